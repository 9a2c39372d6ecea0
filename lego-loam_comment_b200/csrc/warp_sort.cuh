// Warp-level sorting pieces shared by fa_extract_kernel and the self tests: a stable LSD radix sort over 16-bit
// entries with 32-bit keys, and the exact reproduction of libstdc++'s std::sort permutation by one warp.
#pragma once
#include <climits>
#include "kernel_utils.cuh"
#include "std_sort_replay.cuh"

namespace lego {

constexpr int WS_MAX_SORT = 13 * 32;   // entries one warp permutes from registers

// ---- stable LSD radix sort pieces (8-bit digits) over lists of 16-bit entries; key(e) is the 32-bit key of entry e.
// One warp owns `n` consecutive entries and 256 private counters; ranks inside a 32-entry chunk come from
// __match_any_sync, so there are no shared-memory atomics and equal digits keep their order.
__device__ __forceinline__ unsigned float_sort_key(float v) {
    const unsigned b = __float_as_uint(v);
    return b ^ ((unsigned)((int)b >> 31) | 0x80000000u);
}
// A batch of up to 32*U consecutive entries of one warp's list, held in registers between the counting and the
// scattering half of a pass: entry, digit and the lanes holding the same digit in the same 32-entry chunk.
template <int U>
struct RadixBatch { unsigned ed[U]; unsigned peers[U]; };   // ed = entry | digit << 16

template <int U, class KeyF>
__device__ __forceinline__ void radix_batch_load(RadixBatch<U>& r, const unsigned short* in, int n, int shift, int lid, KeyF key) {
#pragma unroll
    for (int u = 0; u < U; ++u) { const int q = lid + 32 * u; r.ed[u] = q < n ? (unsigned)in[q] : 0u; }
#pragma unroll
    for (int u = 0; u < U; ++u) { const int q = lid + 32 * u; r.ed[u] |= (q < n ? ((key((unsigned short)r.ed[u]) >> shift) & 255u) : 256u) << 16; }
#pragma unroll
    for (int u = 0; u < U; ++u) r.peers[u] = (32 * u < n) ? __match_any_sync(0xffffffffu, r.ed[u] >> 16) : 0u;
}
// counters are 16 bit, two per 32-bit word, so a chunk's leader can add its group with one shared-memory atomic
template <int U>
__device__ __forceinline__ void radix_batch_count(const RadixBatch<U>& r, int n, unsigned* h32, int lid) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int q = lid + 32 * u;
        const unsigned dg = r.ed[u] >> 16;
        if (q < n && lid == __ffs(r.peers[u]) - 1) atomicAdd(&h32[dg >> 1], (unsigned)__popc(r.peers[u]) << ((dg & 1) * 16));
    }
}
template <int U>
__device__ __forceinline__ void radix_batch_scatter(const RadixBatch<U>& r, int n, unsigned* h32, unsigned short* out, int lid) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (32 * u < n) {   // warp-uniform
            const int q = lid + 32 * u;
            const int leader = __ffs(r.peers[u]) - 1;
            const unsigned dg = r.ed[u] >> 16;
            unsigned old = 0;
            if (q < n && lid == leader) old = atomicAdd(&h32[dg >> 1], (unsigned)__popc(r.peers[u]) << ((dg & 1) * 16));
            __syncwarp();   // chunk u's update is ordered before chunk u+1's
            old = (__shfl_sync(0xffffffffu, old, leader) >> ((dg & 1) * 16)) & 0xffffu;
            if (q < n) out[old + __popc(r.peers[u] & ((1u << lid) - 1u))] = (unsigned short)r.ed[u];
        }
    }
}
// exclusive scan of a warp's 256 counters, in place
__device__ __forceinline__ void warp_scan256(unsigned short* h, int lid) {
    int v[8], s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = h[8 * lid + i]; s += v[i]; }
    int run = warp_inclusive_scan(s) - s;
#pragma unroll
    for (int i = 0; i < 8; ++i) { h[8 * lid + i] = (unsigned short)run; run += v[i]; }
    __syncwarp();
}

// Stable LSD radix sort of one warp's list with 4-bit digits and lane-private counters: lane l owns the l-th block
// of ceil(n/32) consecutive entries and one column of the 16 x 32 counter table, so counting and ranking need no
// votes or atomics (MATCH / VOTE share the ADU pipe, which saturates long before anything else when every warp of
// the SM sorts).  The exclusive scan runs over (digit, lane), which keeps equal digits in list order.
// cnt: 512 counters.  Returns the list that holds the result (in or out).
template <class KeyF>
__device__ __forceinline__ unsigned short* warp_radix_sort_blocked(unsigned short* in, unsigned short* out, int n, int keyBits,
                                                                   unsigned short* cnt, int lid, KeyF key) {
    const int m = (n + 31) >> 5;
    const int b0 = min(lid * m, n), b1 = min(b0 + m, n);
    for (int shift = 0; shift < keyBits; shift += 4) {
#pragma unroll
        for (int b = 0; b < 16; ++b) cnt[b * 32 + lid] = 0;
        for (int q = b0; q < b1; ++q) { const unsigned dg = (key(in[q]) >> shift) & 15u; cnt[dg * 32 + lid] = (unsigned short)(cnt[dg * 32 + lid] + 1); }
        __syncwarp();
        {
            int v[16], s = 0;
#pragma unroll
            for (int i = 0; i < 16; ++i) { v[i] = cnt[16 * lid + i]; s += v[i]; }
            int run = warp_inclusive_scan(s) - s;
#pragma unroll
            for (int i = 0; i < 16; ++i) { cnt[16 * lid + i] = (unsigned short)run; run += v[i]; }
        }
        __syncwarp();
        for (int q = b0; q < b1; ++q) {
            const unsigned short e = in[q];
            const unsigned dg = (key(e) >> shift) & 15u;
            const int pos = cnt[dg * 32 + lid];
            cnt[dg * 32 + lid] = (unsigned short)(pos + 1);
            out[pos] = e;
        }
        __syncwarp();
        unsigned short* t = in; in = out; out = t;
    }
    return in;
}

// The partition phase of libstdc++'s introsort (std_sort_replay.cuh has the single-thread statement) done by one warp.
// Hoare's unguarded partition is sequential as written, but its outcome has a closed form on the arrangement it
// starts from: with L_1 < L_2 < ... the positions whose value is not below the pivot and R_1 > R_2 > ... the
// positions whose value is not above it, the loop swaps the pairs (L_k, R_k) for as long as L_k < R_k (K pairs) and
// ends with cut = min(L_{K+1}, R_K), because both scans only ever walk over slots no swap has touched yet.  So the
// warp compacts the two stopper lists with ballots, counts K, swaps the K pairs in parallel and moves on.  What
// std::sort does afterwards (guarded-16 + unguarded insertion sort) is a STABLE sort of that arrangement, which the
// caller gets from the radix sort.  Checked against std::sort on tie-heavy inputs in tests (lego_selftest_std_sort).
__device__ __forceinline__ void warp_introsort_partitions(ssr::Item* a, int n, unsigned short* Ls, unsigned short* Rs, int lid) {
    int stk[24];
    int sp = 0;
    stk[sp++] = (n << 10) | (2 * ssr::floor_log2(n) << 20);
    const unsigned lt = (1u << lid) - 1u;
    while (sp > 0) {
        const int e = stk[--sp];
        int first = e & 1023, last = (e >> 10) & 1023, depth = e >> 20;
        while (last - first > 16) {
            if (depth == 0) { if (lid == 0) ssr::heap_sort(a + first, last - first); __syncwarp(); break; }
            --depth;
            if (lid == 0) {   // median of (first+1, mid, last-1) moved to first
                const int ia = first + 1, ib = first + (last - first) / 2, ic = last - 1;
                const float va = a[ia].value, vb = a[ib].value, vc = a[ic].value;
                int pick;
                if (va < vb) pick = (vb < vc) ? ib : ((va < vc) ? ic : ia);
                else pick = (va < vc) ? ia : ((vb < vc) ? ic : ib);
                ssr::swap_items(a[first], a[pick]);
            }
            __syncwarp();
            const float p = a[first].value;
            int nL = 0, nR = 0;
            for (int i0 = first + 1; i0 < last; i0 += 32) {
                const int i = i0 + lid;
                const bool v = i < last;
                const float val = v ? a[i].value : 0.f;
                const bool A = v && !(val < p), B = v && !(p < val);
                const unsigned mA = __ballot_sync(0xffffffffu, A), mB = __ballot_sync(0xffffffffu, B);
                if (A) Ls[nL + __popc(mA & lt)] = (unsigned short)i;
                if (B) Rs[nR + __popc(mB & lt)] = (unsigned short)i;
                nL += __popc(mA); nR += __popc(mB);
            }
            __syncwarp();
            const int m = min(nL, nR);
            int K = 0;
            for (int k0 = 0; k0 < m; k0 += 32) {
                const int k = k0 + lid;
                const bool ok = k < m && Ls[k] < Rs[nR - 1 - k];
                const unsigned mk = __ballot_sync(0xffffffffu, ok);
                K += __popc(mk);
                if (mk != 0xffffffffu) break;
            }
            int cut = INT_MAX;
            if (K < nL) cut = Ls[K];
            if (K >= 1) cut = min(cut, (int)Rs[nR - K]);
            for (int k = lid; k < K; k += 32) ssr::swap_items(a[Ls[k]], a[Rs[nR - 1 - k]]);
            __syncwarp();
            stk[sp++] = cut | (last << 10) | (depth << 20);
            last = cut;
        }
    }
}

// std::sort(a, a + n) with the reference's comparator (value only, utility.h:139-148), n <= WS_MAX_SORT, by one warp.
// Distinct values have one sorted order and the radix sort finds it; if equal values exist, introsort's partitions
// are replayed first (permuting a) and the radix sort (stable, like std::sort's closing insertion sort) finishes the
// job.  Returns the sorted order as a list of positions into a: element q of the result is a[order[q]].
// ordA / ordB: n entries of scratch each (the result is one of them), h: 512 counters.
__device__ __forceinline__ const unsigned short* warp_std_sort_order(ssr::Item* a, int n, unsigned short* ordA, unsigned short* ordB,
                                                                     unsigned short* h, int lid) {
    auto key = [&](unsigned short e) -> unsigned { return float_sort_key(a[e].value); };
    unsigned short* in = ordA;
    for (int attempt = 0; attempt < 2; ++attempt) {
        for (int q = lid; q < n; q += 32) ordA[q] = (unsigned short)q;
        __syncwarp();
        in = warp_radix_sort_blocked(ordA, ordB, n, 32, h, lid, key);
        if (attempt == 1) break;
        bool tie = false;
        for (int q = lid; q + 1 < n; q += 32) tie |= a[in[q]].value == a[in[q + 1]].value;
        if (!__any_sync(0xffffffffu, tie)) break;
        warp_introsort_partitions(a, n, ordA, ordB, lid);
    }
    return in;
}

// In-place form (self test).
__device__ __forceinline__ void warp_std_sort(ssr::Item* a, int n, unsigned short* ordA, unsigned short* ordB, unsigned short* h, int lid) {
    if (n <= 1) return;
    const unsigned short* in = warp_std_sort_order(a, n, ordA, ordB, h, lid);
    ssr::Item mine[WS_MAX_SORT / 32];
#pragma unroll
    for (int u = 0; u < WS_MAX_SORT / 32; ++u) { const int q = lid + 32 * u; if (q < n) mine[u] = a[in[q]]; }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < WS_MAX_SORT / 32; ++u) { const int q = lid + 32 * u; if (q < n) a[q] = mine[u]; }
    __syncwarp();
}

}  // namespace lego
