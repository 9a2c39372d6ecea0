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
template <class KeyF>
__device__ __forceinline__ void warp_radix_count(const unsigned short* in, int n, int shift, unsigned short* h, int lid, KeyF key) {
    for (int q0 = 0; q0 < n; q0 += 32) {
        const int q = q0 + lid;
        const bool valid = q < n;
        const unsigned dg = valid ? ((key(in[q]) >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        if (valid && lid == __ffs(peers) - 1) h[dg] = (unsigned short)(h[dg] + __popc(peers));
        __syncwarp();
    }
}
template <class KeyF>
__device__ __forceinline__ void warp_radix_scatter(const unsigned short* in, unsigned short* out, int n, int shift, unsigned short* h, int lid, KeyF key) {
    for (int q0 = 0; q0 < n; q0 += 32) {
        const int q = q0 + lid;
        const bool valid = q < n;
        const unsigned short e = valid ? in[q] : (unsigned short)0;
        const unsigned dg = valid ? ((key(e) >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        const int leader = __ffs(peers) - 1;
        int old = 0;
        if (valid && lid == leader) { old = h[dg]; h[dg] = (unsigned short)(old + __popc(peers)); }
        old = __shfl_sync(0xffffffffu, old, leader);
        if (valid) out[old + __popc(peers & ((1u << lid) - 1u))] = e;
        __syncwarp();
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
// are replayed first and the radix sort (stable, like std::sort's closing insertion sort) finishes the job.
// ordA / ordB: n entries of scratch each, h: 256 counters.
__device__ __forceinline__ void warp_std_sort(ssr::Item* a, int n, unsigned short* ordA, unsigned short* ordB, unsigned short* h, int lid) {
    if (n <= 1) return;
    auto key = [&](unsigned short e) -> unsigned { return float_sort_key(a[e].value); };
    for (int attempt = 0; attempt < 2; ++attempt) {
        unsigned short* in = ordA; unsigned short* out = ordB;
        for (int q = lid; q < n; q += 32) in[q] = (unsigned short)q;
        __syncwarp();
        for (int shift = 0; shift < 32; shift += 8) {
#pragma unroll
            for (int i = 0; i < 8; ++i) h[lid + 32 * i] = 0;
            __syncwarp();
            warp_radix_count(in, n, shift, h, lid, key);
            warp_scan256(h, lid);
            warp_radix_scatter(in, out, n, shift, h, lid, key);
            unsigned short* t = in; in = out; out = t;
        }
        if (attempt == 0) {
            bool tie = false;
            for (int q = lid; q + 1 < n; q += 32) tie |= a[in[q]].value == a[in[q + 1]].value;
            if (__any_sync(0xffffffffu, tie)) { warp_introsort_partitions(a, n, ordA, ordB, lid); continue; }
        }
        ssr::Item mine[WS_MAX_SORT / 32];
#pragma unroll
        for (int u = 0; u < WS_MAX_SORT / 32; ++u) { const int q = lid + 32 * u; if (q < n) mine[u] = a[in[q]]; }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < WS_MAX_SORT / 32; ++u) { const int q = lid + 32 * u; if (q < n) a[q] = mine[u]; }
        __syncwarp();
        return;
    }
}

}  // namespace lego
