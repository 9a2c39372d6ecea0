// extractFeatures (reference featureAssociation.cpp:680-784) as one CTA per (ring, lane).
//
// Rings are independent in the reference's loop (suppression reaches +-5 slots, the ring margins are 5+5 slots;
// validated in SURVEY.md section 8), so each CTA owns one ring and walks its six sectors in order:
//   phase 1  six warps replay libstdc++'s std::sort on their sector's {curvature, index} slots (std_sort_replay.cuh);
//            the sort range is [sp, ep) while the scans below include slot ep (quirk Q10);
//   phase 2  warp 0 does the greedy picks with warp ballots: 32 candidates are tested at once, the first eligible
//            one is taken, its +-5 neighbourhood is suppressed (stopping at column gaps > 10), and the remaining
//            lanes are re-tested -- identical to the sequential scan because suppression only ever removes
//            candidates;
//   phase 3  the ring's less-flat candidates go through pcl::VoxelGrid(0.2) semantics: bounding box, voxel keys,
//            block-wide bitonic sort by (key, position), one thread per voxel averages its members in order.
// cloudNeighborPicked / cloudLabel live in persistent global arrays exactly like the reference's (never cleared);
// the CTA stages its ring's window in shared memory and writes it back.
#include <cfloat>
#include <climits>
#include "engine.h"
#include "kernel_utils.cuh"
#include "parity_math.cuh"
#include "std_sort_replay.cuh"

namespace lego {

struct RingWin {
    int lo, hi;                       // window [lo, hi) of cloud indices staged in shared memory
    int* col; float* curv; signed char* picked; signed char* label; unsigned char* ground;   // shared
    const int* gcol; const float* gcurv; int* gpicked; int* glabel; const uint8_t* gground;  // global (lane base)
};
__device__ __forceinline__ bool in_win(const RingWin& w, int i) { return i >= w.lo && i < w.hi; }
__device__ __forceinline__ int w_col(const RingWin& w, int i) { return in_win(w, i) ? w.col[i - w.lo] : w.gcol[i]; }
__device__ __forceinline__ float w_curv(const RingWin& w, int i) { return in_win(w, i) ? w.curv[i - w.lo] : w.gcurv[i]; }
__device__ __forceinline__ int w_picked(const RingWin& w, int i) { return in_win(w, i) ? (int)w.picked[i - w.lo] : w.gpicked[i]; }
__device__ __forceinline__ bool w_ground(const RingWin& w, int i) { return (in_win(w, i) ? w.ground[i - w.lo] : w.gground[i]) != 0; }
__device__ __forceinline__ void w_set_picked(RingWin& w, int i) { if (in_win(w, i)) w.picked[i - w.lo] = 1; else w.gpicked[i] = 1; }
__device__ __forceinline__ void w_set_label(RingWin& w, int i, int v) { if (in_win(w, i)) w.label[i - w.lo] = (signed char)v; else w.glabel[i] = v; }

// Suppression of the picked point's neighbourhood (featureAssociation.cpp:720-732 and 751-767): the point and up to
// five slots on either side are marked, each walk stopping at the first column gap > 10.  An index below 0 ends the
// backward walk: the reference reads out of bounds there (quirk Q11) and the oracle defines it as "stop".
// The same walk done by a whole warp: lanes 0-4 test the five forward steps, lanes 5-9 the five backward steps, a
// ballot finds where each walk stops, and the lanes before the stop mark their slot.
__device__ __forceinline__ void suppress_warp(RingWin& w, int ind, int lid) {
    bool bad = false; int slot = ind;
    if (lid < 5) {
        const int l = lid + 1;
        slot = ind + l;
        bad = abs(w_col(w, ind + l) - w_col(w, ind + l - 1)) > 10;
    } else if (lid < 10) {
        const int l = -(lid - 4);
        slot = ind + l;
        bad = (ind + l < 0) || abs(w_col(w, ind + l) - w_col(w, ind + l + 1)) > 10;
    }
    const unsigned m = __ballot_sync(0xffffffffu, bad);
    const int nF = (m & 0x1fu) ? __ffs(m & 0x1fu) - 1 : 5;
    const int nB = ((m >> 5) & 0x1fu) ? __ffs((m >> 5) & 0x1fu) - 1 : 5;
    if (lid < nF || (lid >= 5 && lid - 5 < nB) || lid == 10) w_set_picked(w, slot);
}

__device__ __forceinline__ void sector_bounds(int start, int end, int j, int* sp, int* ep) {
    *sp = (start * (6 - j) + end * j) / 6;
    *ep = (start * (5 - j) + end * (j + 1)) / 6 - 1;
}

__global__ void __launch_bounds__(256) fa_extract_kernel(IpDev d, FaDev f, Consts c, float leaf) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = blockIdx.y, ring = blockIdx.x, tid = threadIdx.x;
    const int warp = tid >> 5, lid = tid & 31;
    const size_t base = (size_t)lane * c.N;
    const int start = d.startRing[lane * c.R + ring], end = d.endRing[lane * c.R + ring];
    const int rowStart = start - 4, rowEnd = end + 6;
    int* rc = f.ringCounts + ((size_t)lane * c.R + ring) * 4;
    if (rowEnd - rowStart <= 0) { if (tid < 4) rc[tid] = 0; return; }

    // ---- shared memory carve-up; region A (items | col | curv) is reused for the voxel keys in phase 3 ----
    const int Cw = c.C + 8;
    ssr::Item* items = reinterpret_cast<ssr::Item*>(smem);
    int* s_col = reinterpret_cast<int*>(smem + (size_t)8 * c.C);
    float* s_curv = reinterpret_cast<float*>(smem + (size_t)8 * c.C + (size_t)4 * Cw);
    unsigned long long* vkeys = reinterpret_cast<unsigned long long*>(smem);
    unsigned char* tail = smem + (size_t)8 * c.C + (size_t)8 * Cw;
    signed char* s_picked = reinterpret_cast<signed char*>(tail);
    signed char* s_label = reinterpret_cast<signed char*>(tail + Cw);
    unsigned char* s_ground = tail + 2 * Cw;
    __shared__ int s_sp[6], s_ep[6], s_tie[6];
    __shared__ int s_pickSharp[12], s_pickLess[120], s_pickFlat[24], s_nPick[3];
    __shared__ float s_red[6][8];
    __shared__ int s_scan[8];
    __shared__ int s_carry;

    RingWin w;
    w.lo = max(rowStart - 1, 0); w.hi = rowEnd;
    w.col = s_col; w.curv = s_curv; w.picked = s_picked; w.label = s_label; w.ground = s_ground;
    w.gcol = d.segCol + base; w.gcurv = f.curv + base; w.gpicked = f.picked + base; w.glabel = f.label + base; w.gground = d.segGround + base;

    for (int k = w.lo + tid; k < w.hi; k += blockDim.x) {
        s_col[k - w.lo] = w.gcol[k]; s_curv[k - w.lo] = w.gcurv[k];
        s_picked[k - w.lo] = (signed char)w.gpicked[k]; s_label[k - w.lo] = (signed char)w.glabel[k]; s_ground[k - w.lo] = w.gground[k];
    }
    for (int k = start + tid; k < end; k += blockDim.x) {
        ssr::Item it;
        if (k == 4) { it.value = f.slot4val[lane]; it.ind = f.slot4ind[lane]; }   // stale slot, quirk Q11
        else { it.value = w.gcurv[k]; it.ind = k; }
        items[k - start] = it;
    }
    if (tid < 6) { int sp, ep; sector_bounds(start, end, tid, &sp, &ep); s_sp[tid] = sp; s_ep[tid] = ep; s_tie[tid] = 0; }
    if (tid < 3) s_nPick[tid] = 0;
    __syncthreads();

    // ---- phase 1: std::sort(cloudSmoothness.begin()+sp, cloudSmoothness.begin()+ep) per sector ----
    // A sector whose curvatures are all distinct has exactly one sorted order, so it is sorted in parallel by ranking
    // (every thread counts the smaller keys of its items).  Equal curvatures are common (quirk Q17) and their final
    // order depends on libstdc++'s introsort path: those sectors keep their original arrangement and one thread replays
    // std::sort on it.
    {
        constexpr int MAXI = 9;                     // items per thread: covers Horizon_SCAN <= 2304
        const bool rankable = end - start <= MAXI * 256;
        ssr::Item mine[MAXI]; int dst[MAXI], sec[MAXI];
#pragma unroll
        for (int u = 0; u < MAXI; ++u) {
            const int k = start + tid + u * 256;
            dst[u] = -1; sec[u] = 0;
            if (rankable && k < end) {
                int j = -1;
#pragma unroll
                for (int t = 0; t < 6; ++t) if (k >= s_sp[t] && k < s_ep[t]) j = t;
                if (j >= 0) {
                    const ssr::Item it = items[k - start];
                    const int sp = s_sp[j], ep = s_ep[j];
                    // rank = number of smaller curvatures in the sector, eq = number of equal ones (itself included).
                    // The curvatures are read from the staged cloudCurvature window (contiguous floats, 4 per load);
                    // slot 4 of cloudSmoothness is the one entry that is NOT cloudCurvature (quirk Q11), patched below.
                    const float x = it.value;
                    const float* vp = s_curv + (sp - w.lo);
                    const int n = ep - sp;
                    int rank = 0, eq = 0, q = 0;
                    for (; q < n && ((size_t)(vp + q) & 15); ++q) { const float v = vp[q]; rank += v < x; eq += v == x; }
                    for (; q + 4 <= n; q += 4) {
                        const float4 v = *reinterpret_cast<const float4*>(vp + q);
                        rank += (v.x < x) + (v.y < x) + (v.z < x) + (v.w < x);
                        eq += (v.x == x) + (v.y == x) + (v.z == x) + (v.w == x);
                    }
                    for (; q < n; ++q) { const float v = vp[q]; rank += v < x; eq += v == x; }
                    if (start == 4 && j == 0) {
                        const float c4 = s_curv[4 - w.lo], t4 = items[0].value;
                        rank += (t4 < x) - (c4 < x); eq += (t4 == x) - (c4 == x);
                    }
                    if (eq > 1) s_tie[j] = 1;
                    mine[u] = it; dst[u] = sp - start + rank; sec[u] = j;
                }
            }
        }
        if (!rankable && tid < 6) s_tie[tid] = 1;
        __syncthreads();
#pragma unroll
        for (int u = 0; u < MAXI; ++u) if (dst[u] >= 0 && !s_tie[sec[u]]) items[dst[u]] = mine[u];
        if (lid == 0 && warp < 6 && s_tie[warp]) {
            const int sp = s_sp[warp], ep = s_ep[warp];
            if (sp < ep) ssr::std_sort(items + (sp - start), ep - sp);
        }
    }
    __syncthreads();
    if (tid == 0 && start == 4 && s_sp[0] < s_ep[0]) { f.slot4val[lane] = items[0].value; f.slot4ind[lane] = items[0].ind; }

    // ---- phase 2: greedy selection, sectors in order, one warp ----
    if (warp == 0) {
        int nSharp = 0, nLess = 0, nFlat = 0;
        for (int j = 0; j < 6; ++j) {
            const int sp = s_sp[j], ep = s_ep[j];
            if (sp >= ep) continue;
            // corner pass, k = ep down to sp (featureAssociation.cpp:701-734)
            int picks = 0; bool stop = false;
            for (int top = ep; top >= sp && !stop; top -= 32) {
                const int k = top - lid;
                const bool inr = k >= sp;
                const int ind = inr ? items[k - start].ind : 0;
                unsigned done = 0u;
                while (true) {
                    const bool elig = inr && !((done >> lid) & 1u) && w_picked(w, ind) == 0 && w_curv(w, ind) > c.edgeThreshold && !w_ground(w, ind);
                    const unsigned m = __ballot_sync(0xffffffffu, elig);
                    if (!m) break;
                    const int first = __ffs(m) - 1;
                    const int pind = __shfl_sync(0xffffffffu, ind, first);
                    ++picks;
                    if (picks > 20) { stop = true; break; }
                    if (lid == 0) {   // the points themselves are copied by the whole CTA after the selection
                        if (picks <= 2) { w_set_label(w, pind, 2); s_pickSharp[nSharp] = pind; s_pickLess[nLess] = pind; }
                        else { w_set_label(w, pind, 1); s_pickLess[nLess] = pind; }
                    }
                    suppress_warp(w, pind, lid);
                    if (picks <= 2) ++nSharp;
                    ++nLess;
                    __syncwarp();
                    done |= (first == 31) ? 0xffffffffu : ((2u << first) - 1u);
                }
            }
            // flat pass, k = sp up to ep (featureAssociation.cpp:736-769); the 4th pick breaks BEFORE suppression (Q12)
            picks = 0; stop = false;
            for (int bot = sp; bot <= ep && !stop; bot += 32) {
                const int k = bot + lid;
                const bool inr = k <= ep;
                const int ind = inr ? items[k - start].ind : 0;
                unsigned done = 0u;
                while (true) {
                    const bool elig = inr && !((done >> lid) & 1u) && w_picked(w, ind) == 0 && w_curv(w, ind) < c.surfThreshold && w_ground(w, ind);
                    const unsigned m = __ballot_sync(0xffffffffu, elig);
                    if (!m) break;
                    const int first = __ffs(m) - 1;
                    const int pind = __shfl_sync(0xffffffffu, ind, first);
                    ++picks;
                    if (lid == 0) {
                        w_set_label(w, pind, -1);
                        s_pickFlat[nFlat] = pind;
                    }
                    if (picks < 4) suppress_warp(w, pind, lid);
                    ++nFlat;
                    __syncwarp();
                    if (picks >= 4) { stop = true; break; }
                    done |= (first == 31) ? 0xffffffffu : ((2u << first) - 1u);
                }
            }
        }
        if (lid == 0) { rc[0] = nSharp; rc[1] = nLess; rc[2] = nFlat; s_nPick[0] = nSharp; s_nPick[1] = nLess; s_nPick[2] = nFlat; }
    }
    __syncthreads();
    {
        const float4* adjp = f.segAdj + base;
        float4* oSharp = f.ringSharp + ((size_t)lane * c.R + ring) * 12;
        float4* oLess = f.ringLessSharp + ((size_t)lane * c.R + ring) * 120;
        float4* oFlat = f.ringFlat + ((size_t)lane * c.R + ring) * 24;
        if (tid < s_nPick[0]) oSharp[tid] = adjp[s_pickSharp[tid]];
        if (tid >= 32 && tid - 32 < s_nPick[2]) oFlat[tid - 32] = adjp[s_pickFlat[tid - 32]];
        if (tid >= 64 && tid - 64 < s_nPick[1]) oLess[tid - 64] = adjp[s_pickLess[tid - 64]];
    }

    // ---- write back this ring's share of the persistent arrays ([lo, hi-1): the last slot is never modified here
    //      but may be by the next ring's CTA) ----
    for (int k = w.lo + tid; k < w.hi - 1; k += blockDim.x) { w.gpicked[k] = s_picked[k - w.lo]; w.glabel[k] = s_label[k - w.lo]; }

    // ---- phase 3: less-flat candidates of this ring -> VoxelGrid(leaf) (featureAssociation.cpp:771-782) ----
    const float4* adj = f.segAdj + base;
    auto is_cand = [&](int k) -> bool {
        bool in_sector = false;
#pragma unroll
        for (int j = 0; j < 6; ++j) in_sector |= (s_sp[j] < s_ep[j] && k >= s_sp[j] && k <= s_ep[j]);
        return in_sector && s_label[k - w.lo] <= 0;
    };
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    int myCount = 0;
    for (int k = start + tid; k < end; k += blockDim.x) if (is_cand(k)) {
        const float4 p = adj[k];
        mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
        mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
        ++myCount;
    }
    __syncthreads();   // everyone is done with items / col / curv before region A becomes vkeys
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = warp_min_f(mn[a]); mx[a] = warp_max_f(mx[a]); }
    if (lid == 0) { for (int a = 0; a < 3; ++a) { s_red[a][warp] = mn[a]; s_red[3 + a][warp] = mx[a]; } }
    myCount = warp_sum(myCount);
    if (lid == 0) s_scan[warp] = myCount;
    __syncthreads();
    int nCand = 0;
    for (int q = 0; q < 8; ++q) nCand += s_scan[q];
    if (nCand == 0) { if (tid == 0) rc[3] = 0; return; }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float lo = s_red[a][0], hi = s_red[3 + a][0];
        for (int q = 1; q < 8; ++q) { lo = fminf(lo, s_red[a][q]); hi = fmaxf(hi, s_red[3 + a][q]); }
        mn[a] = lo; mx[a] = hi;
    }
    const float inv = PM_DIV(1.0f, leaf);
    long long dd[3]; int minb[3], divb[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        dd[a] = (long long)PM_MUL(PM_SUB(mx[a], mn[a]), inv) + 1;
        minb[a] = (int)floorf(PM_MUL(mn[a], inv));
        divb[a] = (int)floorf(PM_MUL(mx[a], inv)) - minb[a] + 1;
    }
    const bool overflow = dd[0] * dd[1] * dd[2] > (long long)INT_MAX;
    float4* out = f.ringLessFlat + ((size_t)lane * c.R + ring) * c.C;
    int n2 = 1; while (n2 < nCand) n2 <<= 1;
    __syncthreads();
    // ordered list of candidates with their voxel keys: key << 12 | (k - start)
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int k0 = start; k0 < end; k0 += blockDim.x) {
        const int k = k0 + tid;
        const int flag = (k < end && is_cand(k)) ? 1 : 0;
        const int inc = warp_inclusive_scan(flag);
        if (lid == 31) s_scan[warp] = inc;
        __syncthreads();
        int off = s_carry + inc - flag;
        for (int q = 0; q < warp; ++q) off += s_scan[q];
        if (flag) {
            const float4 p = adj[k];
            const int i0 = (int)PM_SUB(floorf(PM_MUL(p.x, inv)), (float)minb[0]);
            const int i1 = (int)PM_SUB(floorf(PM_MUL(p.y, inv)), (float)minb[1]);
            const int i2 = (int)PM_SUB(floorf(PM_MUL(p.z, inv)), (float)minb[2]);
            const long long key = overflow ? (long long)off : ((long long)i0 + (long long)i1 * divb[0] + (long long)i2 * divb[0] * divb[1]);
            vkeys[off] = ((unsigned long long)key << 12) | (unsigned long long)(k - start);
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int q = 0; q < 8; ++q) s += s_scan[q]; s_carry += s; }
        __syncthreads();
    }
    for (int i = nCand + tid; i < n2; i += blockDim.x) vkeys[i] = ~0ull;
    __syncthreads();
    for (int kk = 2; kk <= n2; kk <<= 1)
        for (int jj = kk >> 1; jj > 0; jj >>= 1) {
            for (int i = tid; i < n2; i += blockDim.x) {
                const int p = i ^ jj;
                if (p > i) {
                    const unsigned long long a = vkeys[i], b = vkeys[p];
                    const bool asc = (i & kk) == 0;
                    if ((a > b) == asc) { vkeys[i] = b; vkeys[p] = a; }
                }
            }
            __syncthreads();
        }
    // one output per distinct key, mean of the members in sorted order (float32 sums, then divide)
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int i0 = 0; i0 < nCand; i0 += blockDim.x) {
        const int i = i0 + tid;
        int head = 0;
        if (i < nCand) head = (i == 0 || (vkeys[i - 1] >> 12) != (vkeys[i] >> 12)) ? 1 : 0;
        const int inc = warp_inclusive_scan(head);
        if (lid == 31) s_scan[warp] = inc;
        __syncthreads();
        int off = s_carry + inc - head;
        for (int q = 0; q < warp; ++q) off += s_scan[q];
        if (head) {
            const unsigned long long key = vkeys[i] >> 12;
            float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f; int cnt = 0;
            for (int m = i; m < nCand && (vkeys[m] >> 12) == key; ++m) {
                const float4 p = adj[start + (int)(vkeys[m] & 0xfffull)];
                sx = PM_ADD(sx, p.x); sy = PM_ADD(sy, p.y); sz = PM_ADD(sz, p.z); sw = PM_ADD(sw, p.w); ++cnt;
            }
            const float fc = (float)cnt;
            out[off] = make_float4(PM_DIV(sx, fc), PM_DIV(sy, fc), PM_DIV(sz, fc), PM_DIV(sw, fc));
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int q = 0; q < 8; ++q) s += s_scan[q]; s_carry += s; }
        __syncthreads();
    }
    if (tid == 0) rc[3] = s_carry;
}

// Concatenate the ring-local lists in ring order (the reference appends ring by ring, featureAssociation.cpp:782):
// one CTA per (ring, lane) sums the counts of the rings before it and copies its own four lists.
__global__ void __launch_bounds__(128) fa_gather_features_kernel(FaDev f, Consts c) {
    const int ring = blockIdx.x, lane = blockIdx.y, tid = threadIdx.x;
    __shared__ int off[4], cnt[4], tot[4];
    if (tid < 4) {
        int s = 0, all = 0;
        for (int r = 0; r < c.R; ++r) {
            const int n = f.ringCounts[((size_t)lane * c.R + r) * 4 + tid];
            if (r < ring) s += n;
            all += n;
        }
        off[tid] = s; tot[tid] = all; cnt[tid] = f.ringCounts[((size_t)lane * c.R + ring) * 4 + tid];
        if (ring == 0) f.featCounts[lane * 4 + tid] = all;
    }
    __syncthreads();
    const int cap[4] = {12, 120, 24, c.C};
    const float4* src[4] = {f.ringSharp + (size_t)lane * c.R * 12, f.ringLessSharp + (size_t)lane * c.R * 120,
                            f.ringFlat + (size_t)lane * c.R * 24, f.ringLessFlat + (size_t)lane * c.R * c.C};
    float4* dst[4] = {f.sharp + (size_t)lane * c.R * 12, f.lessSharp + (size_t)lane * c.R * 120,
                      f.flat + (size_t)lane * c.R * 24, f.lessFlat + (size_t)lane * c.N};
#pragma unroll
    for (int t = 0; t < 4; ++t)
        for (int i = tid; i < cnt[t]; i += blockDim.x) dst[t][off[t] + i] = src[t][(size_t)ring * cap[t] + i];
}

// Kernels defined in feature_association.cu
__global__ void fa_prepare_kernel(IpDev d, FaDev f, Consts c);
__global__ void fa_distort_kernel(IpDev d, FaDev f, Consts c);

void Engine::adjustDistortionCalculateSmoothnessMarkOccluded() {
    cuda_check(cudaMemsetAsync(fa.halfFirst, 0x7F, sizeof(int) * c.L, stream), "memset halfFirst");
    dim3 g((c.N + 255) / 256, c.L);
    { ProfScope ps(this, "fa_prepare_kernel"); fa_prepare_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    { ProfScope ps(this, "fa_distort_kernel"); fa_distort_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    launches += 3;
}

void Engine::extractFeatures() {
    const size_t smem = (size_t)8 * c.C + (size_t)8 * (c.C + 8) + (size_t)3 * (c.C + 8);
    static bool attr_set = false;
    if (!attr_set) {
        cuda_check(cudaFuncSetAttribute(fa_extract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024), "smem attr");
        attr_set = true;
    }
    dim3 g(c.R, c.L);
    { ProfScope ps(this, "fa_extract_kernel"); fa_extract_kernel<<<g, 256, smem, stream>>>(ip, fa, c, 0.2f); }   // downSizeFilter.setLeafSize(0.2,...) featureAssociation.cpp:225
    { ProfScope ps(this, "fa_gather_features_kernel"); fa_gather_features_kernel<<<dim3(c.R, c.L), 128, 0, stream>>>(fa, c); }
    launches += 2;
}

}  // namespace lego
