// extractFeatures (reference featureAssociation.cpp:680-784) as three kernels, each shaped after the parallelism its
// step has.  Rings are independent in the reference's loop (suppression reaches +-5 slots, the ring margins are 5+5
// slots; validated in SURVEY.md section 8) and so are the sorts of the six sectors of a ring.
//   fa_sort_sectors_kernel   one warp per (sector, ring, lane): std::sort of the sector's {curvature, index} slots
//            with exactly libstdc++'s permutation (warp_sort.cuh: radix sort when the curvatures are distinct,
//            introsort partition replay + stable radix sort when some are equal, quirk Q17).  The sort range is
//            [sp, ep) while the scans include slot ep (quirk Q10).  The warp then writes the two lists the greedy
//            selection walks: edge candidates (curvature > edgeThreshold, not ground) in descending order and flat
//            candidates (curvature < surfThreshold, ground) in ascending order, each with the reach of the
//            suppression walk of featureAssociation.cpp:720-732 from that point (steps before a column gap > 10).
//   fa_select_kernel         one warp per (ring, lane): the greedy picks, sectors in order.  32 candidates are tested
//            at once, the first one not yet suppressed is taken, the lanes inside its reach drop out by comparing
//            indices and lanes 0..10 mark cloudNeighborPicked -- identical to the sequential scan because
//            suppression only ever removes candidates.
//   fa_voxel_ring_kernel     one CTA per (ring, lane): the ring's less-flat candidates through pcl::VoxelGrid(0.2)
//            semantics: bounding box, voxel keys, block-wide stable radix sort by key (input order kept inside a
//            voxel), one thread per voxel averages its members in order.
// cloudNeighborPicked / cloudLabel live in persistent global arrays exactly like the reference's (never cleared).
#include <cfloat>
#include <climits>
#include <cstdlib>
#include <stdexcept>
#include "engine.h"
#include "kernel_utils.cuh"
#include "parity_math.cuh"
#include "std_sort_replay.cuh"
#include "warp_sort.cuh"

namespace lego {

// bit 0: may become an edge point (curvature > edgeThreshold, not ground); bit 1: may become a flat point
__device__ __forceinline__ int pick_stat(float curv, bool ground, float edgeT, float surfT) {
    return ((curv > edgeT && !ground) ? 1 : 0) | ((curv < surfT && ground) ? 2 : 0);
}

// Suppression of the picked point's neighbourhood (featureAssociation.cpp:720-732 and 751-767): the point and up to
// five slots on either side are marked, each walk stopping at the first column gap > 10.  An index below 0 ends the
// backward walk: the reference reads out of bounds there (quirk Q11) and the oracle defines it as "stop".
// reach = forward steps | backward steps << 4.
__device__ __forceinline__ int suppression_reach(const int* col, int ind) {
    int nF = 0, nB = 0;
    while (nF < 5 && abs(col[ind + nF + 1] - col[ind + nF]) <= 10) ++nF;
    while (nB < 5 && ind - nB - 1 >= 0 && abs(col[ind - nB - 1] - col[ind - nB]) <= 10) ++nB;
    return nF | (nB << 4);
}

__device__ __forceinline__ void sector_bounds(int start, int end, int j, int* sp, int* ep) {
    *sp = (start * (6 - j) + end * j) / 6;
    *ep = (start * (5 - j) + end * (j + 1)) / 6 - 1;
}

constexpr int FS_WARPS = 4;          // sectors per CTA of fa_sort_sectors_kernel
constexpr int SEL_WARPS = 4;         // rings per CTA of fa_select_kernel
constexpr int FX_MAX_ITERS = 16;     // 256-wide sweeps over one ring (Horizon_SCAN <= 4096)
constexpr int CAND_SLOW = 0xff;      // reach marker: the point is outside its ring's [start, end) (stale slot 4, quirk Q11)

struct SortSmem { ssr::Item a[WS_MAX_SORT]; unsigned short oA[WS_MAX_SORT], oB[WS_MAX_SORT], h[512]; };

// resident CTAs per SM the compiler is asked to leave registers for (compile-time knobs; unset = its own choice)
#ifdef LEGO_SS_MINBLOCKS
#define SS_BOUNDS __launch_bounds__(32 * FS_WARPS, LEGO_SS_MINBLOCKS)
#else
#define SS_BOUNDS __launch_bounds__(32 * FS_WARPS)
#endif
#ifndef LEGO_VR_MINBLOCKS
#define LEGO_VR_MINBLOCKS 6     // 40 registers instead of 54, no spills: 6 instead of 4 CTAs per SM, 11.4 -> 9.3 ms per 25-scan step of 592 lanes
#endif
#define VR_BOUNDS __launch_bounds__(256, LEGO_VR_MINBLOCKS)
__global__ void SS_BOUNDS fa_sort_sectors_kernel(IpDev d, FaDev f, Consts c) {
    __shared__ __align__(16) SortSmem sm[FS_WARPS];
    const int lane = blockIdx.y, warp = threadIdx.x >> 5, lid = threadIdx.x & 31;
    const int sec = blockIdx.x * FS_WARPS + warp;
    if (sec >= c.R * 6) return;
    const int ring = sec / 6, j = sec - ring * 6;
    const size_t base = (size_t)lane * c.N;
    const int start = d.startRing[lane * c.R + ring], end = d.endRing[lane * c.R + ring];
    int* cnt = f.candCounts + (((size_t)lane * c.R + ring) * 6 + j) * 2;
    int sp, ep;
    sector_bounds(start, end, j, &sp, &ep);
    if (end + 6 - (start - 4) <= 0 || sp >= ep) { if (lid < 2) cnt[lid] = 0; return; }   // featureAssociation.cpp:696-697
    const int n = ep - sp;
    const float* gcurv = f.curv + base;
    const int* gcol = d.segCol + base;
    const uint8_t* gground = d.segGround + base;
    SortSmem& S = sm[warp];

    // ---- std::sort(cloudSmoothness.begin()+sp, cloudSmoothness.begin()+ep) ----
    const bool big = n > WS_MAX_SORT;
    ssr::Item* a = big ? reinterpret_cast<ssr::Item*>(f.sortScratch) + base + sp : S.a;
    for (int q = lid; q < n; q += 32) {
        const int k = sp + q;
        ssr::Item it;
        if (k == 4) { it.value = f.slot4val[lane]; it.ind = f.slot4ind[lane]; }   // stale slot, quirk Q11
        else { it.value = gcurv[k]; it.ind = k; }
        a[q] = it;
    }
    __syncwarp();
    const unsigned short* ord = nullptr;
    if (!big) ord = warp_std_sort_order(a, n, S.oA, S.oB, S.h, lid);
    else { if (lid == 0) ssr::std_sort(a, n); __syncwarp(); }
    auto sorted_item = [&](int q) -> ssr::Item { return big ? a[q] : a[ord[q]]; };
    if (sp == 4 && lid == 0) { const ssr::Item it = sorted_item(0); f.slot4val[lane] = it.value; f.slot4ind[lane] = it.ind; }

    // ---- candidate lists of the two scans (featureAssociation.cpp:701-769) ----
    int* cand = f.cand + base;
    const unsigned lt = (1u << lid) - 1u;
    auto packed_of = [&](int ind) -> int {
        const int reach = (ind >= start && ind < end) ? suppression_reach(gcol, ind) : CAND_SLOW;
        return ind | (reach << 24);
    };
    int nEdge = 0, nFlat = 0;
    for (int c0 = 0; c0 <= n; c0 += 32) {          // k = ep down to sp
        const int t = c0 + lid;
        const bool valid = t <= n;
        const int ind = !valid ? 0 : (t == 0 ? ep : sorted_item(n - t).ind);
        const bool is = valid && (pick_stat(gcurv[ind], gground[ind] != 0, c.edgeThreshold, c.surfThreshold) & 1);
        const unsigned m = __ballot_sync(0xffffffffu, is);
        if (is) cand[sp + nEdge + __popc(m & lt)] = packed_of(ind);
        nEdge += __popc(m);
    }
    for (int c0 = 0; c0 <= n; c0 += 32) {          // k = sp up to ep
        const int t = c0 + lid;
        const bool valid = t <= n;
        const int ind = !valid ? 0 : (t == n ? ep : sorted_item(t).ind);
        const bool is = valid && (pick_stat(gcurv[ind], gground[ind] != 0, c.edgeThreshold, c.surfThreshold) & 2);
        const unsigned m = __ballot_sync(0xffffffffu, is);
        if (is) cand[ep - (nFlat + __popc(m & lt))] = packed_of(ind);
        nFlat += __popc(m);
    }
    if (lid == 0) { cnt[0] = nEdge; cnt[1] = nFlat; }
}

// General suppression for a point anywhere in the cloud (only the stale slot-4 entry of quirk Q11 can be outside its
// own ring): lanes 0-4 test the five forward steps, lanes 5-9 the five backward steps, a ballot finds where each
// walk stops, and the lanes before the stop mark their slot.
__device__ __forceinline__ void suppress_warp(const int* gcol, int* gpicked, int ind, int lid, int N) {
    bool bad = false; int slot = ind;
    if (lid < 5) {
        const int l = lid + 1;
        slot = ind + l;
        bad = (ind + l >= N) || abs(gcol[ind + l] - gcol[ind + l - 1]) > 10;   // the end of the array ends the walk too
    } else if (lid < 10) {
        const int l = -(lid - 4);
        slot = ind + l;
        bad = (ind + l < 0) || abs(gcol[ind + l] - gcol[ind + l + 1]) > 10;
    }
    const unsigned m = __ballot_sync(0xffffffffu, bad);
    const int nF = (m & 0x1fu) ? __ffs(m & 0x1fu) - 1 : 5;
    const int nB = ((m >> 5) & 0x1fu) ? __ffs((m >> 5) & 0x1fu) - 1 : 5;
    if (lid < nF || (lid >= 5 && lid - 5 < nB) || lid == 10) gpicked[slot] = 1;
}

__global__ void __launch_bounds__(32 * SEL_WARPS) fa_select_kernel(IpDev d, FaDev f, Consts c) {
    __shared__ int s_pick[SEL_WARPS][12 + 120 + 24];
    const int lane = blockIdx.y, warp = threadIdx.x >> 5, lid = threadIdx.x & 31;
    const int ring = blockIdx.x * SEL_WARPS + warp;
    if (ring >= c.R) return;
    const size_t base = (size_t)lane * c.N;
    const int start = d.startRing[lane * c.R + ring], end = d.endRing[lane * c.R + ring];
    int* rc = f.ringCounts + ((size_t)lane * c.R + ring) * 4;
    if (end + 6 - (start - 4) <= 0) { if (lid < 3) rc[lid] = 0; return; }
    const int* gcol = d.segCol + base;
    int* gpicked = f.picked + base;
    int* glabel = f.label + base;
    const int* cand = f.cand + base;
    const int* cnt = f.candCounts + ((size_t)lane * c.R + ring) * 12;
    int* pickSharp = s_pick[warp]; int* pickLess = pickSharp + 12; int* pickFlat = pickLess + 120;

    int nSharp = 0, nLess = 0, nFlat = 0;
    for (int j = 0; j < 6; ++j) {
        int sp, ep;
        sector_bounds(start, end, j, &sp, &ep);
        if (sp >= ep) continue;
        for (int pass = 0; pass < 2; ++pass) {
            // pass 0: edge points, largest curvature first (featureAssociation.cpp:701-734)
            // pass 1: flat points, smallest curvature first (featureAssociation.cpp:736-769); its 4th pick breaks BEFORE suppression (Q12)
            const int len = cnt[j * 2 + pass];
            int picks = 0; bool stop = false;
            for (int c0 = 0; c0 < len && !stop; c0 += 32) {
                const int t = c0 + lid;
                const bool valid = t < len;
                const int packed = valid ? cand[pass == 0 ? sp + t : ep - t] : 0;
                const int ind = packed & 0xffffff;
                bool pk = valid ? gpicked[ind] != 0 : true;
                while (true) {
                    const unsigned m = __ballot_sync(0xffffffffu, !pk);
                    if (!m) break;
                    const int first = __ffs(m) - 1;
                    const int pp = __shfl_sync(0xffffffffu, packed, first);
                    const int pind = pp & 0xffffff, reach = (pp >> 24) & 0xff;
                    ++picks;
                    if (pass == 0 && picks > 20) { stop = true; break; }
                    if (lid == 0) {   // the points themselves are copied after the selection
                        if (pass == 1) { glabel[pind] = -1; pickFlat[nFlat] = pind; }
                        else if (picks <= 2) { glabel[pind] = 2; pickSharp[nSharp] = pind; pickLess[nLess] = pind; }
                        else { glabel[pind] = 1; pickLess[nLess] = pind; }
                    }
                    if (pass == 1) ++nFlat; else { ++nLess; if (picks <= 2) ++nSharp; }
                    if (pass == 1 && picks >= 4) { stop = true; break; }
                    if (reach != CAND_SLOW) {
                        const int nF = reach & 15, nB = reach >> 4;
                        const int o = lid - 5;
                        if (lid <= 10 && o >= -nB && o <= nF) gpicked[pind + o] = 1;
                        if (ind >= pind - nB && ind <= pind + nF) pk = true;
                    } else {
                        suppress_warp(gcol, gpicked, pind, lid, c.N);
                        __syncwarp();
                        pk = valid ? gpicked[ind] != 0 : true;
                    }
                }
                __syncwarp();   // marks of this chunk are visible to the loads of the next one
            }
        }
    }
    if (lid == 0) { rc[0] = nSharp; rc[1] = nLess; rc[2] = nFlat; }
    __syncwarp();
    const float4* adj = f.segAdj + base;
    float4* oSharp = f.ringSharp + ((size_t)lane * c.R + ring) * 12;
    float4* oLess = f.ringLessSharp + ((size_t)lane * c.R + ring) * 120;
    float4* oFlat = f.ringFlat + ((size_t)lane * c.R + ring) * 24;
    if (lid < nSharp) oSharp[lid] = adj[pickSharp[lid]];
    if (lid < nFlat) oFlat[lid] = adj[pickFlat[lid]];
    for (int i = lid; i < nLess; i += 32) oLess[i] = adj[pickLess[i]];
}

// Less-flat candidates of one ring -> pcl::VoxelGrid(leaf) (featureAssociation.cpp:771-782).
__global__ void VR_BOUNDS fa_voxel_ring_kernel(IpDev d, FaDev f, Consts c, float leaf) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = blockIdx.y, ring = blockIdx.x, tid = threadIdx.x;
    const int warp = tid >> 5, lid = tid & 31;
    const size_t base = (size_t)lane * c.N;
    const int start = d.startRing[lane * c.R + ring], end = d.endRing[lane * c.R + ring];
    int* rc = f.ringCounts + ((size_t)lane * c.R + ring) * 4;
    if (end + 6 - (start - 4) <= 0) { if (tid == 0) rc[3] = 0; return; }

    // shared memory: voxel index and ring position per candidate, two index lists for the sort
    const int Cw = c.C + 8;
    unsigned* vkey = reinterpret_cast<unsigned*>(smem);
    unsigned short* cpos = reinterpret_cast<unsigned short*>(smem + (size_t)4 * c.C);
    unsigned short* ordA = reinterpret_cast<unsigned short*>(smem + (size_t)6 * c.C);
    unsigned short* ordB = ordA + Cw;
    __shared__ __align__(16) unsigned short s_hist[8 * 256];
    __shared__ int s_sp[6], s_ep[6];
    __shared__ float s_red[6][8];
    __shared__ int s_scan[8];
    __shared__ int s_cnt[FX_MAX_ITERS * 8 + 1];
    if (tid < 6) { int sp, ep; sector_bounds(start, end, tid, &sp, &ep); s_sp[tid] = sp; s_ep[tid] = ep; }
    __syncthreads();

    const float4* adj = f.segAdj + base;
    const int* glabel = f.label + base;
    auto is_cand = [&](int k) -> bool {
        bool in_sector = false;
#pragma unroll
        for (int j = 0; j < 6; ++j) in_sector |= (s_sp[j] < s_ep[j] && k >= s_sp[j] && k <= s_ep[j]);
        return in_sector && glabel[k] <= 0;
    };
    // exclusive scan of the per-(sweep, warp) counts in s_cnt[0 .. nSlots) by warp 0; the total lands in s_cnt[nSlots]
    auto scan_counts = [&](int nSlots) {
        if (warp == 0) {
            int v[FX_MAX_ITERS * 8 / 32], s = 0;
#pragma unroll
            for (int i = 0; i < FX_MAX_ITERS * 8 / 32; ++i) { const int q = lid * (FX_MAX_ITERS * 8 / 32) + i; v[i] = q < nSlots ? s_cnt[q] : 0; s += v[i]; }
            int run = warp_inclusive_scan(s) - s;
#pragma unroll
            for (int i = 0; i < FX_MAX_ITERS * 8 / 32; ++i) { const int q = lid * (FX_MAX_ITERS * 8 / 32) + i; if (q < nSlots) s_cnt[q] = run; run += v[i]; }
            if (lid == 31) s_cnt[nSlots] = run;
        }
    };
    const int nSweep = (end - start + 255) >> 8;      // <= FX_MAX_ITERS (checked on the host)
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int it = 0; it < nSweep; ++it) {
        const int k = start + it * 256 + tid;
        const bool flag = k < end && is_cand(k);
        if (flag) {
            const float4 p = adj[k];
            mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
            mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        if (lid == 0) s_cnt[it * 8 + warp] = __popc(bal);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = warp_min_f(mn[a]); mx[a] = warp_max_f(mx[a]); }
    if (lid == 0) { for (int a = 0; a < 3; ++a) { s_red[a][warp] = mn[a]; s_red[3 + a][warp] = mx[a]; } }
    __syncthreads();
    scan_counts(nSweep * 8);
    __syncthreads();
    const int nCand = s_cnt[nSweep * 8];
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
    // pcl::VoxelGrid refuses a grid whose index would overflow an int and passes the cloud through unfiltered
    const bool overflow = dd[0] * dd[1] * dd[2] > (long long)INT_MAX;
    const long long cells = (long long)divb[0] * divb[1] * divb[2];
    const int keyBits = (overflow || cells <= 1) ? 0 : min(32, 64 - __clzll(cells - 1));
    float4* out = f.ringLessFlat + ((size_t)lane * c.R + ring) * c.C;
    // ordered list of candidates: voxel index, position in the ring, identity permutation
    for (int it = 0; it < nSweep; ++it) {
        const int k = start + it * 256 + tid;
        const bool flag = k < end && is_cand(k);
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        if (flag) {
            const int off = s_cnt[it * 8 + warp] + __popc(bal & ((1u << lid) - 1u));
            const float4 p = adj[k];
            const int i0 = (int)PM_SUB(floorf(PM_MUL(p.x, inv)), (float)minb[0]);
            const int i1 = (int)PM_SUB(floorf(PM_MUL(p.y, inv)), (float)minb[1]);
            const int i2 = (int)PM_SUB(floorf(PM_MUL(p.z, inv)), (float)minb[2]);
            vkey[off] = overflow ? (unsigned)off : (unsigned)((long long)i0 + (long long)i1 * divb[0] + (long long)i2 * divb[0] * divb[1]);
            cpos[off] = (unsigned short)(k - start);
            ordA[off] = (unsigned short)off;
        }
    }
    __syncthreads();
    // block-wide stable radix sort of the permutation by voxel index: warp w owns a contiguous share of the list
    unsigned short* ord = ordA;
    {
        unsigned short* in = ordA; unsigned short* outl = ordB;
        const int per = ((nCand + 255) >> 8) << 5;
        const int wb = min(warp * per, nCand), nw = min(wb + per, nCand) - wb;
        constexpr int U3 = 5;   // entries per lane held in registers at a time
        unsigned* h32 = reinterpret_cast<unsigned*>(s_hist + warp * 256);
        auto key = [&](unsigned short e) -> unsigned { return vkey[e]; };
        for (int shift = 0; shift < keyBits; shift += 8) {
#pragma unroll
            for (int q = 0; q < 8; ++q) s_hist[q * 256 + tid] = 0;
            __syncthreads();
            RadixBatch<U3> r;
            for (int q0 = 0; q0 < nw; q0 += 32 * U3) {
                radix_batch_load(r, in + wb + q0, min(nw - q0, 32 * U3), shift, lid, key);
                radix_batch_count(r, min(nw - q0, 32 * U3), h32, lid);
            }
            __syncthreads();
            {   // digit-major, warp-minor exclusive scan of the 8 x 256 counters: thread = digit
                int run = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q) { const int t = s_hist[q * 256 + tid]; s_hist[q * 256 + tid] = (unsigned short)run; run += t; }
                const int inc = warp_inclusive_scan(run);
                if (lid == 31) s_scan[warp] = inc;
                __syncthreads();
                int basep = inc - run;
                for (int q = 0; q < warp; ++q) basep += s_scan[q];
#pragma unroll
                for (int q = 0; q < 8; ++q) s_hist[q * 256 + tid] = (unsigned short)(s_hist[q * 256 + tid] + basep);
            }
            __syncthreads();
            for (int q0 = 0; q0 < nw; q0 += 32 * U3) {
                if (nw > 32 * U3) radix_batch_load(r, in + wb + q0, min(nw - q0, 32 * U3), shift, lid, key);
                radix_batch_scatter(r, min(nw - q0, 32 * U3), h32, outl, lid);
            }
            __syncthreads();
            unsigned short* t = in; in = outl; outl = t;
        }
        ord = in;
    }
    // one output per distinct voxel, mean of the members in sorted order (float32 sums, then divide)
    const int nHeadSweep = (nCand + 255) >> 8;
    for (int it = 0; it < nHeadSweep; ++it) {
        const int i = it * 256 + tid;
        const bool head = i < nCand && (i == 0 || vkey[ord[i - 1]] != vkey[ord[i]]);
        const unsigned bal = __ballot_sync(0xffffffffu, head);
        if (lid == 0) s_cnt[it * 8 + warp] = __popc(bal);
    }
    __syncthreads();
    scan_counts(nHeadSweep * 8);
    __syncthreads();
    for (int it = 0; it < nHeadSweep; ++it) {
        const int i = it * 256 + tid;
        const bool head = i < nCand && (i == 0 || vkey[ord[i - 1]] != vkey[ord[i]]);
        const unsigned bal = __ballot_sync(0xffffffffu, head);
        if (head) {
            const int off = s_cnt[it * 8 + warp] + __popc(bal & ((1u << lid) - 1u));
            const unsigned key = vkey[ord[i]];
            float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f; int cnt = 0;
            for (int m = i; m < nCand; ++m) {
                const int e = ord[m];
                if (vkey[e] != key) break;
                const float4 p = adj[start + (int)cpos[e]];
                sx = PM_ADD(sx, p.x); sy = PM_ADD(sy, p.y); sz = PM_ADD(sz, p.z); sw = PM_ADD(sw, p.w); ++cnt;
            }
            const float fc = (float)cnt;
            out[off] = make_float4(PM_DIV(sx, fc), PM_DIV(sy, fc), PM_DIV(sz, fc), PM_DIV(sw, fc));
        }
    }
    if (tid == 0) rc[3] = s_cnt[nHeadSweep * 8];
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
__global__ void fa_imu_start_kernel(IpDev d, FaDev f, Consts c);

void Engine::adjustDistortionCalculateSmoothnessMarkOccluded(const double* stamps) {
    cuda_check(cudaMemsetAsync(fa.halfFirst, 0x7F, sizeof(int) * c.L, stream), "memset halfFirst");
    dim3 g((c.N + 255) / 256, c.L);
    if (anyImu && stamps) {
        uploadImu(stamps);
        ProfScope ps(this, "fa_imu_start_kernel"); fa_imu_start_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(ip, fa, c);
        launches += 1;
    }
    { ProfScope ps(this, "fa_prepare_kernel"); fa_prepare_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    { ProfScope ps(this, "fa_distort_kernel"); fa_distort_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    launches += 3;
}

void Engine::extractFeatures() {
    if (c.C > 4096 - 8) { set_last_error("Horizon_SCAN > 4088 is not supported by fa_voxel_ring_kernel"); throw std::length_error("capacity"); }
    const size_t smem = (size_t)6 * c.C + (size_t)4 * (c.C + 8);   // vkey | cpos | two index lists
    static bool attr_set = false;
    if (!attr_set) {
        cuda_check(cudaFuncSetAttribute(fa_voxel_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024), "smem attr");
        attr_set = true;
    }
    { ProfScope ps(this, "fa_sort_sectors_kernel"); fa_sort_sectors_kernel<<<dim3((c.R * 6 + FS_WARPS - 1) / FS_WARPS, c.L), 32 * FS_WARPS, 0, stream>>>(ip, fa, c); }
    { ProfScope ps(this, "fa_select_kernel"); fa_select_kernel<<<dim3((c.R + SEL_WARPS - 1) / SEL_WARPS, c.L), 32 * SEL_WARPS, 0, stream>>>(ip, fa, c); }
    { ProfScope ps(this, "fa_voxel_ring_kernel"); fa_voxel_ring_kernel<<<dim3(c.R, c.L), 256, smem, stream>>>(ip, fa, c, 0.2f); }   // downSizeFilter.setLeafSize(0.2,...) featureAssociation.cpp:225
    { ProfScope ps(this, "fa_gather_features_kernel"); fa_gather_features_kernel<<<dim3(c.R, c.L), 128, 0, stream>>>(fa, c); }
    launches += 4;
}

}  // namespace lego
