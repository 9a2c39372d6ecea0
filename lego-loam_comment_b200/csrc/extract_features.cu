// extractFeatures (reference featureAssociation.cpp:680-784) as one CTA per (ring, lane).
//
// Rings are independent in the reference's loop (suppression reaches +-5 slots, the ring margins are 5+5 slots;
// validated in SURVEY.md section 8), so each CTA owns one ring and walks its six sectors in order:
//   phase 0  the ring's window of cloudNeighborPicked / cloudLabel is staged in shared memory together with two
//            derived bytes per point: the static part of the pick conditions (curvature vs threshold, ground flag)
//            and how far the +-5 suppression walk of featureAssociation.cpp:720-732 gets before a column gap > 10;
//   phase 1  six warps sort their sector's {curvature, index} slots: a warp-local stable LSD radix sort gives the
//            order whenever the curvatures are distinct; a sector that holds equal curvatures keeps its arrangement
//            and one thread replays libstdc++'s std::sort on it (std_sort_replay.cuh, quirk Q17);
//            the sort range is [sp, ep) while the scans below include slot ep (quirk Q10);
//   phase 2  warp 0 does the greedy picks with warp ballots: 32 candidates are tested at once, the first eligible
//            one is taken, its neighbourhood is suppressed in registers (and in the shared window for later chunks)
//            and the remaining lanes are re-tested -- identical to the sequential scan because suppression only
//            ever removes candidates;
//   phase 3  the ring's less-flat candidates go through pcl::VoxelGrid(0.2) semantics: bounding box, voxel keys,
//            block-wide stable radix sort by key (input order kept inside a voxel), one thread per voxel averages
//            its members in order.
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

struct RingWin {
    int lo, hi;                       // window [lo, hi) of cloud indices staged in shared memory
    signed char* picked; signed char* label; unsigned char* stat;                            // shared
    const int* gcol; const float* gcurv; int* gpicked; int* glabel; const uint8_t* gground;  // global (lane base)
    float edgeThreshold, surfThreshold;
};
__device__ __forceinline__ bool in_win(const RingWin& w, int i) { return i >= w.lo && i < w.hi; }
// bit 0: may become an edge point (curvature > edgeThreshold, not ground); bit 1: may become a flat point
__device__ __forceinline__ unsigned char pick_stat(float curv, bool ground, float edgeT, float surfT) {
    return (unsigned char)(((curv > edgeT && !ground) ? 1 : 0) | ((curv < surfT && ground) ? 2 : 0));
}
__device__ __forceinline__ int w_stat(const RingWin& w, int i) {
    return in_win(w, i) ? (int)w.stat[i - w.lo] : (int)pick_stat(w.gcurv[i], w.gground[i] != 0, w.edgeThreshold, w.surfThreshold);
}
__device__ __forceinline__ int w_picked(const RingWin& w, int i) { return in_win(w, i) ? (int)w.picked[i - w.lo] : w.gpicked[i]; }
__device__ __forceinline__ void w_set_picked(RingWin& w, int i) { if (in_win(w, i)) w.picked[i - w.lo] = 1; else w.gpicked[i] = 1; }
__device__ __forceinline__ void w_set_label(RingWin& w, int i, int v) { if (in_win(w, i)) w.label[i - w.lo] = (signed char)v; else w.glabel[i] = v; }

// Suppression of the picked point's neighbourhood (featureAssociation.cpp:720-732 and 751-767): the point and up to
// five slots on either side are marked, each walk stopping at the first column gap > 10.  An index below 0 ends the
// backward walk: the reference reads out of bounds there (quirk Q11) and the oracle defines it as "stop".
// General version for a point anywhere in the cloud (only the stale slot-4 entry of quirk Q11 can be outside its own
// ring): lanes 0-4 test the five forward steps, lanes 5-9 the five backward steps, a ballot finds where each walk
// stops, and the lanes before the stop mark their slot.
__device__ __forceinline__ void suppress_warp(RingWin& w, int ind, int lid) {
    bool bad = false; int slot = ind;
    if (lid < 5) {
        const int l = lid + 1;
        slot = ind + l;
        bad = abs(w.gcol[ind + l] - w.gcol[ind + l - 1]) > 10;
    } else if (lid < 10) {
        const int l = -(lid - 4);
        slot = ind + l;
        bad = (ind + l < 0) || abs(w.gcol[ind + l] - w.gcol[ind + l + 1]) > 10;
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

constexpr int FX_MAX_SECTOR = WS_MAX_SORT; // sector sizes the warp sort takes (Horizon_SCAN <= 2496); larger ones use the serial replay
constexpr int FX_MAX_ITERS = 16;         // 256-wide sweeps over one ring (Horizon_SCAN <= 4096)

__global__ void __launch_bounds__(256) fa_extract_kernel(IpDev d, FaDev f, Consts c, float leaf, int dbg) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = blockIdx.y, ring = blockIdx.x, tid = threadIdx.x;
    const int warp = tid >> 5, lid = tid & 31;
    const size_t base = (size_t)lane * c.N;
    const int start = d.startRing[lane * c.R + ring], end = d.endRing[lane * c.R + ring];
    const int rowStart = start - 4, rowEnd = end + 6;
    int* rc = f.ringCounts + ((size_t)lane * c.R + ring) * 4;
    if (rowEnd - rowStart <= 0) { if (tid < 4) rc[tid] = 0; return; }

    // ---- shared memory carve-up ----
    //   region A  8*C bytes   cloudSmoothness slots of the ring (phases 1-2), then voxel keys + positions (phase 3)
    //   region B  4*Cw bytes  column indices while the derived bytes are computed, then the two sort index lists
    //   region C  4*Cw bytes  picked | label | stat | reach, one byte per window slot each
    const int Cw = c.C + 8;
    ssr::Item* items = reinterpret_cast<ssr::Item*>(smem);
    unsigned* vkey = reinterpret_cast<unsigned*>(smem);
    unsigned short* cpos = reinterpret_cast<unsigned short*>(smem + (size_t)4 * c.C);
    int* s_col = reinterpret_cast<int*>(smem + (size_t)8 * c.C);
    unsigned short* ordA = reinterpret_cast<unsigned short*>(smem + (size_t)8 * c.C);
    unsigned short* ordB = ordA + Cw;
    unsigned char* tail = smem + (size_t)8 * c.C + (size_t)4 * Cw;
    signed char* s_picked = reinterpret_cast<signed char*>(tail);
    signed char* s_label = reinterpret_cast<signed char*>(tail + Cw);
    unsigned char* s_stat = tail + 2 * Cw;
    unsigned char* s_reach = tail + 3 * Cw;   // forward steps | backward steps << 4 of the suppression walk
    __shared__ unsigned short s_hist[8 * 256];
    __shared__ int s_sp[6], s_ep[6];
    __shared__ int s_pickSharp[12], s_pickLess[120], s_pickFlat[24], s_nPick[3];
    __shared__ float s_red[6][8];
    __shared__ int s_scan[8];
    __shared__ int s_cnt[FX_MAX_ITERS * 8 + 1];

    RingWin w;
    w.lo = max(rowStart - 1, 0); w.hi = rowEnd;
    w.picked = s_picked; w.label = s_label; w.stat = s_stat;
    w.gcol = d.segCol + base; w.gcurv = f.curv + base; w.gpicked = f.picked + base; w.glabel = f.label + base; w.gground = d.segGround + base;
    w.edgeThreshold = c.edgeThreshold; w.surfThreshold = c.surfThreshold;

    // ---- phase 0: stage the window ----
    for (int k = w.lo + tid; k < w.hi; k += blockDim.x) {
        s_col[k - w.lo] = w.gcol[k];
        s_picked[k - w.lo] = (signed char)w.gpicked[k]; s_label[k - w.lo] = (signed char)w.glabel[k];
        s_stat[k - w.lo] = pick_stat(w.gcurv[k], w.gground[k] != 0, c.edgeThreshold, c.surfThreshold);
    }
    for (int k = start + tid; k < end; k += blockDim.x) {
        ssr::Item it;
        if (k == 4) { it.value = f.slot4val[lane]; it.ind = f.slot4ind[lane]; }   // stale slot, quirk Q11
        else { it.value = w.gcurv[k]; it.ind = k; }
        items[k - start] = it;
    }
    if (tid < 6) { int sp, ep; sector_bounds(start, end, tid, &sp, &ep); s_sp[tid] = sp; s_ep[tid] = ep; }
    if (tid < 3) s_nPick[tid] = 0;
    __syncthreads();
    // reach of the suppression walk from every slot a sector can pick, [start, end): slots k-5 .. k+5 are inside the window
    for (int k = start + tid; k < end; k += blockDim.x) {
        const int* cc = s_col + (k - w.lo);
        int nF = 0, nB = 0;
        while (nF < 5 && abs(cc[nF + 1] - cc[nF]) <= 10) ++nF;
        while (nB < 5 && k - nB - 1 >= 0 && abs(cc[-nB - 1] - cc[-nB]) <= 10) ++nB;
        s_reach[k - w.lo] = (unsigned char)(nF | (nB << 4));
    }
    __syncthreads();   // region B becomes the sort index lists

    // ---- phase 1: std::sort(cloudSmoothness.begin()+sp, cloudSmoothness.begin()+ep) per sector, one warp each ----
    // A sector whose curvatures are all distinct has exactly one sorted order, which the radix sort finds.  Equal
    // curvatures do occur (quirk Q17) and their final order depends on libstdc++'s introsort path: for such a sector
    // the warp replays introsort's partitions first (warp_sort.cuh).
    if (warp < 6) {
        const int sp = s_sp[warp], ep = s_ep[warp];
        const int n = ep - sp, b = sp - start;
        if (n > FX_MAX_SECTOR) {
            if (lid == 0) ssr::std_sort(items + b, n);
        } else if (!(dbg & 1)) {
            warp_std_sort(items + b, n, ordA + b, ordB + b, s_hist + warp * 256, lid);
        }
    }
    __syncthreads();
    if (tid == 0 && start == 4 && s_sp[0] < s_ep[0]) { f.slot4val[lane] = items[0].value; f.slot4ind[lane] = items[0].ind; }

    // ---- phase 2: greedy selection, sectors in order, one warp ----
    // Every lane keeps its candidate's state in registers: index + reach (packed), static eligibility, picked flag.
    // A pick is broadcast with one shuffle; the lanes inside its reach drop out by comparing indices, and lanes
    // 0..10 mark the shared window for the chunks that follow.
    if (warp == 0) {
        int nSharp = 0, nLess = 0, nFlat = 0;
        for (int j = 0; j < ((dbg & 2) ? 0 : 6); ++j) {
            const int sp = s_sp[j], ep = s_ep[j];
            if (sp >= ep) continue;
            for (int pass = 0; pass < 2; ++pass) {
                // pass 0: corner points, k = ep down to sp (featureAssociation.cpp:701-734)
                // pass 1: flat points, k = sp up to ep (featureAssociation.cpp:736-769); its 4th pick breaks BEFORE suppression (Q12)
                int picks = 0; bool stop = false;
                for (int c0 = 0; c0 <= ep - sp && !stop; c0 += 32) {
                    const int k = pass == 0 ? ep - c0 - lid : sp + c0 + lid;
                    const bool inr = k >= sp && k <= ep;
                    const int ind = inr ? items[k - start].ind : 0;
                    const bool fastp = inr && ind >= start && ind < end;
                    const int packed = ind | ((fastp ? (int)s_reach[ind - w.lo] : 0xff) << 24);
                    const bool cand = inr && ((w_stat(w, ind) >> pass) & 1);
                    bool pk = inr ? w_picked(w, ind) != 0 : true;
                    while (true) {
                        const unsigned m = __ballot_sync(0xffffffffu, cand && !pk);
                        if (!m) break;
                        const int first = __ffs(m) - 1;
                        const int pp = __shfl_sync(0xffffffffu, packed, first);
                        const int pind = pp & 0xffffff, reach = (pp >> 24) & 0xff;
                        ++picks;
                        if (pass == 0 && picks > 20) { stop = true; break; }
                        if (lid == 0) {   // the points themselves are copied by the whole CTA after the selection
                            if (pass == 1) { w_set_label(w, pind, -1); s_pickFlat[nFlat] = pind; }
                            else if (picks <= 2) { w_set_label(w, pind, 2); s_pickSharp[nSharp] = pind; s_pickLess[nLess] = pind; }
                            else { w_set_label(w, pind, 1); s_pickLess[nLess] = pind; }
                        }
                        if (pass == 1) ++nFlat; else { ++nLess; if (picks <= 2) ++nSharp; }
                        if (pass == 1 && picks >= 4) { stop = true; break; }
                        if (reach != 0xff) {
                            const int nF = reach & 15, nB = reach >> 4;
                            const int o = lid - 5;
                            if (lid <= 10 && o >= -nB && o <= nF) s_picked[pind + o - w.lo] = 1;
                            if (ind >= pind - nB && ind <= pind + nF) pk = true;
                        } else {
                            suppress_warp(w, pind, lid);
                            __syncwarp();
                            pk = inr ? w_picked(w, ind) != 0 : true;
                        }
                    }
                    __syncwarp();   // marks of this chunk are visible to the loads of the next one
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
    if (dbg & 4) { if (tid == 0) rc[3] = 0; return; }
    const float4* adj = f.segAdj + base;
    auto is_cand = [&](int k) -> bool {
        bool in_sector = false;
#pragma unroll
        for (int j = 0; j < 6; ++j) in_sector |= (s_sp[j] < s_ep[j] && k >= s_sp[j] && k <= s_ep[j]);
        return in_sector && s_label[k - w.lo] <= 0;
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
    __syncthreads();   // also: everyone is done with the cloudSmoothness slots before region A becomes vkey / cpos
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
        unsigned short* h = s_hist + warp * 256;
        auto key = [&](unsigned short e) -> unsigned { return vkey[e]; };
        for (int shift = 0; shift < keyBits; shift += 8) {
#pragma unroll
            for (int q = 0; q < 8; ++q) s_hist[q * 256 + tid] = 0;
            __syncthreads();
            warp_radix_count(in + wb, nw, shift, h, lid, key);
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
            warp_radix_scatter(in + wb, outl, nw, shift, h, lid, key);
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

void Engine::adjustDistortionCalculateSmoothnessMarkOccluded() {
    cuda_check(cudaMemsetAsync(fa.halfFirst, 0x7F, sizeof(int) * c.L, stream), "memset halfFirst");
    dim3 g((c.N + 255) / 256, c.L);
    { ProfScope ps(this, "fa_prepare_kernel"); fa_prepare_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    { ProfScope ps(this, "fa_distort_kernel"); fa_distort_kernel<<<g, 256, 0, stream>>>(ip, fa, c); }
    launches += 3;
}

void Engine::extractFeatures() {
    const size_t smem = (size_t)8 * c.C + (size_t)8 * (c.C + 8);   // regions A, B, C of fa_extract_kernel
    if (c.C > 4096 - 8) { set_last_error("Horizon_SCAN > 4088 is not supported by fa_extract_kernel"); throw std::length_error("capacity"); }
    static bool attr_set = false;
    if (!attr_set) {
        cuda_check(cudaFuncSetAttribute(fa_extract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024), "smem attr");
        attr_set = true;
    }
    static const int dbg = getenv("LEGO_DBG_SKIP") ? atoi(getenv("LEGO_DBG_SKIP")) : 0;
    dim3 g(c.R, c.L);
    { ProfScope ps(this, "fa_extract_kernel"); fa_extract_kernel<<<g, 256, smem, stream>>>(ip, fa, c, 0.2f, dbg); }   // downSizeFilter.setLeafSize(0.2,...) featureAssociation.cpp:225
    { ProfScope ps(this, "fa_gather_features_kernel"); fa_gather_features_kernel<<<dim3(c.R, c.L), 128, 0, stream>>>(fa, c); }
    launches += 2;
}

}  // namespace lego
