// GPU voxel hash for exact nearest-neighbour queries: the replacement for pcl::KdTreeFLANN in the reference's
// correspondence searches (featureAssociation.cpp:1054,1165 scan-to-scan 1-NN; mapOptmization.cpp:1099,1181
// scan-to-map 5-NN).  Points are binned into cubic cells of edge `cell` metres; an open-addressing table maps the
// packed cell coordinate to a contiguous run of a cell-sorted copy of the cloud (xyz + original index in w), so a
// query walks the 27 (or, for the 5 m odometry search, growing shells of) neighbouring cells and streams their
// points from consecutive addresses.
//
// Exactness: every point closer than k*cell lies in a cell whose Chebyshev distance to the query's cell is <= k,
// so a shell search that stops once its current best is <= k*cell is exact; the acceptance thresholds of the
// reference (5th neighbour < 1 m for mapping, nearest < 5 m for odometry) bound k.  Equal distances are resolved
// by the lowest original index, which makes results independent of the order points land inside a cell.
//
// Build = 3 launches over all lanes (blockIdx.y = lane): insert+count, per-lane exclusive scan of the slot
// counts, scatter.  Tables are sized to >= 2x the capacity of the cloud, so probe sequences stay short.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "kernel_utils.cuh"

namespace lego {

struct VoxelHash {
    unsigned long long* keys;   // [L][H]  packed cell coordinate, VH_EMPTY = free
    int* count;                 // [L][H]  points in the cell
    int* start;                 // [L][H]  END of the cell's run in `sorted` (the scatter pass advances it from the run's first position)
    int* slotOf;                // [L][cap] table slot of every input point
    float4* sorted;             // [L][cap] cell-sorted copy, w = original index (int bits)
    int H;                      // slots ALLOCATED per lane, power of two; a build uses only vh_table_size(n, H) of them
    const int* n;               // [lane * nStride] live size of the indexed cloud (device counter)
    int nStride;
    int minH;                   // lower limit of the table size in use (= H switches the live sizing off)
    int cap;                    // points per lane
    float invCell;              // 1 / cell edge
};

#define VH_EMPTY 0xFFFFFFFFFFFFFFFFull

// the 27 cells around a query ordered by box distance class: centre, 6 faces, 12 edges, 8 corners;
// entry = (dx+1) | (dy+1) << 2 | (dz+1) << 4
#define VH_C(dx, dy, dz) (((dx) + 1) | (((dy) + 1) << 2) | (((dz) + 1) << 4))
static __device__ __constant__ unsigned char VH_ORDER27[27] = {
    VH_C(0, 0, 0),
    VH_C(-1, 0, 0), VH_C(1, 0, 0), VH_C(0, -1, 0), VH_C(0, 1, 0), VH_C(0, 0, -1), VH_C(0, 0, 1),
    VH_C(-1, -1, 0), VH_C(1, -1, 0), VH_C(-1, 1, 0), VH_C(1, 1, 0), VH_C(-1, 0, -1), VH_C(1, 0, -1), VH_C(-1, 0, 1), VH_C(1, 0, 1),
    VH_C(0, -1, -1), VH_C(0, 1, -1), VH_C(0, -1, 1), VH_C(0, 1, 1),
    VH_C(-1, -1, -1), VH_C(1, -1, -1), VH_C(-1, 1, -1), VH_C(1, 1, -1), VH_C(-1, -1, 1), VH_C(1, -1, 1), VH_C(-1, 1, 1), VH_C(1, 1, 1)};

__host__ __device__ __forceinline__ unsigned long long vh_pack(int ix, int iy, int iz) {
    return ((unsigned long long)(unsigned)(ix + (1 << 20)) & 0x1FFFFF) |
           (((unsigned long long)(unsigned)(iy + (1 << 20)) & 0x1FFFFF) << 21) |
           (((unsigned long long)(unsigned)(iz + (1 << 20)) & 0x1FFFFF) << 42);
}
__host__ __device__ __forceinline__ unsigned vh_hash(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return (unsigned)k;
}
// Table size actually used for a cloud of n points: the smallest power of two >= 2n (at least 1024, at most the
// allocation).  Sizing by the live count instead of the capacity keeps the clear / scan passes proportional to the
// data.  The table holds CELLS, and a 1 m cell of these clouds holds several points, so the load factor stays
// far below its worst case of 1/2 and look-ups of EMPTY neighbour cells -- the common case of the shell searches --
// end after ~1.1 probes.
__host__ __device__ __forceinline__ int vh_table_size(int n, int Hmax, int Hmin = 1024) {
    int H = Hmin;
    while (H < 2 * n && H < Hmax) H <<= 1;
    return H;
}
__device__ __forceinline__ int vh_cell(float v, float invCell) { return (int)floorf(v * invCell); }

// slot holding `key`, or -1
__device__ __forceinline__ int vh_find(const unsigned long long* __restrict__ keys, int H, unsigned long long key) {
    unsigned s = vh_hash(key) & (unsigned)(H - 1);
    while (true) {
        const unsigned long long k = keys[s];
        if (k == key) return (int)s;
        if (k == VH_EMPTY) return -1;
        s = (s + 1) & (unsigned)(H - 1);
    }
}

// ---- build (kernels live in voxel_hash.cu); n[lane * nStride] = number of points of the lane's cloud ----
__global__ void vh_clear_kernel(VoxelHash h);
__global__ void vh_insert_kernel(VoxelHash h, const float4* __restrict__ pts, size_t ptsStride);
__global__ void vh_scan_kernel(VoxelHash h);
__global__ void vh_scatter_kernel(VoxelHash h, const float4* __restrict__ pts, size_t ptsStride);
__global__ void vh_recount_kernel(VoxelHash h);
// Enqueues the five build launches for all lanes (cloud sizes come from h.n); returns the number of launches.
int vh_build(const VoxelHash& h, int lanes, const float4* pts, size_t ptsStride, cudaStream_t stream);

// ---- query helpers ----
struct VhLane { const unsigned long long* keys; const int* start; const int* count; const float4* sorted; int H; float invCell; };
__device__ __forceinline__ VhLane vh_lane(const VoxelHash& h, int lane) {
    VhLane v;
    v.keys = h.keys + (size_t)lane * h.H; v.start = h.start + (size_t)lane * h.H; v.count = h.count + (size_t)lane * h.H;
    v.sorted = h.sorted + (size_t)lane * h.cap; v.H = vh_table_size(min(h.n[lane * h.nStride], h.cap), h.H, h.minH); v.invCell = h.invCell;
    return v;
}

// squared distance exactly as the reference / PCL evaluate it in float32: ((dx*dx + dy*dy) + dz*dz)
__device__ __forceinline__ float vh_dist2(float4 a, float x, float y, float z) {
    const float dx = __fsub_rn(a.x, x), dy = __fsub_rn(a.y, y), dz = __fsub_rn(a.z, z);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Exact nearest neighbour within sqrt(maxD2): returns the original index or -1; *d2out = its squared distance.
// Shells of cells are visited outwards; shell k only adds the cells at Chebyshev distance exactly k.
__device__ inline int vh_nearest(const VhLane& v, float x, float y, float z, float maxD2, float cellEdge, float* d2out) {
    const int cx = vh_cell(x, v.invCell), cy = vh_cell(y, v.invCell), cz = vh_cell(z, v.invCell);
    float best = maxD2; int bestI = -1;
    const int maxShell = (int)ceilf(sqrtf(maxD2) * v.invCell);
    for (int k = 0; k <= maxShell; ++k) {
        for (int dz = -k; dz <= k; ++dz)
            for (int dy = -k; dy <= k; ++dy) {
                const bool face = (abs(dz) == k) || (abs(dy) == k);
                const int step = face ? 1 : (k == 0 ? 1 : 2 * k);
                for (int dx = -k; dx <= k; dx += step) {
                    const int s = vh_find(v.keys, v.H, vh_pack(cx + dx, cy + dy, cz + dz));
                    if (s < 0) continue;
                    const int e = v.start[s], b = e - v.count[s];
                    for (int t = b; t < e; ++t) {
                        const float4 a = v.sorted[t];
                        const float d = vh_dist2(a, x, y, z);
                        const int ai = __float_as_int(a.w);
                        if (d < best || (d == best && bestI >= 0 && ai < bestI)) { best = d; bestI = ai; }
                    }
                }
            }
        // every unvisited point is farther than k * cellEdge from the query
        const float reach = (float)k * cellEdge;
        if (bestI >= 0 && best <= reach * reach) break;
    }
    *d2out = best;
    return bestI;
}

// Warp-cooperative form of vh_nearest: all 32 lanes of a warp call it with the SAME query; the cells of each shell are
// dealt out to the lanes, so the 27-cell common case costs one probe per lane and the rare 5 m miss (1331 cells) 42.
// Same result as vh_nearest: minimum of (distance, original index) in lexicographic order.
__device__ inline int vh_nearest_warp(const VhLane& v, float x, float y, float z, float maxD2, float cellEdge, float* d2out) {
    const int lid = threadIdx.x & 31;
    const float ux = x * v.invCell, uy = y * v.invCell, uz = z * v.invCell;
    const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
    const int cx = (int)flx, cy = (int)fly, cz = (int)flz;
    // distance from the query to the lower / upper face of its own cell, per axis
    const float lo0 = (ux - flx) * cellEdge, lo1 = (uy - fly) * cellEdge, lo2 = (uz - flz) * cellEdge;
    const float hi0 = cellEdge - lo0, hi1 = cellEdge - lo1, hi2 = cellEdge - lo2;
    float best = maxD2; int bestI = 0x7fffffff;
    auto offer = [&](const float4& a) {
        const float d = vh_dist2(a, x, y, z);
        const int ai = __float_as_int(a.w);
        if (d < best || (d == best && bestI != 0x7fffffff && ai < bestI)) { best = d; bestI = ai; }
    };
    auto warp_min = [&]() {   // warp minimum of (best, bestI); lanes that found nothing carry (maxD2, INT_MAX)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bestI, o);
            if (od < best || (od == best && oi < bestI)) { best = od; bestI = oi; }
        }
    };
    // the query's own cell first, its points dealt out to the lanes: the hit it yields prunes most of the shell
    {
        const int s = vh_find(v.keys, v.H, vh_pack(cx, cy, cz));
        if (s >= 0) {
            const int e = v.start[s], b = e - v.count[s];
            for (int t = b + lid; t < e; t += 32) offer(v.sorted[t]);
        }
        warp_min();
    }
    const int maxShell = (int)ceilf(sqrtf(maxD2) * v.invCell);
    for (int k = 1; k <= max(maxShell, 1); ++k) {
        const int side = 2 * k + 1, ncell = side * side * side;
        for (int c = lid; c < ncell; c += 32) {
            const int dz = c / (side * side) - k, r = c % (side * side), dy = r / side - k, dx = r % side - k;
            // shell k adds the cells at Chebyshev distance exactly k
            if (max(abs(dx), max(abs(dy), abs(dz))) < k) continue;
            // lower bound on the distance to anything inside that cell (with a margin for rounding)
            const float gx = dx < 0 ? lo0 + (float)(-dx - 1) * cellEdge : (dx > 0 ? hi0 + (float)(dx - 1) * cellEdge : 0.f);
            const float gy = dy < 0 ? lo1 + (float)(-dy - 1) * cellEdge : (dy > 0 ? hi1 + (float)(dy - 1) * cellEdge : 0.f);
            const float gz = dz < 0 ? lo2 + (float)(-dz - 1) * cellEdge : (dz > 0 ? hi2 + (float)(dz - 1) * cellEdge : 0.f);
            if (gx * gx + gy * gy + gz * gz > best * 1.0001f + 1e-5f) continue;
            const int s = vh_find(v.keys, v.H, vh_pack(cx + dx, cy + dy, cz + dz));
            if (s < 0) continue;
            const int e = v.start[s], b = e - v.count[s];
            int t = b;
            for (; t + 3 < e; t += 4) {   // four loads in flight
                const float4 a0 = v.sorted[t], a1 = v.sorted[t + 1], a2 = v.sorted[t + 2], a3 = v.sorted[t + 3];
                offer(a0); offer(a1); offer(a2); offer(a3);
            }
            for (; t < e; ++t) offer(v.sorted[t]);
        }
        warp_min();
        const float reach = (float)k * cellEdge;
        if (bestI != 0x7fffffff && best <= reach * reach) break;   // every unvisited point is farther than k cells
    }
    *d2out = best;
    return bestI == 0x7fffffff ? -1 : bestI;
}

// Exact 5 nearest neighbours among the 27 cells around the query, restricted to squared distances <= bound (callers
// pass the reference's acceptance limit: the 5th neighbour must be closer than 1 m, mapOptmization.cpp:1101, 1183, so
// farther points can never matter and one cell edge of reach is exact).  d2[] ascending, ties by original index.
// The centre cell is scanned first; a neighbouring cell is skipped when the distance from the query to its box already
// exceeds the current 5th best (a lower bound on every point inside, taken with a safety margin for rounding), which
// leaves 2-4 of the 27 cells in the common case.  Returns the number of candidates kept (<= 5).
__device__ inline int vh_knn5(const VhLane& v, float x, float y, float z, float bound, float* d2, int* idx) {
    const float ux = x * v.invCell, uy = y * v.invCell, uz = z * v.invCell;
    const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
    const int cx = (int)flx, cy = (int)fly, cz = (int)flz;
    const float cell = 1.0f / v.invCell;
    // distance from the query to the lower / upper face of its own cell, per axis
    const float lo[3] = {(ux - flx) * cell, (uy - fly) * cell, (uz - flz) * cell};
    const float hi[3] = {cell - lo[0], cell - lo[1], cell - lo[2]};
    int found = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) { d2[k] = bound; idx[k] = 0x7fffffff; }
    // visiting order: centre, 6 faces, 12 edges, 8 corners (nearest boxes first, so the bound tightens early)
    for (int o = 0; o < 27; ++o) {
            const int code = VH_ORDER27[o];
            const int dx = (code & 3) - 1, dy = ((code >> 2) & 3) - 1, dz = ((code >> 4) & 3) - 1;
            const float gx = dx < 0 ? lo[0] : (dx > 0 ? hi[0] : 0.f);
            const float gy = dy < 0 ? lo[1] : (dy > 0 ? hi[1] : 0.f);
            const float gz = dz < 0 ? lo[2] : (dz > 0 ? hi[2] : 0.f);
            const float boxD2 = gx * gx + gy * gy + gz * gz;
            if (boxD2 > d2[4] * 1.0001f + 1e-5f) continue;
            const int s = vh_find(v.keys, v.H, vh_pack(cx + dx, cy + dy, cz + dz));
            if (s < 0) continue;
            const int e = v.start[s], b = e - v.count[s];
            for (int t = b; t < e; ++t) {
                const float4 a = v.sorted[t];
                float d = vh_dist2(a, x, y, z);
                int ai = __float_as_int(a.w);
                if (d > d2[4] || (d == d2[4] && ai > idx[4])) continue;
                ++found;
                // insertion into the sorted list of 5 (registers, fully unrolled)
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const bool lt = d < d2[k] || (d == d2[k] && ai < idx[k]);
                    if (lt) { const float td = d2[k]; const int ti = idx[k]; d2[k] = d; idx[k] = ai; d = td; ai = ti; }
                }
            }
        }
    return found < 5 ? found : 5;
}

// ---- 5-NN with a candidate cache for the later LM iterations of one mapping call ---------------------------------
// scan2MapOptimization repeats the same queries up to 10 times while the pose moves by millimetres.  The first search
// of a query (vh_knn5_fill) keeps its VH_CACHE_K nearest candidates, looking VH_CACHE_MARGIN beyond the acceptance
// bound, together with the query position and the squared radius rc2 inside which that list is COMPLETE: the K-th
// distance when K points were found, else the widened bound -- and never more than the distance to the nearest map
// region outside the 27 visited cells.  A later iteration re-ranks the cached candidates at the new position
// (vh_knn5_cached).  A point outside the list was at least sqrt(rc2) from the old position, hence at least
// sqrt(rc2) - delta from the new one (delta = how far the query moved); when the new 5th distance -- or, with fewer
// than five candidates inside the bound, the bound itself -- is smaller than that (with a millimetre of slack for
// float rounding), the re-ranked list IS what vh_knn5 would return: same distances, same tie order, same count.
// Otherwise the caller searches again, so the result never differs from vh_knn5.
#ifndef LEGO_KNN_K
#define LEGO_KNN_K 8
#endif
constexpr int VH_CACHE_K = LEGO_KNN_K;   // multiple of 4 (stored as int4s)
constexpr float VH_CACHE_MARGIN = 0.1f;

// Exact top VH_CACHE_K inside (sqrt(bound) + margin)^2, ties by original index.  d2 / idx: what vh_knn5(bound) returns.
// cand[VH_CACHE_K]: original indices (-1 = none).
__device__ inline int vh_knn5_fill(const VhLane& v, float x, float y, float z, float bound, float* d2, int* idx, int* cand, float* rc2) {
    const float ux = x * v.invCell, uy = y * v.invCell, uz = z * v.invCell;
    const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
    const int cx = (int)flx, cy = (int)fly, cz = (int)flz;
    const float cell = 1.0f / v.invCell;
    const float lo[3] = {(ux - flx) * cell, (uy - fly) * cell, (uz - flz) * cell};
    const float hi[3] = {cell - lo[0], cell - lo[1], cell - lo[2]};
    const float wideR = sqrtf(bound) + VH_CACHE_MARGIN, wide = wideR * wideR;
    float e2[VH_CACHE_K]; int ei[VH_CACHE_K];
    int found = 0;
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) { e2[k] = wide; ei[k] = 0x7fffffff; }
    for (int o = 0; o < 27; ++o) {
        const int code = VH_ORDER27[o];
        const int dx = (code & 3) - 1, dy = ((code >> 2) & 3) - 1, dz = ((code >> 4) & 3) - 1;
        const float gx = dx < 0 ? lo[0] : (dx > 0 ? hi[0] : 0.f);
        const float gy = dy < 0 ? lo[1] : (dy > 0 ? hi[1] : 0.f);
        const float gz = dz < 0 ? lo[2] : (dz > 0 ? hi[2] : 0.f);
        const float boxD2 = gx * gx + gy * gy + gz * gz;
        if (boxD2 > e2[VH_CACHE_K - 1] * 1.0001f + 1e-5f) continue;
        const int s = vh_find(v.keys, v.H, vh_pack(cx + dx, cy + dy, cz + dz));
        if (s < 0) continue;
        const int e = v.start[s], b = e - v.count[s];
        for (int t = b; t < e; ++t) {
            const float4 a = v.sorted[t];
            float d = vh_dist2(a, x, y, z);
            int ai = __float_as_int(a.w);
            if (d > e2[VH_CACHE_K - 1] || (d == e2[VH_CACHE_K - 1] && ai > ei[VH_CACHE_K - 1])) continue;
            ++found;
#pragma unroll
            for (int k = 0; k < VH_CACHE_K; ++k) {
                const bool lt = d < e2[k] || (d == e2[k] && ai < ei[k]);
                if (lt) { const float td = e2[k]; const int ti = ei[k]; e2[k] = d; ei[k] = ai; d = td; ai = ti; }
            }
        }
    }
    int found5 = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const bool in = ei[k] != 0x7fffffff && e2[k] <= bound;
        d2[k] = in ? e2[k] : bound; idx[k] = in ? ei[k] : 0x7fffffff;
        found5 += in ? 1 : 0;
    }
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) cand[k] = ei[k] == 0x7fffffff ? -1 : ei[k];
    // everything nearer than `reach` lies inside the 27 visited cells
    const float reach = cell + fminf(fminf(fminf(lo[0], hi[0]), fminf(lo[1], hi[1])), fminf(lo[2], hi[2]));
    const float rc = fminf(found >= VH_CACHE_K ? sqrtf(e2[VH_CACHE_K - 1]) : wideR, reach);
    *rc2 = rc * rc;
    return found5;
}

// ---- flat form of vh_knn5_fill --------------------------------------------------------------------------------
// Same result as vh_knn5_fill (the top list is the K smallest (distance, index) pairs inside the widened bound, whatever
// order the points are met in), shaped for SIMT execution: ncu showed the nested version running with 9.75 of 32 threads
// active (profiles/r02_ncu_full_top7_vlp16_l296.txt) -- every lane sits in a different cell / point of its own nested loops.
// Here a thread (1) scans its own cell, (2) walks the other 26 cells in ONE uniform loop that only prunes by box distance
// and probes the hash, noting the non-empty cells in a per-thread list in shared memory, (3) runs ONE flat loop over all
// listed points; switching to the next cell costs two loads and a re-check of the (tighter) bound, no probe.
// (distance, index) travel as one 64-bit key -- float bits of a non-negative distance order like the float -- so an
// insertion slot is a min / max pair.  `list` = VH_LIST_CELLS ints of shared memory owned by this thread, stride `stride`.
constexpr int VH_LIST_CELLS = 26;
__device__ __forceinline__ unsigned long long vh_key64(float d, int i) { return ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)i; }

__device__ inline int vh_knn5_fill_flat(const VhLane& v, float x, float y, float z, float bound, float* d2, int* idx, int* cand, float* rc2,
                                        int* list, int stride) {
    const float ux = x * v.invCell, uy = y * v.invCell, uz = z * v.invCell;
    const float flx = floorf(ux), fly = floorf(uy), flz = floorf(uz);
    const int cx = (int)flx, cy = (int)fly, cz = (int)flz;
    const float cell = 1.0f / v.invCell;
    const float lo[3] = {(ux - flx) * cell, (uy - fly) * cell, (uz - flz) * cell};
    const float hi[3] = {cell - lo[0], cell - lo[1], cell - lo[2]};
    const float wideR = sqrtf(bound) + VH_CACHE_MARGIN, wide = wideR * wideR;
    unsigned long long e[VH_CACHE_K];
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) e[k] = vh_key64(wide, 0x7fffffff);
    auto offer = [&](const float4& a) {
        unsigned long long key = vh_key64(vh_dist2(a, x, y, z), __float_as_int(a.w));
        if (key > e[VH_CACHE_K - 1]) return;
#pragma unroll
        for (int k = 0; k < VH_CACHE_K; ++k) { const unsigned long long t = e[k]; e[k] = key < t ? key : t; key = key < t ? t : key; }
    };
    // (1) the query's own cell
    {
        const int s = vh_find(v.keys, v.H, vh_pack(cx, cy, cz));
        if (s >= 0) {
            const int en = v.start[s];
            for (int t = en - v.count[s]; t < en; ++t) offer(v.sorted[t]);
        }
    }
    // (2) the 26 neighbours, nearest boxes first: prune by box distance against the bound so far, probe, note
    int nList = 0;
    for (int o = 1; o < 27; ++o) {
        const int code = VH_ORDER27[o];
        const int dx = (code & 3) - 1, dy = ((code >> 2) & 3) - 1, dz = ((code >> 4) & 3) - 1;
        const float gx = dx < 0 ? lo[0] : (dx > 0 ? hi[0] : 0.f);
        const float gy = dy < 0 ? lo[1] : (dy > 0 ? hi[1] : 0.f);
        const float gz = dz < 0 ? lo[2] : (dz > 0 ? hi[2] : 0.f);
        const float boxD2 = gx * gx + gy * gy + gz * gz;
        if (boxD2 > __uint_as_float((unsigned)(e[VH_CACHE_K - 1] >> 32)) * 1.0001f + 1e-5f) continue;
        const int s = vh_find(v.keys, v.H, vh_pack(cx + dx, cy + dy, cz + dz));
        if (s < 0) continue;
        list[nList * stride] = s | (o << 26);      // H <= 2^26 slots
        ++nList;
    }
    // (3) every listed point in one loop
    int li = 0, t = 0, en = 0;
    while (true) {
        if (t == en) {
            bool got = false;
            while (li < nList) {
                const int packed = list[li * stride]; ++li;
                const int code = VH_ORDER27[packed >> 26];
                const int dx = (code & 3) - 1, dy = ((code >> 2) & 3) - 1, dz = ((code >> 4) & 3) - 1;
                const float gx = dx < 0 ? lo[0] : (dx > 0 ? hi[0] : 0.f);
                const float gy = dy < 0 ? lo[1] : (dy > 0 ? hi[1] : 0.f);
                const float gz = dz < 0 ? lo[2] : (dz > 0 ? hi[2] : 0.f);
                if (gx * gx + gy * gy + gz * gz > __uint_as_float((unsigned)(e[VH_CACHE_K - 1] >> 32)) * 1.0001f + 1e-5f) continue;
                const int s = packed & ((1 << 26) - 1);
                en = v.start[s]; t = en - v.count[s];
                if (t < en) { got = true; break; }
            }
            if (!got) break;
        }
        offer(v.sorted[t]); ++t;
    }
    int found5 = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const float dk = __uint_as_float((unsigned)(e[k] >> 32)); const int ik = (int)(unsigned)e[k];
        const bool in = ik != 0x7fffffff && dk <= bound;
        d2[k] = in ? dk : bound; idx[k] = in ? ik : 0x7fffffff;
        found5 += in ? 1 : 0;
    }
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) { const int ik = (int)(unsigned)e[k]; cand[k] = ik == 0x7fffffff ? -1 : ik; }
    // everything nearer than `reach` lies inside the 27 visited cells
    const float reach = cell + fminf(fminf(fminf(lo[0], hi[0]), fminf(lo[1], hi[1])), fminf(lo[2], hi[2]));
    const bool full = (int)(unsigned)e[VH_CACHE_K - 1] != 0x7fffffff;    // K points inside the widened bound exist
    const float rc = fminf(full ? sqrtf(__uint_as_float((unsigned)(e[VH_CACHE_K - 1] >> 32))) : wideR, reach);
    *rc2 = rc * rc;
    return found5;
}

// Re-ranks the cached candidates at the new query position.  Returns true when the outcome is proven to be what
// vh_knn5(bound) returns there: d2 / idx hold it and *found5 = min(neighbours inside the bound, 5).
__device__ inline bool vh_knn5_cached(const float4* __restrict__ cloud, const int* cand, float ox, float oy, float oz, float rc2,
                                      float x, float y, float z, float bound, float* d2, int* idx, int* found5) {
    int found = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) { d2[k] = bound; idx[k] = 0x7fffffff; }
    float4 pts[VH_CACHE_K];
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) pts[k] = cand[k] >= 0 ? cloud[cand[k]] : make_float4(0.f, 0.f, 0.f, 0.f);   // all gathers in flight
#pragma unroll
    for (int k = 0; k < VH_CACHE_K; ++k) {
        const int ci = cand[k];
        if (ci < 0) continue;
        float d = vh_dist2(pts[k], x, y, z);
        int ai = ci;
        if (d > d2[4] || (d == d2[4] && ai > idx[4])) continue;
        ++found;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const bool lt = d < d2[j] || (d == d2[j] && ai < idx[j]);
            if (lt) { const float td = d2[j]; const int ti = idx[j]; d2[j] = d; idx[j] = ai; d = td; ai = ti; }
        }
    }
    *found5 = found < 5 ? found : 5;
    const float mx = x - ox, my = y - oy, mz = z - oz;
    const float delta = sqrtf(mx * mx + my * my + mz * mz);
    // d2[4] is the 5th distance, or still `bound` when fewer than five candidates lie inside the bound
    return sqrtf(d2[4]) + delta + 1e-3f < sqrtf(rc2);
}

}  // namespace lego
