// ImageProjection stage as sm_100a kernels: range-image projection, ground removal, connected-component
// segmentation and the ordered compaction into the segmented cloud + cloud_info.
// Replaces the body of ImageProjection::cloudHandler (reference LeGO-LOAM/src/imageProjection.cpp:181-197).
// One launch covers every lane (independent sequence): blockIdx.y = lane.
#include <cfloat>
#include <math_constants.h>
#include <cstdlib>
#include "engine.h"
#include "parity_math.cuh"
#include "kernel_utils.cuh"

namespace lego {

// ------------------------------------------------------------------------------------------------------------
// projectPointCloud (imageProjection.cpp:211-257) + findStartEndAngle (imageProjection.cpp:199-209).
// The reference walks the cloud in input order and lets later points overwrite earlier ones in the same cell
// (quirk Q3); here each point does atomicMax(owner[cell], pointIndex) and a gather pass reads the winner.
// Arithmetic mirrors the C++ expression types: atan2f * 180 is a FLOAT product, the division by M_PI and the
// column computation are double, the result is narrowed to float on assignment.
// An input point: 16-byte x y z intensity, or 12-byte packed x y z (the intensity is never read, imageProjection.cpp:248-255).
__device__ __forceinline__ float4 load_input_point(const float4* __restrict__ in, int i, int floats) {
    if (floats == 4) return in[i];
    const float* q = reinterpret_cast<const float*>(in) + 3 * (size_t)i;
    return make_float4(q[0], q[1], q[2], 0.f);
}

__global__ void __launch_bounds__(256) ip_project_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int P = d.npts[lane];
    const float4* __restrict__ in = d.in_xyzi[lane];
    if (i == 0) {
        float* o = d.orient + lane * 4;
        if (P > 0) {
            const float4 a = load_input_point(in, 0, d.in_floats), b = load_input_point(in, P - 1, d.in_floats);
            const float so = -pm::atan2f_glibc(a.y, a.x);
            float eo = (float)((double)(-pm::atan2f_glibc(b.y, b.x)) + 2.0 * M_PI);
            const float diff = PM_SUB(eo, so);
            if ((double)diff > 3.0 * M_PI) eo = (float)((double)eo - 2.0 * M_PI);
            else if ((double)diff < M_PI) eo = (float)((double)eo + 2.0 * M_PI);
            o[0] = so; o[1] = eo; o[2] = PM_SUB(eo, so);
        }
    }
    if (i >= P) return;
    const float4 p = load_input_point(in, i, d.in_floats);
    long long row;
    if (c.useRing) {
        row = d.in_ring[lane][i];
    } else {
        const float xy = PM_SQRT(PM_ADD(PM_MUL(p.x, p.x), PM_MUL(p.y, p.y)));
        const float va = (float)((double)PM_MUL(pm::atan2f_glibc(p.z, xy), 180.0f) / M_PI);
        const float rf = PM_DIV(PM_ADD(va, c.ang_bottom), c.ang_res_y);
        if (!(rf > -9.0e18f && rf < 9.0e18f)) return;
        row = (long long)rf;   // size_t conversion of the reference truncates toward zero (quirk Q2)
    }
    if (row < 0 || row >= c.R) return;
    const float ha = (float)((double)PM_MUL(pm::atan2f_glibc(p.x, p.y), 180.0f) / M_PI);
    const double colf = -round(((double)ha - 90.0) / (double)c.ang_res_x) + (double)(c.C / 2);
    if (!(colf > -9.0e18 && colf < 9.0e18)) return;
    long long col = (long long)colf;
    if (col >= c.C) col -= c.C;
    if (col < 0 || col >= c.C) return;
    const float range = PM_SQRT(PM_ADD(PM_ADD(PM_MUL(p.x, p.x), PM_MUL(p.y, p.y)), PM_MUL(p.z, p.z)));
    if (range < c.sensorMinimumRange) return;
    atomicMax(d.owner + (size_t)lane * c.N + row * c.C + col, i);
}

// Gather pass: one thread per image cell.  Writes rangeMat / fullCloud exactly as the sequential loop would
// have left them, and resets the per-cell segmentation scratch (union-find parent, component statistics).
__global__ void __launch_bounds__(256) ip_gather_kernel(IpDev d, Consts c, int resetCcl) {
    const int lane = blockIdx.y;
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= c.N) return;
    const size_t g = (size_t)lane * c.N + cell;
    const int o = d.owner[g];
    float4 out; float range;
    if (o < 0) {
        out = make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, -1.0f);
        range = FLT_MAX;
    } else {
        const float4 p = load_input_point(d.in_xyzi[lane], o, d.in_floats);
        range = PM_SQRT(PM_ADD(PM_ADD(PM_MUL(p.x, p.x), PM_MUL(p.y, p.y)), PM_MUL(p.z, p.z)));
        const int row = cell / c.C, col = cell - row * c.C;
        const float inten = (float)((double)(float)row + (double)(float)col / 10000.0);
        out = make_float4(p.x, p.y, p.z, inten);
    }
    d.fullCloud[g] = out;
    d.rangeMat[g] = range;
    d.groundMat[g] = 0;
    if (resetCcl) {   // (the tiled connected-components path writes what it reads itself)
        d.parent[g] = cell;
        d.compSize[g] = 0;
        for (int w = 0; w < c.rowWords; ++w) d.compRows[g * c.rowWords + w] = 0u;
    }
}

// groundRemoval (imageProjection.cpp:260-291): the only sequential dependence is along a column (a -1 written
// for pair i may overwrite the 1 written by pair i-1 and vice versa, quirk Q4), so one thread walks one column.
__global__ void __launch_bounds__(128) ip_ground_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= c.C) return;
    const size_t base = (size_t)lane * c.N;
    const float4* fc = d.fullCloud + base;
    int8_t* gm = d.groundMat + base;
    float4 lower = fc[j];
    for (int i = 0; i < c.G; ++i) {
        if (i + 1 >= c.R) break;
        const float4 upper = fc[j + (i + 1) * c.C];
        if (lower.w == -1.0f || upper.w == -1.0f) {
            gm[j + i * c.C] = -1;
        } else {
            const float dx = PM_SUB(upper.x, lower.x), dy = PM_SUB(upper.y, lower.y), dz = PM_SUB(upper.z, lower.z);
            const float hyp = PM_SQRT(PM_ADD(PM_MUL(dx, dx), PM_MUL(dy, dy)));
            const float angle = (float)((double)PM_MUL(pm::atan2f_glibc(dz, hyp), 180.0f) / M_PI);
            if (fabsf(PM_SUB(angle, c.sensorMountAngle)) <= 10.0f) {
                gm[j + i * c.C] = 1;
                gm[j + (i + 1) * c.C] = 1;
            }
        }
        lower = upper;
    }
}

// A cell takes part in segmentation iff the reference would have initialised its label to 0:
// not ground and holding a return (imageProjection.cpp:295-301).
__device__ __forceinline__ bool seg_cell(const IpDev& d, size_t g) { return d.owner[g] >= 0 && d.groundMat[g] != 1; }

__device__ __forceinline__ int uf_find(int* parent, int i) {
    volatile int* vp = parent;
    int p;
    while ((p = vp[i]) != i) i = p;
    return i;
}
// link by smaller index: the root of a finished component is its smallest cell index = the BFS seed the
// reference's row-major driver would have started from (imageProjection.cpp:314-317).
__device__ __forceinline__ void uf_union(int* parent, int a, int b) {
    while (true) {
        a = uf_find(parent, a); b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }
        const int old = atomicMin(parent + a, b);
        if (old == a) return;
        a = old;
    }
}

// Edge test of labelComponents (imageProjection.cpp:411-423): symmetric in the two cells, so the BFS flood fill
// is plain connected components of the graph these edges define (validated in SURVEY.md section 8).
__device__ __forceinline__ bool seg_edge(float r1, float r2, float sinA, float cosA, const Consts& c) {
    const float d1 = fmaxf(r1, r2), d2 = fminf(r1, r2);
    const float y = PM_MUL(d2, sinA), x = PM_SUB(d1, PM_MUL(d2, cosA));
    // atan2f(y, x) > theta is y / x > tan(theta) in the first quadrant: a ratio 1e-4 clear of tan(theta) decides without the
    // arc tangent (the restated glibc atan2f is exact to the last bit but ~150 instructions, and two edge tests per cell were
    // most of the segmentation's time); only the band around the threshold, 4e-5 rad wide against 1e-7 of float rounding,
    // evaluates it.  segTanHi = 0 (theta outside (0, pi/2)) switches the shortcut off.
    if (c.segTanHi > 0.0f && x > 0.0f && y >= 0.0f && x < 1e30f && y < 1e30f) {
        if (y > x * c.segTanHi) return true;
        if (y < x * c.segTanLo) return false;
    }
    return pm::atan2f_glibc(y, x) > c.segmentTheta;
}

// Connected components in three passes.  (1) ip_ccl_runs_kernel: horizontal runs.  Every cell tests the edge to its
// left neighbour once; a run of connected cells inside a 256-cell block gets its leftmost cell as parent by a ballot
// scan, without atomics, and its last cell remembers the run length.  (2) ip_ccl_union_kernel: the remaining edges --
// vertical ones, the column wrap, and the horizontal edges a block boundary cut -- are merged with the lock-free
// min-index union.  (3) ip_ccl_stats_kernel: every cell is pointed at its root, and every RUN (not every cell) adds
// its length and row bit to the root's statistics, which takes the atomic traffic on large components down by the
// run length.
constexpr int RUN_LEFT_BIT = 1 << 30;

__global__ void __launch_bounds__(256) ip_ccl_runs_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y, tid = threadIdx.x, lid = tid & 31, warp = tid >> 5;
    const int cell = blockIdx.x * blockDim.x + tid;
    const size_t base = (size_t)lane * c.N;
    const bool valid = cell < c.N;
    const int col = valid ? cell % c.C : 0;
    // loads first, decisions after: one DRAM round trip instead of three dependent ones (the left neighbour's values sit in the
    // same lines as this thread's)
    const size_t g = base + (valid ? cell : 0), gl = valid && cell > 0 ? g - 1 : g;
    const int own = d.owner[g], ownL = d.owner[gl];
    const unsigned char gm = d.groundMat[g], gmL = d.groundMat[gl];
    const float r = d.rangeMat[g], rL = d.rangeMat[gl];
    const bool seg = valid && own >= 0 && gm != 1;
    bool leftConn = false;
    if (seg && tid > 0 && col > 0 && ownL >= 0 && gmL != 1)
        leftConn = seg_edge(r, rL, c.sinAlphaX, c.cosAlphaX, c);
    const bool brk = !leftConn;
    __shared__ int s_lastBrk[8];
    __shared__ unsigned char s_brk[257];
    const unsigned mask = __ballot_sync(0xffffffffu, brk);
    if (lid == 0) s_lastBrk[warp] = mask ? (warp << 5) + 31 - __clz(mask) : -1;
    s_brk[tid] = brk ? 1 : 0;
    if (tid == 0) s_brk[256] = 1;
    __syncthreads();
    if (!valid) return;
    const unsigned lower = mask & (lid == 31 ? 0xffffffffu : ((2u << lid) - 1u));
    int leaderT;
    if (lower) leaderT = (warp << 5) + 31 - __clz(lower);
    else { leaderT = 0; for (int w = warp - 1; w >= 0; --w) if (s_lastBrk[w] >= 0) { leaderT = s_lastBrk[w]; break; } }
    const int leader = cell - tid + leaderT;
    const bool isLast = seg && (s_brk[tid + 1] != 0);
    if (seg) d.parent[base + cell] = leader;
    // run length at the last cell; bit 30 tells the union pass that this cell hangs on its left neighbour
    d.runLen[base + cell] = (isLast ? cell - leader + 1 : 0) | (leftConn ? RUN_LEFT_BIT : 0);
}

__global__ void __launch_bounds__(256) ip_ccl_union_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y, tid = threadIdx.x;
    const int cell = blockIdx.x * blockDim.x + tid;
    const size_t base = (size_t)lane * c.N;
    const bool in = cell < c.N;
    const int row = in ? cell / c.C : 0, col = cell - row * c.C;
    // Everything the vertical edge needs is loaded up front and unconditionally: the kernel is a chain of dependent DRAM
    // round trips otherwise (cell class -> class of the cell below -> its range -> the run bits; L1 hit rate 1 %), and with all
    // loads in flight at once it is one.
    const bool below = in && row + 1 < c.R;
    const size_t g = base + (in ? cell : 0), gd = below ? g + c.C : g;
    const int own = d.owner[g], ownD = d.owner[gd];
    const unsigned char gm = d.groundMat[g], gmD = d.groundMat[gd];
    const float r = d.rangeMat[g], rD = d.rangeMat[gd];
    const int rl = d.runLen[g], rlD = d.runLen[gd];
    const bool seg = in && own >= 0 && gm != 1;
    int* parent = d.parent + base;
    // vertical neighbour, rows do not wrap (imageProjection.cpp:400-401)
    bool vert = false, hang = false;
    if (seg && below && ownD >= 0 && gmD != 1) {
        vert = seg_edge(r, rD, c.sinAlphaY, c.cosAlphaY, c);
        // both this cell and the one below hang on their left neighbours: if the left pair is joined
        // vertically too, this edge joins the same two runs again
        hang = (rl & RUN_LEFT_BIT) && (rlD & RUN_LEFT_BIT);
    }
    __shared__ unsigned char s_vert[256];
    s_vert[tid] = vert ? 1 : 0;
    __syncthreads();
    if (!seg) return;
    if (vert && !(hang && tid > 0 && s_vert[tid - 1])) uf_union(parent, cell, cell + c.C);
    // horizontal edge a block boundary cut out of the run pass
    if (tid == 0 && col > 0 && seg_cell(d, base + cell - 1) &&
        seg_edge(r, d.rangeMat[base + cell - 1], c.sinAlphaX, c.cosAlphaX, c))
        uf_union(parent, cell, cell - 1);
    // columns wrap (imageProjection.cpp:403-406)
    if (col == c.C - 1 && c.C > 1) {
        const int rc = row * c.C;
        if (seg_cell(d, base + rc) && seg_edge(r, d.rangeMat[base + rc], c.sinAlphaX, c.cosAlphaX, c))
            uf_union(parent, cell, rc);
    }
}

// Component statistics: size and the set of rows among NON-seed members.  The reference sets lineCountFlag only
// for cells it pushes, never for the seed itself (imageProjection.cpp:376-384 vs 431, quirk Q5).  The seed is the
// smallest index of the component, i.e. the leader of its own run.
__global__ void __launch_bounds__(256) ip_ccl_stats_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y;
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= c.N) return;
    const size_t base = (size_t)lane * c.N;
    int* parent = d.parent + base;
    // independent loads first (see ip_ccl_union_kernel); any value on a cell's parent chain leads to its root
    const int own = d.owner[base + cell];
    const unsigned char gm = d.groundMat[base + cell];
    const int p0 = parent[cell];
    const int rl = d.runLen[base + cell];
    if (!(own >= 0 && gm != 1)) return;
    const int root = p0 == cell ? cell : uf_find(parent, p0);
    if (root != cell) parent[cell] = root;
    const int len = rl & ~RUN_LEFT_BIT;
    if (len > 0) {
        const int leader = cell - len + 1;
        const int n = len - (leader == root ? 1 : 0);
        if (n > 0) {
            atomicAdd(d.compSize + base + root, n);
            const int row = cell / c.C;
            atomicOr(d.compRows + (base + root) * c.rowWords + (row >> 5), 1u << (row & 31));
        }
    }
}

// ---- tiled connected components ------------------------------------------------------------------------------------------
// The three passes above send every vertical edge through an atomicMin chain in L2 (ip_ccl_union_kernel: L1 hit rate 1 %).
// Here a CTA owns a tile of ALL rows x TW columns (<= 4096 cells) and resolves it in shared memory:
//   ip_ccl_tile_kernel    ranges staged in shared memory, both edge tests per cell kept as ballot words, horizontal runs get
//                         their first cell as parent from those words (no atomics), vertical edges are lock-free min-index
//                         unions on the shared-memory forest of run heads (both ends pointed at the surviving root), then per
//                         tile component (rooted at its smallest cell) the number of members besides that root and the row
//                         bits of those members.  Row-major order inside a tile is the order of the image, so "smallest local
//                         index" = "smallest cell index".
//   ip_ccl_border_kernel  the edges between neighbouring tiles and the column wrap (R per border) joined on the global forest
//                         the tile roots form: a few hundred unions per image instead of one per run.
//   ip_ccl_finish_kernel  every cell is pointed at its final root; every tile root that is not the final root hands over its
//                         statistics plus itself: the seed never counts (quirk Q5), every other cell does exactly once.
// Results are the same forest roots and the same statistics at the roots as the three-pass form (LEGO_CCL_TILE=0), which the
// tests compare cell by cell.  cellClass doubles as the pass-to-pass marker: 0 = no segmentation cell, 1 = member, 2 = tile root.
constexpr int CCL_MAX_TILE = 4096;
constexpr int CCL_THREADS = 256;

__device__ __forceinline__ int s_find(const int* lab, int i) {
    const volatile int* vl = lab;
    int p;
    while ((p = vl[i]) != i) i = p;
    return i;
}
// Parents only ever move to smaller indices of the same set, so lowering a parent to any known ancestor with atomicMin is
// always valid: both ends of a union are pointed at the surviving root (without it the row-by-row links of a component
// spanning the whole tile form chains that every later find walks: 40 % of the kernel's instructions in the first version).
__device__ __forceinline__ void s_union(int* lab, int a0, int b0) {
    int a = a0, b = b0;
    while (true) {
        a = s_find(lab, a); b = s_find(lab, b);
        if (a == b) break;
        if (a < b) { const int t = a; a = b; b = t; }
        const int old = atomicMin(lab + a, b);
        if (old == a) { a = b; break; }
        a = old;
    }
    if (a0 != a) atomicMin(lab + a0, a);
    if (b0 != a) atomicMin(lab + b0, a);
}

#ifndef LEGO_CCL_MINBLOCKS
#define LEGO_CCL_MINBLOCKS 8    // 32 registers, no spills: the kernel waits at its barriers, more resident tiles fill the gaps (9.1 -> 7.7 ms)
#endif
#define CCL_BOUNDS __launch_bounds__(CCL_THREADS, LEGO_CCL_MINBLOCKS)
__global__ void CCL_BOUNDS ip_ccl_tile_kernel(IpDev d, Consts c, int TW, int twShift) {
    extern __shared__ int cclSmem[];
    const int T = c.R * TW;                               // cells of this tile (multiple of 32)
    int* lab = cclSmem;                                   // [T]      union-find forest, local indices
    unsigned* sz2 = reinterpret_cast<unsigned*>(lab + T); // [T / 2]  members besides the root, two 16-bit counters per word
    unsigned* rows = sz2 + T / 2;                         // [T][rowWords] row bits of those members; first holds the staged ranges
    float* rng = reinterpret_cast<float*>(rows);          // [T]      range of a segmentation cell, -1 elsewhere
    __shared__ unsigned leftM[CCL_MAX_TILE / 32], downM[CCL_MAX_TILE / 32];   // edge bits, one word per warp step
    const int lane = blockIdx.y, tid = threadIdx.x, lid = tid & 31;
    const int col0 = blockIdx.x * TW;
    const size_t base = (size_t)lane * c.N;
    unsigned segBits = 0;
    // four cells per trip with their twelve loads in flight together: the tile starts with a cold DRAM round trip per load
    for (int k0 = 0; k0 * CCL_THREADS < T; k0 += 4) {
        int own[4]; unsigned char gm[4]; float rv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = (k0 + u) * CCL_THREADS + tid;
            const int row = i >> twShift, col = col0 + (i & (TW - 1));
            const bool in = i < T && col < c.C;
            const size_t g = base + (in ? (size_t)row * c.C + col : 0);
            own[u] = in ? d.owner[g] : -1; gm[u] = in ? d.groundMat[g] : (unsigned char)1; rv[u] = in ? d.rangeMat[g] : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = (k0 + u) * CCL_THREADS + tid;
            if (i >= T) break;
            const bool seg = own[u] >= 0 && gm[u] != 1;
            if (seg) segBits |= 1u << (k0 + u);
            rng[i] = seg ? rv[u] : -1.0f;
            if ((i & 1) == 0) sz2[i >> 1] = 0u;
        }
    }
    __syncthreads();
    // both edge tests of every cell (imageProjection.cpp:411-423), kept as one bit per cell
    for (int k = 0; k * CCL_THREADS < T; ++k) {
        const int i = k * CCL_THREADS + tid;
        if (i >= T) break;                                // (warp-uniform: T is a multiple of 32)
        bool left = false, down = false;
        if ((segBits >> k) & 1u) {
            const float r = rng[i];
            if ((i & (TW - 1)) > 0) {
                const float rl = rng[i - 1];
                left = rl >= 0.0f && seg_edge(r, rl, c.sinAlphaX, c.cosAlphaX, c);
            }
            if (i + TW < T) {                             // rows do not wrap (imageProjection.cpp:400-401)
                const float rd = rng[i + TW];
                down = rd >= 0.0f && seg_edge(r, rd, c.sinAlphaY, c.cosAlphaY, c);
            }
        }
        const unsigned lm = __ballot_sync(0xffffffffu, left), dm = __ballot_sync(0xffffffffu, down);
        if (lid == 0) { leftM[i >> 5] = lm; downM[i >> 5] = dm; }
    }
    __syncthreads();
    for (int q = tid; q < T * c.rowWords; q += CCL_THREADS) rows[q] = 0u;   // (the ranges are dead from here on)
    // horizontal runs without atomics: a cell's parent is the first cell of its run = the nearest cell at or before it that
    // does not hang on its left neighbour (column 0 of the tile never does, so the walk stays inside the row)
    for (int k = 0; k * CCL_THREADS < T; ++k) {
        const int i = k * CCL_THREADS + tid;
        if (i >= T) break;
        int w = i >> 5;
        unsigned m = ~leftM[w] & (lid == 31 ? 0xffffffffu : ((2u << lid) - 1u));
        while (!m) m = ~leftM[--w];
        lab[i] = (w << 5) + 31 - __clz(m);
    }
    __syncthreads();
    // vertical edges on the shared-memory forest; an edge whose two cells both hang on their left neighbours is skipped when
    // the left pair is joined vertically too (it joins the same two runs again)
    for (int k = 0; k * CCL_THREADS < T; ++k) {
        const int i = k * CCL_THREADS + tid;
        if (i >= T) break;
        if (!((downM[i >> 5] >> lid) & 1u)) continue;
        const int below = i + TW;
        const bool hang = ((leftM[i >> 5] >> lid) & 1u) && ((leftM[below >> 5] >> (below & 31)) & 1u) && ((downM[(i - 1) >> 5] >> ((i - 1) & 31)) & 1u);
        if (!hang) s_union(lab, lab[i], lab[below]);   // from the runs' first cells: pointing THEM at the root shortens every member's walk
    }
    __syncthreads();
    // roots, then the statistics: the 32 cells of a warp step lie in one row, mostly in one component -> one atomic per group
    int rootOf[CCL_MAX_TILE / CCL_THREADS];
#pragma unroll
    for (int k = 0; k < CCL_MAX_TILE / CCL_THREADS; ++k) {
        const int i = k * CCL_THREADS + tid;
        rootOf[k] = -1;
        if (i < T && ((segBits >> k) & 1u)) rootOf[k] = s_find(lab, i);
    }
#pragma unroll
    for (int k = 0; k < CCL_MAX_TILE / CCL_THREADS; ++k) {
        const int i = k * CCL_THREADS + tid;
        if (i >= T) break;                                // warp-uniform
        const int root = rootOf[k];
        const unsigned grp = __match_any_sync(0xffffffffu, root);
        const unsigned self = __ballot_sync(0xffffffffu, root == i);
        if (root >= 0 && lid == __ffs(grp) - 1) {
            const int n = __popc(grp) - ((grp & self) ? 1 : 0);
            if (n > 0) {
                const int row = i >> twShift;
                atomicAdd(sz2 + (root >> 1), (unsigned)n << ((root & 1) * 16));
                atomicOr(rows + root * c.rowWords + (row >> 5), 1u << (row & 31));
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < CCL_MAX_TILE / CCL_THREADS; ++k) {
        const int i = k * CCL_THREADS + tid;
        if (i >= T) break;
        const int row = i >> twShift, col = col0 + (i & (TW - 1));
        if (col >= c.C) continue;
        const size_t g = base + (size_t)row * c.C + col;
        const int root = rootOf[k];
        d.cellClass[g] = root < 0 ? 0 : (root == i ? 2 : 1);
        if (root < 0) continue;
        d.parent[g] = (root >> twShift) * c.C + col0 + (root & (TW - 1));
        if (root == i) {
            d.compSize[g] = (int)((sz2[i >> 1] >> ((i & 1) * 16)) & 0xffffu);
            for (int w = 0; w < c.rowWords; ++w) d.compRows[g * c.rowWords + w] = rows[i * c.rowWords + w];
        }
    }
}

__global__ void __launch_bounds__(256) ip_ccl_border_kernel(IpDev d, Consts c, int TW, int nTiles) {
    const int lane = blockIdx.x;
    const size_t base = (size_t)lane * c.N;
    int* parent = d.parent + base;
    for (int e = threadIdx.x; e < nTiles * c.R; e += blockDim.x) {
        const int t = e / c.R, row = e - t * c.R;
        // border t >= 1: last column of tile t-1 | first column of tile t; t == 0: the column wrap (imageProjection.cpp:403-406)
        if (t == 0 && c.C <= 1) continue;
        const int colB = t * TW, colA = t == 0 ? c.C - 1 : colB - 1;
        const int a = row * c.C + colA, b = row * c.C + colB;
        if (!d.cellClass[base + a] || !d.cellClass[base + b]) continue;
        if (seg_edge(d.rangeMat[base + a], d.rangeMat[base + b], c.sinAlphaX, c.cosAlphaX, c)) uf_union(parent, a, b);
    }
}

__global__ void __launch_bounds__(256) ip_ccl_finish_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y;
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= c.N) return;
    const size_t base = (size_t)lane * c.N;
    const int mark = d.cellClass[base + cell];
    const int p0 = d.parent[base + cell];
    if (!mark) return;
    int* parent = d.parent + base;
    const int root = p0 == cell ? cell : uf_find(parent, p0);
    if (root == cell) return;
    parent[cell] = root;
    if (mark == 2) {   // a tile root that lost to a smaller cell: its members and itself join the final root's statistics
        atomicAdd(d.compSize + base + root, d.compSize[base + cell] + 1);
        const int row = cell / c.C;
        for (int w = 0; w < c.rowWords; ++w) {
            const unsigned bits = d.compRows[(base + cell) * c.rowWords + w] | (w == (row >> 5) ? 1u << (row & 31) : 0u);
            if (bits) atomicOr(d.compRows + (base + root) * c.rowWords + w, bits);
        }
    }
}

// feasibleSegment of imageProjection.cpp:441-451 for the component rooted at `root`.
__device__ __forceinline__ bool comp_valid(const IpDev& d, const Consts& c, size_t base, int root) {
    const int size = d.compSize[base + root] + 1;
    if (size >= 30) return true;
    if (size < c.segmentValidPointNum) return false;
    int lines = 0;
    for (int w = 0; w < c.rowWords; ++w) lines += __popc(d.compRows[(base + root) * c.rowWords + w]);
    return lines >= c.segmentValidLineNum;
}

// Per-cell classification shared by the counting and the writing pass (imageProjection.cpp:326-351).
// returns bit0 = cell heads a valid segment, bit1 = cell goes to segmentedCloud, bit2 = cell goes to outlierCloud,
// bit3 = cell heads a segment that is NOT valid (its label becomes 999999), bit4 = ground cell
__device__ __forceinline__ int classify_cell(const IpDev& d, const Consts& c, size_t base, int row, int col) {
    const int cell = row * c.C + col;
    const size_t g = base + cell;
    const int own = d.owner[g];
    const bool ground = d.groundMat[g] == 1;
    const int root = d.parent[g];          // loaded with the other two, used only for segmentation cells
    int f = ground ? 16 : 0;
    if (own >= 0 && !ground) {
        const bool valid = comp_valid(d, c, base, root);
        if (valid) { if (root == cell) f |= 1; f |= 2; }
        else {
            if (root == cell) f |= 8;
            if (row > c.G && col % 5 == 0) f |= 4;
        }
    } else if (ground) {
        if (!(col % 5 != 0 && col > 5 && col < c.C - 5)) f |= 2;
    }
    return f;
}

__global__ void __launch_bounds__(256) ip_row_count_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y, row = blockIdx.x;
    const size_t base = (size_t)lane * c.N;
    int n0 = 0, n1 = 0, n2 = 0;
    for (int col = threadIdx.x; col < c.C; col += blockDim.x) {
        const int f = classify_cell(d, c, base, row, col);
        d.cellClass[base + row * c.C + col] = (unsigned char)f;   // read back by ip_compact_kernel
        n0 += f & 1; n1 += (f >> 1) & 1; n2 += (f >> 2) & 1;
    }
    __shared__ int sh[3][8];
    n0 = warp_sum(n0); n1 = warp_sum(n1); n2 = warp_sum(n2);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sh[0][w] = n0; sh[1][w] = n1; sh[2][w] = n2; }
    __syncthreads();
    if (threadIdx.x < 3) {
        int s = 0;
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += sh[threadIdx.x][k];
        d.rowCounts[((size_t)lane * c.R + row) * 4 + threadIdx.x] = s;
    }
}

// Ordered compaction (imageProjection.cpp:319-355): row-major order is kept by giving every row block the
// prefix of the counts of the rows before it and ranking inside the row: a first sweep counts the three classes per
// (256-column sweep, warp) with ballots, one warp per class scans those counts, a second sweep writes.  Also numbers
// the valid segments 1..K in seed order (imageProjection.cpp:453-454) and marks the rest 999999 at their root cells.
constexpr int COMPACT_MAX_SWEEPS = 16;   // Horizon_SCAN <= 4096
__global__ void __launch_bounds__(256) ip_compact_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y, row = blockIdx.x, tid = threadIdx.x;
    const size_t base = (size_t)lane * c.N;
    __shared__ int pre[3];
    __shared__ int cnt[3][COMPACT_MAX_SWEEPS * 8 + 1];
    const int w = tid >> 5, l = tid & 31;
    const int nSweep = (c.C + 255) >> 8;
    if (tid < 3) {
        int s = 0;
        for (int r = 0; r < row; ++r) s += d.rowCounts[((size_t)lane * c.R + r) * 4 + tid];
        pre[tid] = s;
    }
    const unsigned char* cls = d.cellClass + base + row * c.C;
    for (int it = 0; it < nSweep; ++it) {
        const int col = it * 256 + tid;
        const int f = col < c.C ? cls[col] : 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const unsigned bal = __ballot_sync(0xffffffffu, (f >> k) & 1);
            if (l == 0) cnt[k][it * 8 + w] = __popc(bal);
        }
    }
    __syncthreads();
    if (w < 3) {   // exclusive scan of class w's counts, total in the last slot
        const int nSlots = nSweep * 8;
        int v[COMPACT_MAX_SWEEPS * 8 / 32], s = 0;
#pragma unroll
        for (int i = 0; i < COMPACT_MAX_SWEEPS * 8 / 32; ++i) { const int q = l * (COMPACT_MAX_SWEEPS * 8 / 32) + i; v[i] = q < nSlots ? cnt[w][q] : 0; s += v[i]; }
        int run = warp_inclusive_scan(s) - s;
#pragma unroll
        for (int i = 0; i < COMPACT_MAX_SWEEPS * 8 / 32; ++i) { const int q = l * (COMPACT_MAX_SWEEPS * 8 / 32) + i; if (q < nSlots) cnt[w][q] = run; run += v[i]; }
        if (l == 31) cnt[w][nSlots] = run;
    }
    __syncthreads();
    if (tid == 0) d.startRing[lane * c.R + row] = pre[1] - 1 + 5;
    const unsigned lt = (1u << l) - 1u;
    for (int it = 0; it < nSweep; ++it) {
        const int col = it * 256 + tid;
        const int f = col < c.C ? cls[col] : 0;
        int off[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const unsigned bal = __ballot_sync(0xffffffffu, (f >> k) & 1);
            off[k] = pre[k] + cnt[k][it * 8 + w] + __popc(bal & lt);
        }
        if (col < c.C) {
            const size_t g = base + row * c.C + col;
            if (f & 1) d.labelMat[g] = off[0] + 1;
            else if (f & 8) d.labelMat[g] = 999999;
            if (f & 2) {
                const size_t o = base + off[1];
                d.segCloud[o] = d.fullCloud[g];
                d.segGround[o] = (f & 16) ? 1 : 0;
                d.segCol[o] = col;
                d.segRange[o] = d.rangeMat[g];
            }
            if ((f & 4) && off[2] < c.Ocap) d.outlier[(size_t)lane * c.Ocap + off[2]] = d.fullCloud[g];
        }
    }
    if (tid == 0) {
        const int nSlots = nSweep * 8;
        d.endRing[lane * c.R + row] = pre[1] + cnt[1][nSlots] - 1 - 5;
        if (row == c.R - 1) {
            d.counts[lane * 8 + 0] = pre[1] + cnt[1][nSlots];
            d.counts[lane * 8 + 1] = min(pre[2] + cnt[2][nSlots], c.Ocap);
            d.counts[lane * 8 + 2] = pre[0] + cnt[0][nSlots];
        }
    }
}

// labelMat for every cell (only needed when someone looks at it: parity tests, segmentedCloudPure).
__global__ void __launch_bounds__(256) ip_label_finalize_kernel(IpDev d, Consts c) {
    const int lane = blockIdx.y;
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= c.N) return;
    const size_t base = (size_t)lane * c.N;
    const size_t g = base + cell;
    if (!(d.owner[g] >= 0 && d.groundMat[g] != 1)) { d.labelMat[g] = -1; return; }
    const int root = d.parent[g];
    if (root != cell) d.labelMat[g] = d.labelMat[base + root];
}

// ------------------------------------------------------------------------------------------------------------
void Engine::projectPointCloud() {
    projectPointCloudLaunch();
    inputConsumed();
}

void Engine::projectPointCloudLaunch() {
    { ProfScope ps(this, "memset_owner"); cuda_check(cudaMemsetAsync(ip.owner, 0xFF, sizeof(int) * (size_t)c.L * c.N, stream), "memset owner"); }
    dim3 g1((c.Pcap + 255) / 256, c.L);
    { ProfScope ps(this, "ip_project_kernel"); ip_project_kernel<<<g1, 256, 0, stream>>>(ip, c); }
    dim3 g2((c.N + 255) / 256, c.L);
    { ProfScope ps(this, "ip_gather_kernel"); ip_gather_kernel<<<g2, 256, 0, stream>>>(ip, c, cclTileW > 0 ? 0 : 1); }
    launches += 3;
}

void Engine::groundRemoval() {
    dim3 g((c.C + 127) / 128, c.L);
    { ProfScope ps(this, "ip_ground_kernel"); ip_ground_kernel<<<g, 128, 0, stream>>>(ip, c); }
    launches += 1;
}

// Tile width of the shared-memory connected components: all rows x TW columns <= CCL_MAX_TILE cells, TW a power of two >= 32;
// 0 = the three-pass form (images with more than 128 rows, or LEGO_CCL_TILE=0).  128 columns measured best for 16 rows
// (tile + border + finish 11.8 / 11.1 / 11.1 ms per 25-scan step of 592 lanes at 256 / 128 / 64: smaller tiles wait less at
// their barriers, narrower ones pay for more border unions); LEGO_CCL_TW overrides.
int ccl_tile_width(int R) {
    const char* e = std::getenv("LEGO_CCL_TILE");
    if (e && e[0] == '0') return 0;
    int tw = 128;
    if (const char* w = std::getenv("LEGO_CCL_TW")) { const int v = std::atoi(w); if (v == 32 || v == 64 || v == 128 || v == 256) tw = v; }
    while (tw > 32 && R * tw > CCL_MAX_TILE) tw >>= 1;
    return R * tw <= CCL_MAX_TILE ? tw : 0;
}
size_t ccl_tile_smem(int R, int tw, int rowWords) { return (size_t)R * tw * (4 + 2 + 4 * (size_t)rowWords); }
cudaError_t ccl_tile_set_smem(size_t bytes) { return cudaFuncSetAttribute(ip_ccl_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes); }

void Engine::cloudSegmentation() {
    dim3 g((c.N + 255) / 256, c.L);
    if (cclTileW > 0) {
        const int nTiles = (c.C + cclTileW - 1) / cclTileW;
        int sh = 0; while ((1 << sh) < cclTileW) ++sh;
        { ProfScope ps(this, "ip_ccl_tile_kernel"); ip_ccl_tile_kernel<<<dim3(nTiles, c.L), CCL_THREADS, ccl_tile_smem(c.R, cclTileW, c.rowWords), stream>>>(ip, c, cclTileW, sh); }
        { ProfScope ps(this, "ip_ccl_border_kernel"); ip_ccl_border_kernel<<<c.L, 256, 0, stream>>>(ip, c, cclTileW, nTiles); }
        { ProfScope ps(this, "ip_ccl_finish_kernel"); ip_ccl_finish_kernel<<<g, 256, 0, stream>>>(ip, c); }
    } else {
    { ProfScope ps(this, "ip_ccl_runs_kernel"); ip_ccl_runs_kernel<<<g, 256, 0, stream>>>(ip, c); }
    { ProfScope ps(this, "ip_ccl_union_kernel"); ip_ccl_union_kernel<<<g, 256, 0, stream>>>(ip, c); }
    { ProfScope ps(this, "ip_ccl_stats_kernel"); ip_ccl_stats_kernel<<<g, 256, 0, stream>>>(ip, c); }
    }
    dim3 gr(c.R, c.L);
    { ProfScope ps(this, "ip_row_count_kernel"); ip_row_count_kernel<<<gr, 256, 0, stream>>>(ip, c); }
    { ProfScope ps(this, "ip_compact_kernel"); ip_compact_kernel<<<gr, 256, 0, stream>>>(ip, c); }
    launches += 5;
}

void Engine::finalizeLabelMat() {
    dim3 g((c.N + 255) / 256, c.L);
    ip_label_finalize_kernel<<<g, 256, 0, stream>>>(ip, c);
    launches += 1;
}

}  // namespace lego
