// Incremental local map: laserCloudCornerFromMapDS / laserCloudSurfFromMapDS of extractSurroundingKeyFrames
// (mapOptmization.cpp:1013-1063) without re-sorting the whole local map in every mapping call.
//
// What the reference computes: the clouds of the key frames in surroundingExistingKeyPosesID, concatenated in list
// order, through pcl::VoxelGrid (leaf 0.2 corner / 0.4 surf).  VoxelGrid's output is one point per occupied voxel in
// ascending order of the voxel's linear index idx = i + j*dx + k*dx*dy relative to the cloud's bounding box, each the
// float32 mean of its members summed in input order (oracle/thirdparty/tp_voxel_grid.h, csrc/voxel_grid.cu).
//
// Two facts make an incremental form exact:
//   * i, j, k are the ABSOLUTE cell coordinates I = floor(x * inv) ... minus the box minimum, and 0 <= i < dx etc., so
//     ascending idx is the lexicographic order of (K, J, I): neither membership nor order depend on the bounding box.
//   * the list changes by erasing the frames that left the neighbourhood and appending the new ones at the END
//     (mapOptmization.cpp:1013-1049).  The members of a voxel therefore keep their relative order from call to call,
//     and new members come after all old ones.
// So the device keeps every lane's transformed key-frame points sorted by (K, J, I | list position, point index).
// A mapping call (a) stably removes the points of departed frames, (b) sorts only the ARRIVING frames' points with the
// radix machinery of voxel_grid.cu, (c) merges them in with "new after old on equal voxels", (d) re-reduces the runs of
// equal voxels reading the sorted points front to back.  Sums run over the same values in the same order as the full
// filter's, so the result is bit-identical (tests/test_gpu_map_optimization.py compares the two paths); traffic drops from
// ~4 radix passes over the whole map to one read and one write of it.
//
// Not reproduced: PCL returns the input unchanged when dx*dy*dz exceeds INT_MAX.  The full filter mirrors that; here
// the extent of the whole map is never formed, only the arriving frames' (error bit 8 when THEIR grid overflows).  A local
// map of key frames within 50 m would need > 250 m of extent in every axis at leaf 0.2 to get there.
#include <climits>
#include "engine.h"
#include "parity_math.cuh"

namespace lego {

__device__ __forceinline__ unsigned long long im_key(float4 p, float inv) {
    const int I = (int)floorf(PM_MUL(p.x, inv)), J = (int)floorf(PM_MUL(p.y, inv)), K = (int)floorf(PM_MUL(p.z, inv));
    return vh_pack(I, J, K);   // x in the low 21 bits, z in the high ones: ascending = (K, J, I) lexicographic
}

constexpr int IM_THREADS = 256;   // every thread owns IM_TILE / IM_THREADS = 4 consecutive points of a tile
constexpr int IM_PER = IM_TILE / IM_THREADS;

__device__ __forceinline__ int block_sum(int v, int* ws) {
    const int lid = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lid == 0) ws[w] = v;
    __syncthreads();
    int t = 0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) t += ws[q];
    return t;
}

// exclusive prefix of `v` over the block + the block total
__device__ __forceinline__ int block_excl_scan(int v, int* ws, int* total) {
    const int lid = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int inc = warp_inclusive_scan(v);
    __syncthreads();
    if (lid == 31) ws[w] = inc;
    __syncthreads();
    int off = inc - v, t = 0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) { const int x = ws[q]; if (q < w) off += x; t += x; }
    *total = t;
    return off;
}

// which arriving key frame does raw point `i` belong to (entries are laid out back to back in list order)
__device__ __forceinline__ int im_frame_of(const MoKfEntry* en, int nE, int type, int i) {
    int lo = 0, hi = nE - 1;
    while (lo < hi) {   // last entry whose first point is <= i
        const int mid = (lo + hi + 1) >> 1;
        const int first = type == 0 ? en[mid].dstC : en[mid].dstS;
        if (first <= i) lo = mid; else hi = mid - 1;
    }
    return en[lo].id;
}

// (b) the arriving points in sorted order: into the other ping-pong buffer when the lane rebuilds, else into newPts.
__global__ void __launch_bounds__(IM_THREADS) im_gather_new_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl& ctl = m.imCtl[lane];
    if (!ctl.changed) return;
    const VgSeg& sorter = *s.sorter;
    const int n = sorter.state[9];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        s.state[3] = n;
        if (sorter.state[10] || sorter.state[12] || (!ctl.rebuild && n > s.newCap)) atomicOr(s.errors, 8);
    }
    if (n == 0) return;
    const unsigned* vals = vg_sorted_vals(sorter);
    const MoKfEntry* en = m.entries + (size_t)lane * m.maxSel;
    const int nE = m.nEntries[lane];
    const float inv = PM_DIV(1.0f, s.leaf);
    const int next = s.state[0] ^ 1;
    for (int j = blockIdx.x * IM_THREADS + threadIdx.x; j < n; j += gridDim.x * IM_THREADS) {
        const int src = (int)vals[j];
        const float4 p = s.raw[src];
        const int f = im_frame_of(en, nE, s.type, src);
        if (ctl.rebuild) { s.pts[next][j] = p; s.frame[next][j] = f; }
        else if (j < s.newCap) { s.newPts[j] = p; s.newFrame[j] = f; s.newKey[j] = im_key(p, inv); }
    }
}

__device__ __forceinline__ bool im_departed(const ImCtl& ctl, int f) {
    bool d = false;
    for (int q = 0; q < ctl.nDeparted; ++q) d |= (ctl.departed[q] == f);
    return d;
}

// (a) survivors per tile of the current sorted cloud (reads only the frame ids)
__global__ void __launch_bounds__(IM_THREADS) im_mark_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    __shared__ int ws[32];
    const int n = s.state[1];
    const int* __restrict__ fr = s.frame[s.state[0]];
    int* tileKept = s.tileKept;
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        const int i0 = tile * IM_TILE + threadIdx.x * IM_PER;
        int keep = 0;
        if (i0 + IM_PER <= n) {
            const int4 f = *reinterpret_cast<const int4*>(fr + i0);
            keep = (im_departed(ctl, f.x) ? 0 : 1) + (im_departed(ctl, f.y) ? 0 : 1) + (im_departed(ctl, f.z) ? 0 : 1) + (im_departed(ctl, f.w) ? 0 : 1);
        } else {
            for (int e = 0; e < IM_PER; ++e) if (i0 + e < n && !im_departed(ctl, fr[i0 + e])) ++keep;
        }
        const int t = block_sum(keep, ws);
        if (threadIdx.x == 0) tileKept[tile] = t;
    }
}

// (c) survivors move to rank + (arrivals with a smaller voxel).  A thread owns 4 consecutive points: one binary search
// over the arrivals for the first, the others gallop on from there (voxels ascend, arrivals are ~30x sparser).  The rank of
// every 32nd point is kept for the arrivals' pass.
#ifndef LEGO_IMO_MINBLOCKS
#define LEGO_IMO_MINBLOCKS 6    // 40 registers (44 B spilled) instead of 63: 6 CTAs per SM, 6.35 -> 5.87 ms per 25-scan step of 592 lanes
#endif
#define IMO_BOUNDS __launch_bounds__(IM_THREADS, LEGO_IMO_MINBLOCKS)
__global__ void IMO_BOUNDS im_merge_old_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    __shared__ int ws[32];
    const int n = s.state[1], nNew = min(s.state[3], s.newCap);
    const int cur = s.state[0];
    const float4* __restrict__ src = s.pts[cur]; const int* __restrict__ fr = s.frame[cur];
    float4* __restrict__ dst = s.pts[cur ^ 1]; int* __restrict__ dfr = s.frame[cur ^ 1];
    const unsigned long long* __restrict__ nk = s.newKey;
    const int* tileKept = s.tileKept; int* rank32 = s.rank; int* state = s.state;
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        int pre = 0;
        for (int k = threadIdx.x; k < tile; k += IM_THREADS) pre += tileKept[k];
        pre = block_sum(pre, ws);
        const int i0 = tile * IM_TILE + threadIdx.x * IM_PER;
        float4 p[IM_PER]; int f[IM_PER]; int keep[IM_PER]; int mine = 0;
#pragma unroll
        for (int e = 0; e < IM_PER; ++e) {
            keep[e] = 0; f[e] = -1; p[e] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i0 + e < n) { p[e] = src[i0 + e]; f[e] = fr[i0 + e]; keep[e] = im_departed(ctl, f[e]) ? 0 : 1; }
            mine += keep[e];
        }
        int total;
        int r = pre + block_excl_scan(mine, ws, &total);
        if ((threadIdx.x & 7) == 0 && i0 < n) rank32[i0 >> 5] = r;
        int lo = -1;
#pragma unroll
        for (int e = 0; e < IM_PER; ++e) {
            if (!keep[e]) continue;
            const unsigned long long key = im_key(p[e], inv);
            if (lo < 0) {                     // arrivals with key < this point's key (equal voxels: old before new)
                int a = 0, z = nNew;
                while (a < z) { const int mid = (a + z) >> 1; if (nk[mid] < key) a = mid + 1; else z = mid; }
                lo = a;
            } else {
                int step = 0;
                while (lo < nNew && nk[lo] < key && step < 8) { ++lo; ++step; }
                if (lo < nNew && nk[lo] < key) {   // a long stretch of arrivals in between: search the rest
                    int a = lo, z = nNew;
                    while (a < z) { const int mid = (a + z) >> 1; if (nk[mid] < key) a = mid + 1; else z = mid; }
                    lo = a;
                }
            }
            dst[r + lo] = p[e]; dfr[r + lo] = f[e];
            ++r;
        }
        if (threadIdx.x == 0 && (tile + 1) * IM_TILE >= n) state[2] = pre + total;
    }
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) state[2] = 0;
}

// (c) arrivals move to their index + (survivors with a voxel <= theirs)
__global__ void __launch_bounds__(256) im_merge_new_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    const int n = s.state[1], nNew = min(s.state[3], s.newCap), kept = s.state[2];
    const int cur = s.state[0];
    const float4* __restrict__ old = s.pts[cur]; const int* __restrict__ fr = s.frame[cur];
    float4* __restrict__ dst = s.pts[cur ^ 1]; int* __restrict__ dfr = s.frame[cur ^ 1];
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nNew; j += gridDim.x * blockDim.x) {
        const unsigned long long key = s.newKey[j];
        int lo = 0, hi = n;                   // first old point with a larger voxel
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (im_key(old[mid], inv) <= key) lo = mid + 1; else hi = mid; }
        int before = kept;
        if (lo < n) {                         // survivors in front of old[lo]: the recorded rank of its group of 32 + the group's head
            before = s.rank[lo >> 5];
            for (int q = lo & ~31; q < lo; ++q) before += im_departed(ctl, fr[q]) ? 0 : 1;
        }
        dst[j + before] = s.newPts[j];
        dfr[j + before] = s.newFrame[j];
    }
}

__global__ void im_commit_kernel(ImBatch b, MoDev m) {
    const int si = blockIdx.x * blockDim.x + threadIdx.x;
    if (si >= b.nSeg) return;
    const ImSeg& s = b.segs[si];
    const int lane = si >> 1;
    if (!m.active[lane]) return;
    const ImCtl& ctl = m.imCtl[lane];
    if (!ctl.changed) return;
    int n = ctl.rebuild ? s.state[3] : s.state[2] + min(s.state[3], s.newCap);
    if (n > s.cap) { n = s.cap; atomicOr(s.errors, 4); }
    s.state[0] ^= 1; s.state[1] = n;
    *s.nTotal = n;
}

// (d) voxel runs of the merged cloud.  A tile's points are staged in shared memory (one coalesced read); the thread that
// owns the first point of a run folds the run front to back -- the float32 sums of pcl::VoxelGrid: members in input order,
// starting from 0 -- continuing in global memory when the run leaves the tile.  Results go to tile-local slots of the idle
// ping-pong buffer; im_compact_kernel moves them to their final positions once every tile's count is known.
__global__ void __launch_bounds__(IM_THREADS) im_reduce_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane] || !m.imCtl[lane].changed) return;
    __shared__ int ws[32];
    __shared__ float4 sp[IM_TILE];
    __shared__ unsigned long long sk[IM_TILE];
    __shared__ unsigned short sh[IM_TILE];
    const int n = s.state[1];
    const float4* __restrict__ pts = s.pts[s.state[0]];
    float4* __restrict__ stage = s.pts[s.state[0] ^ 1];
    int* tileHeads = s.tileHeads;
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        const int t0 = tile * IM_TILE, t1 = min(t0 + IM_TILE, n);
        __syncthreads();
        for (int k = threadIdx.x; k < t1 - t0; k += IM_THREADS) { const float4 p = pts[t0 + k]; sp[k] = p; sk[k] = im_key(p, inv); }
        __syncthreads();
        const int l0 = threadIdx.x * IM_PER;
        unsigned long long prev = 0ull; bool hasPrev = false;
        if (t0 + l0 < t1 && t0 + l0 > 0) { prev = l0 > 0 ? sk[l0 - 1] : im_key(pts[t0 - 1], inv); hasPrev = true; }
        int head[IM_PER]; int mine = 0;
#pragma unroll
        for (int e = 0; e < IM_PER; ++e) {
            head[e] = 0;
            if (t0 + l0 + e < t1) {
                const unsigned long long k = sk[l0 + e];
                head[e] = (!hasPrev || k != prev) ? 1 : 0;
                prev = k; hasPrev = true;
            }
            mine += head[e];
        }
        int total;
        int off = block_excl_scan(mine, ws, &total);
        // the tile's run heads, compacted: afterwards thread h folds run h, so every active thread has exactly one run
        // (one thread walking its up-to-4 heads one after the other ran with 12 of 32 threads active)
#pragma unroll
        for (int e = 0; e < IM_PER; ++e) if (head[e]) sh[off++] = (unsigned short)(l0 + e);
        __syncthreads();
        for (int h = threadIdx.x; h < total; h += IM_THREADS) {
            const int l = sh[h];
            const unsigned long long key = sk[l];
            const float4 p = sp[l];
            float sx = PM_ADD(0.f, p.x), sy = PM_ADD(0.f, p.y), sz = PM_ADD(0.f, p.z), sw = PM_ADD(0.f, p.w); int cnt = 1;
            int q = l + 1;
            for (; q < t1 - t0 && sk[q] == key; ++q) {
                const float4 a = sp[q];
                sx = PM_ADD(sx, a.x); sy = PM_ADD(sy, a.y); sz = PM_ADD(sz, a.z); sw = PM_ADD(sw, a.w); ++cnt;
            }
            if (q == t1 - t0) {               // the run may go on in the next tiles
                for (int g = t1; g < n; ++g) {
                    const float4 a = pts[g];
                    if (im_key(a, inv) != key) break;
                    sx = PM_ADD(sx, a.x); sy = PM_ADD(sy, a.y); sz = PM_ADD(sz, a.z); sw = PM_ADD(sw, a.w); ++cnt;
                }
            }
            const float fc = (float)cnt;
            stage[t0 + h] = make_float4(PM_DIV(sx, fc), PM_DIV(sy, fc), PM_DIV(sz, fc), PM_DIV(sw, fc));
        }
        if (threadIdx.x == 0) tileHeads[tile] = total;
    }
}

__global__ void __launch_bounds__(IM_THREADS) im_compact_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane] || !m.imCtl[lane].changed) return;
    __shared__ int ws[32];
    const int n = s.state[1];
    if (n == 0) { if (blockIdx.x == 0 && threadIdx.x == 0) *s.nOut = 0; return; }
    const float4* __restrict__ stage = s.pts[s.state[0] ^ 1];
    float4* __restrict__ out = s.out;
    const int* tileHeads = s.tileHeads;
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        int pre = 0;
        for (int k = threadIdx.x; k < tile; k += IM_THREADS) pre += tileHeads[k];
        pre = block_sum(pre, ws);
        const int cnt = tileHeads[tile];
        for (int k = threadIdx.x; k < cnt; k += IM_THREADS) if (pre + k < s.outCap) out[pre + k] = stage[tile * IM_TILE + k];
        if (threadIdx.x == 0 && (tile + 1) * IM_TILE >= n) {
            int t = pre + cnt;
            if (t > s.outCap) { t = s.outCap; atomicOr(s.errors, 4); }
            *s.nOut = t;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------- host
void Engine::allocIncrementalMap(const std::vector<VgSeg>& mapSegs) {
    const size_t L = c.L;
    const char* env = std::getenv("LEGO_INCR_MAP");
    imEnabled = !loopMode() && !(env && env[0] == '0');
    mo.imCtl = alloc<ImCtl>(L);
    mo.imNewCounts = alloc<int>(L * 4);
    if (!imEnabled) return;
    // the arriving frames are sorted with the local map's own radix scratch (vgMap is not used in this mode)
    std::vector<VgSeg> ns = mapSegs;
    for (size_t l = 0; l < L; ++l) { ns[2 * l].nIn = mo.imNewCounts + l * 4; ns[2 * l + 1].nIn = mo.imNewCounts + l * 4 + 1; }
    vgMapNew = vgMap;
    vgMapNew.segs = alloc<VgSeg>(ns.size());
    cuda_check(cudaMemcpyAsync(vgMapNew.segs, ns.data(), sizeof(VgSeg) * ns.size(), cudaMemcpyHostToDevice, stream), "vg table");
    const int newCap = 2 * c.N;   // arrivals of one call beyond this (about eight key frames) make the lane rebuild
    std::vector<ImSeg> segs(2 * L);
    for (size_t l = 0; l < L; ++l)
        for (int t = 0; t < 2; ++t) {
            ImSeg& g = segs[2 * l + t];
            const int cap = t == 0 ? mo.cornerRawCap : mo.surfRawCap;
            for (int k = 0; k < 2; ++k) { g.pts[k] = alloc<float4>(cap, false); g.frame[k] = alloc<int>(cap, false); }
            g.rank = alloc<int>(cap / 32 + 2, false);
            g.newPts = alloc<float4>(newCap, false); g.newFrame = alloc<int>(newCap, false); g.newKey = alloc<unsigned long long>(newCap, false);
            g.raw = t == 0 ? mo.cornerFromMap + l * mo.cornerRawCap : mo.surfFromMap + l * mo.surfRawCap;
            g.nRaw = mo.imNewCounts + l * 4 + t;
            g.sorter = vgMapNew.segs + 2 * l + t;
            g.out = t == 0 ? mo.cornerFromMapDS + l * mo.cornerDSCap : mo.surfFromMapDS + l * mo.surfDSCap;
            g.nOut = mo.mapCounts + l * 4 + 2 + t;
            g.nTotal = mo.mapCounts + l * 4 + t;
            g.state = alloc<int>(8);
            g.tileKept = alloc<int>(cap / IM_TILE + 2); g.tileHeads = alloc<int>(cap / IM_TILE + 2);
            g.errors = mo.state + l * 16 + 5;
            g.leaf = t == 0 ? 0.2f : 0.4f;
            g.cap = cap; g.newCap = newCap; g.outCap = t == 0 ? mo.cornerDSCap : mo.surfDSCap; g.type = t;
        }
    im.segs = alloc<ImSeg>(segs.size());
    im.nSeg = (int)segs.size();
    im.maxTiles = (mo.surfRawCap + IM_TILE - 1) / IM_TILE;
    im.maxNewTiles = (newCap + IM_TILE - 1) / IM_TILE;
    cuda_check(cudaMemcpyAsync(im.segs, segs.data(), sizeof(ImSeg) * segs.size(), cudaMemcpyHostToDevice, stream), "im table");
    cuda_check(cudaStreamSynchronize(stream), "im table sync");
}

// The cache diff of extractSurroundingKeyFrames (mapOptmization.cpp:1013-1049) against what the device holds: `all` is
// this call's list in order.  Valid increments are "some frames erased, the others in unchanged order, new ones at the
// end"; anything else (first call, a capacity truncation, a burst of arrivals) rebuilds the lane's sorted cloud.
void Engine::diffSurroundingKeyFrames(int lane, const std::vector<MoKfEntry>& all, std::vector<MoKfEntry>& arriving, ImCtl* ctl,
                                      int* nCorner, int* nSurf) {
    MoLaneHost& H = moHost[lane];
    std::vector<int> now; now.reserve(all.size());
    for (const MoKfEntry& e : all) now.push_back(e.id);
    ctl->changed = 1; ctl->rebuild = 1; ctl->nDeparted = 0;
    size_t nKept = 0;
    if (H.imValid) {
        // survivors must be a prefix of the new list, in the old order
        std::vector<int> departed;
        size_t k = 0;
        for (int id : H.imList) { if (k < now.size() && now[k] == id) ++k; else departed.push_back(id); }
        bool ok = departed.size() <= (size_t)IM_MAX_DEPARTED;
        for (size_t q = k; ok && q < now.size(); ++q) ok = std::find(H.imList.begin(), H.imList.end(), now[q]) == H.imList.end();
        for (int id : departed) ok = ok && std::find(now.begin(), now.end(), id) == now.end();
        long arrivingPts[2] = {0, 0};
        for (size_t q = k; q < all.size(); ++q) { arrivingPts[0] += all[q].nC; arrivingPts[1] += all[q].nS + all[q].nO; }
        ok = ok && arrivingPts[0] <= 2L * c.N && arrivingPts[1] <= 2L * c.N;
        if (ok) {
            ctl->rebuild = 0; nKept = k;
            ctl->nDeparted = (int)departed.size();
            for (size_t q = 0; q < departed.size(); ++q) ctl->departed[q] = departed[q];
            ctl->changed = (departed.empty() && k == now.size()) ? 0 : 1;
        }
    }
    // the arriving frames, packed from offset 0 of the raw buffers
    arriving.clear();
    int offC = 0, offS = 0;
    for (size_t q = nKept; q < all.size(); ++q) {
        MoKfEntry e = all[q];
        e.dstC = offC; offC += e.nC;
        e.dstS = offS; offS += e.nS;
        e.dstO = offS; offS += e.nO;
        arriving.push_back(e);
    }
    *nCorner = offC; *nSurf = offS;
    H.imList = now; H.imValid = true;
}

__global__ void mo_assemble_map_kernel(MoDev m);

int Engine::incrementalLocalMap() {
    int n = 0;
    { ProfScope ps(this, "im_sort_arrivals"); n += vg_sort_keys(vgMapNew, stream); }
    dim3 gNew(std::min(im.maxTiles * 4, 32), im.nSeg), gOld(std::min(im.maxTiles, 32), im.nSeg), gB(std::min(im.maxNewTiles * 4, 16), im.nSeg);
    { ProfScope ps(this, "im_gather_new_kernel"); im_gather_new_kernel<<<gNew, IM_THREADS, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_mark_kernel"); im_mark_kernel<<<gOld, IM_THREADS, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_merge_old_kernel"); im_merge_old_kernel<<<gOld, IM_THREADS, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_merge_new_kernel"); im_merge_new_kernel<<<gB, 256, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_commit_kernel"); im_commit_kernel<<<(im.nSeg + 127) / 128, 128, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_reduce_kernel"); im_reduce_kernel<<<gOld, IM_THREADS, 0, stream>>>(im, mo); }
    { ProfScope ps(this, "im_compact_kernel"); im_compact_kernel<<<gOld, IM_THREADS, 0, stream>>>(im, mo); }
    return n + 7;
}

}  // namespace lego
