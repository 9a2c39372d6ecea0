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

__device__ __forceinline__ int block_sum_1024(int v, int* ws) {
    const int lid = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lid == 0) ws[w] = v;
    __syncthreads();
    int t = 0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) t += ws[q];
    return t;
}

// exclusive prefix of `flag` over the block (1024 threads) + the block total
__device__ __forceinline__ int block_excl_scan_1024(int flag, int* ws, int* total) {
    const int lid = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int inc = warp_inclusive_scan(flag);
    __syncthreads();
    if (lid == 31) ws[w] = inc;
    __syncthreads();
    int off = inc - flag, t = 0;
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
__global__ void __launch_bounds__(IM_TILE) im_gather_new_kernel(ImBatch b, MoDev m) {
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
    for (int j = blockIdx.x * IM_TILE + threadIdx.x; j < n; j += gridDim.x * IM_TILE) {
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

// (a) survivors per tile of the current sorted cloud
__global__ void __launch_bounds__(IM_TILE) im_mark_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl& ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    __shared__ int ws[32];
    const int n = s.state[1];
    const int* fr = s.frame[s.state[0]];
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        const int i = tile * IM_TILE + threadIdx.x;
        const int keep = (i < n && !im_departed(ctl, fr[i])) ? 1 : 0;
        const int t = block_sum_1024(keep, ws);
        if (threadIdx.x == 0) s.tileKept[tile] = t;
    }
}

// (c) survivors move to rank + (arrivals with a smaller voxel); every point also records its rank for the arrivals' pass
__global__ void __launch_bounds__(IM_TILE) im_merge_old_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl& ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    __shared__ int ws[32];
    const int n = s.state[1], nNew = min(s.state[3], s.newCap);
    const int cur = s.state[0];
    const float4* __restrict__ src = s.pts[cur]; const int* __restrict__ fr = s.frame[cur];
    float4* __restrict__ dst = s.pts[cur ^ 1]; int* __restrict__ dfr = s.frame[cur ^ 1];
    const unsigned long long* __restrict__ nk = s.newKey;
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        int pre = 0;
        for (int k = threadIdx.x; k < tile; k += IM_TILE) pre += s.tileKept[k];
        pre = block_sum_1024(pre, ws);
        const int i = tile * IM_TILE + threadIdx.x;
        float4 p = make_float4(0.f, 0.f, 0.f, 0.f); int f = -1, keep = 0;
        if (i < n) { p = src[i]; f = fr[i]; keep = im_departed(ctl, f) ? 0 : 1; }
        int total;
        const int r = pre + block_excl_scan_1024(keep, ws, &total);
        if (i < n) s.rank[i] = r;
        if (keep) {
            const unsigned long long key = im_key(p, inv);
            int lo = 0, hi = nNew;            // arrivals with key < this point's key (equal voxels: old before new)
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (nk[mid] < key) lo = mid + 1; else hi = mid; }
            dst[r + lo] = p; dfr[r + lo] = f;
        }
        if (threadIdx.x == 0 && (tile + 1) * IM_TILE >= n) s.state[2] = pre + total;
    }
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) s.state[2] = 0;
}

// (c) arrivals move to their index + (survivors with a voxel <= theirs)
__global__ void __launch_bounds__(256) im_merge_new_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane]) return;
    const ImCtl& ctl = m.imCtl[lane];
    if (!ctl.changed || ctl.rebuild) return;
    const int n = s.state[1], nNew = min(s.state[3], s.newCap), kept = s.state[2];
    const int cur = s.state[0];
    const float4* __restrict__ old = s.pts[cur];
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nNew; j += gridDim.x * blockDim.x) {
        const unsigned long long key = s.newKey[j];
        int lo = 0, hi = n;                   // first old point with a larger voxel
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (im_key(old[mid], inv) <= key) lo = mid + 1; else hi = mid; }
        const int before = lo < n ? s.rank[lo] : kept;
        s.pts[cur ^ 1][j + before] = s.newPts[j];
        s.frame[cur ^ 1][j + before] = s.newFrame[j];
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

// (d) voxel runs of the merged cloud: heads per tile, then one thread per head averages its run front to back
__global__ void __launch_bounds__(IM_TILE) im_heads_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane] || !m.imCtl[lane].changed) return;
    __shared__ int ws[32];
    const int n = s.state[1];
    const float4* __restrict__ pts = s.pts[s.state[0]];
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        const int i = tile * IM_TILE + threadIdx.x;
        int head = 0;
        if (i < n) head = (i == 0 || im_key(pts[i - 1], inv) != im_key(pts[i], inv)) ? 1 : 0;
        const int t = block_sum_1024(head, ws);
        if (threadIdx.x == 0) s.tileHeads[tile] = t;
    }
}

__global__ void __launch_bounds__(IM_TILE) im_reduce_kernel(ImBatch b, MoDev m) {
    const ImSeg& s = b.segs[blockIdx.y];
    const int lane = blockIdx.y >> 1;
    if (!m.active[lane] || !m.imCtl[lane].changed) return;
    __shared__ int ws[32];
    const int n = s.state[1];
    if (n == 0) { if (blockIdx.x == 0 && threadIdx.x == 0) *s.nOut = 0; return; }
    const float4* __restrict__ pts = s.pts[s.state[0]];
    const float inv = PM_DIV(1.0f, s.leaf);
    for (int tile = blockIdx.x; tile * IM_TILE < n; tile += gridDim.x) {
        int pre = 0;
        for (int k = threadIdx.x; k < tile; k += IM_TILE) pre += s.tileHeads[k];
        pre = block_sum_1024(pre, ws);
        const int i = tile * IM_TILE + threadIdx.x;
        unsigned long long key = 0ull; int head = 0; float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n) { p = pts[i]; key = im_key(p, inv); head = (i == 0 || im_key(pts[i - 1], inv) != key) ? 1 : 0; }
        int total;
        const int off = pre + block_excl_scan_1024(head, ws, &total);
        if (head) {
            // the float32 sums of pcl::VoxelGrid: members in input order, starting from 0
            float sx = PM_ADD(0.f, p.x), sy = PM_ADD(0.f, p.y), sz = PM_ADD(0.f, p.z), sw = PM_ADD(0.f, p.w); int cnt = 1;
            for (int q = i + 1; q < n; ++q) {
                const float4 a = pts[q];
                if (im_key(a, inv) != key) break;
                sx = PM_ADD(sx, a.x); sy = PM_ADD(sy, a.y); sz = PM_ADD(sz, a.z); sw = PM_ADD(sw, a.w); ++cnt;
            }
            const float fc = (float)cnt;
            if (off < s.outCap) s.out[off] = make_float4(PM_DIV(sx, fc), PM_DIV(sy, fc), PM_DIV(sz, fc), PM_DIV(sw, fc));
        }
        if (threadIdx.x == 0 && (tile + 1) * IM_TILE >= n) {
            int t = pre + total;
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
            g.rank = alloc<int>(cap, false);
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
    ProfScope ps(this, "im_merge_reduce");
    dim3 gNew(std::min(im.maxTiles, 16), im.nSeg), gOld(std::min(im.maxTiles, 32), im.nSeg), gB(std::min(im.maxNewTiles * 4, 16), im.nSeg);
    im_gather_new_kernel<<<gNew, IM_TILE, 0, stream>>>(im, mo);
    im_mark_kernel<<<gOld, IM_TILE, 0, stream>>>(im, mo);
    im_merge_old_kernel<<<gOld, IM_TILE, 0, stream>>>(im, mo);
    im_merge_new_kernel<<<gB, 256, 0, stream>>>(im, mo);
    im_commit_kernel<<<(im.nSeg + 127) / 128, 128, 0, stream>>>(im, mo);
    im_heads_kernel<<<gOld, IM_TILE, 0, stream>>>(im, mo);
    im_reduce_kernel<<<gOld, IM_TILE, 0, stream>>>(im, mo);
    return n + 7;
}

}  // namespace lego
