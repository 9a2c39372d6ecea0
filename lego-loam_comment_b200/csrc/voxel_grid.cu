// Kernels of the batched VoxelGrid (see voxel_grid.cuh).  blockIdx.y = segment; a CTA takes the tiles blockIdx.x,
// blockIdx.x + gridDim.x, ... of its segment.
#include <cfloat>
#include <climits>
#include "kernel_utils.cuh"
#include "parity_math.cuh"
#include "voxel_grid.cuh"

namespace lego {

// order-preserving float <-> int map, so min/max of floats can use integer atomics
__device__ __forceinline__ int f2ord(float f) { const int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7fffffff); }

__device__ __forceinline__ bool vg_live(const VgSeg& s) { return s.active == nullptr || *s.active != 0; }

__global__ void vg_init_kernel(VgBatch b) {
    const int si = blockIdx.x * blockDim.x + threadIdx.x;
    if (si >= b.nSeg) return;
    const VgSeg& s = b.segs[si];
    if (!vg_live(s)) { s.state[9] = 0; return; }
    for (int a = 0; a < 3; ++a) { s.state[a] = INT_MAX; s.state[3 + a] = INT_MIN; }
    s.state[6] = 0; s.state[7] = 0; s.state[8] = 0; s.state[10] = 0; s.state[12] = 0;
    int n = *s.nIn;
    if (n > s.cap) { n = s.cap; s.state[10] = 1; }
    s.state[9] = n;
}

__global__ void __launch_bounds__(VG_THREADS) vg_bbox_kernel(VgBatch b) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {   // the grid covers a typical cloud, larger ones loop
    const int t0 = tile * VG_TILE;
    const int t1 = min(t0 + VG_TILE, n);
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = t0 + threadIdx.x; i < t1; i += VG_THREADS) {
        const float4 p = s.in[i];
        mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
        mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = warp_min_f(mn[a]); mx[a] = warp_max_f(mx[a]); }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(s.state + a, f2ord(mn[a])); atomicMax(s.state + 3 + a, f2ord(mx[a])); }
    }
    }
}

// voxel index of every point: ijk = int(floor(p * inv) - float(min_b)), idx = i + j*dx + k*dx*dy  (PCL voxel_grid.hpp)
__global__ void __launch_bounds__(VG_THREADS) vg_keys_kernel(VgBatch b) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    if (blockIdx.x * VG_TILE >= n) return;
    const float inv = PM_DIV(1.0f, s.leaf);
    int minb[3], divb[3]; long long dd[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = ord2f(s.state[a]), hi = ord2f(s.state[3 + a]);
        dd[a] = (long long)PM_MUL(PM_SUB(hi, lo), inv) + 1;
        minb[a] = (int)floorf(PM_MUL(lo, inv));
        divb[a] = (int)floorf(PM_MUL(hi, inv)) - minb[a] + 1;
    }
    // PCL refuses grids with more than INT_MAX voxels and returns the input unchanged: identity keys do the same
    const bool overflow = dd[0] * dd[1] * dd[2] > (long long)INT_MAX;
    if (overflow && blockIdx.x == 0 && threadIdx.x == 0) s.state[12] = 1;
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {
    const int t0 = tile * VG_TILE, t1 = min(t0 + VG_TILE, n);
    for (int i = t0 + threadIdx.x; i < t1; i += VG_THREADS) {
        const float4 p = s.in[i];
        unsigned key;
        if (overflow) key = (unsigned)i;
        else {
            const int i0 = (int)PM_SUB(floorf(PM_MUL(p.x, inv)), (float)minb[0]);
            const int i1 = (int)PM_SUB(floorf(PM_MUL(p.y, inv)), (float)minb[1]);
            const int i2 = (int)PM_SUB(floorf(PM_MUL(p.z, inv)), (float)minb[2]);
            key = (unsigned)(i0 + i1 * divb[0] + i2 * divb[0] * divb[1]);
        }
        s.keys[0][i] = key;
        s.vals[0][i] = (unsigned)i;
    }
    }
}

// ---- LSD radix sort, one 8-bit digit per pass -----------------------------------------------------------------
__global__ void __launch_bounds__(VG_THREADS) rs_hist_kernel(VgBatch b, int shift) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    __shared__ int hist[256];
    const unsigned* src = s.keys[s.state[6] ^ s.state[7]];
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {
    const int t0 = tile * VG_TILE, t1 = min(t0 + VG_TILE, n);
    hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = t0 + threadIdx.x; i < t1; i += VG_THREADS) atomicAdd(&hist[(src[i] >> shift) & 255u], 1);
    __syncthreads();
    s.blockHist[threadIdx.x * b.maxBlocks + tile] = hist[threadIdx.x];
    __syncthreads();
    }
}

// One CTA per segment: applies the buffer flip left pending by the previous pass, turns the block histograms into
// global destination offsets and detects a pass whose digit is the same for every key (nothing to move).
__global__ void __launch_bounds__(256) rs_scan_kernel(VgBatch b) {
    const VgSeg& s = b.segs[blockIdx.x];
    const int n = s.state[9];
    const int d = threadIdx.x;
    __shared__ int tot[256];
    __shared__ int skip;
    if (d == 0) { s.state[6] ^= s.state[7]; s.state[7] = 0; skip = 0; }
    if (n == 0) { if (d == 0) s.state[8] = 1; return; }
    const int nb = (n + VG_TILE - 1) / VG_TILE;
    int* bh = s.blockHist + d * b.maxBlocks;
    int run = 0;
    for (int k = 0; k < nb; ++k) { const int t = bh[k]; bh[k] = run; run += t; }
    tot[d] = run;
    __syncthreads();
    if (run == n) skip = 1;
    // exclusive scan of the 256 digit totals (8 warps, then the warp sums)
    const int lid = d & 31, w = d >> 5;
    const int inc = warp_inclusive_scan(run);
    __shared__ int wsum[8];
    if (lid == 31) wsum[w] = inc;
    __syncthreads();
    int base = inc - run;
    for (int q = 0; q < w; ++q) base += wsum[q];
    for (int k = 0; k < nb; ++k) bh[k] += base;
    __syncthreads();
    if (d == 0) { s.state[8] = skip; s.state[7] = skip ? 0 : 1; }
}

#ifdef LEGO_RS_MINBLOCKS
#define RS_BOUNDS __launch_bounds__(VG_THREADS, LEGO_RS_MINBLOCKS)
#else
#define RS_BOUNDS __launch_bounds__(VG_THREADS)
#endif
__global__ void RS_BOUNDS rs_scatter_kernel(VgBatch b, int shift) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    if (s.state[8]) return;
    const int cur = s.state[6];
    const unsigned* __restrict__ srcK = s.keys[cur]; const unsigned* __restrict__ srcV = s.vals[cur];
    unsigned* __restrict__ dstK = s.keys[cur ^ 1]; unsigned* __restrict__ dstV = s.vals[cur ^ 1];
    __shared__ int base[256];
    __shared__ int warpCnt[8][257];
    const int tid = threadIdx.x, lid = tid & 31, w = tid >> 5;
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {
    const int t0 = tile * VG_TILE, t1 = min(t0 + VG_TILE, n);
    base[tid] = s.blockHist[tid * b.maxBlocks + tile];
#pragma unroll
    for (int q = 0; q < 8; ++q) warpCnt[q][tid] = 0;
    __syncthreads();
    for (int c0 = t0; c0 < t1; c0 += 4 * VG_THREADS) {
        unsigned k[4], v[4]; int dig[4], rank[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = c0 + w * 128 + j * 32 + lid;
            const bool valid = i < t1;
            k[j] = valid ? srcK[i] : 0u; v[j] = valid ? srcV[i] : 0u;
            dig[j] = valid ? (int)((k[j] >> shift) & 255u) : 256;
            const unsigned m = __match_any_sync(0xffffffffu, dig[j]);
            const int lower = __popc(m & ((1u << lid) - 1u));
            const int old = valid ? warpCnt[w][dig[j]] : 0;
            __syncwarp();
            if (valid && lower == 0) warpCnt[w][dig[j]] = old + __popc(m);
            __syncwarp();
            rank[j] = old + lower;
        }
        __syncthreads();
        {
            int run = base[tid];
#pragma unroll
            for (int q = 0; q < 8; ++q) { const int t = warpCnt[q][tid]; warpCnt[q][tid] = run; run += t; }
            base[tid] = run;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) if (dig[j] < 256) {
            const int dst = warpCnt[w][dig[j]] + rank[j];
            dstK[dst] = k[j]; dstV[dst] = v[j];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 8; ++q) warpCnt[q][tid] = 0;
        __syncthreads();
    }
    }
}

// ---- reduction of equal-key runs --------------------------------------------------------------------------------
__global__ void __launch_bounds__(VG_THREADS) vg_heads_kernel(VgBatch b) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    __shared__ int ws[8];
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {
    const int t0 = tile * VG_TILE, t1 = min(t0 + VG_TILE, n);
    const unsigned* keys = s.keys[s.state[6] ^ s.state[7]];
    int c = 0;
    for (int i = t0 + threadIdx.x; i < t1; i += VG_THREADS) c += (i == 0 || keys[i - 1] != keys[i]) ? 1 : 0;
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int q = 0; q < 8; ++q) t += ws[q]; s.blockHeads[tile] = t; }
    __syncthreads();
    }
}

__global__ void __launch_bounds__(VG_THREADS) vg_reduce_kernel(VgBatch b) {
    const VgSeg& s = b.segs[blockIdx.y];
    const int n = s.state[9];
    if (n == 0) { if (blockIdx.x == 0 && threadIdx.x == 0 && vg_live(s)) *s.nOut = 0; return; }
    const int bufi = s.state[6] ^ s.state[7];
    const unsigned* __restrict__ keys = s.keys[bufi]; const unsigned* __restrict__ vals = s.vals[bufi];
    const int tid = threadIdx.x, lid = tid & 31, w = tid >> 5;
    __shared__ int ws[8];
    __shared__ int carry;
    for (int tile = blockIdx.x; tile * VG_TILE < n; tile += gridDim.x) {
    const int t0 = tile * VG_TILE, t1 = min(t0 + VG_TILE, n);
    // heads in the tiles before this one
    int pre = 0;
    for (int k = tid; k < tile; k += VG_THREADS) pre += s.blockHeads[k];
    pre = warp_sum(pre);
    if (lid == 0) ws[w] = pre;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 8; ++q) t += ws[q]; carry = t; }
    __syncthreads();
    for (int c0 = t0; c0 < t1; c0 += VG_THREADS) {
        const int i = c0 + tid;
        const int head = (i < t1 && (i == 0 || keys[i - 1] != keys[i])) ? 1 : 0;
        const int inc = warp_inclusive_scan(head);
        if (lid == 31) ws[w] = inc;
        __syncthreads();
        int off = carry + inc - head;
        for (int q = 0; q < w; ++q) off += ws[q];
        if (head) {
            const unsigned key = keys[i];
            float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f; int cnt = 0;
            for (int m = i; m < n && keys[m] == key; ++m) {
                const float4 p = s.in[vals[m]];
                sx = PM_ADD(sx, p.x); sy = PM_ADD(sy, p.y); sz = PM_ADD(sz, p.z); sw = PM_ADD(sw, p.w); ++cnt;
            }
            const float fc = (float)cnt;
            if (off < s.outCap) s.out[off] = make_float4(PM_DIV(sx, fc), PM_DIV(sy, fc), PM_DIV(sz, fc), PM_DIV(sw, fc));
        }
        __syncthreads();
        if (tid == 0) { int t = 0; for (int q = 0; q < 8; ++q) t += ws[q]; carry += t; }
        __syncthreads();
    }
    if (t1 == n && tid == 0) {
        int total = carry;
        if (total > s.outCap) { total = s.outCap; s.state[10] = 1; }
        *s.nOut = total;
    }
    __syncthreads();
    }
}

int vg_sort_keys(const VgBatch& b, cudaStream_t stream) {
    if (b.nSeg <= 0) return 0;
    dim3 g(b.maxBlocks < 8 ? b.maxBlocks : 8, b.nSeg);
    vg_init_kernel<<<(b.nSeg + 127) / 128, 128, 0, stream>>>(b);
    vg_bbox_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    vg_keys_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    for (int pass = 0; pass < 4; ++pass) {
        rs_hist_kernel<<<g, VG_THREADS, 0, stream>>>(b, pass * 8);
        rs_scan_kernel<<<b.nSeg, 256, 0, stream>>>(b);
        rs_scatter_kernel<<<g, VG_THREADS, 0, stream>>>(b, pass * 8);
    }
    return 15;
}

int vg_filter(const VgBatch& b, cudaStream_t stream) {
    if (b.nSeg <= 0) return 0;
    // tiles per segment in flight: 8 x 4096 points cover every cloud of the 16/32-ring sensors in one trip; the kernels
    // stride over the live tiles, so a grid sized by the allocation (hundreds of empty CTAs per segment) is avoided
    dim3 g(b.maxBlocks < 8 ? b.maxBlocks : 8, b.nSeg);
    vg_init_kernel<<<(b.nSeg + 127) / 128, 128, 0, stream>>>(b);
    vg_bbox_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    vg_keys_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    for (int pass = 0; pass < 4; ++pass) {
        rs_hist_kernel<<<g, VG_THREADS, 0, stream>>>(b, pass * 8);
        rs_scan_kernel<<<b.nSeg, 256, 0, stream>>>(b);
        rs_scatter_kernel<<<g, VG_THREADS, 0, stream>>>(b, pass * 8);
    }
    vg_heads_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    vg_reduce_kernel<<<g, VG_THREADS, 0, stream>>>(b);
    return 17;
}

}  // namespace lego
