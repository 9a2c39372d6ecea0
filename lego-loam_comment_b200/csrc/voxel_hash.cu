// Build kernels of the voxel hash (see voxel_hash.cuh).
#include <algorithm>
#include "voxel_hash.cuh"

namespace lego {

// ---- build kernels; n[lane * nStride] = number of points of the lane's cloud ----
__global__ void __launch_bounds__(256) vh_clear_kernel(VoxelHash h) {
    const int lane = blockIdx.y;
    const int Hd = vh_table_size(min(h.n[lane * h.nStride], h.cap), h.H, h.minH);
    // grid-stride over the LIVE table: the grid is sized for a typical table, not for the allocation
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Hd; i += gridDim.x * blockDim.x) {
        h.keys[(size_t)lane * h.H + i] = VH_EMPTY;
        h.count[(size_t)lane * h.H + i] = 0;
    }
}

__global__ void __launch_bounds__(256) vh_insert_kernel(VoxelHash h, const float4* __restrict__ pts, size_t ptsStride) {
    const int lane = blockIdx.y;
    const int M = min(h.n[lane * h.nStride], h.cap);
    const unsigned mask = (unsigned)(vh_table_size(M, h.H, h.minH) - 1);
    unsigned long long* keys = h.keys + (size_t)lane * h.H;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M; i += gridDim.x * blockDim.x) {
        const float4 p = pts[(size_t)lane * ptsStride + i];
        const unsigned long long key = vh_pack(vh_cell(p.x, h.invCell), vh_cell(p.y, h.invCell), vh_cell(p.z, h.invCell));
        unsigned s = vh_hash(key) & mask;
        while (true) {
            const unsigned long long prev = atomicCAS(keys + s, VH_EMPTY, key);
            if (prev == VH_EMPTY || prev == key) break;
            s = (s + 1) & mask;
        }
        atomicAdd(h.count + (size_t)lane * h.H + s, 1);
        h.slotOf[(size_t)lane * h.cap + i] = (int)s;
    }
}

// exclusive scan of count[] into start[], one CTA of 1024 threads per lane
__global__ void __launch_bounds__(1024) vh_scan_kernel(VoxelHash h) {
    const int lane = blockIdx.x, tid = threadIdx.x, lid = tid & 31, w = tid >> 5;
    const int* cnt = h.count + (size_t)lane * h.H;
    int* st = h.start + (size_t)lane * h.H;
    const int Hd = vh_table_size(min(h.n[lane * h.nStride], h.cap), h.H, h.minH);
    __shared__ int wsum[32];
    __shared__ int carry;
    if (tid == 0) carry = 0;
    __syncthreads();
    // each thread owns 4 consecutive slots per round -> 4096 slots per round
    for (int base = 0; base < Hd; base += 4096) {
        const int i0 = base + tid * 4;
        int v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (i0 + k < Hd) ? cnt[i0 + k] : 0;
        const int mine = v[0] + v[1] + v[2] + v[3];
        const int inc = warp_inclusive_scan(mine);
        if (lid == 31) wsum[w] = inc;
        __syncthreads();
        if (w == 0) { const int s = warp_inclusive_scan(wsum[lid]); wsum[lid] = s; }
        __syncthreads();
        int off = carry + inc - mine + (w > 0 ? wsum[w - 1] : 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) { if (i0 + k < Hd) st[i0 + k] = off; off += v[k]; }
        __syncthreads();
        if (tid == 0) carry += wsum[31];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) vh_scatter_kernel(VoxelHash h, const float4* __restrict__ pts, size_t ptsStride) {
    const int lane = blockIdx.y;
    const int M = min(h.n[lane * h.nStride], h.cap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M; i += gridDim.x * blockDim.x) {
        const float4 p = pts[(size_t)lane * ptsStride + i];
        const int s = h.slotOf[(size_t)lane * h.cap + i];
        // start[] doubles as the cursor: it ends up at the END of the run, which is what the look-ups read
        const int pos = atomicAdd(h.start + (size_t)lane * h.H + s, 1);
        h.sorted[(size_t)lane * h.cap + pos] = make_float4(p.x, p.y, p.z, __int_as_float(i));
    }
}

int vh_build(const VoxelHash& h, int lanes, const float4* pts, size_t ptsStride, cudaStream_t stream) {
    dim3 gH(std::min((h.H + 255) / 256, 16), lanes);
    const int pb = std::min((h.cap + 255) / 256, 16);   // the insert / scatter kernels stride over the points
    dim3 gP(pb < 1 ? 1 : pb, lanes);
    vh_clear_kernel<<<gH, 256, 0, stream>>>(h);
    vh_insert_kernel<<<gP, 256, 0, stream>>>(h, pts, ptsStride);
    vh_scan_kernel<<<lanes, 1024, 0, stream>>>(h);
    vh_scatter_kernel<<<gP, 256, 0, stream>>>(h, pts, ptsStride);
    return 4;
}

}  // namespace lego
