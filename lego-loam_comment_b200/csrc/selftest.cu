// Device self-tests exported through the C ABI so the parity suite can check the two building blocks whose
// bit-exactness everything else rests on, directly on the GPU: the glibc atan2f model and the std::sort replay.
#include <algorithm>
#include <cstring>
#include <vector>
#include "engine.h"
#include "parity_math.cuh"
#include "std_sort_replay.cuh"
#include "warp_sort.cuh"
#include "voxel_grid.cuh"

namespace lego {

__global__ void selftest_atan2f_kernel(const float* y, const float* x, float* out, long n) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = pm::atan2f_glibc(y[i], x[i]);
}

// one thread per segment sorts its slice in place (global memory) with the replayed std::sort
__global__ void selftest_sort_kernel(ssr::Item* items, const int* offsets, int nseg) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nseg) ssr::std_sort(items + offsets[s], offsets[s + 1] - offsets[s]);
}

// one warp per segment (<= WS_MAX_SORT entries): slice staged in shared memory, sorted with warp_std_sort
__global__ void __launch_bounds__(32) selftest_warp_sort_kernel(ssr::Item* items, const int* offsets, int nseg) {
    __shared__ ssr::Item a[WS_MAX_SORT];
    __shared__ unsigned short oA[WS_MAX_SORT], oB[WS_MAX_SORT], h[512];
    const int s = blockIdx.x, lid = threadIdx.x;
    const int n = offsets[s + 1] - offsets[s];
    ssr::Item* g = items + offsets[s];
    for (int i = lid; i < n; i += 32) a[i] = g[i];
    __syncwarp();
    warp_std_sort(a, n, oA, oB, h, lid);
    __syncwarp();
    for (int i = lid; i < n; i += 32) g[i] = a[i];
}

}  // namespace lego

namespace lego {
__global__ void selftest_sincosf_kernel(const float* x, float* s, float* c, long n) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { s[i] = pm::sinf_glibc(x[i]); c[i] = pm::cosf_glibc(x[i]); }
}
}  // namespace lego

extern "C" {

int lego_selftest_sincosf(const float* x, float* sin_out, float* cos_out, long n) {
    try {
        float *dx, *ds, *dc;
        lego::cuda_check(cudaMalloc(&dx, n * 4), "malloc"); lego::cuda_check(cudaMalloc(&ds, n * 4), "malloc"); lego::cuda_check(cudaMalloc(&dc, n * 4), "malloc");
        cudaMemcpy(dx, x, n * 4, cudaMemcpyHostToDevice);
        lego::selftest_sincosf_kernel<<<(unsigned)((n + 255) / 256), 256>>>(dx, ds, dc, n);
        lego::cuda_check(cudaMemcpy(sin_out, ds, n * 4, cudaMemcpyDeviceToHost), "selftest copy");
        lego::cuda_check(cudaMemcpy(cos_out, dc, n * 4, cudaMemcpyDeviceToHost), "selftest copy");
        cudaFree(dx); cudaFree(ds); cudaFree(dc);
        return LEGO_OK;
    } catch (...) { return LEGO_ERR_CUDA; }
}

int lego_selftest_atan2f(const float* y, const float* x, float* out, long n) {
    try {
        float *dy, *dx, *dout;
        lego::cuda_check(cudaMalloc(&dy, n * 4), "malloc"); lego::cuda_check(cudaMalloc(&dx, n * 4), "malloc"); lego::cuda_check(cudaMalloc(&dout, n * 4), "malloc");
        cudaMemcpy(dy, y, n * 4, cudaMemcpyHostToDevice); cudaMemcpy(dx, x, n * 4, cudaMemcpyHostToDevice);
        lego::selftest_atan2f_kernel<<<(unsigned)((n + 255) / 256), 256>>>(dy, dx, dout, n);
        lego::cuda_check(cudaMemcpy(out, dout, n * 4, cudaMemcpyDeviceToHost), "selftest copy");
        cudaFree(dy); cudaFree(dx); cudaFree(dout);
        return LEGO_OK;
    } catch (...) { return LEGO_ERR_CUDA; }
}

// values: total floats; offsets: nseg+1 ints; inds_out: total ints = for every segment the original positions
// (relative to the segment) in sorted order.
int lego_selftest_std_sort(const float* values, const int* offsets, int nseg, int* inds_out) {
    try {
        const int total = offsets[nseg];
        std::vector<ssr::Item> h(total);
        for (int s = 0; s < nseg; ++s) for (int i = offsets[s]; i < offsets[s + 1]; ++i) { h[i].value = values[i]; h[i].ind = i - offsets[s]; }
        ssr::Item* d; int* doff;
        lego::cuda_check(cudaMalloc(&d, sizeof(ssr::Item) * (total + 1)), "malloc"); lego::cuda_check(cudaMalloc(&doff, 4 * (nseg + 1)), "malloc");
        cudaMemcpy(d, h.data(), sizeof(ssr::Item) * total, cudaMemcpyHostToDevice); cudaMemcpy(doff, offsets, 4 * (nseg + 1), cudaMemcpyHostToDevice);
        lego::selftest_sort_kernel<<<(nseg + 63) / 64, 64>>>(d, doff, nseg);
        lego::cuda_check(cudaMemcpy(h.data(), d, sizeof(ssr::Item) * total, cudaMemcpyDeviceToHost), "selftest copy");
        cudaFree(d); cudaFree(doff);
        for (int i = 0; i < total; ++i) inds_out[i] = h[i].ind;
        return LEGO_OK;
    } catch (...) { return LEGO_ERR_CUDA; }
}

// Same contract as lego_selftest_std_sort, sorted by one warp per segment (csrc/warp_sort.cuh); every segment must
// hold at most 416 values.
int lego_selftest_warp_sort(const float* values, const int* offsets, int nseg, int* inds_out) {
    try {
        const int total = offsets[nseg];
        for (int s = 0; s < nseg; ++s) if (offsets[s + 1] - offsets[s] > lego::WS_MAX_SORT) { lego::set_last_error("segment larger than 416"); return LEGO_ERR_ARG; }
        std::vector<ssr::Item> h(total);
        for (int s = 0; s < nseg; ++s) for (int i = offsets[s]; i < offsets[s + 1]; ++i) { h[i].value = values[i]; h[i].ind = i - offsets[s]; }
        ssr::Item* d; int* doff;
        lego::cuda_check(cudaMalloc(&d, sizeof(ssr::Item) * (total + 1)), "malloc"); lego::cuda_check(cudaMalloc(&doff, 4 * (nseg + 1)), "malloc");
        cudaMemcpy(d, h.data(), sizeof(ssr::Item) * total, cudaMemcpyHostToDevice); cudaMemcpy(doff, offsets, 4 * (nseg + 1), cudaMemcpyHostToDevice);
        if (nseg > 0) lego::selftest_warp_sort_kernel<<<nseg, 32>>>(d, doff, nseg);
        lego::cuda_check(cudaMemcpy(h.data(), d, sizeof(ssr::Item) * total, cudaMemcpyDeviceToHost), "selftest copy");
        cudaFree(d); cudaFree(doff);
        for (int i = 0; i < total; ++i) inds_out[i] = h[i].ind;
        return LEGO_OK;
    } catch (...) { return LEGO_ERR_CUDA; }
}

// Batched VoxelGrid on nseg independent clouds (offsets has nseg+1 entries, in points): out gets the filtered
// clouds back to back in segment order, out_counts[s] their sizes.
int lego_selftest_voxel_grid(const float* xyzi, const int* offsets, int nseg, float leaf, float* out, int* out_counts) {
    try {
        using lego::VgSeg;
        const int total = offsets[nseg];
        int maxCap = 1;
        for (int s = 0; s < nseg; ++s) maxCap = std::max(maxCap, offsets[s + 1] - offsets[s]);
        lego::VgBatch b; b.nSeg = nseg; b.maxBlocks = (maxCap + lego::VG_TILE - 1) / lego::VG_TILE;
        float4 *din, *dout; int *dn, *dno, *dhist, *dheads, *dstate; unsigned* dscratch;
        const size_t tot = (size_t)std::max(total, 1);
        lego::cuda_check(cudaMalloc(&din, tot * 16), "malloc"); lego::cuda_check(cudaMalloc(&dout, tot * 16), "malloc");
        lego::cuda_check(cudaMalloc(&dn, nseg * 4), "malloc"); lego::cuda_check(cudaMalloc(&dno, nseg * 4), "malloc");
        lego::cuda_check(cudaMalloc(&dhist, (size_t)nseg * 256 * b.maxBlocks * 4), "malloc");
        lego::cuda_check(cudaMalloc(&dheads, (size_t)nseg * b.maxBlocks * 4), "malloc");
        lego::cuda_check(cudaMalloc(&dstate, (size_t)nseg * 16 * 4), "malloc");
        lego::cuda_check(cudaMalloc(&dscratch, tot * 4 * 4), "malloc");
        lego::cuda_check(cudaMalloc(&b.segs, sizeof(VgSeg) * nseg), "malloc");
        std::vector<VgSeg> segs(nseg); std::vector<int> n(nseg);
        for (int s = 0; s < nseg; ++s) {
            VgSeg& g = segs[s];
            n[s] = offsets[s + 1] - offsets[s];
            g.in = din + offsets[s]; g.nIn = dn + s; g.out = dout + offsets[s]; g.nOut = dno + s; g.active = nullptr;
            g.leaf = leaf; g.cap = n[s]; g.outCap = n[s];
            g.keys[0] = dscratch + offsets[s]; g.keys[1] = dscratch + tot + offsets[s];
            g.vals[0] = dscratch + 2 * tot + offsets[s]; g.vals[1] = dscratch + 3 * tot + offsets[s];
            g.blockHist = dhist + (size_t)s * 256 * b.maxBlocks; g.blockHeads = dheads + (size_t)s * b.maxBlocks; g.state = dstate + s * 16;
        }
        cudaMemcpy(din, xyzi, (size_t)total * 16, cudaMemcpyHostToDevice);
        cudaMemcpy(dn, n.data(), nseg * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(b.segs, segs.data(), sizeof(VgSeg) * nseg, cudaMemcpyHostToDevice);
        lego::vg_filter(b, 0);
        lego::cuda_check(cudaDeviceSynchronize(), "voxel grid selftest");
        lego::cuda_check(cudaMemcpy(out_counts, dno, nseg * 4, cudaMemcpyDeviceToHost), "selftest copy");
        std::vector<float> tmp((size_t)tot * 4);
        cudaMemcpy(tmp.data(), dout, (size_t)total * 16, cudaMemcpyDeviceToHost);
        size_t w = 0;
        for (int s = 0; s < nseg; ++s) { std::memcpy(out + w * 4, tmp.data() + (size_t)offsets[s] * 4, (size_t)out_counts[s] * 16); w += out_counts[s]; }
        cudaFree(din); cudaFree(dout); cudaFree(dn); cudaFree(dno); cudaFree(dhist); cudaFree(dheads); cudaFree(dstate); cudaFree(dscratch); cudaFree(b.segs);
        return LEGO_OK;
    } catch (...) { return LEGO_ERR_CUDA; }
}

}  // extern "C"
