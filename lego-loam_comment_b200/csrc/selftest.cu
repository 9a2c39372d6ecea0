// Device self-tests exported through the C ABI so the parity suite can check the two building blocks whose
// bit-exactness everything else rests on, directly on the GPU: the glibc atan2f model and the std::sort replay.
#include <vector>
#include "engine.h"
#include "parity_math.cuh"
#include "std_sort_replay.cuh"

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

}  // namespace lego

extern "C" {

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

}  // extern "C"
