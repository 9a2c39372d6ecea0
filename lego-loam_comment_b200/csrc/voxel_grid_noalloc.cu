// lego_voxel_grid_noalloc -- pcl::VoxelGrid<PointXYZI>::filter on ONE device-resident cloud without any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It runs the same kernels
// as every other VoxelGrid call of the library (voxel_grid.cu: bounding box -> voxel keys -> stable LSD radix sort ->
// run heads -> float32 means in input order); only the SCRATCH differs: instead of buffers owned by a handle
// (Engine::makeVgBatch) or allocated per call (lego_selftest_voxel_grid), the sort's ping-pong key / value arrays, the
// block histograms, the per-segment state and the one-entry segment table live in the file-scope device array below.
//
//   * limit: LEGO_VG_NOALLOC_MAX_POINTS input points (262 144); scratch = 4.3 MB that every GPU which loads the
//     library holds until the process exits.
//   * every call shares that one array per device and finds the previous call's contents in it: the set-up kernel
//     rewrites the segment table and clears histograms, heads and state; keys / values are fully rewritten before
//     they are read.  Two calls that overlap on the same GPU (different streams, or different host threads) RACE.
//   * nothing here calls cudaMalloc / cudaFree or an allocating library routine, and nothing synchronises: the
//     launches are enqueued on the caller's stream and the call returns.
#include <mutex>
#include <string>
#include "../../include/lego_b200.h"
#include "engine.h"
#include "voxel_grid.cuh"

namespace {

constexpr int MAXN = LEGO_VG_NOALLOC_MAX_POINTS;
constexpr int MAXB = (MAXN + lego::VG_TILE - 1) / lego::VG_TILE;    // tiles of the largest cloud
static_assert(MAXN % 64 == 0, "sections of the scratch array stay 256-byte aligned");

// section offsets in 4-byte words
constexpr size_t OFF_KEYS0 = 0, OFF_KEYS1 = OFF_KEYS0 + MAXN, OFF_VALS0 = OFF_KEYS1 + MAXN, OFF_VALS1 = OFF_VALS0 + MAXN;
constexpr size_t OFF_HIST = OFF_VALS1 + MAXN;                        // [256][MAXB]
constexpr size_t OFF_HEADS = OFF_HIST + (size_t)256 * MAXB;          // [MAXB] (padded to 64 words)
constexpr size_t OFF_STATE = OFF_HEADS + ((MAXB + 63) / 64) * 64;    // [16]   (padded to 64 words)
constexpr size_t OFF_NIN = OFF_STATE + 64;                           // [1]    (padded to 64 words)
constexpr size_t OFF_SEG = OFF_NIN + 64;                             // one VgSeg (padded to 64 words = 256 B)
constexpr size_t SCRATCH_WORDS = OFF_SEG + 64;
static_assert(sizeof(lego::VgSeg) <= 256, "segment table entry fits its slot");

// In HOST code this name is only a stand-in for the symbol (not a device address): the host side below asks the runtime
// for the address once per device and hands that pointer to the kernels.
static __device__ __align__(256) unsigned g_vg_scratch[SCRATCH_WORDS];

// Builds the one-entry segment table inside the scratch array and clears what must start at zero.
__global__ void vg_noalloc_setup_kernel(unsigned* scratch, const float4* in, int n, float leaf, float4* out, int outCap, int* nOut) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (size_t i = OFF_HIST + tid; i < OFF_SEG; i += nth) scratch[i] = 0u;     // histograms, heads, state, nIn
    if (tid == 0) {
        lego::VgSeg g;
        g.in = in; g.nIn = reinterpret_cast<const int*>(scratch + OFF_NIN); g.out = out; g.nOut = nOut; g.active = nullptr;
        g.leaf = leaf; g.cap = MAXN; g.outCap = outCap;
        g.keys[0] = scratch + OFF_KEYS0; g.keys[1] = scratch + OFF_KEYS1;
        g.vals[0] = scratch + OFF_VALS0; g.vals[1] = scratch + OFF_VALS1;
        g.blockHist = reinterpret_cast<int*>(scratch + OFF_HIST); g.blockHeads = reinterpret_cast<int*>(scratch + OFF_HEADS);
        g.state = reinterpret_cast<int*>(scratch + OFF_STATE);
        *reinterpret_cast<lego::VgSeg*>(scratch + OFF_SEG) = g;
    }
}
// nIn is written by a second tiny kernel so that it cannot race the clearing loop above (different threads touch it there)
__global__ void vg_noalloc_count_kernel(unsigned* scratch, int n) { *reinterpret_cast<int*>(scratch + OFF_NIN) = n; }

// device address of g_vg_scratch, looked up once per device with that device current
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned* g_ptr[MAX_DEVICES] = {};

int scratch_for_current_device(unsigned** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { lego::set_last_error(std::string("cudaGetDevice: ") + cudaGetErrorString(e)); return LEGO_ERR_CUDA; }
    if (dev < 0 || dev >= MAX_DEVICES) { lego::set_last_error("device ordinal out of range"); return LEGO_ERR_ARG; }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_vg_scratch);
        if (e != cudaSuccess || !p) {
            lego::set_last_error(std::string("cudaGetSymbolAddress(g_vg_scratch): ") + cudaGetErrorString(e));
            return LEGO_ERR_CUDA;
        }
        g_ptr[dev] = static_cast<unsigned*>(p);
    }
    *out = g_ptr[dev];
    return LEGO_OK;
}

}  // namespace

extern "C" {

int lego_voxel_grid_noalloc_max_points(void) { return MAXN; }

int lego_voxel_grid_noalloc(const float* d_xyzi, int n, float leaf, float* d_out, int out_cap, int* d_n_out, void* cuda_stream) {
    if (!d_xyzi || !d_out || !d_n_out || !(leaf > 0.f) || out_cap < 0) {
        lego::set_last_error("lego_voxel_grid_noalloc: null pointer, leaf <= 0 or negative out_cap");
        return LEGO_ERR_ARG;
    }
    if (n < 0 || n > MAXN) {
        lego::set_last_error("lego_voxel_grid_noalloc: " + std::to_string(n) + " points, the static scratch holds at most " +
                             std::to_string(MAXN) + " (LEGO_VG_NOALLOC_MAX_POINTS)");
        return LEGO_ERR_CAPACITY;
    }
    unsigned* scratch = nullptr;
    const int rc = scratch_for_current_device(&scratch);
    if (rc != LEGO_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    vg_noalloc_setup_kernel<<<32, 256, 0, stream>>>(scratch, reinterpret_cast<const float4*>(d_xyzi), n, leaf,
                                                    reinterpret_cast<float4*>(d_out), out_cap, d_n_out);
    vg_noalloc_count_kernel<<<1, 1, 0, stream>>>(scratch, n);
    lego::VgBatch b;
    b.segs = reinterpret_cast<lego::VgSeg*>(scratch + OFF_SEG);
    b.nSeg = 1;
    b.maxBlocks = n > 0 ? (n + lego::VG_TILE - 1) / lego::VG_TILE : 1;
    lego::vg_filter(b, stream);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { lego::set_last_error(std::string("lego_voxel_grid_noalloc launch: ") + cudaGetErrorString(e)); return LEGO_ERR_CUDA; }
    return LEGO_OK;
}

}  // extern "C"
