// FeatureAssociation, odometry half: scan-to-scan two-step LM (featureAssociation.cpp:1044-1478, 1666-1695),
// pose integration (featureAssociation.cpp:1697-1725) and the end-of-sweep re-projection + "last" cloud swap
// (featureAssociation.cpp:1759-1815).  blockIdx.y / blockIdx.x = lane; the whole LM loop of one lane runs inside
// one persistent CTA so there is no host round trip between iterations.
#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include "engine.h"
#include "imu_kernels.cuh"
#include "kernel_utils.cuh"
#include "odometry_kernels.cuh"

namespace lego {

__global__ void fa_init_state_kernel(FaDev f, int L);

VoxelHash Engine::allocHash(int cap, float cell, const int* n, int nStride, int lanes) {
    VoxelHash h;
    const size_t nl = lanes > 0 ? lanes : c.L;
    int H = 1024;
    while (H < 2 * cap) H <<= 1;
    h.H = H; h.cap = cap; h.invCell = 1.0f / cell; h.n = n; h.nStride = nStride;
    { const char* e = std::getenv("LEGO_VH_MIN"); h.minH = e && *e ? std::min(H, std::max(1024, std::atoi(e))) : 1024; }
    h.keys = alloc<unsigned long long>(nl * H); h.count = alloc<int>(nl * H); h.start = alloc<int>(nl * H);
    h.slotOf = alloc<int>(nl * cap); h.sorted = alloc<float4>(nl * cap);
    return h;
}

void Engine::allocOdometry() {
    const size_t L = c.L, N = c.N, R = c.R;
    // CTAs per lane of the scan-to-scan LM kernel: a cluster of up to 16 (beyond the portable 8: opt-in attribute) while that
    // still leaves every lane its own SMs (the B200 has 148); LEGO_ODO_CLUSTER = 1 / 2 / 4 / 8 / 16 forces a size
    {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, params.device);
        const char* e = std::getenv("LEGO_ODO_CLUSTER");
        int want = e && *e ? std::atoi(e) : 0;
        if (want != 1 && want != 2 && want != 4 && want != 8 && want != 16) { want = 16; while (want > 1 && (int)L * want > sms) want >>= 1; }
        odoCluster = want;
    }
    for (int b = 0; b < 2; ++b) { fa.cornerLast[b] = alloc<float4>(L * R * 120); fa.surfLast[b] = alloc<float4>(L * N); }
    fa.lastCounts = alloc<int>(L * 4);
    fa.transformCur = alloc<float>(L * 6); fa.transformSum = alloc<float>(L * 6); fa.matP = alloc<float>(L * 9);
    fa.lmState = alloc<int>(L * 8);
    fa.cornerInd = alloc<int>(L * R * 12 * 2); fa.surfInd = alloc<int>(L * R * 24 * 3);
    fa.partials = alloc<double>(L * 16); fa.tickets = alloc<unsigned int>(L);
    fa.outlierLast = alloc<float4>(L * c.Ocap); fa.flags = alloc<int>(L);
    fa.hashCorner = allocHash(c.R * 120, ODO_CORNER_CELL, fa.lastCounts + 0, 4); fa.hashSurf = allocHash(c.N, 1.0f, fa.lastCounts + 1, 4);
    fa_init_state_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(fa, c.L);
}

bool Engine::lookupOdometry(int lane, const std::string& n, void** ptr, int* elem, long* count) {
    const size_t N = c.N, R = c.R, l = lane;
    auto set = [&](void* p, int e, long cnt) { *ptr = p; *elem = e; *count = cnt; return true; };
    auto cnt_of = [&](int* base, int stride, int which) {
        int v = 0;
        cuda_check(cudaMemcpyAsync(&v, base + l * stride + which, sizeof(int), cudaMemcpyDeviceToHost, stream), "count");
        cuda_check(cudaStreamSynchronize(stream), "count sync");
        return (long)v;
    };
    if (n == "transformCur") return set(fa.transformCur + l * 6, 4, 6);
    if (n == "transformSum") return set(fa.transformSum + l * 6, 4, 6);
    if (n == "matP") return set(fa.matP + l * 9, 4, 9);
    if (n == "lmState") return set(fa.lmState + l * 8, 4, 8);
    if (n == "lastCounts") return set(fa.lastCounts + l * 4, 4, 4);
    if (n == "faFlags") return set(fa.flags + l, 4, 1);
    if (n == "laserCloudCornerLast") return set(fa.cornerLast[faBuf] + l * R * 120, 16, cnt_of(fa.lastCounts, 4, 0));
    if (n == "laserCloudSurfLast") return set(fa.surfLast[faBuf] + l * N, 16, cnt_of(fa.lastCounts, 4, 1));
    if (n == "outlierCloudLast") return set(fa.outlierLast + l * c.Ocap, 16, cnt_of(ip.counts, 8, 1));
    return lookupMapOptimization(lane, n, ptr, elem, count);
}

// ------------------------------------------------------------------------------------------------------------
template <bool CL>
__device__ __forceinline__ void odometry_body(IpDev& d, FaDev& f, Consts& c, int buf, int lane, bool leader) {
    extern __shared__ __align__(16) unsigned char odo_smem[];
    __shared__ OdoShared sh;
    float4* sel = reinterpret_cast<float4*>(odo_smem);
    const int tid = threadIdx.x;
    if (!f.lmState[lane * 8 + 0]) return;                       // first scan: checkSystemInitialization only
    const int nCornerLast = f.lastCounts[lane * 4 + 0], nSurfLast = f.lastCounts[lane * 4 + 1];
    if (nCornerLast < 10 || nSurfLast < 100) return;            // featureAssociation.cpp:1668-1669
    if (tid < 6) sh.T[tid] = f.transformCur[lane * 6 + tid];
    if (tid < 9) sh.matP[tid] = f.matP[lane * 9 + tid];
    if (tid == 0) sh.isDegenerate = f.lmState[lane * 8 + 1];
    __syncthreads();
    const size_t R = c.R, N = c.N;
    int itS = 0, itC = 0;
    {
        const VhLane vh = vh_lane(f.hashSurf, lane);
        itS = odo_phase<true, CL>(sh, sel, f.flat + lane * R * 24, f.featCounts[lane * 4 + 2], f.surfLast[buf] + lane * N, nSurfLast, vh,
                        f.surfInd + lane * R * 24 * 3, c);
    }
    {
        const VhLane vh = vh_lane(f.hashCorner, lane);
        itC = odo_phase<false, CL>(sh, sel, f.sharp + lane * R * 12, f.featCounts[lane * 4 + 0], f.cornerLast[buf] + lane * R * 120, nCornerLast, vh,
                         f.cornerInd + lane * R * 12 * 2, c);
    }
    if (!leader) return;        // every CTA of a cluster ends with the same state; the first one stores it
    if (tid < 6) f.transformCur[lane * 6 + tid] = sh.T[tid];
    if (tid < 9) f.matP[lane * 9 + tid] = sh.matP[tid];
    if (tid == 0) { f.lmState[lane * 8 + 1] = sh.isDegenerate; f.lmState[lane * 8 + 6] = itS; f.lmState[lane * 8 + 7] = itC; }
}

// 256 threads with registers left for 5 CTAs per SM: 740 resident CTAs hold all 592 lanes of the default bench in ONE wave
// (512 threads x 2 per SM = 296 ran two), 15.9 -> 13.8 ms per 25-scan step although the kernel spills ~1 KB per thread;
// 512 x 1 (114 registers, no spills) 19.6 ms, 512 x 3 16.2 ms, 384 x 2 16.6 ms, 256 x 6 14.2 ms (profiles/README.md)
#ifndef LEGO_ODO_MINBLOCKS
#define LEGO_ODO_MINBLOCKS 5
#endif
#define ODO_BOUNDS __launch_bounds__(ODO_THREADS, LEGO_ODO_MINBLOCKS)
__global__ void ODO_BOUNDS fa_odometry_kernel(IpDev d, FaDev f, Consts c, int buf) {
    odometry_body<false>(d, f, c, buf, blockIdx.x, true);
}

// The same kernel for few lanes: one thread-block cluster (2, 4 or 8 CTAs on neighbouring SMs, launch attribute) per lane.
// One CTA per lane leaves a GPU with fewer lanes than SMs mostly idle while each lane's ~130 searches per refresh queue up
// behind 16 warps.
__global__ void __launch_bounds__(ODO_THREADS) fa_odometry_cluster_kernel(IpDev d, FaDev f, Consts c, int buf) {
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
    odometry_body<true>(d, f, c, buf, blockIdx.x / cluster.num_blocks(), cluster.block_rank() == 0);
}

// publishCloudsLast (featureAssociation.cpp:1759-1789) / checkSystemInitialization (featureAssociation.cpp:1605-1637):
// the less-sharp and less-flat clouds become the next "last" clouds, re-projected to the end of the sweep except on
// the very first scan; adjustOutlierCloud (featureAssociation.cpp:1746-1757) swaps the outliers into camera axes.
__global__ void __launch_bounds__(256) fa_clouds_last_kernel(IpDev d, FaDev f, Consts c, int nb) {
    const int lane = blockIdx.y;
    const size_t R = c.R, N = c.N;
    const bool inited = f.lmState[lane * 8 + 0] != 0;
    float T[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) T[k] = f.transformCur[lane * 6 + k];
    const int nC = f.featCounts[lane * 4 + 1], nS = f.featCounts[lane * 4 + 3], nO = d.counts[lane * 8 + 1];
    const int t0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    const bool imu = f.imu.ptr[lane * 2] >= 0;   // the IMU tail of TransformToEnd is the identity without IMU data
    const float* st = f.imu.state + lane * IMU_STATE;
    auto to_end = [&](float4 p) { const float4 q = pose::transform_to_end(p, T); return imu ? imu_transform_to_end_tail(st, q) : q; };
    const float4* ls = f.lessSharp + lane * R * 120; float4* cl = f.cornerLast[nb] + lane * R * 120;
    for (int i = t0; i < nC; i += stride) { const float4 p = ls[i]; cl[i] = inited ? to_end(p) : p; }
    const float4* lf = f.lessFlat + lane * N; float4* sl = f.surfLast[nb] + lane * N;
    for (int i = t0; i < nS; i += stride) { const float4 p = lf[i]; sl[i] = inited ? to_end(p) : p; }
    const float4* oc = d.outlier + (size_t)lane * c.Ocap; float4* ol = f.outlierLast + (size_t)lane * c.Ocap;
    for (int i = t0; i < nO; i += stride) { const float4 p = oc[i]; ol[i] = make_float4(p.y, p.z, p.x, p.w); }
}

// integrateTransformation (featureAssociation.cpp:1697-1725) and the publish bookkeeping of publishCloudsLast
// (frameCount / skipFrameNum, featureAssociation.cpp:1790-1794), one thread per lane.
__global__ void fa_finish_scan_kernel(IpDev d, FaDev f, Consts c) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= c.L) return;
    int flags = 1;
    const float* st = f.imu.state + lane * IMU_STATE;
    if (!f.lmState[lane * 8 + 0]) {
        f.lmState[lane * 8 + 0] = 1;
        flags |= 16;   // corner/surf "last" clouds are published once at initialisation, without odometry
        // checkSystemInitialization (featureAssociation.cpp:1632-1633); both are 0 without IMU data
        f.transformSum[lane * 6 + 0] += st[IS_PITCH_START];
        f.transformSum[lane * 6 + 2] += st[IS_ROLL_START];
    } else {
        float S[6], T[6];
        for (int k = 0; k < 6; ++k) { S[k] = f.transformSum[lane * 6 + k]; T[k] = f.transformCur[lane * 6 + k]; }
        // PluginIMURotation(rx, ry, rz, imuPitchStart, imuYawStart, imuRollStart, imuPitchLast, imuYawLast, imuRollLast, ...)
        const float imuRot[6] = {st[IS_PITCH_START], st[IS_YAW_START], st[IS_ROLL_START], st[IS_RPY_LAST + 1], st[IS_RPY_LAST + 2], st[IS_RPY_LAST + 0]};
        pose::integrate_transformation(S, T, imuRot);
        for (int k = 0; k < 6; ++k) f.transformSum[lane * 6 + k] = S[k];
        flags |= 2;
        int fc = f.lastCounts[lane * 4 + 3] + 1;
        if (fc >= 2) { fc = 0; flags |= 4; }   // skipFrameNum = 1 (featureAssociation.cpp:284)
        f.lastCounts[lane * 4 + 3] = fc;
    }
    f.lastCounts[lane * 4 + 0] = f.featCounts[lane * 4 + 1];
    f.lastCounts[lane * 4 + 1] = f.featCounts[lane * 4 + 3];
    f.flags[lane] = flags;
}

__global__ void fa_init_state_kernel(FaDev f, int L) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane < L) f.lastCounts[lane * 4 + 3] = 1;   // frameCount = skipFrameNum (featureAssociation.cpp:314)
}

// updateInitialGuess (featureAssociation.cpp:1639-1664), one thread per lane; a no-op without IMU data (everything is 0)
__global__ void fa_imu_initial_guess_kernel(FaDev f, Consts c) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= c.L || !f.lmState[lane * 8 + 0]) return;   // the first scan only runs checkSystemInitialization
    float* st = f.imu.state + lane * IMU_STATE;
    float* T = f.transformCur + lane * 6;
    for (int k = 0; k < 3; ++k) { st[IS_RPY_LAST + k] = st[IS_RPY_CUR + k]; st[IS_VELO_FROM_START + k] = st[IS_VELO_FROM_START_CUR + k]; }
    const float ax = st[IS_ANG_FROM_START], ay = st[IS_ANG_FROM_START + 1], az = st[IS_ANG_FROM_START + 2];
    if (ax != 0 || ay != 0 || az != 0) { T[0] = -ay; T[1] = -az; T[2] = -ax; }
    const float vx = st[IS_VELO_FROM_START], vy = st[IS_VELO_FROM_START + 1], vz = st[IS_VELO_FROM_START + 2];
    if (vx != 0 || vy != 0 || vz != 0) { T[3] -= vx * c.scanPeriod; T[4] -= vy * c.scanPeriod; T[5] -= vz * c.scanPeriod; }
}

void Engine::updateTransformation() {
    if (anyImu) { ProfScope ps(this, "fa_imu_initial_guess_kernel"); fa_imu_initial_guess_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(fa, c); launches += 1; }
    const size_t smem = sizeof(float4) * 24 * (size_t)c.R;
    static bool attr_set = false;
    if (!attr_set) {
        cuda_check(cudaFuncSetAttribute(fa_odometry_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024), "smem attr");
        cuda_check(cudaFuncSetAttribute(fa_odometry_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024), "smem attr");
        cuda_check(cudaFuncSetAttribute(fa_odometry_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1), "cluster attr");
        attr_set = true;
    }
    if (odoCluster > 1) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(c.L * odoCluster)); cfg.blockDim = dim3(ODO_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)odoCluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        ProfScope ps(this, "fa_odometry_kernel");
        cuda_check(cudaLaunchKernelEx(&cfg, fa_odometry_cluster_kernel, ip, fa, c, faBuf), "cluster launch");
    } else {
        ProfScope ps(this, "fa_odometry_kernel"); fa_odometry_kernel<<<c.L, ODO_THREADS, smem, stream>>>(ip, fa, c, faBuf);
    }
    launches += 1;
}

void Engine::integrateTransformationPublishCloudsLast() {
    integrateTransformationPublishCloudsLastLaunch();
    faBuf ^= 1;
}

void Engine::integrateTransformationPublishCloudsLastLaunch() {
    const int nb = faBuf ^ 1;
    dim3 g(ODO_TE_BLOCKS, c.L);
    { ProfScope ps(this, "fa_clouds_last_kernel"); fa_clouds_last_kernel<<<g, 256, 0, stream>>>(ip, fa, c, nb); }
    { ProfScope ps(this, "fa_finish_scan_kernel"); fa_finish_scan_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(ip, fa, c); }
    // the two kd-tree rebuilds of featureAssociation.cpp:1786-1787 become two voxel-hash builds
    launches += 2;
    { ProfScope ps(this, "vh_build(cornerLast)"); launches += vh_build(fa.hashCorner, c.L, fa.cornerLast[nb], (size_t)c.R * 120, stream); }
    { ProfScope ps(this, "vh_build(surfLast)"); launches += vh_build(fa.hashSurf, c.L, fa.surfLast[nb], (size_t)c.N, stream); }
}

}  // namespace lego

#ifdef LEGO_ODO_PROF
extern "C" int lego_debug_odo_prof(unsigned long long* out8) {
    return cudaMemcpyFromSymbol(out8, lego::g_odo_prof, sizeof(unsigned long long) * 8) == cudaSuccess ? 0 : -1;
}
extern "C" int lego_debug_odo_prof2(unsigned long long* out8) {
    return cudaMemcpyFromSymbol(out8, lego::g_odo_prof2, sizeof(unsigned long long) * 8) == cudaSuccess ? 0 : -1;
}
#endif
