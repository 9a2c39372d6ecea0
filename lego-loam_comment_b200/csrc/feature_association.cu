// FeatureAssociation, feature half: de-skew (adjustDistortion), curvature (calculateSmoothness), occlusion masks
// (markOccludedPoints) and the sorted greedy corner / flat selection with per-ring VoxelGrid (extractFeatures).
// Replaces featureAssociation.cpp:491-784 of the reference.  blockIdx.y = lane everywhere.
#include <cfloat>
#include <climits>
#include "engine.h"
#include "imu_kernels.cuh"
#include "kernel_utils.cuh"
#include "parity_math.cuh"
#include "std_sort_replay.cuh"

namespace lego {

// ------------------------------------------------------------------------------------------------------------
void Engine::allocFeatureAssociation() {
    const size_t L = c.L, N = c.N, R = c.R, C = c.C;
    fa.segAdj = alloc<float4>(L * N);
    // the reference allocates these with new[] and reads parts of them before ever writing (quirks Q8/Q11);
    // the oracle zero-fills them, and so does this allocation.  They persist across scans, never cleared.
    fa.curv = alloc<float>(L * N); fa.picked = alloc<int>(L * N); fa.label = alloc<int>(L * N);
    fa.slot4val = alloc<float>(L); fa.slot4ind = alloc<int>(L); fa.halfFirst = alloc<int>(L);
    fa.ringSharp = alloc<float4>(L * R * 12); fa.ringLessSharp = alloc<float4>(L * R * 120);
    fa.ringFlat = alloc<float4>(L * R * 24); fa.ringLessFlat = alloc<float4>(L * R * C);
    fa.ringCounts = alloc<int>(L * R * 4);
    fa.cand = alloc<int>(L * N); fa.candCounts = alloc<int>(L * R * 12);
    fa.sortScratch = C > 6 * 416 ? (void*)alloc<double>(L * N) : nullptr;
    fa.sharp = alloc<float4>(L * R * 12); fa.lessSharp = alloc<float4>(L * R * 120);
    fa.flat = alloc<float4>(L * R * 24); fa.lessFlat = alloc<float4>(L * N);
    fa.featCounts = alloc<int>(L * 4);
    fa.imu.time = alloc<double>(L * IMU_QUE); fa.imu.ring = alloc<float>(L * IMU_QUE * IMU_CH);
    fa.imu.ptr = alloc<int>(L * 2); fa.imu.scanTime = alloc<double>(L); fa.imu.state = alloc<float>(L * IMU_STATE);
    {   // imuPointerLast = -1 (featureAssociation.cpp:251); cos(0) = 1 for the start angles (updateImuRollPitchYawStartSinCos
        // has not run yet in the reference either, where these members are zero-initialised: sin = cos = 0 -- but they are
        // only ever read after an IMU message arrived, so the value before that is immaterial)
        std::vector<int> ptr(L * 2, 0);
        for (size_t l = 0; l < L; ++l) ptr[2 * l] = -1;
        cuda_check(cudaMemcpyAsync(fa.imu.ptr, ptr.data(), sizeof(int) * ptr.size(), cudaMemcpyHostToDevice, stream), "imu ptr init");
        cuda_check(cudaStreamSynchronize(stream), "imu ptr init sync");
    }
    allocOdometry();
}

bool Engine::lookupFeatureAssociation(int lane, const std::string& n, void** ptr, int* elem, long* count) {
    const size_t N = c.N, R = c.R, l = lane;
    auto set = [&](void* p, int e, long cnt) { *ptr = p; *elem = e; *count = cnt; return true; };
    auto cnt_of = [&](int* base, int stride, int which) {
        int v = 0;
        cuda_check(cudaMemcpyAsync(&v, base + l * stride + which, sizeof(int), cudaMemcpyDeviceToHost, stream), "count");
        cuda_check(cudaStreamSynchronize(stream), "count sync");
        return (long)v;
    };
    if (n == "segmentedCloudAdj") return set(fa.segAdj + l * N, 16, cnt_of(ip.counts, 8, 0));
    if (n == "cloudCurvature") return set(fa.curv + l * N, 4, cnt_of(ip.counts, 8, 0));
    if (n == "cloudNeighborPicked") return set(fa.picked + l * N, 4, cnt_of(ip.counts, 8, 0));
    if (n == "cloudLabel") return set(fa.label + l * N, 4, cnt_of(ip.counts, 8, 0));
    if (n == "slot4val") return set(fa.slot4val + l, 4, 1);
    if (n == "slot4ind") return set(fa.slot4ind + l, 4, 1);
    if (n == "halfFirst") return set(fa.halfFirst + l, 4, 1);
    if (n == "cornerPointsSharp") return set(fa.sharp + l * R * 12, 16, cnt_of(fa.featCounts, 4, 0));
    if (n == "cornerPointsLessSharp") return set(fa.lessSharp + l * R * 120, 16, cnt_of(fa.featCounts, 4, 1));
    if (n == "surfPointsFlat") return set(fa.flat + l * R * 24, 16, cnt_of(fa.featCounts, 4, 2));
    if (n == "surfPointsLessFlat") return set(fa.lessFlat + l * N, 16, cnt_of(fa.featCounts, 4, 3));
    if (n == "featCounts") return set(fa.featCounts + l * 4, 4, 4);
    if (n == "imuState") return set(fa.imu.state + l * IMU_STATE, 4, IMU_STATE);
    return lookupOdometry(lane, n, ptr, elem, count);
}

// ------------------------------------------------------------------------------------------------------------
// Orientation of a segmented point for the first half of the sweep ("branch A", featureAssociation.cpp:504-512):
// returns the unwrapped ori and whether this point would flip halfPassed.
__device__ __forceinline__ float ori_branch_a(float4 p, float startOri, bool* trigger) {
    // camera axes: point.x = lidar y, point.z = lidar x (featureAssociation.cpp:500-502)
    float ori = -pm::atan2f_glibc(p.y, p.x);
    if ((double)ori < (double)startOri - M_PI / 2) ori = (float)((double)ori + 2 * M_PI);
    else if ((double)ori > (double)startOri + M_PI * 3 / 2) ori = (float)((double)ori - 2 * M_PI);
    *trigger = (double)PM_SUB(ori, startOri) > M_PI;
    return ori;
}
__device__ __forceinline__ float ori_branch_b(float4 p, float endOri) {
    float ori = -pm::atan2f_glibc(p.y, p.x);
    ori = (float)((double)ori + 2 * M_PI);
    if ((double)ori < (double)endOri - M_PI * 3 / 2) ori = (float)((double)ori + 2 * M_PI);
    else if ((double)ori > (double)endOri + M_PI / 2) ori = (float)((double)ori - 2 * M_PI);
    return ori;
}

// Occlusion rule of one loop iteration k of markOccludedPoints (featureAssociation.cpp:647-670), valid for k in [5, S-6):
// column gap < 10 and r[k] - r[k+1] > 0.3 -> type 1 (marks k-5..k); r[k+1] - r[k] > 0.3 -> type 2 (marks k+1..k+6); else 0.

// calculateSmoothness + markOccludedPoints as one gather stencil per point, plus the halfPassed search of
// adjustDistortion.  The reference's scatter writes (picked[i-5..i] = 1 ...) are all "set to 1", so gathering the
// union of the rules that can reach index i is order-free and exact.  Entries the reference does not reset
// (i < 5, i >= S-5) keep their previous value OR-ed with this scan's marks, exactly as its persistent arrays do.
__global__ void __launch_bounds__(256) fa_prepare_kernel(IpDev d, FaDev f, Consts c) {
    // The block's 256 points plus a halo of 6 + 7 slots are staged once; the occlusion rule of every k in reach is then
    // evaluated ONCE per block (it was evaluated by each of the 12 points it can reach) and read back as a byte.
    constexpr int HB = 6, HA = 7, W = 256 + HB + HA;
    __shared__ float sr[W];
    __shared__ int scol[W];
    __shared__ unsigned char socc[W];
    const int lane = blockIdx.y;
    const int i0 = blockIdx.x * blockDim.x, tid = threadIdx.x, i = i0 + tid;
    const int S = d.counts[lane * 8 + 0];
    if (i0 >= S) return;
    const size_t base = (size_t)lane * c.N;
    const float* r = d.segRange + base;
    const int* col = d.segCol + base;
    for (int q = tid; q < W; q += 256) {
        const int k = i0 - HB + q;
        const bool in = k >= 0 && k < S;
        sr[q] = in ? r[k] : 0.0f;
        scol[q] = in ? col[k] : 0;
    }
    __syncthreads();
    for (int q = tid; q < W - 1; q += 256) {      // rule of loop iteration k = i0 - 6 + q (reads k and k + 1)
        const int k = i0 - HB + q;
        int t = 0;
        if (k >= 5 && k < S - 6 && abs(scol[q + 1] - scol[q]) < 10) {
            if ((double)PM_SUB(sr[q], sr[q + 1]) > 0.3) t = 1;
            else if ((double)PM_SUB(sr[q + 1], sr[q]) > 0.3) t = 2;
        }
        socc[q] = (unsigned char)t;
    }
    __syncthreads();
    if (i >= S) return;
    const int q = tid + HB;                        // this point's slot in the staged arrays
    int pk;
    if (i >= 5 && i < S - 5) {
        float s = PM_ADD(sr[q - 5], sr[q - 4]);
        s = PM_ADD(s, sr[q - 3]); s = PM_ADD(s, sr[q - 2]); s = PM_ADD(s, sr[q - 1]);
        s = PM_SUB(s, PM_MUL(sr[q], 10.0f));
        s = PM_ADD(s, sr[q + 1]); s = PM_ADD(s, sr[q + 2]); s = PM_ADD(s, sr[q + 3]); s = PM_ADD(s, sr[q + 4]); s = PM_ADD(s, sr[q + 5]);
        f.curv[base + i] = PM_MUL(s, s);
        f.label[base + i] = 0;
        pk = 0;
    } else {
        pk = f.picked[base + i];
    }
#pragma unroll
    for (int k = 0; k <= 5; ++k) if (socc[q + k] == 1) pk = 1;       // iterations i .. i+5 mark k-5 .. k
#pragma unroll
    for (int k = 1; k <= 6; ++k) if (socc[q - k] == 2) pk = 1;       // iterations i-6 .. i-1 mark k+1 .. k+6
    if (i >= 5 && i < S - 6) {
        const float d1 = fabsf(PM_SUB(sr[q - 1], sr[q])), d2 = fabsf(PM_SUB(sr[q + 1], sr[q]));
        if ((double)d1 > 0.02 * (double)sr[q] && (double)d2 > 0.02 * (double)sr[q]) pk = 1;
    }
    f.picked[base + i] = pk;
    // halfPassed: first index whose branch-A orientation is more than pi past the start (quirk Q9)
    bool trig;
    ori_branch_a(d.segCloud[base + i], d.orient[lane * 4 + 0], &trig);
    if (trig) atomicMin(f.halfFirst + lane, i);
}

// The i == 0 branch of adjustDistortion's IMU block (featureAssociation.cpp:570-610), one thread per lane: the IMU state at
// the first point of the sweep becomes the "start" state, and the integrated angular rotation since the previous sweep's
// start becomes imuAngularFromStart (the rotation prior of updateInitialGuess).
__global__ void fa_imu_start_kernel(IpDev d, FaDev f, Consts c) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= c.L || f.imu.ptr[lane * 2] < 0 || d.counts[lane * 8 + 0] < 1) return;
    const size_t base = (size_t)lane * c.N;
    const float so = d.orient[lane * 4 + 0], od = d.orient[lane * 4 + 2];
    bool trig;
    const float ori = ori_branch_a(d.segCloud[base], so, &trig);
    const float relTime = PM_DIV(PM_SUB(ori, so), od);
    const float pointTime = PM_MUL(relTime, c.scanPeriod);
    const ImuSample s = imu_interpolate(f.imu, lane, f.imu.scanTime[lane], pointTime);
    float* st = f.imu.state + lane * IMU_STATE;
    for (int k = 0; k < 3; ++k) {
        st[IS_ROLL_START + k] = s.rpy[k]; st[IS_VELO_START + k] = s.velo[k]; st[IS_SHIFT_START + k] = s.shift[k];
        st[IS_RPY_CUR + k] = s.rpy[k];
        st[IS_ANG_FROM_START + k] = s.ang[k] - st[IS_ANG_LAST + k];
        st[IS_ANG_LAST + k] = s.ang[k];
    }
    st[IS_COS_ROLL] = pm::cosf_glibc(s.rpy[0]); st[IS_COS_PITCH] = pm::cosf_glibc(s.rpy[1]); st[IS_COS_YAW] = pm::cosf_glibc(s.rpy[2]);
    st[IS_SIN_ROLL] = pm::sinf_glibc(s.rpy[0]); st[IS_SIN_PITCH] = pm::sinf_glibc(s.rpy[1]); st[IS_SIN_YAW] = pm::sinf_glibc(s.rpy[2]);
}

// adjustDistortion (featureAssociation.cpp:491-619): axis swap and intensity = ring + 0.1*relTime; with IMU data also the
// per-point de-skew TransformToStartIMU (the last point leaves its interpolated IMU state behind, as the loop does).
__global__ void __launch_bounds__(256) fa_distort_kernel(IpDev d, FaDev f, Consts c) {
    const int lane = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int S = d.counts[lane * 8 + 0];
    if (i >= S) return;
    const size_t base = (size_t)lane * c.N;
    const float4 p = d.segCloud[base + i];
    const float so = d.orient[lane * 4 + 0], eo = d.orient[lane * 4 + 1], od = d.orient[lane * 4 + 2];
    bool trig;
    float ori = (i <= f.halfFirst[lane]) ? ori_branch_a(p, so, &trig) : ori_branch_b(p, eo);
    const float relTime = PM_DIV(PM_SUB(ori, so), od);
    const float inten = PM_ADD((float)(int)p.w, PM_MUL(c.scanPeriod, relTime));
    float px = p.y, py = p.z, pz = p.x;
    if (f.imu.ptr[lane * 2] >= 0 && i > 0) {
        const float pointTime = PM_MUL(relTime, c.scanPeriod);
        const ImuSample s = imu_interpolate(f.imu, lane, f.imu.scanTime[lane], pointTime);
        float* st = f.imu.state + lane * IMU_STATE;
        imu_transform_to_start(st, s.rpy, px, py, pz);
        if (i == S - 1) {
            float v[3]; imu_velo_to_start(st, s.velo, v);
            for (int k = 0; k < 3; ++k) { st[IS_RPY_CUR + k] = s.rpy[k]; st[IS_VELO_FROM_START_CUR + k] = v[k]; }
        }
    }
    f.segAdj[base + i] = make_float4(px, py, pz, inten);
}

}  // namespace lego
