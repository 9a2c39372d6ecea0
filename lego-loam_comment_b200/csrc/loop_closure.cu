// Loop closure of mapOptimization (reference LeGO-LOAM/src/mapOptmization.cpp:802-945, 1456-1478):
//   detectLoopClosure      host: radius search over a few hundred key POSES (814-873), like the reference
//   lc_assemble_kernel     latestSurfKeyFrameCloud / nearHistorySurfKeyFrameCloud: key-frame clouds out of the device
//                          pool, transformed by their key poses (840-858, 577-606)
//   vg_filter              downSizeFilterHistoryKeyFrames, leaf 0.4 (860-861)
//   vh_build + lc_icp_kernel   pcl::IterativeClosestPoint::align as configured at 892-905: one persistent CTA per loop
//                          candidate runs the whole ICP (exact voxel-hash nearest neighbour, cross-covariance reduced in
//                          double, 3x3 SVD, convergence tests, fitness score) without returning to the host
//   addLoopConstraint      host: the BetweenFactor of 915-942 on the lane's pose graph (pose_graph.h)
//   correctPoses           host: key poses rewritten from the graph estimate (1456-1478); the clouds stay in the pool in
//                          the sensor frame, so nothing moves on the device
// PCL and GTSAM are not vendored in the reference; the ICP and graph semantics follow their published algorithms
// (oracle/thirdparty/tp_icp.h and tp_pose_graph.h restate the same for the parity tests).
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <atomic>
#include <thread>
#include "engine.h"
#include "map_kernels.cuh"

namespace lego {

constexpr int LC_THREADS = 1024;
constexpr int LC_SUMS = 17;          // n, sum s (3), sum d (3), sum d s^T (9), sum d2
constexpr float LC_SHELL_D2 = 9.0f;  // the shell search is exact for neighbours closer than 3 m; farther ones take the scan path

__global__ void __launch_bounds__(256) lc_assemble_kernel(MoDev m, LcDev lc) {
    const int slot = blockIdx.y;
    if (!lc.slotActive[slot]) return;
    const int lane = lc.slotLane[slot];
    const int nE = lc.counts[slot * 4 + 3];
    const float4* pool = m.pool + (size_t)lane * m.poolCap;
    for (int e = blockIdx.x; e < nE; e += gridDim.x) {
        const LcEntry en = lc.entries[(size_t)slot * lc.maxEntries + e];
        MapSinCos t;
        t.cRoll = en.ctRoll; t.sRoll = en.stRoll; t.cPitch = en.ctPitch; t.sPitch = en.stPitch; t.cYaw = en.ctYaw; t.sYaw = en.stYaw;
        t.tX = en.tInX; t.tY = en.tInY; t.tZ = en.tInZ;
        float4* dst = en.which == 0 ? lc.src + (size_t)slot * lc.srcCap : lc.tgtRaw + (size_t)slot * lc.tgtRawCap;
        for (int i = threadIdx.x; i < en.n; i += blockDim.x) dst[en.dst + i] = map_transform(t, pool[en.src + i]);
    }
}

// y = M [p 1], float, left to right (pcl::transformPointCloud with a Matrix4f)
__device__ __forceinline__ float4 lc_apply(const float* M, float4 p) {
    float4 o;
    o.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[0], p.x), __fmul_rn(M[1], p.y)), __fmul_rn(M[2], p.z)), M[3]);
    o.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[4], p.x), __fmul_rn(M[5], p.y)), __fmul_rn(M[6], p.z)), M[7]);
    o.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[8], p.x), __fmul_rn(M[9], p.y)), __fmul_rn(M[10], p.z)), M[11]);
    o.w = p.w;
    return o;
}

__device__ inline double lc_det3(const double* M) {
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// A = U diag(s) V^T by one-sided Jacobi rotations of the columns of A (double); s descending.
__device__ inline void lc_svd3(const double* A, double* U, double* s, double* V) {
    double B[9];
    for (int i = 0; i < 9; ++i) { B[i] = A[i]; V[i] = (i % 4 == 0) ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < 3; ++r) { a += B[3 * r + p] * B[3 * r + p]; b += B[3 * r + q] * B[3 * r + q]; c += B[3 * r + p] * B[3 * r + q]; }
                off = fmax(off, fabs(c) / (sqrt(a * b) + 1e-300));
                if (fabs(c) <= 1e-300) continue;
                const double zeta = (b - a) / (2.0 * c);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int r = 0; r < 3; ++r) {
                    const double bp = B[3 * r + p], bq = B[3 * r + q];
                    B[3 * r + p] = cs * bp - sn * bq; B[3 * r + q] = sn * bp + cs * bq;
                    const double vp = V[3 * r + p], vq = V[3 * r + q];
                    V[3 * r + p] = cs * vp - sn * vq; V[3 * r + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
    int ord[3] = {0, 1, 2};
    double n[3];
    for (int j = 0; j < 3; ++j) n[j] = sqrt(B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j]);
    for (int i = 0; i < 2; ++i) for (int j = i + 1; j < 3; ++j) if (n[ord[j]] > n[ord[i]]) { const int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    double Vs[9];
    for (int j = 0; j < 3; ++j) {
        const int o = ord[j];
        s[j] = n[o];
        for (int r = 0; r < 3; ++r) { Vs[3 * r + j] = V[3 * r + o]; U[3 * r + j] = n[o] > 1e-300 ? B[3 * r + o] / n[o] : 0.0; }
    }
    for (int i = 0; i < 9; ++i) V[i] = Vs[i];
    if (s[2] <= 1e-300 * s[0] || s[2] == 0.0) {   // rank 2: complete U with u0 x u1
        U[2] = U[3] * U[7] - U[6] * U[4]; U[5] = U[6] * U[1] - U[0] * U[7]; U[8] = U[0] * U[4] - U[3] * U[1];
    }
}

// Eigen::umeyama(src, dst, false) from the raw moments of the pairs: acc = n, sum s, sum d, sum d s^T (row-major d x s)
__device__ inline void lc_umeyama(const double* acc, float* T) {
    const double n = acc[0];
    double ms[3], md[3], S[9];
    for (int a = 0; a < 3; ++a) { ms[a] = acc[1 + a] / n; md[a] = acc[4 + a] / n; }
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) S[3 * r + c] = acc[7 + 3 * r + c] / n - md[r] * ms[c];
    double U[9], d[3], V[9];
    lc_svd3(S, U, d, V);
    double sg[3] = {1.0, 1.0, 1.0};
    if (lc_det3(S) < 0) sg[2] = -1.0;
    int rank = 0;
    for (int i = 0; i < 3; ++i) if (!(fabs(d[i]) <= fabs(d[0]) * 1e-12)) ++rank;
    if (rank == 2) sg[2] = (lc_det3(U) * lc_det3(V) > 0) ? 1.0 : -1.0;
    double R[9];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { double a = 0; for (int k = 0; k < 3; ++k) a += U[3 * r + k] * sg[k] * V[3 * c + k]; R[3 * r + c] = a; }
    for (int i = 0; i < 16; ++i) T[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) T[4 * r + c] = (float)R[3 * r + c];
        T[4 * r + 3] = (float)(md[r] - (R[3 * r] * ms[0] + R[3 * r + 1] * ms[1] + R[3 * r + 2] * ms[2]));
    }
}

// One CTA per slot: the complete ICP of one loop candidate.
__global__ void __launch_bounds__(LC_THREADS) lc_icp_kernel(LcDev lc, LcIcpParams prm) {
    const int slot = blockIdx.x, tid = threadIdx.x, lid = tid & 31, warp = tid >> 5;
    if (!lc.slotActive[slot]) return;
    const int ns = min(lc.counts[slot * 4 + 0], lc.srcCap), nt = min(lc.counts[slot * 4 + 2], lc.tgtDSCap);
    const float4* src = lc.src + (size_t)slot * lc.srcCap;
    float4* work = lc.work + (size_t)slot * lc.srcCap;
    const float4* tgt = lc.tgtDS + (size_t)slot * lc.tgtDSCap;
    int* miss = lc.miss + (size_t)slot * lc.srcCap;
    int* nnIdx = lc.nnIdx + (size_t)slot * lc.srcCap;
    float* nnD2 = lc.nnD2 + (size_t)slot * lc.srcCap;
    LcResult& res = lc.results[slot];
    __shared__ double red[32][LC_SUMS];
    __shared__ double sTot[LC_SUMS];
    __shared__ float sT[16], sF[16];
    __shared__ int sMiss, sNext, sStop, sConverged, sIter;
    __shared__ double sPrevMSE;
    if (tid == 0) {
        for (int i = 0; i < 16; ++i) { sF[i] = (i % 5 == 0) ? 1.f : 0.f; sT[i] = sF[i]; }
        sStop = 0; sConverged = 0; sIter = 0; sPrevMSE = DBL_MAX;
    }
    __syncthreads();
    if (ns == 0 || nt == 0) {
        if (tid == 0) { res.converged = 0; res.iterations = 0; res.pairs = 0; res.nSrc = ns; res.fitness = 1e30; for (int i = 0; i < 16; ++i) res.T[i] = sF[i]; }
        return;
    }
    const VhLane vh = vh_lane(lc.hash, slot);
    for (int i = tid; i < ns; i += LC_THREADS) work[i] = src[i];
    __syncthreads();

    // exact nearest target point of every source point: voxel-hash shells out to 3 m per thread, then the few points
    // with nothing that close are scanned against the whole target, one warp per point (ties: lowest target index)
    auto nearest_all = [&](const float4* pts, bool withFinal) {
        if (tid == 0) { sMiss = 0; sNext = 0; }
        __syncthreads();
        // points are handed out one at a time: the shell search takes 1 to 125 cells depending on where the neighbour sits
        for (int i = atomicAdd(&sNext, 1); i < ns; i = atomicAdd(&sNext, 1)) {
            float4 p = pts[i];
            if (withFinal) p = lc_apply(sF, p);
            float d2;
            const int j = vh_nearest(vh, p.x, p.y, p.z, LC_SHELL_D2, 1.0f, &d2);
            if (j < 0) miss[atomicAdd(&sMiss, 1)] = i;
            nnIdx[i] = j; nnD2[i] = d2;
        }
        __syncthreads();
        const int nm = sMiss;
        for (int k = warp; k < nm; k += LC_THREADS / 32) {
            const int i = miss[k];
            float4 p = pts[i];
            if (withFinal) p = lc_apply(sF, p);
            float best = FLT_MAX; int bi = INT_MAX;
            for (int t = lid; t < nt; t += 32) {
                const float d = vh_dist2(tgt[t], p.x, p.y, p.z);
                if (d < best || (d == best && t < bi)) { best = d; bi = t; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float od = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (od < best || (od == best && oi < bi)) { best = od; bi = oi; }
            }
            if (lid == 0) { nnIdx[i] = bi; nnD2[i] = best; }
        }
        __syncthreads();
    };
    auto block_sum = [&](double* acc) {   // -> sTot, valid for every thread after the call
#pragma unroll
        for (int k = 0; k < LC_SUMS; ++k) {
            double v = acc[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lid == 0) red[warp][k] = v;
        }
        __syncthreads();
        if (tid < LC_SUMS) { double t = 0.0; for (int w = 0; w < LC_THREADS / 32; ++w) t += red[w][tid]; sTot[tid] = t; }
        __syncthreads();
    };

    while (true) {
        nearest_all(work, false);
        double acc[LC_SUMS];
#pragma unroll
        for (int k = 0; k < LC_SUMS; ++k) acc[k] = 0.0;
        for (int i = tid; i < ns; i += LC_THREADS) {
            const int j = nnIdx[i];
            const float d2 = nnD2[i];
            if (j < 0 || j == INT_MAX || d2 > prm.maxDist2) continue;
            const float4 s = work[i], d = tgt[j];
            const double sv[3] = {(double)s.x, (double)s.y, (double)s.z}, dv[3] = {(double)d.x, (double)d.y, (double)d.z};
            acc[0] += 1.0;
#pragma unroll
            for (int a = 0; a < 3; ++a) { acc[1 + a] += sv[a]; acc[4 + a] += dv[a]; }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) acc[7 + 3 * r + c] += dv[r] * sv[c];
            acc[16] += (double)d2;
        }
        block_sum(acc);
        if (tid == 0) {
            res.pairs = (int)sTot[0];
            if (sTot[0] < 3.0) { sStop = 1; sConverged = 0; }
            else lc_umeyama(sTot, sT);
        }
        __syncthreads();
        if (sStop) break;
        for (int i = tid; i < ns; i += LC_THREADS) work[i] = lc_apply(sT, work[i]);
        if (tid == 0) {
            float F[16];   // final = T * final, float, inner sums left to right from zero
            for (int r = 0; r < 4; ++r)
                for (int c = 0; c < 4; ++c) {
                    float s = 0.f;
                    for (int k = 0; k < 4; ++k) s = __fadd_rn(s, __fmul_rn(sT[4 * r + k], sF[4 * k + c]));
                    F[4 * r + c] = s;
                }
            for (int i = 0; i < 16; ++i) sF[i] = F[i];
            const int it = ++sIter;
            const double mse = sTot[16] / sTot[0];
            // pcl::registration::DefaultConvergenceCriteria::hasConverged
            const double cosAngle = 0.5 * ((double)sT[0] + (double)sT[5] + (double)sT[10] - 1.0);
            const double trSqr = (double)sT[3] * sT[3] + (double)sT[7] * sT[7] + (double)sT[11] * sT[11];
            if (it >= prm.maxIter) { sStop = 1; sConverged = 1; }
            else if (cosAngle >= 1.0 - prm.epsTransform && trSqr <= prm.epsTransform) { sStop = 1; sConverged = 1; }
            else if (fabs(mse - sPrevMSE) < 1e-12) { sStop = 1; sConverged = 1; }
            else if (fabs(mse - sPrevMSE) / sPrevMSE < prm.epsFitness) { sStop = 1; sConverged = 1; }
            else sPrevMSE = mse;
        }
        __syncthreads();
        if (sStop) break;
    }
    // Registration::getFitnessScore(): mean squared distance of the source under the FINAL transform to its nearest target point
    nearest_all(src, true);
    double acc[LC_SUMS];
#pragma unroll
    for (int k = 0; k < LC_SUMS; ++k) acc[k] = 0.0;
    for (int i = tid; i < ns; i += LC_THREADS) { acc[0] += 1.0; acc[16] += (double)nnD2[i]; }
    block_sum(acc);
    if (tid == 0) {
        res.converged = sConverged; res.iterations = sIter; res.nSrc = ns;
        res.fitness = sTot[0] > 0 ? sTot[16] / sTot[0] : DBL_MAX;
        for (int i = 0; i < 16; ++i) res.T[i] = sF[i];
    }
}

// ------------------------------------------------------------------------------------------------ device memory
template <class A> static void lc_alloc(A& a, LcDev& lc, int slots, int srcCap, int tgtRawCap, int tgtDSCap, int maxEntries) {
    lc.slots = slots; lc.srcCap = srcCap; lc.tgtRawCap = tgtRawCap; lc.tgtDSCap = tgtDSCap; lc.maxEntries = maxEntries;
    const size_t S = slots;
    lc.slotLane = a.template alloc<int>(S); lc.slotActive = a.template alloc<int>(S); lc.counts = a.template alloc<int>(S * 4);
    lc.entries = a.template alloc<LcEntry>(S * maxEntries);
    lc.src = a.template alloc<float4>(S * srcCap, false); lc.work = a.template alloc<float4>(S * srcCap, false);
    lc.tgtRaw = a.template alloc<float4>(S * tgtRawCap, false); lc.tgtDS = a.template alloc<float4>(S * tgtDSCap, false);
    lc.miss = a.template alloc<int>(S * srcCap, false); lc.nnIdx = a.template alloc<int>(S * srcCap, false); lc.nnD2 = a.template alloc<float>(S * srcCap, false);
    lc.results = a.template alloc<LcResult>(S);
}

void Engine::allocLoopClosure() {
    if (!loopMode()) return;
    const int hist = std::max(0, params.historyKeyframeSearchNum);
    const int perKf = (int)std::min<size_t>((size_t)c.N / 4, 16384);
    const int slots = std::min(c.L, 16);
    const int tgtRaw = (2 * hist + 1) * perKf;
    lc_alloc(*this, lc, slots, mo.scanCornerCap + mo.scanSurfCap, tgtRaw, std::max(tgtRaw / 2, perKf), 2 + 2 * (2 * hist + 1));
    lc.hash = allocHash(lc.tgtDSCap, 1.0f, lc.counts + 2, 4, slots);
    std::vector<VgSeg> segs;
    for (int s = 0; s < slots; ++s)
        segs.push_back(makeVgSeg(lc.tgtRaw + (size_t)s * lc.tgtRawCap, lc.counts + s * 4 + 1, lc.tgtDS + (size_t)s * lc.tgtDSCap, lc.counts + s * 4 + 2,
                                 lc.slotActive + s, 0.4f, lc.tgtRawCap, lc.tgtDSCap));
    vgLoop = makeVgBatch(segs, lc.tgtRawCap);
    const size_t bytes = sizeof(LcEntry) * (size_t)slots * lc.maxEntries + sizeof(int) * (size_t)slots * 6 + sizeof(LcResult) * (size_t)slots;
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&lcPinned), bytes), "cudaMallocHost");
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&moOverridesHost), sizeof(MoOverride) * c.L), "cudaMallocHost");
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&moAftHost), sizeof(float) * 6 * c.L), "cudaMallocHost");
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&moScanCountsHost), sizeof(int) * 8 * c.L), "cudaMallocHost");
    std::memset(moOverridesHost, 0, sizeof(MoOverride) * c.L);
}

// ------------------------------------------------------------------------------------------------ host bookkeeping
// extractSurroundingKeyFrames with loopClosureEnableFlag (mapOptmization.cpp:960-1000): the local map is the queue of
// the surroundingKeyframeSearchNum most recent key frames.  The reference caches each frame's TRANSFORMED cloud at push
// time; key poses only change in correctPoses, which clears the queue, so transforming at gather time is the same.
void Engine::selectRecentKeyFrames(int lane, std::vector<MoKfEntry>& out, int* nCorner, int* nSurf) {
    MoLaneHost& H = moHost[lane];
    out.clear(); *nCorner = 0; *nSurf = 0;
    if (H.keyFrames.empty()) return;
    const int want = params.surroundingKeyframeSearchNum, numPoses = (int)H.keyFrames.size();
    if ((int)H.recent.size() < want) {   // queue not full: rebuild it from the newest key frames
        H.recent.clear();
        for (int i = numPoses - 1; i >= 0; --i) {
            H.recent.push_front(i);
            if ((int)H.recent.size() >= want) break;
        }
    } else if (H.latestFrameID != numPoses - 1) {   // (latestFrameID is only ever updated here: mapOptmization.cpp:981-987)
        H.recent.pop_front();
        H.latestFrameID = numPoses - 1;
        H.recent.push_back(H.latestFrameID);
    }
    int offC = 0, offS = 0;
    for (int id : H.recent) {
        const MoKeyFrame& k = H.keyFrames[id];
        MoKfEntry e;
        e.srcC = k.poolOff; e.nC = k.nC; e.dstC = offC; offC += k.nC;
        e.srcS = k.poolOff + k.nC; e.nS = k.nS; e.dstS = offS; offS += k.nS;
        e.srcO = k.poolOff + k.nC + k.nS; e.nO = k.nO; e.dstO = offS; offS += k.nO;
        e.ctRoll = cosf(k.roll); e.stRoll = sinf(k.roll); e.ctPitch = cosf(k.pitch); e.stPitch = sinf(k.pitch);
        e.ctYaw = cosf(k.yaw); e.stYaw = sinf(k.yaw); e.tInX = k.x; e.tInY = k.y; e.tInZ = k.z;
        if (offC > mo.cornerRawCap || offS > mo.surfRawCap || (int)out.size() >= mo.maxSel) { H.lastError |= 4; offC -= k.nC; offS -= k.nS + k.nO; break; }
        out.push_back(e);
    }
    *nCorner = offC; *nSurf = offS;
}

static pg::Pose pose_from_transform(const float* t) {   // Pose3(Rot3::RzRyRx(t[2], t[0], t[1]), Point3(t[5], t[3], t[4]))
    return pg::make_pose((double)t[2], (double)t[0], (double)t[1], (double)t[5], (double)t[3], (double)t[4]);
}
static const double kOdometryVariances[6] = {1e-6, 1e-6, 1e-6, 1e-8, 1e-8, 1e-6};   // mapOptmization.cpp:346-350

// latestEstimate -> thisPose6D and transformAftMapped (mapOptmization.cpp:1411-1440)
static void estimate_to_floats(const pg::Pose& p, float* pose6D, float* aft) {
    double roll, pitch, yaw; pg::rpy(p.R, &roll, &pitch, &yaw);
    pose6D[0] = (float)p.t[1]; pose6D[1] = (float)p.t[2]; pose6D[2] = (float)p.t[0];
    pose6D[3] = (float)pitch; pose6D[4] = (float)yaw; pose6D[5] = (float)roll;
    if (aft) { aft[0] = pose6D[3]; aft[1] = pose6D[4]; aft[2] = pose6D[5]; aft[3] = pose6D[0]; aft[4] = pose6D[1]; aft[5] = pose6D[2]; }
}

// saveKeyFramesAndFactor (mapOptmization.cpp:1353-1440) on the host for the lanes whose graph holds a loop factor; the
// other lanes keep the device-only path (their estimate is the inserted value).
// Runs body(i) for i in [0, n) on up to hardware_concurrency host threads.  Used for the per-lane pose-graph updates: every
// lane owns its graph, key-frame list and override slot, and a Gauss-Newton step over a few hundred poses takes
// milliseconds to tenths of a second, so lanes are worth running side by side (the reference's iSAM2 is host code too).
template <class F> static void parallel_lanes(int n, F body) {
    const int nt = std::min<int>(n, std::max(1u, std::thread::hardware_concurrency()));
    if (nt <= 1) { for (int i = 0; i < n; ++i) body(i); return; }
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t) pool.emplace_back([&]() { for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) body(i); });
    for (auto& th : pool) th.join();
}

void Engine::saveKeyFramesHostGraph(const std::vector<int>& active) {
    bool any = false;
    for (int l = 0; l < c.L; ++l) { moOverridesHost[l].use = 0; moOverridesHost[l].save = 0; if (active[l] && moHost[l].graph.hasLoop) any = true; }
    if (any) {
        cuda_check(cudaMemcpyAsync(moAftHost, mo.transformAftMapped, sizeof(float) * 6 * c.L, cudaMemcpyDeviceToHost, stream), "aft");
        cuda_check(cudaMemcpyAsync(moScanCountsHost, mo.scanCounts, sizeof(int) * 8 * c.L, cudaMemcpyDeviceToHost, stream), "scanCounts");
        cuda_check(cudaStreamSynchronize(stream), "graph sync");
        parallel_lanes(c.L, [&](int l) {
            MoLaneHost& H = moHost[l];
            if (!active[l] || !H.graph.hasLoop) return;
            MoOverride& ov = moOverridesHost[l];
            ov.use = 1;
            const float* Aft = moAftHost + l * 6;
            const float cx = Aft[3], cy = Aft[4], cz = Aft[5];
            const float dist = sqrtf((H.prevRobotPos[0] - cx) * (H.prevRobotPos[0] - cx) + (H.prevRobotPos[1] - cy) * (H.prevRobotPos[1] - cy) +
                                     (H.prevRobotPos[2] - cz) * (H.prevRobotPos[2] - cz));
            if ((double)dist < 0.3) return;
            const int n = (int)H.keyFrames.size();
            const MoKeyFrame& last = H.keyFrames.back();
            const int used = last.poolOff + last.nC + last.nS + last.nO;
            const int* sc = moScanCountsHost + l * 8;
            if (n >= mo.maxKeyFrames || used + sc[0] + sc[1] + sc[2] > mo.poolCap) { H.lastError |= 1; return; }
            H.prevRobotPos[0] = cx; H.prevRobotPos[1] = cy; H.prevRobotPos[2] = cz;
            const pg::Pose from = pose_from_transform(H.transformLast), to = pose_from_transform(Aft);
            H.graph.addBetween(n - 1, n, pg::between(from, to), kOdometryVariances);
            H.graph.insert(n, to);
            H.graph.update(); H.graph.update();      // isam->update(graph, initialEstimate); isam->update();
            H.estimate = H.graph.x;                   // isamCurrentEstimate
            estimate_to_floats(H.graph.x[n], ov.pose6D, ov.aft);
            for (int i = 0; i < 6; ++i) H.transformLast[i] = ov.aft[i];
            ov.save = 1;
        });
    }
    cuda_check(cudaMemcpyAsync(mo.overrides, moOverridesHost, sizeof(MoOverride) * c.L, cudaMemcpyHostToDevice, stream), "overrides");
}

// correctPoses (mapOptmization.cpp:1456-1478).  isamCurrentEstimate is the snapshot taken at the last key-frame save,
// which may predate the loop factor when no key frame was saved since (the reference behaves the same).
void Engine::correctPoses(int lane) {
    MoLaneHost& H = moHost[lane];
    if (!H.aLoopIsClosed) return;
    H.recent.clear();
    const size_t n = std::min(H.estimate.size(), H.keyFrames.size());
    for (size_t i = 0; i < n; ++i) {
        float p[6]; estimate_to_floats(H.estimate[i], p, nullptr);
        MoKeyFrame& k = H.keyFrames[i];
        k.x = p[0]; k.y = p[1]; k.z = p[2]; k.roll = p[3]; k.pitch = p[4]; k.yaw = p[5];
    }
    H.aLoopIsClosed = false;
}

// detectLoopClosure (mapOptmization.cpp:814-873), the key-POSE part
bool Engine::detectLoopClosure(int lane, int* closest, int* latest) {
    MoLaneHost& H = moHost[lane];
    const float r = params.historyKeyframeSearchRadius, r2 = r * r;
    std::vector<std::pair<float, int> > hits;
    for (size_t i = 0; i < H.keyFrames.size(); ++i) {
        const float dx = H.keyFrames[i].x - H.currentRobotPos[0], dy = H.keyFrames[i].y - H.currentRobotPos[1], dz = H.keyFrames[i].z - H.currentRobotPos[2];
        const float d = dx * dx + dy * dy + dz * dz;
        if (d <= r2) hits.push_back(std::make_pair(d, (int)i));
    }
    std::sort(hits.begin(), hits.end());
    *closest = -1;
    for (auto& h : hits)
        if (std::abs(H.keyFrames[h.second].time - H.timeOdometry) > 30.0) { *closest = h.second; break; }
    if (*closest < 0) return false;
    *latest = (int)H.keyFrames.size() - 1;
    return true;
}

// pcl::getTransformation / pcl::getTranslationAndEulerAngles (common/impl/eigen.hpp), float
static void pcl_get_transformation(float x, float y, float z, float roll, float pitch, float yaw, float* t) {
    const float A = cosf(yaw), B = sinf(yaw), C = cosf(pitch), D = sinf(pitch), E = cosf(roll), F = sinf(roll), DE = D * E, DF = D * F;
    t[0] = A * C; t[1] = A * DF - B * E; t[2] = B * F + A * DE; t[3] = x;
    t[4] = B * C; t[5] = A * E + B * DF; t[6] = B * DE - A * F; t[7] = y;
    t[8] = -D;    t[9] = C * F;          t[10] = C * E;         t[11] = z;
    t[12] = 0.f; t[13] = 0.f; t[14] = 0.f; t[15] = 1.f;
}
static void pcl_get_translation_and_euler(const float* t, float& x, float& y, float& z, float& roll, float& pitch, float& yaw) {
    x = t[3]; y = t[7]; z = t[11];
    roll = atan2f(t[9], t[10]); pitch = asinf(-t[8]); yaw = atan2f(t[4], t[0]);
}

// the pose constraint of performLoopClosure (mapOptmization.cpp:915-944)
void Engine::addLoopConstraint(int lane, int closest, int latest, const LcResult& r) {
    MoLaneHost& H = moHost[lane];
    float x, y, z, roll, pitch, yaw;
    pcl_get_translation_and_euler(r.T, x, y, z, roll, pitch, yaw);                       // correction in the camera frame
    float corrLidar[16]; pcl_get_transformation(z, x, y, yaw, roll, pitch, corrLidar);
    const MoKeyFrame& kl = H.keyFrames[latest];
    float tWrong[16]; pcl_get_transformation(kl.z, kl.x, kl.y, kl.yaw, kl.roll, kl.pitch, tWrong);   // pclPointToAffine3fCameraToLidar
    float tCorrect[16];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { float s = 0.f; for (int k = 0; k < 4; ++k) s += corrLidar[4 * i + k] * tWrong[4 * k + j]; tCorrect[4 * i + j] = s; }
    pcl_get_translation_and_euler(tCorrect, x, y, z, roll, pitch, yaw);
    const pg::Pose poseFrom = pg::make_pose((double)roll, (double)pitch, (double)yaw, (double)x, (double)y, (double)z);
    const MoKeyFrame& kc = H.keyFrames[closest];
    const pg::Pose poseTo = pg::make_pose((double)kc.yaw, (double)kc.roll, (double)kc.pitch, (double)kc.z, (double)kc.x, (double)kc.y);   // pclPointTogtsamPose3
    const float noiseScore = (float)r.fitness;
    const double var[6] = {noiseScore, noiseScore, noiseScore, noiseScore, noiseScore, noiseScore};
    H.graph.addBetween(latest, closest, pg::between(poseFrom, poseTo), var);
    H.graph.update(); H.graph.update();   // isam->update(gtSAMgraph); isam->update();
    H.aLoopIsClosed = true;
    H.loopsClosed += 1;
}

// latestSurfKeyFrameCloud / nearHistorySurfKeyFrameCloudDS + ICP for up to lc.slots candidates at a time
void Engine::runIcpBatch(const std::vector<int>& lanes, const std::vector<int>& closest, const std::vector<int>& latest) {
    const int S = lc.slots;
    LcEntry* hE = reinterpret_cast<LcEntry*>(lcPinned);
    int* hCounts = reinterpret_cast<int*>(lcPinned + sizeof(LcEntry) * (size_t)S * lc.maxEntries);   // counts[S][4] | slotActive[S] | slotLane[S]
    int* hActive = hCounts + 4 * S; int* hLane = hActive + S;
    LcResult* hRes = reinterpret_cast<LcResult*>(hLane + S);
    const int hist = params.historyKeyframeSearchNum;
    LcIcpParams prm; prm.maxDist2 = 100.0f * 100.0f; prm.maxIter = 100; prm.epsTransform = 1e-6; prm.epsFitness = 1e-6;   // mapOptmization.cpp:894-898
    for (size_t base = 0; base < lanes.size(); base += S) {
        const int nb = (int)std::min<size_t>(S, lanes.size() - base);
        int maxE = 0;
        for (int s = 0; s < S; ++s) {
            hActive[s] = 0; hLane[s] = -1; for (int k = 0; k < 4; ++k) hCounts[4 * s + k] = 0;
            if (s >= nb) continue;
            const int l = lanes[base + s];
            MoLaneHost& H = moHost[l];
            auto entry = [&](int id, bool corner, int which, int dst) {
                const MoKeyFrame& k = H.keyFrames[id];
                LcEntry e;
                e.src = corner ? k.poolOff : k.poolOff + k.nC; e.n = corner ? k.nC : k.nS; e.dst = dst; e.which = which;
                e.ctRoll = cosf(k.roll); e.stRoll = sinf(k.roll); e.ctPitch = cosf(k.pitch); e.stPitch = sinf(k.pitch);
                e.ctYaw = cosf(k.yaw); e.stYaw = sinf(k.yaw); e.tInX = k.x; e.tInY = k.y; e.tInZ = k.z;
                return e;
            };
            LcEntry* out = hE + (size_t)s * lc.maxEntries;
            int nE = 0, nSrc = 0, nTgt = 0;
            const int lt = latest[base + s], cl = closest[base + s];
            out[nE] = entry(lt, true, 0, nSrc); nSrc += out[nE].n; ++nE;
            out[nE] = entry(lt, false, 0, nSrc); nSrc += out[nE].n; ++nE;
            bool fits = nSrc <= lc.srcCap;
            for (int j = -hist; j <= hist && fits; ++j) {
                const int id = cl + j;
                if (id < 0 || id > lt) continue;
                const MoKeyFrame& k = H.keyFrames[id];
                if (nTgt + k.nC + k.nS > lc.tgtRawCap || nE + 2 > lc.maxEntries) { fits = false; break; }
                out[nE] = entry(id, true, 1, nTgt); nTgt += k.nC; ++nE;
                out[nE] = entry(id, false, 1, nTgt); nTgt += k.nS; ++nE;
            }
            if (!fits) { H.lastError |= 8; continue; }
            hActive[s] = 1; hLane[s] = l; hCounts[4 * s + 0] = nSrc; hCounts[4 * s + 1] = nTgt; hCounts[4 * s + 3] = nE;
            maxE = std::max(maxE, nE);
        }
        cuda_check(cudaMemcpyAsync(lc.entries, hE, sizeof(LcEntry) * (size_t)S * lc.maxEntries, cudaMemcpyHostToDevice, stream), "lc entries");
        cuda_check(cudaMemcpyAsync(lc.counts, hCounts, sizeof(int) * 4 * S, cudaMemcpyHostToDevice, stream), "lc counts");
        cuda_check(cudaMemcpyAsync(lc.slotActive, hActive, sizeof(int) * S, cudaMemcpyHostToDevice, stream), "lc active");
        cuda_check(cudaMemcpyAsync(lc.slotLane, hLane, sizeof(int) * S, cudaMemcpyHostToDevice, stream), "lc lanes");
        if (maxE > 0) {
            { ProfScope ps(this, "lc_assemble_kernel"); dim3 g(std::min(maxE, 64), S); lc_assemble_kernel<<<g, 256, 0, stream>>>(mo, lc); }
            { ProfScope ps(this, "vg_filter(historyKeyFrames)"); launches += vg_filter(vgLoop, stream); }
            { ProfScope ps(this, "vh_build(historyKeyFrames)"); launches += vh_build(lc.hash, S, lc.tgtDS, (size_t)lc.tgtDSCap, stream); }
            { ProfScope ps(this, "lc_icp_kernel"); lc_icp_kernel<<<S, LC_THREADS, 0, stream>>>(lc, prm); }
            launches += 2;
        }
        cuda_check(cudaMemcpyAsync(hRes, lc.results, sizeof(LcResult) * S, cudaMemcpyDeviceToHost, stream), "lc results");
        cuda_check(cudaMemcpyAsync(hCounts, lc.counts, sizeof(int) * 4 * S, cudaMemcpyDeviceToHost, stream), "lc counts back");
        cuda_check(cudaStreamSynchronize(stream), "icp sync");
        parallel_lanes(nb, [&](int s) {   // (the slots of one batch belong to different lanes)
            if (!hActive[s]) return;
            const int l = lanes[base + s];
            MoLaneHost& H = moHost[l];
            const LcResult& r = hRes[s];
            H.lastLoop[1] = 1; H.lastLoop[3] = closest[base + s]; H.lastLoop[4] = latest[base + s];
            H.lastLoop[5] = r.converged; H.lastLoop[6] = r.iterations; H.lastLoop[7] = r.nSrc; H.lastLoop[8] = hCounts[4 * s + 2];
            std::memcpy(H.lastIcpT, r.T, sizeof(r.T)); H.lastIcpFitness = r.fitness;
            // icp.hasConverged() == false || icp.getFitnessScore() > historyKeyframeFitnessScore -> no constraint
            if (!r.converged || r.fitness > (double)params.historyKeyframeFitnessScore) return;
            addLoopConstraint(l, closest[base + s], latest[base + s], r);
            H.lastLoop[2] = 1;
        });
    }
}

// performLoopClosure (mapOptmization.cpp:875-945) for every lane that is due.  The reference calls it from a 1 Hz side
// thread; here it runs after a mapping call once >= 1.0 s of scan time passed since the lane's last call.
void Engine::performLoopClosure(const std::vector<int>& active) {
    std::vector<int> lanes, closest, latest;
    for (int l = 0; l < c.L; ++l) {
        MoLaneHost& H = moHost[l];
        if (!active[l] || !(H.timeOdometry - H.timeLastLoop >= 1.0)) continue;
        H.timeLastLoop = H.timeOdometry;
        H.loopCalls += 1;
        const int init[9] = {1, 0, 0, -1, -1, 0, 0, 0, 0};
        std::memcpy(H.lastLoop, init, sizeof(init));
        if (H.keyFrames.empty()) continue;
        int cl, lt;
        if (!detectLoopClosure(l, &cl, &lt)) continue;
        lanes.push_back(l); closest.push_back(cl); latest.push_back(lt);
    }
    if (!lanes.empty()) runIcpBatch(lanes, closest, latest);
}

}  // namespace lego

// ------------------------------------------------------------------------------------------------ C ABI
using lego::DevArena;

struct lego_pg { lego::pg::Graph g; };

extern "C" {

int lego_icp_align(int device, const float* source, int n_source, const float* target, int n_target,
                   double max_correspondence_distance, int max_iterations, double transformation_epsilon,
                   double euclidean_fitness_epsilon, float T_out[16], double* fitness_out, int* converged_out, int* iterations_out) {
    if (!source || !target || n_source < 0 || n_target < 0 || !T_out) return LEGO_ERR_ARG;
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); lego::set_last_error("no CUDA device"); return LEGO_ERR_CUDA; }
        lego::cuda_check(cudaSetDevice(device), "cudaSetDevice");
        DevArena a;
        lego::LcDev lc = {};
        lego::lc_alloc(a, lc, 1, std::max(n_source, 1), 1, std::max(n_target, 1), 1);
        lego::VoxelHash h;
        int H = 1024; while (H < 2 * lc.tgtDSCap) H <<= 1;
        h.H = H; h.cap = lc.tgtDSCap; h.invCell = 1.0f; h.n = lc.counts + 2; h.nStride = 4; h.minH = 1024;
        h.keys = a.alloc<unsigned long long>(H); h.count = a.alloc<int>(H); h.start = a.alloc<int>(H);
        h.slotOf = a.alloc<int>(lc.tgtDSCap); h.sorted = a.alloc<float4>(lc.tgtDSCap);
        lc.hash = h;
        const int counts[4] = {n_source, 0, n_target, 0}, one = 1, zero = 0;
        lego::cuda_check(cudaMemcpy(lc.counts, counts, sizeof(counts), cudaMemcpyHostToDevice), "counts");
        lego::cuda_check(cudaMemcpy(lc.slotActive, &one, sizeof(int), cudaMemcpyHostToDevice), "active");
        lego::cuda_check(cudaMemcpy(lc.slotLane, &zero, sizeof(int), cudaMemcpyHostToDevice), "lane");
        if (n_source) lego::cuda_check(cudaMemcpy(lc.src, source, sizeof(float4) * (size_t)n_source, cudaMemcpyHostToDevice), "source");
        if (n_target) lego::cuda_check(cudaMemcpy(lc.tgtDS, target, sizeof(float4) * (size_t)n_target, cudaMemcpyHostToDevice), "target");
        lego::vh_build(lc.hash, 1, lc.tgtDS, (size_t)lc.tgtDSCap, nullptr);
        lego::LcIcpParams prm;
        prm.maxDist2 = (float)(max_correspondence_distance * max_correspondence_distance); prm.maxIter = max_iterations;
        prm.epsTransform = transformation_epsilon; prm.epsFitness = euclidean_fitness_epsilon;
        lego::lc_icp_kernel<<<1, lego::LC_THREADS>>>(lc, prm);
        lego::cuda_check(cudaGetLastError(), "lc_icp_kernel launch");
        lego::LcResult r;
        lego::cuda_check(cudaMemcpy(&r, lc.results, sizeof(r), cudaMemcpyDeviceToHost), "result");
        std::memcpy(T_out, r.T, sizeof(r.T));
        if (fitness_out) *fitness_out = r.fitness;
        if (converged_out) *converged_out = r.converged;
        if (iterations_out) *iterations_out = r.iterations;
        return LEGO_OK;
    } catch (const std::exception& ex) {
        if (lego::last_error().empty()) lego::set_last_error(ex.what());
        return LEGO_ERR_CUDA;
    }
}

int lego_pg_create(lego_pg** out) { if (!out) return LEGO_ERR_ARG; *out = new lego_pg(); return LEGO_OK; }
int lego_pg_destroy(lego_pg* h) { if (!h) return LEGO_ERR_ARG; delete h; return LEGO_OK; }
int lego_pg_add_prior(lego_pg* h, int key, const double pose[6], const double variances[6]) {
    if (!h || key < 0 || !pose || !variances) return LEGO_ERR_ARG;
    h->g.addPrior(key, lego::pg::make_pose(pose[0], pose[1], pose[2], pose[3], pose[4], pose[5]), variances);
    return LEGO_OK;
}
int lego_pg_add_between(lego_pg* h, int key1, int key2, const double from[6], const double to[6], const double variances[6]) {
    if (!h || key1 < 0 || key2 < 0 || !from || !to || !variances) return LEGO_ERR_ARG;
    const lego::pg::Pose a = lego::pg::make_pose(from[0], from[1], from[2], from[3], from[4], from[5]);
    const lego::pg::Pose b = lego::pg::make_pose(to[0], to[1], to[2], to[3], to[4], to[5]);
    h->g.addBetween(key1, key2, lego::pg::between(a, b), variances);
    return LEGO_OK;
}
int lego_pg_insert(lego_pg* h, int key, const double pose[6]) {
    if (!h || key < 0 || !pose) return LEGO_ERR_ARG;
    h->g.insert(key, lego::pg::make_pose(pose[0], pose[1], pose[2], pose[3], pose[4], pose[5]));
    return LEGO_OK;
}
int lego_pg_update(lego_pg* h) { if (!h) return LEGO_ERR_ARG; h->g.update(); return LEGO_OK; }
int lego_pg_size(lego_pg* h) { return h ? (int)h->g.x.size() : LEGO_ERR_ARG; }
int lego_pg_estimate(lego_pg* h, int key, double o[6]) {
    if (!h || key < 0 || key >= (int)h->g.x.size() || !o) return LEGO_ERR_ARG;
    const lego::pg::Pose& p = h->g.x[key];
    o[0] = p.t[0]; o[1] = p.t[1]; o[2] = p.t[2];
    lego::pg::rpy(p.R, &o[3], &o[4], &o[5]);
    return LEGO_OK;
}

}  // extern "C"
