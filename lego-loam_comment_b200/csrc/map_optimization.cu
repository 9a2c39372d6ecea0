// mapOptimization scan-to-map as sm_100a kernels (reference LeGO-LOAM/src/mapOptmization.cpp:1487-1522, the body of
// mapOptimization::run with loop closure off).  blockIdx.y (or blockIdx.x for the per-lane kernels) = lane; lanes whose
// gates did not open this step (mo.active[lane] == 0) return at once.
//   mo_begin_kernel        laserOdometryHandler + transformAssociateToMap                      (608-621, 376-461)
//   mo_assemble_map_kernel extractSurroundingKeyFrames: transform + concatenate key-frame clouds (1041-1056, 545-575)
//   vg_filter              downSizeFilterCorner/Surf/Outlier on the local map and the current scan (1058-1089)
//   mo_concat_kernel       laserCloudSurfTotalLast = surfLastDS + outlierLastDS                   (1083-1086)
//   vh_build               the two kd-tree builds of scan2MapOptimization                         (1333-1334)
//   mo_residual_kernel /   <= 10 x {cornerOptimization + surfOptimization rows over the whole GPU, LMOptimization in one
//   mo_solve_kernel        warp per lane}, convergence decided on the device, then transformUpdate   (1093-1350, 463-496)
//   mo_save_keyframe_kernel saveKeyFramesAndFactor: 0.3 m gate, key pose, clouds into the pool     (1353-1454)
#include <cfloat>
#include "engine.h"
#include "map_kernels.cuh"

namespace lego {

__global__ void mo_begin_kernel(FaDev f, MoDev m, int L) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= L) return;
    MoRecord& rec = m.records[lane];
    rec.ran = 0; rec.saved = 0; rec.error = 0; rec.iters = 0;
    if (!m.active[lane]) return;
    rec.ran = 1;
    float S[6], T[6], I[6];
    odometry_message_round_trip(f.transformSum + lane * 6, S);
    for (int k = 0; k < 6; ++k) { m.transformSum[lane * 6 + k] = S[k]; T[k] = m.transformTobeMapped[lane * 6 + k]; I[k] = m.transformIncre[lane * 6 + k]; }
    transform_associate_to_map(S, m.transformBefMapped + lane * 6, m.transformAftMapped + lane * 6, I, T);
    for (int k = 0; k < 6; ++k) { m.transformTobeMapped[lane * 6 + k] = T[k]; m.transformIncre[lane * 6 + k] = I[k]; }
}

// One CTA per (entry slice, lane): entry e = one surrounding key frame; its corner cloud goes to laserCloudCornerFromMap,
// its surf and outlier clouds to laserCloudSurfFromMap, transformed by the key pose (mapOptmization.cpp:1041-1056).
__global__ void __launch_bounds__(256) mo_assemble_map_kernel(MoDev m) {
    const int lane = blockIdx.y;
    if (!m.active[lane]) return;
    const int nE = m.nEntries[lane];
    const float4* pool = m.pool + (size_t)lane * m.poolCap;
    float4* cm = m.cornerFromMap + (size_t)lane * m.cornerRawCap;
    float4* sm_ = m.surfFromMap + (size_t)lane * m.surfRawCap;
    for (int e = blockIdx.x; e < nE; e += gridDim.x) {
        const MoKfEntry en = m.entries[(size_t)lane * m.maxSel + e];
        MapSinCos t;
        t.cRoll = en.ctRoll; t.sRoll = en.stRoll; t.cPitch = en.ctPitch; t.sPitch = en.stPitch; t.cYaw = en.ctYaw; t.sYaw = en.stYaw;
        t.tX = en.tInX; t.tY = en.tInY; t.tZ = en.tInZ;
        for (int i = threadIdx.x; i < en.nC; i += blockDim.x) cm[en.dstC + i] = map_transform(t, pool[en.srcC + i]);
        for (int i = threadIdx.x; i < en.nS; i += blockDim.x) sm_[en.dstS + i] = map_transform(t, pool[en.srcS + i]);
        for (int i = threadIdx.x; i < en.nO; i += blockDim.x) sm_[en.dstO + i] = map_transform(t, pool[en.srcO + i]);
    }
}

__global__ void __launch_bounds__(256) mo_concat_kernel(MoDev m, int surfCap, int outlierCap, int totalCap) {
    const int lane = blockIdx.y;
    if (!m.active[lane]) return;
    const int nS = m.scanCounts[lane * 8 + 1], nO = m.scanCounts[lane * 8 + 2];
    const float4* s = m.surfLastDS + (size_t)lane * surfCap;
    const float4* o = m.outlierLastDS + (size_t)lane * outlierCap;
    float4* t = m.surfTotalLast + (size_t)lane * totalCap;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nS + nO; i += gridDim.x * blockDim.x) t[i] = i < nS ? s[i] : o[i - nS];
    if (blockIdx.x == 0 && threadIdx.x == 0) m.scanCounts[lane * 8 + 3] = nS + nO;
}

__device__ __forceinline__ void mo_accumulate(double* acc, const MapSinCos& sc, float4 o, float cx, float cy, float cz, float cd) {
    const float srx = sc.sRoll, crx = sc.cRoll, sry = sc.sPitch, cry = sc.cPitch, srz = sc.sYaw, crz = sc.cYaw;
    const float arx = (crx * sry * srz * o.x + crx * crz * sry * o.y - srx * sry * o.z) * cx
                    + (-srx * srz * o.x - crz * srx * o.y - crx * o.z) * cy
                    + (crx * cry * srz * o.x + crx * cry * crz * o.y - cry * srx * o.z) * cz;
    const float ary = ((cry * srx * srz - crz * sry) * o.x + (sry * srz + cry * crz * srx) * o.y + crx * cry * o.z) * cx
                    + ((-cry * crz - srx * sry * srz) * o.x + (cry * srz - crz * srx * sry) * o.y - crx * sry * o.z) * cz;
    const float arz = ((crz * srx * sry - cry * srz) * o.x + (-cry * crz - srx * sry * srz) * o.y) * cx
                    + (crx * crz * o.x - crx * srz * o.y) * cy
                    + ((sry * srz + cry * crz * srx) * o.x + (crz * sry - cry * srx * srz) * o.y) * cz;
    const double A[6] = {(double)arx, (double)ary, (double)arz, (double)cx, (double)cy, (double)cz};
    const double B = (double)(-cd);
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = i; j < 6; ++j) acc[k++] += A[i] * A[j];
#pragma unroll
    for (int i = 0; i < 6; ++i) acc[21 + i] += A[i] * B;
    acc[27] += 1.0;
}

// cornerOptimization for one point (mapOptmization.cpp:1096-1172): returns true and the weighted coefficients when
// the point yields a residual row.
__device__ inline bool mo_corner_residual(const float4* __restrict__ map, const float4 ps, int found, float d5, const int* idx, float* cx_, float* cy_, float* cz_, float* cd_) {
    if (found < 5 || !((double)d5 < 1.0)) return false;
    float px[5], py[5], pz[5];
    float cx = 0.f, cy = 0.f, cz = 0.f;
#pragma unroll
    for (int j = 0; j < 5; ++j) { const float4 q = map[idx[j]]; px[j] = q.x; py[j] = q.y; pz[j] = q.z; cx += q.x; cy += q.y; cz += q.z; }
    cx /= 5; cy /= 5; cz /= 5;
    float a11 = 0, a12 = 0, a13 = 0, a22 = 0, a23 = 0, a33 = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const float ax = px[j] - cx, ay = py[j] - cy, az = pz[j] - cz;
        a11 += ax * ax; a12 += ax * ay; a13 += ax * az; a22 += ay * ay; a23 += ay * az; a33 += az * az;
    }
    a11 /= 5; a12 /= 5; a13 /= 5; a22 /= 5; a23 /= 5; a33 /= 5;
    const float A1[9] = {a11, a12, a13, a12, a22, a23, a13, a23, a33};
    float D1[3], V1[9];
    sm::eigen_sym<3>(A1, D1, V1);
    if (!(D1[0] > 3 * D1[1])) return false;
    const float x0 = ps.x, y0 = ps.y, z0 = ps.z;
    const float x1 = (float)((double)cx + 0.1 * (double)V1[0]), y1 = (float)((double)cy + 0.1 * (double)V1[1]), z1 = (float)((double)cz + 0.1 * (double)V1[2]);
    const float x2 = (float)((double)cx - 0.1 * (double)V1[0]), y2 = (float)((double)cy - 0.1 * (double)V1[1]), z2 = (float)((double)cz - 0.1 * (double)V1[2]);
    const float m11 = (x0 - x1) * (y0 - y2) - (x0 - x2) * (y0 - y1);
    const float m22 = (x0 - x1) * (z0 - z2) - (x0 - x2) * (z0 - z1);
    const float m33 = (y0 - y1) * (z0 - z2) - (y0 - y2) * (z0 - z1);
    const float a012 = sqrtf(m11 * m11 + m22 * m22 + m33 * m33);
    const float l12 = sqrtf((x1 - x2) * (x1 - x2) + (y1 - y2) * (y1 - y2) + (z1 - z2) * (z1 - z2));
    const float la = ((y1 - y2) * m11 + (z1 - z2) * m22) / a012 / l12;
    const float lb = -((x1 - x2) * m11 - (z1 - z2) * m33) / a012 / l12;
    const float lc = -((x1 - x2) * m22 + (y1 - y2) * m33) / a012 / l12;
    const float ld2 = a012 / l12;
    const float s = (float)(1.0 - 0.9 * (double)fabsf(ld2));
    if (!((double)s > 0.1)) return false;
    *cx_ = s * la; *cy_ = s * lb; *cz_ = s * lc; *cd_ = s * ld2;
    return true;
}

// surfOptimization for one point (mapOptmization.cpp:1179-1225)
__device__ inline bool mo_surf_residual(const float4* __restrict__ map, const float4 ps, int found, float d5, const int* idx, float* cx_, float* cy_, float* cz_, float* cd_) {
    if (found < 5 || !((double)d5 < 1.0)) return false;
    float A0[15];
#pragma unroll
    for (int j = 0; j < 5; ++j) { const float4 q = map[idx[j]]; A0[j * 3] = q.x; A0[j * 3 + 1] = q.y; A0[j * 3 + 2] = q.z; }
    const float B0[5] = {-1.f, -1.f, -1.f, -1.f, -1.f};
    float X0[3];
    sm::solve_qr<5, 3>(A0, B0, X0);
    float pa = X0[0], pb = X0[1], pc = X0[2], pd = 1.f;
    const float ps_ = sqrtf(pa * pa + pb * pb + pc * pc);
    pa /= ps_; pb /= ps_; pc /= ps_; pd /= ps_;
#pragma unroll
    for (int j = 0; j < 5; ++j)
        if ((double)fabsf(pa * A0[j * 3] + pb * A0[j * 3 + 1] + pc * A0[j * 3 + 2] + pd) > 0.2) return false;
    const float pd2 = pa * ps.x + pb * ps.y + pc * ps.z + pd;
    const float s = (float)(1.0 - 0.9 * (double)fabsf(pd2) / (double)sqrtf(sqrtf(ps.x * ps.x + ps.y * ps.y + ps.z * ps.z)));
    if (!((double)s > 0.1)) return false;
    *cx_ = s * pa; *cy_ = s * pb; *cz_ = s * pc; *cd_ = s * pd2;
    return true;
}

// ---- scan2MapOptimization (mapOptmization.cpp:1329-1350) as <= 10 x {mo_residual_kernel, mo_solve_kernel} -------------
// A lane's residual rows are independent, so they are spread over as many CTAs as there are queries (grid.x covers the
// capacity, CTAs past the live query count leave at once); every CTA writes its 28 partial sums and a one-warp kernel
// per lane adds them in block order (deterministic), solves the 6x6 system and decides convergence on the device.
// state[6] = 1 while the lane is iterating; a finished lane makes the remaining launches of the call no-ops.
__global__ void mo_lm_begin_kernel(MoDev m, int L) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= L) return;
    int go = 0;
    if (m.active[lane]) {
        const int nCornerMap = m.mapCounts[lane * 4 + 2], nSurfMap = m.mapCounts[lane * 4 + 3];
        go = (nCornerMap > 10 && nSurfMap > 100) ? 1 : 0;   // guard of mapOptmization.cpp:1331
        m.state[lane * 16 + 3] = 0; m.state[lane * 16 + 4] = 0;
    }
    m.state[lane * 16 + 6] = go;
    m.knnMissCount[lane] = 0;
}

// One LM iteration's residual rows in three launches, so that each phase runs dense:
//   mo_rerank_kernel  (iterations >= 1) re-ranks every query's cached neighbour candidates at the new pose; a proven
//                     answer (voxel_hash.cuh) goes to knnRes, an unproven query onto the lane's miss list
//   mo_search_kernel  one thread per query (iteration 0) or per miss: voxel-hash search, refills the cache, writes knnRes
//   mo_residual_kernel  line / plane fit of every query from knnRes, rows reduced into per-warp partial sums
// A search is a chain of dependent loads (hash probe -> cell header -> points); mixed into one kernel with the re-rank,
// the slowest query of a CTA set its lifetime and the cache saved nothing (profiles/README.md, session 20).
// knnRes: [2 x int4] idx0..3 | idx4, d5 bits, found, -.  The miss list's order depends on atomics; the results do not.
__device__ __forceinline__ void mo_store_result(MoDev& m, size_t slot, int found, float d5, const int* idx) {
    m.knnRes[slot * 2] = make_int4(idx[0], idx[1], idx[2], idx[3]);
    m.knnRes[slot * 2 + 1] = make_int4(idx[4], __float_as_int(d5), found, 0);
}

__global__ void __launch_bounds__(256) mo_rerank_kernel(MoDev m) {
    const int lane = blockIdx.y, tid = threadIdx.x, lid = tid & 31;
    if (!m.state[lane * 16 + 6]) return;
    const int nCorner = m.scanCounts[lane * 8 + 0], nQ = nCorner + m.scanCounts[lane * 8 + 4];
    const MapSinCos sc = map_sincos_cta(m.transformTobeMapped + lane * 6);
    const float4* mapC = m.cornerFromMapDS + (size_t)lane * m.cornerDSCap;
    const float4* mapS = m.surfFromMapDS + (size_t)lane * m.surfDSCap;
    const float4* qC = m.cornerLastDS + (size_t)lane * m.scanCornerCap;
    const float4* qS = m.surfTotalLastDS + (size_t)lane * (m.scanSurfCap + m.scanOutlierCap);
    const size_t slotBase = (size_t)lane * m.resBlocks * MO_RES_THREADS;
    for (int base = blockIdx.x * 256; base < nQ; base += gridDim.x * 256) {
        const int q = base + tid;
        bool miss = false;
        if (q < nQ) {
            const bool corner = q < nCorner;
            const float4 ps = map_transform(sc, corner ? qC[q] : qS[q - nCorner]);
            const float4 meta = m.knnMeta[slotBase + q];
            int cand[VH_CACHE_K];
#pragma unroll
            for (int j = 0; j < VH_CACHE_K / 4; ++j) {
                const int4 c4 = m.knnCand[(slotBase + q) * (VH_CACHE_K / 4) + j];
                cand[4 * j] = c4.x; cand[4 * j + 1] = c4.y; cand[4 * j + 2] = c4.z; cand[4 * j + 3] = c4.w;
            }
            float d2[5]; int idx[5]; int found;
            if (vh_knn5_cached(corner ? mapC : mapS, cand, meta.x, meta.y, meta.z, meta.w, ps.x, ps.y, ps.z, 1.0f, d2, idx, &found))
                mo_store_result(m, slotBase + q, found, d2[4], idx);
            else miss = true;
        }
        // one atomic per warp
        const unsigned bal = __ballot_sync(0xffffffffu, miss);
        if (bal) {
            int pos = 0;
            if (lid == 0) pos = atomicAdd(&m.knnMissCount[lane], __popc(bal));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (miss) m.knnMissList[slotBase + pos + __popc(bal & ((1u << lid) - 1))] = q;
        }
    }
}

__global__ void __launch_bounds__(256, 4) mo_search_kernel(MoDev m, int iter) {
    const int lane = blockIdx.y, tid = threadIdx.x;
    if (!m.state[lane * 16 + 6]) return;
    const int nCorner = m.scanCounts[lane * 8 + 0], nQ = nCorner + m.scanCounts[lane * 8 + 4];
    const int n = iter == 0 ? nQ : m.knnMissCount[lane];
    const MapSinCos sc = map_sincos_cta(m.transformTobeMapped + lane * 6);
    const VhLane vhC = vh_lane(m.hashCornerMap, lane), vhS = vh_lane(m.hashSurfMap, lane);
    const float4* qC = m.cornerLastDS + (size_t)lane * m.scanCornerCap;
    const float4* qS = m.surfTotalLastDS + (size_t)lane * (m.scanSurfCap + m.scanOutlierCap);
    const size_t slotBase = (size_t)lane * m.resBlocks * MO_RES_THREADS;
    __shared__ int cellList[VH_LIST_CELLS * 256];   // per thread: the non-empty cells around its query (column tid: no bank conflicts)
    for (int i = blockIdx.x * 256 + tid; i < n; i += gridDim.x * 256) {
        const int q = iter == 0 ? i : m.knnMissList[slotBase + i];
        const bool corner = q < nCorner;
        const float4 ps = map_transform(sc, corner ? qC[q] : qS[q - nCorner]);
        float d2[5]; int idx[5]; int cand[VH_CACHE_K]; float rc2;
#ifdef LEGO_KNN_NESTED
        const int found = vh_knn5_fill(corner ? vhC : vhS, ps.x, ps.y, ps.z, 1.0f, d2, idx, cand, &rc2);
#else
        const int found = vh_knn5_fill_flat(corner ? vhC : vhS, ps.x, ps.y, ps.z, 1.0f, d2, idx, cand, &rc2, cellList + tid, 256);
#endif
#pragma unroll
        for (int j = 0; j < VH_CACHE_K / 4; ++j)
            m.knnCand[(slotBase + q) * (VH_CACHE_K / 4) + j] = make_int4(cand[4 * j], cand[4 * j + 1], cand[4 * j + 2], cand[4 * j + 3]);
        m.knnMeta[slotBase + q] = make_float4(ps.x, ps.y, ps.z, rc2);
        mo_store_result(m, slotBase + q, found, d2[4], idx);
    }
}

__global__ void __launch_bounds__(MO_RES_THREADS, 4) mo_residual_kernel(MoDev m) {
    const int lane = blockIdx.y, tid = threadIdx.x, lid = tid & 31, warp = tid >> 5;
    if (!m.state[lane * 16 + 6]) return;
    const int nCorner = m.scanCounts[lane * 8 + 0], nSurf = m.scanCounts[lane * 8 + 4];
    const MapSinCos sc = map_sincos_cta(m.transformTobeMapped + lane * 6);
    const float4* mapC = m.cornerFromMapDS + (size_t)lane * m.cornerDSCap;
    const float4* mapS = m.surfFromMapDS + (size_t)lane * m.surfDSCap;
    const size_t slotBase = (size_t)lane * m.resBlocks * MO_RES_THREADS;
    // a CTA takes the 256-query blocks blk = blockIdx.x, blockIdx.x + gridDim.x, ...; the partial sums stay per block
    // and warp (not per CTA) so the solve kernel adds them in the same order whatever the grid size
    for (int blk = blockIdx.x; blk * MO_RES_THREADS < nCorner + nSurf; blk += gridDim.x) {
        const int q = blk * MO_RES_THREADS + tid;
        double acc[MO_ACC];
#pragma unroll
        for (int k = 0; k < MO_ACC; ++k) acc[k] = 0.0;
        if (q < nCorner + nSurf) {
            const bool corner = q < nCorner;
            const float4 o = corner ? m.cornerLastDS[(size_t)lane * m.scanCornerCap + q]
                                    : m.surfTotalLastDS[(size_t)lane * (m.scanSurfCap + m.scanOutlierCap) + (q - nCorner)];
            const float4 ps = map_transform(sc, o);
            const int4 r0 = m.knnRes[(slotBase + q) * 2], r1 = m.knnRes[(slotBase + q) * 2 + 1];
            const int idx[5] = {r0.x, r0.y, r0.z, r0.w, r1.x};
            float cx, cy, cz, cd;
            const bool ok = corner ? mo_corner_residual(mapC, ps, r1.z, __int_as_float(r1.y), idx, &cx, &cy, &cz, &cd)
                                   : mo_surf_residual(mapS, ps, r1.z, __int_as_float(r1.y), idx, &cx, &cy, &cz, &cd);
            if (ok) mo_accumulate(acc, sc, o, cx, cy, cz, cd);
        }
        // every warp writes its own 28 sums
        double mine = 0.0;
#pragma unroll
        for (int k = 0; k < MO_ACC; ++k) {
            double v = acc[k];
#pragma unroll
            for (int o2 = 16; o2 > 0; o2 >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o2);
            if (lid == k) mine = v;
        }
        if (lid < MO_ACC) m.partials[(((size_t)lane * m.resBlocks + blk) * (MO_RES_THREADS / 32) + warp) * MO_ACC + lid] = mine;
    }
}

// LMOptimization (mapOptmization.cpp:1229-1327) + the loop control of scan2MapOptimization + transformUpdate (463-496)
// The solve of one LM iteration for one lane from the 28 summed accumulators s (LMOptimization, loop control of
// scan2MapOptimization, transformUpdate).  One thread.  Returns 1 when the lane is finished.
__device__ inline int mo_solve_lane(MoDev& m, int lane, int iter, const double* s) {
    float* T = m.transformTobeMapped + lane * 6;
    float* matP = m.matP + lane * 36;
    const int rows = (int)s[27];
    int done = 0;
    if (rows < 50) {
        done = 1;   // LMOptimization returns false without touching the pose: every further iteration would repeat this one
    } else {
        float AtA[36], AtB[6], X[6];
        int k = 0;
        for (int i = 0; i < 6; ++i) for (int j = i; j < 6; ++j) { AtA[i * 6 + j] = (float)s[k]; AtA[j * 6 + i] = (float)s[k]; ++k; }
        for (int i = 0; i < 6; ++i) AtB[i] = (float)s[21 + i];
        sm::solve_qr<6, 6>(AtA, AtB, X);
        int degenerate = m.state[lane * 16 + 0];
        if (iter == 0) {   // eigen decomposition only when the cheap bound cannot rule degeneracy out
            degenerate = sm::min_eigen_certainly_above<6>(AtA, 100.f) ? 0 : (sm::degeneracy_projector<6>(AtA, 100.f, matP) ? 1 : 0);
            m.state[lane * 16 + 0] = degenerate;
        }
        if (degenerate) {
            float X2[6];
            for (int i = 0; i < 6; ++i) { double t = 0.0; for (int j = 0; j < 6; ++j) t += (double)matP[i * 6 + j] * (double)X[j]; X2[i] = (float)t; }
            for (int i = 0; i < 6; ++i) X[i] = X2[i];
        }
        for (int i = 0; i < 6; ++i) T[i] += X[i];
        const float r2d = 57.29577951308232f;   // pcl::rad2deg(float)
        double dr = 0.0, dt = 0.0;
        for (int i = 0; i < 3; ++i) { const double a = (double)(X[i] * r2d), b = (double)(X[3 + i] * 100); dr += a * a; dt += b * b; }
        const float deltaR = (float)sqrt(dr), deltaT = (float)sqrt(dt);
        if ((double)deltaR < 0.05 && (double)deltaT < 0.05) done = 1;
    }
    m.knnMissCount[lane] = 0;
    m.state[lane * 16 + 3] = iter + 1; m.state[lane * 16 + 4] = rows;
    const int finished = (done || iter == 9) ? 1 : 0;
    if (finished) {
        m.state[lane * 16 + 6] = 0;
        m.records[lane].iters = iter + 1;
        // transformUpdate (mapOptmization.cpp:463-496); the host interpolated the IMU roll / pitch at the end of the sweep
        const float* bl = m.imuBlend + lane * 4;
        if (bl[0] != 0.f) {
            T[0] = (float)(0.998 * (double)T[0] + 0.002 * (double)bl[2]);
            T[2] = (float)(0.998 * (double)T[2] + 0.002 * (double)bl[1]);
        }
        for (int i = 0; i < 6; ++i) { m.transformBefMapped[lane * 6 + i] = m.transformSum[lane * 6 + i]; m.transformAftMapped[lane * 6 + i] = T[i]; }
    }
    return finished;
}

__global__ void __launch_bounds__(32) mo_solve_kernel(MoDev m, int iter) {
    const int lane = blockIdx.x, lid = threadIdx.x;
    if (!m.state[lane * 16 + 6]) return;
    __shared__ double s[MO_ACC];
    const int nQ = m.scanCounts[lane * 8 + 0] + m.scanCounts[lane * 8 + 4];
    const int nb = (nQ + MO_RES_THREADS - 1) / MO_RES_THREADS;
    if (lid < MO_ACC) {
        double t = 0.0;
        const double* p = m.partials + (size_t)lane * m.resBlocks * (MO_RES_THREADS / 32) * MO_ACC + lid;
        for (int b = 0; b < nb * (MO_RES_THREADS / 32); ++b) t += p[(size_t)b * MO_ACC];
        s[lid] = t;
    }
    __syncwarp();
    if (lid != 0) return;
    mo_solve_lane(m, lane, iter, s);
}

// scan2MapOptimization (mapOptmization.cpp:1329-1350) as ONE persistent CTA per lane: all <= 10 LM iterations -- re-rank the
// cached neighbours, search the misses, fit the rows, solve -- run back to back with block barriers instead of kernel
// boundaries.  The per-iteration kernels above sweep over every lane in every launch: each lane's map (hash + points,
// ~1.6 MB) is evicted between a lane's iterations and fetched from DRAM again (ncu: iteration 1 reads as many bytes as
// iteration 0), and a mapping call costs 40 launches of which the tail is empty.  Here a lane's iterations touch its map while it
// is still in L2 / L1, a converged lane frees its SM at once, and the call is two launches.  Same device functions, and the
// rows are summed over the same 32-query slices in the same order as mo_residual_kernel / mo_solve_kernel do
// (partials[slice] by the same shuffle butterfly, slices added up in index order), so the poses are bit-identical.
#ifndef LEGO_MO_LM_THREADS
#define LEGO_MO_LM_THREADS 512
#endif
#ifndef LEGO_MO_LM_MINBLOCKS
#define LEGO_MO_LM_MINBLOCKS (1024 / LEGO_MO_LM_THREADS)
#endif
constexpr int MO_LM_THREADS = LEGO_MO_LM_THREADS;
__global__ void __launch_bounds__(MO_LM_THREADS, LEGO_MO_LM_MINBLOCKS) mo_lm_persistent_kernel(MoDev m, int useCache) {
    extern __shared__ int cellList[];   // [VH_LIST_CELLS][MO_LM_THREADS]: per thread, the non-empty cells around its query
    __shared__ float sTrig[6];
    __shared__ int sDone, sMiss, sNext;
    __shared__ double sSum[MO_ACC];
    const int lane = blockIdx.x, tid = threadIdx.x, lid = tid & 31, warp = tid >> 5;
    if (!m.state[lane * 16 + 6]) return;
    const int nCorner = m.scanCounts[lane * 8 + 0], nSurf = m.scanCounts[lane * 8 + 4], nQ = nCorner + nSurf;
    const float* T = m.transformTobeMapped + lane * 6;
    const VhLane vhC = vh_lane(m.hashCornerMap, lane), vhS = vh_lane(m.hashSurfMap, lane);
    const float4* mapC = m.cornerFromMapDS + (size_t)lane * m.cornerDSCap;
    const float4* mapS = m.surfFromMapDS + (size_t)lane * m.surfDSCap;
    const float4* qC = m.cornerLastDS + (size_t)lane * m.scanCornerCap;
    const float4* qS = m.surfTotalLastDS + (size_t)lane * (m.scanSurfCap + m.scanOutlierCap);
    const size_t slotBase = (size_t)lane * m.resBlocks * MO_RES_THREADS;
    const int nSlices = (nQ + 31) / 32;
    for (int iter = 0; iter < 10; ++iter) {
        if (tid < 6) sTrig[tid] = pm::sincosf_glibc(T[tid >> 1], (tid & 1) ^ 1);   // cos, sin, cos, sin, ... of the pose of this iteration
        if (tid == 0) { sMiss = 0; sNext = 0; }
        __syncthreads();
        MapSinCos sc;
        sc.cRoll = sTrig[0]; sc.sRoll = sTrig[1]; sc.cPitch = sTrig[2]; sc.sPitch = sTrig[3]; sc.cYaw = sTrig[4]; sc.sYaw = sTrig[5];
        sc.tX = T[3]; sc.tY = T[4]; sc.tZ = T[5];
        const bool all = iter == 0 || !useCache;
        if (!all) {   // re-rank the cached candidates; a query whose cache cannot prove the answer goes onto the miss list
            for (int base = 0; base < nQ; base += MO_LM_THREADS) {
                const int q = base + tid;
                bool miss = false;
                if (q < nQ) {
                    const bool corner = q < nCorner;
                    const float4 ps = map_transform(sc, corner ? qC[q] : qS[q - nCorner]);
                    const float4 meta = m.knnMeta[slotBase + q];
                    int cand[VH_CACHE_K];
#pragma unroll
                    for (int j = 0; j < VH_CACHE_K / 4; ++j) {
                        const int4 c4 = m.knnCand[(slotBase + q) * (VH_CACHE_K / 4) + j];
                        cand[4 * j] = c4.x; cand[4 * j + 1] = c4.y; cand[4 * j + 2] = c4.z; cand[4 * j + 3] = c4.w;
                    }
                    float d2[5]; int idx[5]; int found;
                    if (vh_knn5_cached(corner ? mapC : mapS, cand, meta.x, meta.y, meta.z, meta.w, ps.x, ps.y, ps.z, 1.0f, d2, idx, &found))
                        mo_store_result(m, slotBase + q, found, d2[4], idx);
                    else miss = true;
                }
                const unsigned bal = __ballot_sync(0xffffffffu, miss);
                if (bal) {
                    int pos = 0;
                    if (lid == 0) pos = atomicAdd(&sMiss, __popc(bal));
                    pos = __shfl_sync(0xffffffffu, pos, 0);
                    if (miss) m.knnMissList[slotBase + pos + __popc(bal & ((1u << lid) - 1))] = q;
                }
            }
            __syncthreads();
        }
        // searches, handed out one at a time: they differ a lot in length
        const int nSearch = all ? nQ : sMiss;
        for (int i = atomicAdd(&sNext, 1); i < nSearch; i = atomicAdd(&sNext, 1)) {
            const int q = all ? i : m.knnMissList[slotBase + i];
            const bool corner = q < nCorner;
            const float4 ps = map_transform(sc, corner ? qC[q] : qS[q - nCorner]);
            float d2[5]; int idx[5]; int cand[VH_CACHE_K]; float rc2;
            const int found = vh_knn5_fill_flat(corner ? vhC : vhS, ps.x, ps.y, ps.z, 1.0f, d2, idx, cand, &rc2, cellList + tid, MO_LM_THREADS);
#pragma unroll
            for (int j = 0; j < VH_CACHE_K / 4; ++j)
                m.knnCand[(slotBase + q) * (VH_CACHE_K / 4) + j] = make_int4(cand[4 * j], cand[4 * j + 1], cand[4 * j + 2], cand[4 * j + 3]);
            m.knnMeta[slotBase + q] = make_float4(ps.x, ps.y, ps.z, rc2);
            mo_store_result(m, slotBase + q, found, d2[4], idx);
        }
        __syncthreads();
        // rows: one warp per 32-query slice, the slice's sums exactly as mo_residual_kernel forms them
        for (int sl = warp; sl < nSlices; sl += MO_LM_THREADS / 32) {
            const int q = sl * 32 + lid;
            double acc[MO_ACC];
#pragma unroll
            for (int k = 0; k < MO_ACC; ++k) acc[k] = 0.0;
            if (q < nQ) {
                const bool corner = q < nCorner;
                const float4 o = corner ? qC[q] : qS[q - nCorner];
                const float4 ps = map_transform(sc, o);
                const int4 r0 = m.knnRes[(slotBase + q) * 2], r1 = m.knnRes[(slotBase + q) * 2 + 1];
                const int idx[5] = {r0.x, r0.y, r0.z, r0.w, r1.x};
                float cx, cy, cz, cd;
                const bool ok = corner ? mo_corner_residual(mapC, ps, r1.z, __int_as_float(r1.y), idx, &cx, &cy, &cz, &cd)
                                       : mo_surf_residual(mapS, ps, r1.z, __int_as_float(r1.y), idx, &cx, &cy, &cz, &cd);
                if (ok) mo_accumulate(acc, sc, o, cx, cy, cz, cd);
            }
            double mine = 0.0;
#pragma unroll
            for (int k = 0; k < MO_ACC; ++k) {
                double v = acc[k];
#pragma unroll
                for (int o2 = 16; o2 > 0; o2 >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o2);
                if (lid == k) mine = v;
            }
            if (lid < MO_ACC) m.partials[((size_t)lane * m.resBlocks * (MO_RES_THREADS / 32) + sl) * MO_ACC + lid] = mine;
        }
        __syncthreads();
        if (warp == 0) {
            if (lid < MO_ACC) {   // (the per-iteration kernels also add the all-zero slices that pad the last 256-query block: x + 0.0 = x)
                double t = 0.0;
                const double* p = m.partials + (size_t)lane * m.resBlocks * (MO_RES_THREADS / 32) * MO_ACC + lid;
                for (int b = 0; b < nSlices; ++b) t += p[(size_t)b * MO_ACC];
                sSum[lid] = t;
            }
            __syncwarp();
            if (lid == 0) sDone = mo_solve_lane(m, lane, iter, sSum);
        }
        __syncthreads();
        if (sDone) break;
    }
}

int mo_lm_threads() { return MO_LM_THREADS; }
cudaError_t mo_lm_set_smem(size_t bytes) { return cudaFuncSetAttribute(mo_lm_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes); }

// Rot3::RzRyRx(x, y, z) then .roll() / .pitch() / .yaw() (the GTSAM calls of mapOptmization.cpp:1376-1440), in double.
__device__ inline void rzryrx_round_trip(double x, double y, double z, double* roll, double* pitch, double* yaw) {
    const double cx = cos(x), sx = sin(x), cy = cos(y), sy = sin(y), cz = cos(z), sz = sin(z);
    const double r00 = cz * cy, r10 = sz * cy, r20 = -sy, r21 = cy * sx, r22 = cy * cx;
    *roll = atan2(r21, r22);
    *pitch = atan2(-r20, sqrt(r21 * r21 + r22 * r22));
    *yaw = atan2(r10, r00);
}

// saveKeyFramesAndFactor (mapOptmization.cpp:1353-1454) with a chain-only pose graph: without loop closures the iSAM2
// estimate of the newest pose is the value that was inserted, so the update reduces to the Rot3 round trip.
__global__ void __launch_bounds__(256) mo_save_keyframe_kernel(MoDev m) {
    const int lane = blockIdx.x, tid = threadIdx.x;
    if (!m.active[lane]) return;
    __shared__ int s_save, s_off, s_err;
    MoRecord& rec = m.records[lane];
    const int nC = m.scanCounts[lane * 8 + 0], nS = m.scanCounts[lane * 8 + 1], nO = m.scanCounts[lane * 8 + 2];
    if (tid == 0) {
        float* Aft = m.transformAftMapped + lane * 6; float* T = m.transformTobeMapped + lane * 6; float* Last = m.transformLast + lane * 6;
        float* prev = m.prevRobotPos + lane * 4;
        const float cx = Aft[3], cy = Aft[4], cz = Aft[5];
        rec.robotPos[0] = cx; rec.robotPos[1] = cy; rec.robotPos[2] = cz;
        const int nKF = m.state[lane * 16 + 1];
        const int used = m.state[lane * 16 + 2];
        const MoOverride ov = m.overrides[lane];
        int err = 0;
        s_save = 0;
        if (ov.use) {
            // the host ran the gate and the pose-graph update (a loop factor is in the graph): apply its decision
            if (ov.save) {
                prev[0] = cx; prev[1] = cy; prev[2] = cz;
                for (int i = 0; i < 6; ++i) { rec.aftRaw[i] = Aft[i]; rec.pose6D[i] = ov.pose6D[i]; Aft[i] = ov.aft[i]; Last[i] = ov.aft[i]; T[i] = ov.aft[i]; }
                m.state[lane * 16 + 1] = nKF + 1;
                m.state[lane * 16 + 2] = used + nC + nS + nO;
                s_save = 1; s_off = used;
                rec.saved = 1; rec.poolOff = used; rec.nC = nC; rec.nS = nS; rec.nO = nO;
            }
        } else {
            const float dist = sqrtf((prev[0] - cx) * (prev[0] - cx) + (prev[1] - cy) * (prev[1] - cy) + (prev[2] - cz) * (prev[2] - cz));
            const bool save = !((double)dist < 0.3);
            if (save || nKF == 0) {
                if (nKF >= m.maxKeyFrames || used + nC + nS + nO > m.poolCap) err = 1;
                else {
                    prev[0] = cx; prev[1] = cy; prev[2] = cz;
                    const float* P = nKF == 0 ? T : Aft;   // first key frame is built from transformTobeMapped (1375-1382)
                    if (nKF == 0) for (int i = 0; i < 6; ++i) Last[i] = T[i];
                    for (int i = 0; i < 6; ++i) rec.aftRaw[i] = P[i];
                    double roll, pitch, yaw;
                    rzryrx_round_trip((double)P[2], (double)P[0], (double)P[1], &roll, &pitch, &yaw);
                    const float kx = P[3], ky = P[4], kz = P[5];   // translation (z, x, y) -> thisPose3D (y, z, x) of it
                    const float kroll = (float)pitch, kpitch = (float)yaw, kyaw = (float)roll;
                    rec.pose6D[0] = kx; rec.pose6D[1] = ky; rec.pose6D[2] = kz; rec.pose6D[3] = kroll; rec.pose6D[4] = kpitch; rec.pose6D[5] = kyaw;
                    if (nKF + 1 > 1) {
                        Aft[0] = kroll; Aft[1] = kpitch; Aft[2] = kyaw; Aft[3] = kx; Aft[4] = ky; Aft[5] = kz;
                        for (int i = 0; i < 6; ++i) { Last[i] = Aft[i]; T[i] = Aft[i]; }
                    }
                    m.state[lane * 16 + 1] = nKF + 1;
                    m.state[lane * 16 + 2] = used + nC + nS + nO;
                    s_save = 1; s_off = used;
                    rec.saved = 1; rec.poolOff = used; rec.nC = nC; rec.nS = nS; rec.nO = nO;
                }
            }
        }
        s_err = err;
        if (err) { m.state[lane * 16 + 5] |= err; rec.error |= err; }
    }
    __syncthreads();
    if (!s_save) return;
    float4* dst = m.pool + (size_t)lane * m.poolCap + s_off;
    const float4* c = m.cornerLastDS + (size_t)lane * m.scanCornerCap;
    const float4* s = m.surfLastDS + (size_t)lane * m.scanSurfCap;
    const float4* o = m.outlierLastDS + (size_t)lane * m.scanOutlierCap;
    for (int i = tid; i < nC; i += blockDim.x) dst[i] = c[i];
    for (int i = tid; i < nS; i += blockDim.x) dst[nC + i] = s[i];
    for (int i = tid; i < nO; i += blockDim.x) dst[nC + nS + i] = o[i];
}

}  // namespace lego
