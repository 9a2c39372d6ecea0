// Device side of the scan-to-scan odometry (see odometry.cu).  One persistent CTA per lane runs both LM phases of
// updateTransformation (featureAssociation.cpp:1666-1695) start to finish: TransformToStart of every feature,
// correspondence refresh every 5th iteration, residuals + 3x3 normal equations reduced in double, the tiny solve
// and the convergence test -- all without leaving the kernel.
#pragma once
#include <cfloat>
#include <cooperative_groups.h>
#include "engine.h"
#include "kernel_utils.cuh"
#include "pose_math.cuh"
#include "smallmat.cuh"
#include "voxel_hash.cuh"

namespace lego {

#ifndef LEGO_ODO_THREADS
#define LEGO_ODO_THREADS 256
#endif
constexpr int ODO_THREADS = LEGO_ODO_THREADS;
constexpr int ODO_WARPS = ODO_THREADS / 32;
constexpr int ODO_TE_BLOCKS = 24;
#ifndef LEGO_ODO_CORNER_CELL
#define LEGO_ODO_CORNER_CELL 4.0f
#endif
constexpr float ODO_CORNER_CELL = LEGO_ODO_CORNER_CELL;   // voxel-hash cell edge of the (sparse) corner cloud, metres

struct OdoShared {
    float T[6];
    float matP[9];
    int isDegenerate;
    int keepGoing;
    int ringOff[136];          // firstGE of build_ring_offsets
    int ringLast[136];         // lastLE
    int nextQuery;             // dynamic distribution of the correspondence searches over the warps
    float trig[6];             // sin / cos of transformCur's three angles for the current iteration (srx crx sry cry srz crz)
    double red[ODO_WARPS][10];
};

// (distance, visiting order) candidate of the adjacent-ring scans; the reference keeps the FIRST strictly smaller
// distance it meets, i.e. the lexicographic minimum of (distance, order).
struct Cand { float d; int ord; int j; };
__device__ __forceinline__ void cand_offer(Cand& b, float d, int ord, int j, float limit) {
    if (d < limit && (d < b.d || (d == b.d && ord < b.ord))) { b.d = d; b.ord = ord; b.j = j; }
}
__device__ __forceinline__ Cand cand_warp_min(Cand b) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float d = __shfl_xor_sync(0xffffffffu, b.d, o);
        const int ord = __shfl_xor_sync(0xffffffffu, b.ord, o);
        const int j = __shfl_xor_sync(0xffffffffu, b.j, o);
        if (d < b.d || (d == b.d && ord < b.ord)) { b.d = d; b.ord = ord; b.j = j; }
    }
    return b;
}

// Exact nearest neighbour of a small cloud by one warp: every lane strides over the points (coalesced, four loads in
// flight), then the warp takes the minimum of (distance, index) -- the rule of vh_nearest / vh_nearest_warp.  Measured
// (clock64 per phase, -DLEGO_ODO_PROF): the 1-NN shell walk is ~4000 cycles per query and 52 % of this kernel; on the ~500-point
// corner cloud of a 16-ring scan the scan takes ~2000.  That shortens a lone handle's step (94.2 k -> 96.3 k scans/s with 296
// lanes) but costs more instructions, and with four handles saturating the GPU the headline drops (119.1 k -> 117.3 k, twice
// each): off by default, -DLEGO_ODO_BRUTE_MAX=1024 (LEGO_NVCC_DEFS) turns it on for latency-bound use (few lanes).
#ifndef LEGO_ODO_BRUTE_MAX
#define LEGO_ODO_BRUTE_MAX 0
#endif
constexpr int ODO_BRUTE_MAX = LEGO_ODO_BRUTE_MAX;
__device__ inline int nearest_brute_warp(const float4* __restrict__ cloud, int M, float x, float y, float z, float maxD2) {
    const int lid = threadIdx.x & 31;
    float best = maxD2; int bestI = 0x7fffffff;
    auto offer = [&](const float4& a, int j) {
        const float d = vh_dist2(a, x, y, z);
        if (d < best || (d == best && bestI != 0x7fffffff && j < bestI)) { best = d; bestI = j; }
    };
    int j = lid;
    for (; j + 96 < M; j += 128) {
        const float4 a0 = cloud[j], a1 = cloud[j + 32], a2 = cloud[j + 64], a3 = cloud[j + 96];
        offer(a0, j); offer(a1, j + 32); offer(a2, j + 64); offer(a3, j + 96);
    }
    for (; j < M; j += 32) offer(cloud[j], j);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bestI, o);
        if (od < best || (od == best && oi < bestI)) { best = od; bestI = oi; }
    }
    return bestI == 0x7fffffff ? -1 : bestI;
}

// Bounds of the reference's adjacent-ring scans (featureAssociation.cpp:1062-1110, 1173-1226).  A scan walks up from
// closest + 1 until the first point whose label int(intensity) exceeds closestPointScan + 2.5, and down from closest - 1
// until the first label below closestPointScan - 2.5.  The label is NOT always the ring: adjustDistortion stores
// ring + 0.1 * relTime, and a slightly negative relTime (common at high yaw rates) makes int() give ring - 1, so inside a
// ring's stretch the labels r - 1 and r interleave.  Only "label(j) in {ring(j) - 1, ring(j)}, rings ascending" holds,
// which is enough for the break positions to be independent of where the scan starts:
//   firstGE[v] = min { j : label(j) >= v }   (upward scans stop there),   lastLE[v] = max { j : label(j) <= v }   (downward).
// An earlier version took "the index where the label steps up to v" with plain stores, which is the same thing only for
// sorted labels -- with interleaved labels several threads stored different indices (a race, visible as run-to-run
// differences of transformCur) and the bounds could cut a scan short.
__device__ inline void build_ring_offsets(const float4* __restrict__ cloud, int M, int R, int* firstGE, int* lastLE) {
    for (int r = threadIdx.x; r <= R + 3; r += blockDim.x) { firstGE[r] = M; lastLE[r] = -1; }
    __syncthreads();
    for (int j = threadIdx.x; j < M; j += blockDim.x) {
        const int v = min(max((int)cloud[j].w, 0), R + 2);
        const int vp = j > 0 ? min(max((int)cloud[j - 1].w, 0), R + 2) : -1;
        const int vn = j + 1 < M ? min(max((int)cloud[j + 1].w, 0), R + 2) : -1;
        if (v != vp) atomicMin(&firstGE[v], j);      // first / last point of every stretch of equal labels
        if (v != vn) atomicMax(&lastLE[v], j);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int r = R + 2; r >= 0; --r) firstGE[r] = min(firstGE[r], firstGE[r + 1]);
        for (int r = 1; r <= R + 3; ++r) lastLE[r] = max(lastLE[r], lastLE[r - 1]);
    }
    __syncthreads();
}

// One LM phase.  SURF: features = surfPointsFlat, unknowns [rx, rz, ty] (featureAssociation.cpp:1155-1377);
// otherwise features = cornerPointsSharp, unknowns [ry, tx, tz] (featureAssociation.cpp:1044-1153, 1379-1478).
#ifdef LEGO_ODO_PROF
static __device__ unsigned long long g_odo_prof2[8];   // per warp: nearest cycles, scan cycles, queries, misses, surf nearest, surf scan
static __device__ unsigned long long g_odo_prof[8];   // ring offsets, transform, search, rows + reduce, solve, searches, iterations
#define ODO_T(i) if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_odo_prof[i], (unsigned long long)(t_ - t0_)); t0_ = t_; }
#else
#define ODO_T(i)
#endif
// CL: the lane is run by a thread-block CLUSTER.  Every CTA of the cluster carries the whole LM state and repeats the cheap parts
// (TransformToStart of the features, rows, reduction, solve: same instructions on the same data, so all copies of T agree bit
// for bit), and the correspondence searches -- 3/4 of the kernel's cycles -- are handed out over the warps of ALL its CTAs
// through one counter in the shared memory of CTA 0 (distributed shared memory).  Results go to `ind` in global memory; the
// cluster barrier behind the searches orders them before the reads of the other CTAs.
template <bool CL> __device__ __forceinline__ int odo_ind(const int* ind, int i) { return CL ? __ldcg(ind + i) : ind[i]; }
template <bool SURF, bool CL>
__device__ int odo_phase(OdoShared& sh, float4* sel, const float4* __restrict__ feats, int Q,
                          const float4* __restrict__ last, int M, const VhLane& vh, int* ind, const Consts& c) {
    constexpr int K = SURF ? 3 : 2;
    const int tid = threadIdx.x, lid = tid & 31, warp = tid >> 5;
    int* nextQuery = &sh.nextQuery;
    bool leader = true;
    if (CL) {
        cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
        nextQuery = cluster.map_shared_rank(&sh.nextQuery, 0);
        leader = cluster.block_rank() == 0;
    }
    const float maxD2 = c.nearestFeatureSearchSqDist;
#ifdef LEGO_ODO_PROF
    long long t0_ = clock64();
#endif
    build_ring_offsets(last, M, c.R, sh.ringOff, sh.ringLast);
    ODO_T(0)
    int itersDone = 0;
    for (int iter = 0; iter < 25; ++iter) {
        itersDone = iter + 1;
        for (int q = tid; q < Q; q += ODO_THREADS) sel[q] = pose::transform_to_start(feats[q], sh.T);
        if (tid == 0 && leader) sh.nextQuery = 0;
        // six otherwise idle threads evaluate the iteration's sin / cos (double-precision restatements of glibc) once for the CTA
        if (tid >= ODO_THREADS - 6) { const int k = tid - (ODO_THREADS - 6); sh.trig[k] = pm::sincosf_glibc(sh.T[k >> 1], k & 1); }
        if (CL && iter % 5 == 0) cooperative_groups::this_cluster().sync(); else __syncthreads();
        ODO_T(1)
        if (iter % 5 == 0) {
            // one query per warp at a time, handed out dynamically: the searches differ a lot in length
            while (true) {
                int q = 0;
                if (lid == 0) q = atomicAdd(nextQuery, 1);
                q = __shfl_sync(0xffffffffu, q, 0);
                if (q >= Q) break;
                float d2near;
                const float4 pq = sel[q];
#ifdef LEGO_ODO_PROF
                const long long w0_ = clock64();
#endif
                const int closest = M <= ODO_BRUTE_MAX ? nearest_brute_warp(last, M, pq.x, pq.y, pq.z, maxD2)
                                                       : vh_nearest_warp(vh, pq.x, pq.y, pq.z, maxD2, SURF ? 1.0f : ODO_CORNER_CELL, &d2near);
#ifdef LEGO_ODO_PROF
                const long long w1_ = clock64();
#endif
                if (lid == 0) ind[q * K] = closest;
                Cand b2 = {maxD2, 0x7fffffff, -1}, b3 = {maxD2, 0x7fffffff, -1};
                if (closest >= 0) {
                    const float4 p = sel[q];
                    const int scan = (int)last[closest].w;
                    // forward scan: the reference bounds it by the CURRENT feature count (quirk Q13)
                    const int hiF = min(min(Q, M), sh.ringOff[min(scan + 3, c.R + 3)]);
                    const int nF = max(hiF - (closest + 1), 0);
                    auto offerF = [&](const float4& a, int j) {
                        const float d = vh_dist2(a, p.x, p.y, p.z);
                        const int ring = (int)a.w, ord = j - (closest + 1);
                        if (SURF) { if (ring <= scan) cand_offer(b2, d, ord, j, maxD2); else cand_offer(b3, d, ord, j, maxD2); }
                        else if (ring > scan) cand_offer(b2, d, ord, j, maxD2);
                    };
                    auto offerB = [&](const float4& a, int j) {
                        const float d = vh_dist2(a, p.x, p.y, p.z);
                        const int ring = (int)a.w, ord = nF + (closest - 1 - j);
                        if (SURF) { if (ring >= scan) cand_offer(b2, d, ord, j, maxD2); else cand_offer(b3, d, ord, j, maxD2); }
                        else if (ring < scan) cand_offer(b2, d, ord, j, maxD2);
                    };
                    // four loads in flight per lane; cand_offer keeps the lexicographic minimum of (distance, order),
                    // so the order in which a lane meets its candidates does not matter
                    int j = closest + 1 + lid;
                    for (; j + 96 < hiF; j += 128) {
                        const float4 a0 = last[j], a1 = last[j + 32], a2 = last[j + 64], a3 = last[j + 96];
                        offerF(a0, j); offerF(a1, j + 32); offerF(a2, j + 64); offerF(a3, j + 96);
                    }
                    for (; j < hiF; j += 32) offerF(last[j], j);
                    const int loB = scan >= 3 ? sh.ringLast[min(scan - 3, c.R + 3)] + 1 : 0;
                    j = closest - 1 - lid;
                    for (; j - 96 >= loB; j -= 128) {
                        const float4 a0 = last[j], a1 = last[j - 32], a2 = last[j - 64], a3 = last[j - 96];
                        offerB(a0, j); offerB(a1, j - 32); offerB(a2, j - 64); offerB(a3, j - 96);
                    }
                    for (; j >= loB; j -= 32) offerB(last[j], j);
                }
                b2 = cand_warp_min(b2);
                if (SURF) b3 = cand_warp_min(b3);
                if (lid == 0) { ind[q * K + 1] = b2.j; if (SURF) ind[q * K + 2] = b3.j; }
#ifdef LEGO_ODO_PROF
                if (lid == 0) { const long long w2_ = clock64(); atomicAdd(&g_odo_prof2[SURF ? 4 : 0], (unsigned long long)(w1_ - w0_)); atomicAdd(&g_odo_prof2[SURF ? 5 : 1], (unsigned long long)(w2_ - w1_)); atomicAdd(&g_odo_prof2[SURF ? 6 : 2], 1ull); if (closest < 0) atomicAdd(&g_odo_prof2[3], 1ull); }
#endif
            }
            if (CL) cooperative_groups::this_cluster().sync(); else __syncthreads();
            ODO_T(2)
#ifdef LEGO_ODO_PROF
            if (threadIdx.x == 0) atomicAdd(&g_odo_prof[5], 1ull);
#endif
        }
        // ---- residuals and normal equations ----
        const float srx = sh.trig[0], crx = sh.trig[1], sry = sh.trig[2], cry = sh.trig[3], srz = sh.trig[4], crz = sh.trig[5];
        const float tx = sh.T[3], ty = sh.T[4], tz = sh.T[5];
        double acc[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) acc[k] = 0.0;
        for (int q = tid; q < Q; q += ODO_THREADS) {
            const float4 p = sel[q];
            const float4 o = feats[q];
            float cx, cy, cz, cd; bool ok = false;
            if (SURF) {
                const int i1 = odo_ind<CL>(ind, q * 3), i2 = odo_ind<CL>(ind, q * 3 + 1), i3 = odo_ind<CL>(ind, q * 3 + 2);
                if (i2 >= 0 && i3 >= 0) {
                    const float4 t1 = last[i1], t2 = last[i2], t3 = last[i3];
                    float pa = (t2.y - t1.y) * (t3.z - t1.z) - (t3.y - t1.y) * (t2.z - t1.z);
                    float pb = (t2.z - t1.z) * (t3.x - t1.x) - (t3.z - t1.z) * (t2.x - t1.x);
                    float pc = (t2.x - t1.x) * (t3.y - t1.y) - (t3.x - t1.x) * (t2.y - t1.y);
                    float pd = -(pa * t1.x + pb * t1.y + pc * t1.z);
                    const float ps = sqrtf(pa * pa + pb * pb + pc * pc);
                    pa /= ps; pb /= ps; pc /= ps; pd /= ps;
                    const float pd2 = pa * p.x + pb * p.y + pc * p.z + pd;
                    float s = 1.f;
                    if (iter >= 5) s = (float)(1.0 - 1.8 * (double)fabsf(pd2) / (double)sqrtf(sqrtf(p.x * p.x + p.y * p.y + p.z * p.z)));
                    if ((double)s > 0.1 && pd2 != 0.f) { cx = s * pa; cy = s * pb; cz = s * pc; cd = s * pd2; ok = true; }
                }
            } else {
                const int i1 = odo_ind<CL>(ind, q * 2), i2 = odo_ind<CL>(ind, q * 2 + 1);
                if (i2 >= 0) {
                    const float4 t1 = last[i1], t2 = last[i2];
                    const float x0 = p.x, y0 = p.y, z0 = p.z, x1 = t1.x, y1 = t1.y, z1 = t1.z, x2 = t2.x, y2 = t2.y, z2 = t2.z;
                    const float m11 = (x0 - x1) * (y0 - y2) - (x0 - x2) * (y0 - y1);
                    const float m22 = (x0 - x1) * (z0 - z2) - (x0 - x2) * (z0 - z1);
                    const float m33 = (y0 - y1) * (z0 - z2) - (y0 - y2) * (z0 - z1);
                    const float a012 = sqrtf(m11 * m11 + m22 * m22 + m33 * m33);
                    const float l12 = sqrtf((x1 - x2) * (x1 - x2) + (y1 - y2) * (y1 - y2) + (z1 - z2) * (z1 - z2));
                    const float la = ((y1 - y2) * m11 + (z1 - z2) * m22) / a012 / l12;
                    const float lb = -((x1 - x2) * m11 - (z1 - z2) * m33) / a012 / l12;
                    const float lc = -((x1 - x2) * m22 + (y1 - y2) * m33) / a012 / l12;
                    const float ld2 = a012 / l12;
                    float s = 1.f;
                    if (iter >= 5) s = (float)(1.0 - 1.8 * (double)fabsf(ld2));
                    if ((double)s > 0.1 && ld2 != 0.f) { cx = s * la; cy = s * lb; cz = s * lc; cd = s * ld2; ok = true; }
                }
            }
            if (!ok) continue;
            float j0, j1, j2;
            if (SURF) {
                const float a1 = crx * sry * srz, a2 = crx * crz * sry, a3 = srx * sry, a4 = tx * a1 - ty * a2 - tz * a3;
                const float a5 = srx * srz, a6 = crz * srx, a7 = ty * a6 - tz * crx - tx * a5;
                const float a8 = crx * cry * srz, a9 = crx * cry * crz, a10 = cry * srx, a11 = tz * a10 + ty * a9 - tx * a8;
                const float b1 = -crz * sry - cry * srx * srz, b2 = cry * crz * srx - sry * srz;
                const float b5 = cry * crz - srx * sry * srz, b6 = cry * srz + crz * srx * sry;
                const float c1 = -b6, c2 = b5, c3 = tx * b6 - ty * b5, c4 = -crx * crz, c5 = crx * srz, c6 = ty * c5 + tx * -c4;
                const float c7 = b2, c8 = -b1, c9 = tx * -b2 - ty * -b1;
                j0 = (-a1 * o.x + a2 * o.y + a3 * o.z + a4) * cx + (a5 * o.x - a6 * o.y + crx * o.z + a7) * cy + (a8 * o.x - a9 * o.y - a10 * o.z + a11) * cz;
                j1 = (c1 * o.x + c2 * o.y + c3) * cx + (c4 * o.x - c5 * o.y + c6) * cy + (c7 * o.x + c8 * o.y + c9) * cz;
                j2 = -b6 * cx + c4 * cy + b2 * cz;
            } else {
                const float b1 = -crz * sry - cry * srx * srz, b2 = cry * crz * srx - sry * srz, b3 = crx * cry, b4 = tx * -b1 + ty * -b2 + tz * b3;
                const float b5 = cry * crz - srx * sry * srz, b6 = cry * srz + crz * srx * sry, b7 = crx * sry, b8 = tz * b7 - ty * b6 - tx * b5;
                const float c5 = crx * srz;
                j0 = (b1 * o.x + b2 * o.y - b3 * o.z + b4) * cx + (b5 * o.x + b6 * o.y - b7 * o.z + b8) * cz;
                j1 = -b5 * cx + c5 * cy + b1 * cz;
                j2 = b7 * cx - srx * cy - b3 * cz;
            }
            const double A0 = j0, A1 = j1, A2 = j2, B = (double)(float)(-0.05 * (double)cd);
            acc[0] += A0 * A0; acc[1] += A0 * A1; acc[2] += A0 * A2; acc[3] += A1 * A1; acc[4] += A1 * A2; acc[5] += A2 * A2;
            acc[6] += A0 * B; acc[7] += A1 * B; acc[8] += A2 * B; acc[9] += 1.0;
        }
#pragma unroll
        for (int k = 0; k < 10; ++k) { const double v = warp_sum_d(acc[k]); if (lid == 0) sh.red[warp][k] = v; }
        __syncthreads();
        ODO_T(3)
        if (tid == 0) {
            double s[10];
            for (int k = 0; k < 10; ++k) { s[k] = 0.0; for (int w = 0; w < ODO_WARPS; ++w) s[k] += sh.red[w][k]; }
            int keep = 1;
            if (s[9] >= 10.0) {   // fewer than 10 rows: `continue` (featureAssociation.cpp:1677, 1690)
                const float AtA[9] = {(float)s[0], (float)s[1], (float)s[2], (float)s[1], (float)s[3], (float)s[4], (float)s[2], (float)s[4], (float)s[5]};
                const float AtB[3] = {(float)s[6], (float)s[7], (float)s[8]};
                float x[3];
                sm::solve_qr<3, 3>(AtA, AtB, x);
                if (iter == 0) sh.isDegenerate = sm::min_eigen_certainly_above<3>(AtA, 10.f) ? 0 : (sm::degeneracy_projector<3>(AtA, 10.f, sh.matP) ? 1 : 0);
                if (sh.isDegenerate) {
                    float x2[3];
                    for (int i = 0; i < 3; ++i) x2[i] = (float)((double)sh.matP[i * 3] * x[0] + (double)sh.matP[i * 3 + 1] * x[1] + (double)sh.matP[i * 3 + 2] * x[2]);
                    x[0] = x2[0]; x[1] = x2[1]; x[2] = x2[2];
                }
                if (SURF) { sh.T[0] += x[0]; sh.T[2] += x[1]; sh.T[4] += x[2]; }
                else { sh.T[1] += x[0]; sh.T[3] += x[1]; sh.T[5] += x[2]; }
                for (int i = 0; i < 6; ++i) if (isnan(sh.T[i])) sh.T[i] = 0.f;
                const float r2d = 57.29577951308232f;
                float deltaR, deltaT;
                if (SURF) { deltaR = sqrtf(x[0] * r2d * x[0] * r2d + x[1] * r2d * x[1] * r2d); deltaT = sqrtf(x[2] * 100.f * x[2] * 100.f); }
                else { deltaR = sqrtf(x[0] * r2d * x[0] * r2d); deltaT = sqrtf(x[1] * 100.f * x[1] * 100.f + x[2] * 100.f * x[2] * 100.f); }
                if ((double)deltaR < 0.1 && (double)deltaT < 0.1) keep = 0;
            }
            sh.keepGoing = keep;
        }
        __syncthreads();
        ODO_T(4)
#ifdef LEGO_ODO_PROF
        if (threadIdx.x == 0) atomicAdd(&g_odo_prof[6], 1ull);
#endif
        if (!sh.keepGoing) break;
    }
    __syncthreads();
    return itersDone;
}

}  // namespace lego
