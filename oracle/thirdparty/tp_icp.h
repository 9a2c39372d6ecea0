// TEST INFRASTRUCTURE ONLY (oracle).
//
// Restatement of pcl::IterativeClosestPoint<PointXYZI, PointXYZI> as the reference configures it in
// mapOptmization.cpp:892-905 (max correspondence distance 100 m, <= 100 iterations, transformation epsilon 1e-6,
// Euclidean fitness epsilon 1e-6, no RANSAC, no correspondence rejectors).  PCL (1.7.2 - 1.8.1 via pcl_ros) is NOT
// vendored in /root/reference and the reference holds no fixture for it: PARITY UNPINNED.  The published algorithm
// (registration/impl/icp.hpp::computeTransformation, default_convergence_criteria.hpp::hasConverged,
// transformation_estimation_svd.hpp -> Eigen::umeyama without scaling, Registration::getFitnessScore):
//
//   working cloud = source;  final = I;  prevMSE = DBL_MAX
//   repeat
//     correspondences = exact 1-NN of every working point in the target, kept when d2 <= maxDist^2
//     fewer than 3 -> not converged, stop
//     T = rigid transform minimising sum |T p - q|^2 over the pairs (Umeyama / Kabsch via a 3x3 SVD)
//     working cloud <- T * working cloud;  final <- T * final;  ++iterations
//     converged when  iterations >= max
//                 or  (trace(R_T) - 1) / 2 >= 1 - eps_T  and  |t_T|^2 <= eps_T
//                 or  |MSE - prevMSE| < 1e-12  or  |MSE - prevMSE| / prevMSE < eps_fit      (MSE = mean pair d2)
//   fitness = mean over source points of the squared distance from final * p to its nearest target point
//
// Deviation that is stated, not hidden: Eigen evaluates the 3x3 cross-covariance with float GEMM kernels whose
// summation order is an implementation detail; here the means and the covariance are accumulated in double (the
// result is then within float rounding of any summation order), and the 3x3 SVD is a double one-sided Jacobi.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>
#include "tp_kdtree.h"

namespace tp {

// y = M * [x 1] in float, left to right: ((m0*x + m1*y) + m2*z) + m3
inline void icp_apply(const float M[16], const float* p, float* o) {
    for (int r = 0; r < 3; ++r) {
        float a = M[4 * r] * p[0], b = M[4 * r + 1] * p[1], c = M[4 * r + 2] * p[2];
        float s = a + b; s = s + c; s = s + M[4 * r + 3];
        o[r] = s;
    }
}

// SVD of a 3x3 double matrix by one-sided Jacobi: A = U diag(s) V^T, s descending (sign of the last column of U
// carries det).  Returns U, s, V.
inline void svd3(const double A[9], double U[9], double s[3], double V[9]) {
    double B[9]; std::memcpy(B, A, sizeof(B));   // columns are rotated until orthogonal: B = A V
    for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < 3; ++r) { a += B[3 * r + p] * B[3 * r + p]; b += B[3 * r + q] * B[3 * r + q]; c += B[3 * r + p] * B[3 * r + q]; }
                off = std::max(off, std::fabs(c) / (std::sqrt(a * b) + 1e-300));
                if (std::fabs(c) <= 1e-300) continue;
                const double zeta = (b - a) / (2.0 * c);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
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
    for (int j = 0; j < 3; ++j) n[j] = std::sqrt(B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j]);
    for (int i = 0; i < 2; ++i) for (int j = i + 1; j < 3; ++j) if (n[ord[j]] > n[ord[i]]) std::swap(ord[i], ord[j]);
    double Vs[9];
    for (int j = 0; j < 3; ++j) {
        const int o = ord[j];
        s[j] = n[o];
        for (int r = 0; r < 3; ++r) { Vs[3 * r + j] = V[3 * r + o]; U[3 * r + j] = n[o] > 1e-300 ? B[3 * r + o] / n[o] : 0.0; }
    }
    std::memcpy(V, Vs, sizeof(Vs));
    // complete U when the smallest singular value vanished: u2 = u0 x u1
    if (s[2] <= 1e-300 * s[0] || s[2] == 0.0) {
        U[2] = U[3] * U[7] - U[6] * U[4]; U[5] = U[6] * U[1] - U[0] * U[7]; U[8] = U[0] * U[4] - U[3] * U[1];
    }
}
inline double det3(const double M[9]) {
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// Eigen::umeyama(src, dst, with_scaling = false) over n pairs (src[i], dst[i]); T row-major 4x4 float.
inline void umeyama_pairs(const std::vector<const float*>& src, const std::vector<const float*>& dst, float T[16]) {
    const size_t n = src.size();
    double ms[3] = {0, 0, 0}, md[3] = {0, 0, 0};
    for (size_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) { ms[a] += src[i][a]; md[a] += dst[i][a]; }
    for (int a = 0; a < 3; ++a) { ms[a] /= (double)n; md[a] /= (double)n; }
    double S[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // sigma = (1/n) sum (dst - md)(src - ms)^T
    for (size_t i = 0; i < n; ++i)
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) S[3 * r + c] += ((double)dst[i][r] - md[r]) * ((double)src[i][c] - ms[c]);
    for (int i = 0; i < 9; ++i) S[i] /= (double)n;
    double U[9], d[3], V[9];
    svd3(S, U, d, V);
    double sg[3] = {1, 1, 1};
    if (det3(S) < 0) sg[2] = -1;
    int rank = 0;
    for (int i = 0; i < 3; ++i) if (!(std::fabs(d[i]) <= std::fabs(d[0]) * 1e-12)) ++rank;   // isMuchSmallerThan
    if (rank == 2) { if (det3(U) * det3(V) > 0) sg[2] = 1; else sg[2] = -1; }
    double R[9];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { double a = 0; for (int k = 0; k < 3; ++k) a += U[3 * r + k] * sg[k] * V[3 * c + k]; R[3 * r + c] = a; }
    for (int i = 0; i < 16; ++i) T[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) T[4 * r + c] = (float)R[3 * r + c];
        T[4 * r + 3] = (float)(md[r] - (R[3 * r] * ms[0] + R[3 * r + 1] * ms[1] + R[3 * r + 2] * ms[2]));
    }
}

struct IcpResult { bool converged = false; int iterations = 0; double fitness = 1e30; float T[16]; int pairsLast = 0; };

// source / target: n x 4 floats (x, y, z, intensity)
inline IcpResult icp_align(const float* source, size_t ns, const float* target, size_t nt, double maxDist, int maxIter,
                           double epsTransform, double epsFitness) {
    IcpResult res;
    for (int i = 0; i < 16; ++i) res.T[i] = (i % 5 == 0) ? 1.f : 0.f;
    if (ns == 0 || nt == 0) return res;
    KdTree3 tree; tree.build(target, nt, 4);
    std::vector<float> work(source, source + 4 * ns);
    const float maxD2 = (float)(maxDist * maxDist);
    double prevMSE = DBL_MAX;
    std::vector<int> ii; std::vector<float> dd;
    std::vector<const float*> ps, pd;
    while (true) {
        ps.clear(); pd.clear();
        double mse = 0;
        for (size_t i = 0; i < ns; ++i) {
            tree.knn(&work[4 * i], 1, ii, dd);
            if (ii.empty() || dd[0] > maxD2) continue;
            ps.push_back(&work[4 * i]); pd.push_back(target + 4 * (size_t)ii[0]); mse += dd[0];
        }
        res.pairsLast = (int)ps.size();
        if (ps.size() < 3) { res.converged = false; break; }
        mse /= (double)ps.size();
        float T[16];
        umeyama_pairs(ps, pd, T);
        for (size_t i = 0; i < ns; ++i) { float o[3]; icp_apply(T, &work[4 * i], o); work[4 * i] = o[0]; work[4 * i + 1] = o[1]; work[4 * i + 2] = o[2]; }
        float F[16];   // final = T * final (float 4x4 product, inner sums left to right)
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) { float s = 0.f; for (int k = 0; k < 4; ++k) s += T[4 * r + k] * res.T[4 * k + c]; F[4 * r + c] = s; }
        std::memcpy(res.T, F, sizeof(F));
        ++res.iterations;
        // DefaultConvergenceCriteria::hasConverged
        if (res.iterations >= maxIter) { res.converged = true; break; }
        const double cosAngle = 0.5 * ((double)T[0] + (double)T[5] + (double)T[10] - 1.0);
        const double trSqr = (double)T[3] * T[3] + (double)T[7] * T[7] + (double)T[11] * T[11];
        if (cosAngle >= 1.0 - epsTransform && trSqr <= epsTransform) { res.converged = true; break; }
        if (std::fabs(mse - prevMSE) < 1e-12) { res.converged = true; break; }
        if (std::fabs(mse - prevMSE) / prevMSE < epsFitness) { res.converged = true; break; }
        prevMSE = mse;
    }
    // Registration::getFitnessScore(max_range = DBL_MAX): source transformed by the FINAL matrix
    double fit = 0; long nr = 0;
    for (size_t i = 0; i < ns; ++i) {
        float o[4]; icp_apply(res.T, source + 4 * i, o);
        tree.knn(o, 1, ii, dd);
        if (ii.empty()) continue;
        fit += dd[0]; ++nr;
    }
    res.fitness = nr > 0 ? fit / (double)nr : DBL_MAX;
    return res;
}

}  // namespace tp
