// TEST INFRASTRUCTURE ONLY.  C entry points over the oracle's restatements of the third-party algorithms the
// reference calls but does not vendor (oracle/thirdparty/: PCL VoxelGrid, FLANN exact k-NN, OpenCV small dense
// algebra).  Builds without /root/reference (oracle/_build/liblego_oracle.so), so kernel-level parity tests can run
// on the GPU box against exactly the code that oracle/_ref links for the same jobs.  Only tests/, smoke() and
// bench.py's cpu_baseline may load this library.
#include <cstring>
#include <vector>
#include "thirdparty/tp_icp.h"
#include "thirdparty/tp_kdtree.h"
#include "thirdparty/tp_pose_graph.h"
#include "thirdparty/tp_smallmat.h"
#include "thirdparty/tp_voxel_grid.h"

extern "C" {

// pcl::VoxelGrid<PointXYZI>::filter with setLeafSize(leaf,leaf,leaf) (featureAssociation.cpp:778-782,
// mapOptmization.cpp:1058-1089).  Returns the number of output points; copies min(m, cap).
long oracle_voxel_grid(const float* xyzi, long n, float leaf, float* out, long cap) {
    std::vector<float> o;
    const long m = static_cast<long>(tp::voxel_grid_filter(xyzi, static_cast<size_t>(n), leaf, o));
    std::memcpy(out, o.data(), sizeof(float) * 4 * static_cast<size_t>(m < cap ? m : cap));
    return m;
}

// pcl::KdTreeFLANN::nearestKSearch for nq queries (mapOptmization.cpp:1099,1181; featureAssociation.cpp:1054,1165).
// idx/d2: nq*k entries, padded with -1 / FLT_MAX when fewer than k points exist.
void oracle_knn(const float* pts, long n, const float* queries, long nq, int k, int* idx, float* d2) {
    tp::KdTree3 tree;
    tree.build(pts, static_cast<size_t>(n), 4);
    std::vector<int> ii; std::vector<float> dd;
    for (long q = 0; q < nq; ++q) {
        tree.knn(queries + 4 * q, k, ii, dd);
        for (int j = 0; j < k; ++j) {
            idx[q * k + j] = j < static_cast<int>(ii.size()) ? ii[j] : -1;
            d2[q * k + j] = j < static_cast<int>(dd.size()) ? dd[j] : FLT_MAX;
        }
    }
}

// cv::solve(A, b, x, DECOMP_QR) for an m x n float system (mapOptmization.cpp:1189, 1276)
int oracle_solve_qr(const float* A, const float* b, float* x, int m, int n) { return tp::solve_qr(A, b, x, m, n) ? 1 : 0; }
// cv::eigen on a symmetric n x n float matrix: descending eigenvalues, eigenvectors in rows (mapOptmization.cpp:1126, 1283)
void oracle_eigen_sym(const float* A, float* evals, float* evecs, int n) { tp::eigen_sym(A, evals, evecs, n); }

// pcl::IterativeClosestPoint::align as configured at mapOptmization.cpp:892-905 (oracle/thirdparty/tp_icp.h)
void oracle_icp(const float* source, long ns, const float* target, long nt, double maxDist, int maxIter, double epsT, double epsF,
                float* T16, double* fitness, int* converged, int* iterations) {
    tp::IcpResult r = tp::icp_align(source, (size_t)ns, target, (size_t)nt, maxDist, maxIter, epsT, epsF);
    std::memcpy(T16, r.T, sizeof(r.T)); *fitness = r.fitness; *converged = r.converged ? 1 : 0; *iterations = r.iterations;
}

// Pose3 graph (oracle/thirdparty/tp_pose_graph.h): poses as (rx, ry, rz, tx, ty, tz) with R = Rot3::RzRyRx(rx, ry, rz).
// factors: nf rows of {i (-1 = prior), j}; meas: nf * 12 doubles = from pose | to pose (prior: pose | unused); var: nf * 6.
// Runs `steps` Gauss-Newton steps and writes the estimate as translation + roll/pitch/yaw per pose.
static tp::PgPose port_pose(const double* p) {
    tp::PgPose o;
    const double cx = std::cos(p[0]), sx = std::sin(p[0]), cy = std::cos(p[1]), sy = std::sin(p[1]), cz = std::cos(p[2]), sz = std::sin(p[2]);
    o.R[0] = cz * cy; o.R[1] = cz * sy * sx - sz * cx; o.R[2] = cz * sy * cx + sz * sx;
    o.R[3] = sz * cy; o.R[4] = sz * sy * sx + cz * cx; o.R[5] = sz * sy * cx - cz * sx;
    o.R[6] = -sy;     o.R[7] = cy * sx;                o.R[8] = cy * cx;
    o.t[0] = p[3]; o.t[1] = p[4]; o.t[2] = p[5];
    return o;
}
void oracle_pose_graph(const double* poses, int n, const int* factors, const double* meas, const double* var, int nf, int steps, double* out) {
    tp::PoseGraph g;
    for (int k = 0; k < n; ++k) g.x.push_back(port_pose(poses + 6 * k));
    for (int f = 0; f < nf; ++f) {
        tp::PgFactor F; F.i = factors[2 * f]; F.j = factors[2 * f + 1];
        F.z = F.i < 0 ? port_pose(meas + 12 * f) : tp::pg_between(port_pose(meas + 12 * f), port_pose(meas + 12 * f + 6));
        for (int a = 0; a < 6; ++a) F.sigma[a] = std::sqrt(var[6 * f + a]);
        g.f.push_back(F);
    }
    for (int s = 0; s < steps; ++s) g.gaussNewtonStep();
    for (int k = 0; k < n; ++k) {
        const tp::PgPose& p = g.x[k];
        out[6 * k] = p.t[0]; out[6 * k + 1] = p.t[1]; out[6 * k + 2] = p.t[2];
        out[6 * k + 3] = std::atan2(p.R[7], p.R[8]);
        out[6 * k + 4] = std::atan2(-p.R[6], std::sqrt(p.R[7] * p.R[7] + p.R[8] * p.R[8]));
        out[6 * k + 5] = std::atan2(p.R[3], p.R[0]);
    }
}

}  // extern "C"
