// TEST INFRASTRUCTURE ONLY.  C entry points over the oracle's restatements of the third-party algorithms the
// reference calls but does not vendor (oracle/thirdparty/: PCL VoxelGrid, FLANN exact k-NN, OpenCV small dense
// algebra).  Builds without /root/reference (oracle/_build/liblego_oracle.so), so kernel-level parity tests can run
// on the GPU box against exactly the code that oracle/_ref links for the same jobs.  Only tests/, smoke() and
// bench.py's cpu_baseline may load this library.
#include <cstring>
#include <vector>
#include "thirdparty/tp_kdtree.h"
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

}  // extern "C"
