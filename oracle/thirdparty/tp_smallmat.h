// TEST INFRASTRUCTURE ONLY (oracle).
//
// Small dense float32 algebra standing in for the OpenCV calls the reference makes (OpenCV is NOT vendored in
// /root/reference): cv::solve(..., DECOMP_QR) (featureAssociation.cpp:1327,1428; mapOptmization.cpp:1189,
// 1276), cv::eigen on symmetric matrices (featureAssociation.cpp:1334,1435; mapOptmization.cpp:1126,1283),
// Mat::inv (featureAssociation.cpp:1349,1450; mapOptmization.cpp:1298) and Mat*Mat.  These restate the
// textbook algorithms OpenCV documents for those entry points (Householder QR least squares, cyclic Jacobi
// with descending eigenvalues / row eigenvectors, LU inverse, GEMM with wide accumulation).  Rounding differs
// from any particular OpenCV build at the 1e-6 level: parity UNPINNED by reference fixtures.
#pragma once
#include <cmath>
#include <cstring>
#include <vector>

namespace tp {

// C(m x n) = A(m x k) * B(k x n), row-major float storage, double accumulation.
inline void matmul(const float* A, const float* B, float* C, int m, int k, int n) {
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) {
            double s = 0.0;
            for (int p = 0; p < k; ++p) s += static_cast<double>(A[i * k + p]) * static_cast<double>(B[p * n + j]);
            C[i * n + j] = static_cast<float>(s);
        }
}

// Least-squares / square solve of A(m x n) X = B(m x 1), m >= n, by Householder reflections in float32.
// Returns false when R is singular to working precision.
inline bool solve_qr(const float* A_in, const float* b_in, float* x, int m, int n) {
    std::vector<float> A(A_in, A_in + m * n), b(b_in, b_in + m), v(m);
    for (int c = 0; c < n; ++c) {
        float nrm2 = 0.f;
        for (int r = c; r < m; ++r) { v[r] = A[r * n + c]; nrm2 += v[r] * v[r]; }
        float head = v[c];
        float alpha = std::sqrt(nrm2);
        v[c] = head + (head >= 0.f ? alpha : -alpha);
        float vn = std::sqrt(nrm2 + v[c] * v[c] - head * head);
        if (vn == 0.f) continue;
        for (int r = c; r < m; ++r) v[r] /= vn;
        for (int j = c; j < n; ++j) {
            float dot = 0.f;
            for (int r = c; r < m; ++r) dot += v[r] * A[r * n + j];
            for (int r = c; r < m; ++r) A[r * n + j] -= 2.f * v[r] * dot;
        }
        float dotb = 0.f;
        for (int r = c; r < m; ++r) dotb += v[r] * b[r];
        for (int r = c; r < m; ++r) b[r] -= 2.f * v[r] * dotb;
    }
    for (int i = n - 1; i >= 0; --i) {
        float s = b[i];
        for (int j = n - 1; j > i; --j) s -= x[j] * A[i * n + j];
        float piv = A[i * n + i];
        if (std::fabs(piv) < 1.1920929e-7f * 100.f) { for (int q = 0; q < n; ++q) x[q] = 0.f; return false; }
        x[i] = s / piv;
    }
    return true;
}

// Symmetric eigen-decomposition by cyclic Jacobi rotations in float32.
// evals: n values, descending.  evecs: n x n, eigenvector k in ROW k (cv::eigen layout).
inline void eigen_sym(const float* A_in, float* evals, float* evecs, int n) {
    std::vector<float> A(A_in, A_in + n * n), V(n * n, 0.f);
    for (int i = 0; i < n; ++i) V[i * n + i] = 1.f;
    for (int sweep = 0; sweep < 60; ++sweep) {
        float off = 0.f;
        for (int p = 0; p < n; ++p) for (int q = p + 1; q < n; ++q) off += A[p * n + q] * A[p * n + q];
        if (off < 1e-30f) break;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) {
                float apq = A[p * n + q];
                if (std::fabs(apq) < 1e-30f) continue;
                float theta = (A[q * n + q] - A[p * n + p]) / (2.f * apq);
                float t = (theta >= 0.f ? 1.f : -1.f) / (std::fabs(theta) + std::sqrt(theta * theta + 1.f));
                float c = 1.f / std::sqrt(t * t + 1.f), s = t * c;
                for (int k = 0; k < n; ++k) {
                    float akp = A[k * n + p], akq = A[k * n + q];
                    A[k * n + p] = c * akp - s * akq; A[k * n + q] = s * akp + c * akq;
                }
                for (int k = 0; k < n; ++k) {
                    float apk = A[p * n + k], aqk = A[q * n + k];
                    A[p * n + k] = c * apk - s * aqk; A[q * n + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < n; ++k) {
                    float vpk = V[p * n + k], vqk = V[q * n + k];
                    V[p * n + k] = c * vpk - s * vqk; V[q * n + k] = s * vpk + c * vqk;
                }
            }
    }
    std::vector<int> ord(n);
    for (int i = 0; i < n; ++i) ord[i] = i;
    for (int i = 0; i < n; ++i) for (int j = i + 1; j < n; ++j) if (A[ord[j] * n + ord[j]] > A[ord[i] * n + ord[i]]) { int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    for (int i = 0; i < n; ++i) {
        evals[i] = A[ord[i] * n + ord[i]];
        for (int k = 0; k < n; ++k) evecs[i * n + k] = V[ord[i] * n + k];
    }
}

// Inverse by Gauss-Jordan with partial pivoting (double work precision, float in/out). False if singular.
inline bool inverse(const float* A_in, float* Ainv, int n) {
    std::vector<double> M(n * 2 * n, 0.0);
    for (int i = 0; i < n; ++i) { for (int j = 0; j < n; ++j) M[i * 2 * n + j] = A_in[i * n + j]; M[i * 2 * n + n + i] = 1.0; }
    for (int c = 0; c < n; ++c) {
        int piv = c;
        for (int r = c + 1; r < n; ++r) if (std::fabs(M[r * 2 * n + c]) > std::fabs(M[piv * 2 * n + c])) piv = r;
        if (std::fabs(M[piv * 2 * n + c]) < 1e-300) { std::memset(Ainv, 0, sizeof(float) * n * n); return false; }
        if (piv != c) for (int j = 0; j < 2 * n; ++j) std::swap(M[c * 2 * n + j], M[piv * 2 * n + j]);
        double d = M[c * 2 * n + c];
        for (int j = 0; j < 2 * n; ++j) M[c * 2 * n + j] /= d;
        for (int r = 0; r < n; ++r) if (r != c) {
            double f = M[r * 2 * n + c];
            if (f != 0.0) for (int j = 0; j < 2 * n; ++j) M[r * 2 * n + j] -= f * M[c * 2 * n + j];
        }
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) Ainv[i * n + j] = static_cast<float>(M[i * 2 * n + n + j]);
    return true;
}

}  // namespace tp
