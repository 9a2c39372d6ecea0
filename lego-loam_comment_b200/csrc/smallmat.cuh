// Small dense float32 algebra that runs inside one thread: the device counterpart of the OpenCV entry points the
// reference calls from its LM solvers -- cv::solve(DECOMP_QR) (featureAssociation.cpp:1327,1428;
// mapOptmization.cpp:1189,1276), cv::eigen on a symmetric matrix (featureAssociation.cpp:1334,1435;
// mapOptmization.cpp:1126,1283) and Mat::inv (featureAssociation.cpp:1349,1450; mapOptmization.cpp:1298).
// Householder least squares, cyclic Jacobi with descending eigenvalues / eigenvectors in rows, Gauss-Jordan
// inverse in double.  Everything is statically sized so the matrices live in registers / local memory.
#pragma once
#include <cuda_runtime.h>

namespace sm {

// Solve A(M x N) x = b in the least-squares sense (M >= N), row-major A.  Returns false (x = 0) when R is singular.
template <int M, int N>
__host__ __device__ inline bool solve_qr(const float* A_in, const float* b_in, float* x) {
    float A[M * N], b[M], v[M];
    for (int i = 0; i < M * N; ++i) A[i] = A_in[i];
    for (int i = 0; i < M; ++i) b[i] = b_in[i];
    for (int c = 0; c < N; ++c) {
        float nrm2 = 0.f;
        for (int r = c; r < M; ++r) { v[r] = A[r * N + c]; nrm2 += v[r] * v[r]; }
        const float head = v[c];
        const float alpha = sqrtf(nrm2);
        v[c] = head + (head >= 0.f ? alpha : -alpha);
        const float vn = sqrtf(nrm2 + v[c] * v[c] - head * head);
        if (vn == 0.f) continue;
        for (int r = c; r < M; ++r) v[r] /= vn;
        for (int j = c; j < N; ++j) {
            float dot = 0.f;
            for (int r = c; r < M; ++r) dot += v[r] * A[r * N + j];
            for (int r = c; r < M; ++r) A[r * N + j] -= 2.f * v[r] * dot;
        }
        float dotb = 0.f;
        for (int r = c; r < M; ++r) dotb += v[r] * b[r];
        for (int r = c; r < M; ++r) b[r] -= 2.f * v[r] * dotb;
    }
    for (int i = N - 1; i >= 0; --i) {
        float s = b[i];
        for (int j = N - 1; j > i; --j) s -= x[j] * A[i * N + j];
        const float piv = A[i * N + i];
        if (fabsf(piv) < 1.1920929e-7f * 100.f) { for (int q = 0; q < N; ++q) x[q] = 0.f; return false; }
        x[i] = s / piv;
    }
    return true;
}

// Symmetric eigen-decomposition: evals descending, eigenvector k in ROW k of evecs (cv::eigen layout).
template <int N>
__host__ __device__ inline void eigen_sym(const float* A_in, float* evals, float* evecs) {
    float A[N * N], V[N * N];
    for (int i = 0; i < N * N; ++i) { A[i] = A_in[i]; V[i] = 0.f; }
    for (int i = 0; i < N; ++i) V[i * N + i] = 1.f;
    for (int sweep = 0; sweep < 60; ++sweep) {
        float off = 0.f;
        for (int p = 0; p < N; ++p) for (int q = p + 1; q < N; ++q) off += A[p * N + q] * A[p * N + q];
        if (off < 1e-30f) break;
        for (int p = 0; p < N; ++p)
            for (int q = p + 1; q < N; ++q) {
                const float apq = A[p * N + q];
                if (fabsf(apq) < 1e-30f) continue;
                const float theta = (A[q * N + q] - A[p * N + p]) / (2.f * apq);
                const float t = (theta >= 0.f ? 1.f : -1.f) / (fabsf(theta) + sqrtf(theta * theta + 1.f));
                const float c = 1.f / sqrtf(t * t + 1.f), s = t * c;
                for (int k = 0; k < N; ++k) {
                    const float akp = A[k * N + p], akq = A[k * N + q];
                    A[k * N + p] = c * akp - s * akq; A[k * N + q] = s * akp + c * akq;
                }
                for (int k = 0; k < N; ++k) {
                    const float apk = A[p * N + k], aqk = A[q * N + k];
                    A[p * N + k] = c * apk - s * aqk; A[q * N + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < N; ++k) {
                    const float vpk = V[p * N + k], vqk = V[q * N + k];
                    V[p * N + k] = c * vpk - s * vqk; V[q * N + k] = s * vpk + c * vqk;
                }
            }
    }
    int ord[N];
    for (int i = 0; i < N; ++i) ord[i] = i;
    for (int i = 0; i < N; ++i)
        for (int j = i + 1; j < N; ++j)
            if (A[ord[j] * N + ord[j]] > A[ord[i] * N + ord[i]]) { const int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    for (int i = 0; i < N; ++i) {
        evals[i] = A[ord[i] * N + ord[i]];
        for (int k = 0; k < N; ++k) evecs[i * N + k] = V[ord[i] * N + k];
    }
}

// Inverse by Gauss-Jordan with partial pivoting, double work precision.  False (zeros) when singular.
template <int N>
__host__ __device__ inline bool inverse(const float* A_in, float* Ainv) {
    double Mx[N * 2 * N];
    for (int i = 0; i < N * 2 * N; ++i) Mx[i] = 0.0;
    for (int i = 0; i < N; ++i) { for (int j = 0; j < N; ++j) Mx[i * 2 * N + j] = A_in[i * N + j]; Mx[i * 2 * N + N + i] = 1.0; }
    for (int c = 0; c < N; ++c) {
        int piv = c;
        for (int r = c + 1; r < N; ++r) if (fabs(Mx[r * 2 * N + c]) > fabs(Mx[piv * 2 * N + c])) piv = r;
        if (fabs(Mx[piv * 2 * N + c]) < 1e-300) { for (int i = 0; i < N * N; ++i) Ainv[i] = 0.f; return false; }
        if (piv != c) for (int j = 0; j < 2 * N; ++j) { const double t = Mx[c * 2 * N + j]; Mx[c * 2 * N + j] = Mx[piv * 2 * N + j]; Mx[piv * 2 * N + j] = t; }
        const double d = Mx[c * 2 * N + c];
        for (int j = 0; j < 2 * N; ++j) Mx[c * 2 * N + j] /= d;
        for (int r = 0; r < N; ++r) if (r != c) {
            const double f = Mx[r * 2 * N + c];
            if (f != 0.0) for (int j = 0; j < 2 * N; ++j) Mx[r * 2 * N + j] -= f * Mx[c * 2 * N + j];
        }
    }
    for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) Ainv[i * N + j] = (float)Mx[i * 2 * N + N + j];
    return true;
}

// C(N x N) = A * B with double accumulation, float storage (cv::Mat operator* on CV_32F).
template <int N>
__host__ __device__ inline void matmul(const float* A, const float* B, float* C) {
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) {
            double s = 0.0;
            for (int p = 0; p < N; ++p) s += (double)A[i * N + p] * (double)B[p * N + j];
            C[i * N + j] = (float)s;
        }
}

// The degeneracy projector of the reference's LM solvers (featureAssociation.cpp:1329-1352, mapOptmization.cpp:
// 1278-1301): eigenvectors whose eigenvalue is below `thresh` (walking up from the smallest) are zeroed,
// matP = V^-1 * V2.  Returns isDegenerate.
template <int N>
__host__ __device__ inline bool degeneracy_projector(const float* AtA, float thresh, float* matP) {
    float E[N], V[N * N], V2[N * N], Vi[N * N];
    eigen_sym<N>(AtA, E, V);
    for (int i = 0; i < N * N; ++i) V2[i] = V[i];
    bool degenerate = false;
    for (int i = N - 1; i >= 0; --i) {
        if (E[i] < thresh) { for (int j = 0; j < N; ++j) V2[i * N + j] = 0.f; degenerate = true; }
        else break;
    }
    inverse<N>(V, Vi);
    matmul<N>(Vi, V2, matP);
    return degenerate;
}

// True when the smallest eigenvalue of the symmetric matrix A is certainly above `thresh` by a margin no float
// eigen solver error can bridge (margin = 1e-5 * trace + 1 % of thresh; the error of a float Jacobi sweep on an N x N
// matrix is a few N * 6e-8 * ||A||).  Test: Cholesky of A - (thresh + margin) I in double succeeds with positive
// pivots.  Callers use it to skip degeneracy_projector when its answer is known to be "not degenerate" -- the
// projector itself is only ever applied when isDegenerate is set (featureAssociation.cpp:1361-1367,
// mapOptmization.cpp:1303-1308).
template <int N>
__host__ __device__ inline bool min_eigen_certainly_above(const float* A, float thresh) {
    double tr = 0.0;
    for (int i = 0; i < N; ++i) tr += (double)A[i * N + i];
    const double shift = (double)thresh + 1e-5 * tr + 0.01 * (double)thresh;
    double Lm[N * N];
    for (int i = 0; i < N; ++i) {
        for (int j = 0; j <= i; ++j) {
            double sum = (double)A[i * N + j] - (i == j ? shift : 0.0);
            for (int k = 0; k < j; ++k) sum -= Lm[i * N + k] * Lm[j * N + k];
            if (i == j) {
                if (!(sum > 1e-9 * (tr + 1.0))) return false;
                Lm[i * N + i] = sqrt(sum);
            } else {
                Lm[i * N + j] = sum / Lm[j * N + j];
            }
        }
    }
    return true;
}

}  // namespace sm
