// TEST INFRASTRUCTURE ONLY (oracle).
//
// Stand-in for what the reference asks of GTSAM 4.0.0-alpha2 (NOT vendored in /root/reference; README.md:18-26):
// a Pose3 graph with one PriorFactor, a chain of BetweenFactors (mapOptmization.cpp:1375-1390) and, once loop closure
// fires, BetweenFactors between non-consecutive key frames (mapOptmization.cpp:939), all with diagonal noise, updated
// through ISAM2::update() and read through calculateEstimate().  PARITY UNPINNED (no GTSAM here, no fixture in the
// reference).  Restated behaviour:
//   * while the graph is a pure chain every BetweenFactor was computed from the very values that were inserted, so its
//     residual is zero and the estimate equals the inserted values: pass-through, no arithmetic;
//   * from the first loop factor on, every ISAM2::update() call is ONE Gauss-Newton step over the whole graph,
//     relinearised at the current estimate (iSAM2 with relinearizeThreshold 0.01 / relinearizeSkip 1 re-solves at
//     every update and differs from a full relinearisation only in variables whose pending delta is below 0.01).
// Conventions (GTSAM Pose3): tangent = [omega (rotation), v (translation)], retraction T <- T * [Exp(omega), v]
// (R <- R Exp(omega), t <- t + R v), factor error = [Log(dR), dt] of  d = measured^-1 * (Ti^-1 Tj)  (prior: measured^-1 * T),
// whitened by 1/sqrt(variance).  Jacobians here are central differences in double (the product uses closed forms).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

namespace tp {

struct PgPose { double R[9]; double t[3]; };

inline void pg_mul(const double A[9], const double B[9], double C[9]) {
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { double s = 0; for (int k = 0; k < 3; ++k) s += A[3 * r + k] * B[3 * k + c]; C[3 * r + c] = s; }
}
inline void pg_exp(const double w[3], double R[9]) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
    double a, b;   // R = I + a K + b K^2
    if (th < 1e-8) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = std::sin(th) / th; b = (1.0 - std::cos(th)) / th2; }
    const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
    double K2[9]; pg_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0 ? 1.0 : 0.0) + a * K[i] + b * K2[i];
}
inline void pg_log(const double R[9], double w[3]) {
    const double tr = R[0] + R[4] + R[8];
    double c = 0.5 * (tr - 1.0); c = c > 1 ? 1 : (c < -1 ? -1 : c);
    const double th = std::acos(c);
    const double v[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    double k;
    if (th < 1e-8) k = 0.5 + th * th / 12.0;
    else if (3.141592653589793 - th < 1e-6) {   // near pi: axis from the diagonal (never reached by these graphs)
        double ax[3]; for (int a = 0; a < 3; ++a) ax[a] = std::sqrt(std::max(0.0, (R[4 * a] + 1.0) * 0.5));
        if (v[0] < 0) ax[0] = -ax[0]; if (v[1] < 0) ax[1] = -ax[1]; if (v[2] < 0) ax[2] = -ax[2];
        for (int a = 0; a < 3; ++a) w[a] = th * ax[a];
        return;
    } else k = th / (2.0 * std::sin(th));
    for (int a = 0; a < 3; ++a) w[a] = k * v[a];
}
inline PgPose pg_between(const PgPose& A, const PgPose& B) {   // A^-1 * B
    PgPose o;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { double s = 0; for (int k = 0; k < 3; ++k) s += A.R[3 * k + r] * B.R[3 * k + c]; o.R[3 * r + c] = s; }
    const double d[3] = {B.t[0] - A.t[0], B.t[1] - A.t[1], B.t[2] - A.t[2]};
    for (int r = 0; r < 3; ++r) o.t[r] = A.R[r] * d[0] + A.R[3 + r] * d[1] + A.R[6 + r] * d[2];
    return o;
}
inline PgPose pg_retract(const PgPose& T, const double xi[6]) {
    PgPose o; double E[9]; pg_exp(xi, E); pg_mul(T.R, E, o.R);
    for (int r = 0; r < 3; ++r) o.t[r] = T.t[r] + T.R[3 * r] * xi[3] + T.R[3 * r + 1] * xi[4] + T.R[3 * r + 2] * xi[5];
    return o;
}
inline void pg_local(const PgPose& D, double e[6]) { pg_log(D.R, e); e[3] = D.t[0]; e[4] = D.t[1]; e[5] = D.t[2]; }

struct PgFactor { int i, j; PgPose z; double sigma[6]; };   // i < 0: prior on j

struct PoseGraph {
    std::vector<PgPose> x;
    std::vector<PgFactor> f;
    bool hasLoop = false;

    void error(const PgFactor& F, const PgPose& Ti, const PgPose& Tj, double e[6]) const {
        PgPose h = F.i < 0 ? Tj : pg_between(Ti, Tj);
        pg_local(pg_between(F.z, h), e);
        for (int a = 0; a < 6; ++a) e[a] /= F.sigma[a];
    }

    // Elimination order: breadth-first levels from the newest pose, reversed (reverse Cuthill-McKee).  Only the cost
    // depends on it (in key-frame order every loop factor stretches the envelope across the chain); the Gauss-Newton step is
    // the same linear solve in exact arithmetic.
    std::vector<int> eliminationOrder() const {
        const int n = (int)x.size();
        std::vector<std::vector<int> > nbr(n);
        for (const PgFactor& F : f) if (F.i >= 0 && F.i != F.j) { nbr[F.i].push_back(F.j); nbr[F.j].push_back(F.i); }
        std::vector<int> seq; std::vector<char> done(n, 0);
        for (int start = n - 1; start >= 0; --start) {
            if (done[start]) continue;
            size_t q = seq.size(); seq.push_back(start); done[start] = 1;
            for (; q < seq.size(); ++q) {
                std::vector<std::pair<size_t, int> > next;
                for (int v : nbr[seq[q]]) if (!done[v]) { done[v] = 1; next.push_back(std::make_pair(nbr[v].size(), v)); }
                std::sort(next.begin(), next.end());
                for (size_t k = 0; k < next.size(); ++k) seq.push_back(next[k].second);
            }
        }
        std::vector<int> where(n);
        for (int k = 0; k < n; ++k) where[seq[n - 1 - k]] = k;
        return where;
    }

    // one Gauss-Newton step: envelope (skyline) Cholesky of the normal equations
    void gaussNewtonStep() {
        const int n = (int)x.size(), D = 6 * n;
        if (n == 0) return;
        const std::vector<int> where = eliminationOrder();
        std::vector<int> first(D);
        for (int r = 0; r < D; ++r) first[r] = r - (r % 6);
        for (const PgFactor& F : f) if (F.i >= 0) { const int lo = std::min(where[F.i], where[F.j]), hi = std::max(where[F.i], where[F.j]); for (int a = 0; a < 6; ++a) first[6 * hi + a] = std::min(first[6 * hi + a], 6 * lo); }
        std::vector<size_t> off(D + 1, 0);
        for (int r = 0; r < D; ++r) off[r + 1] = off[r] + (size_t)(r - first[r] + 1);
        std::vector<double> H(off[D], 0.0), g(D, 0.0);
        auto at = [&](int r, int c) -> double& { return H[off[r] + (size_t)(c - first[r])]; };   // c <= r, c >= first[r]
        const double h = 1e-6;
        for (const PgFactor& F : f) {
            const int nv = F.i < 0 ? 1 : 2;
            const int var[2] = {where[F.i < 0 ? F.j : F.i], where[F.j]};
            double e0[6]; error(F, F.i < 0 ? x[F.j] : x[F.i], x[F.j], e0);
            double J[2][6][6];
            for (int v = 0; v < nv; ++v)
                for (int a = 0; a < 6; ++a) {
                    double xp[6] = {0, 0, 0, 0, 0, 0}, xm[6] = {0, 0, 0, 0, 0, 0}; xp[a] = h; xm[a] = -h;
                    PgPose Ti = F.i < 0 ? x[F.j] : x[F.i], Tj = x[F.j], Ti2 = Ti, Tj2 = Tj;
                    if (F.i < 0 || v == 1) { Tj = pg_retract(x[F.j], xp); Tj2 = pg_retract(x[F.j], xm); }
                    else { Ti = pg_retract(x[F.i], xp); Ti2 = pg_retract(x[F.i], xm); }
                    double ep[6], em[6]; error(F, Ti, Tj, ep); error(F, Ti2, Tj2, em);
                    for (int r = 0; r < 6; ++r) J[v][r][a] = (ep[r] - em[r]) / (2 * h);
                }
            for (int v = 0; v < nv; ++v)
                for (int a = 0; a < 6; ++a) {
                    double s = 0; for (int r = 0; r < 6; ++r) s += J[v][r][a] * e0[r];
                    g[6 * var[v] + a] -= s;
                    for (int w = 0; w < nv; ++w)
                        for (int b = 0; b < 6; ++b) {
                            const int rr = 6 * var[v] + a, cc = 6 * var[w] + b;
                            if (cc > rr) continue;
                            double q = 0; for (int r = 0; r < 6; ++r) q += J[v][r][a] * J[w][r][b];
                            at(rr, cc) += q;
                        }
                }
        }
        // L L^T in place inside the envelope
        for (int r = 0; r < D; ++r) {
            for (int c = first[r]; c <= r; ++c) {
                double s = at(r, c);
                const int k0 = std::max(first[r], first[c]);
                for (int k = k0; k < c; ++k) s -= at(r, k) * at(c, k);
                if (c < r) at(r, c) = s / at(c, c); else at(r, r) = std::sqrt(s > 1e-300 ? s : 1e-300);
            }
        }
        for (int r = 0; r < D; ++r) { double s = g[r]; for (int k = first[r]; k < r; ++k) s -= at(r, k) * g[k]; g[r] = s / at(r, r); }
        for (int r = D - 1; r >= 0; --r) { g[r] /= at(r, r); for (int k = first[r]; k < r; ++k) g[k] -= at(r, k) * g[r]; }
        for (int i = 0; i < n; ++i) x[i] = pg_retract(x[i], &g[6 * where[i]]);
    }
};

}  // namespace tp
