// Host-side pose graph of mapOptimization: what the reference asks of GTSAM's iSAM2 (mapOptmization.cpp:229-232,
// 346-350, 1375-1409 chain of key poses; 929-942 loop-closure constraint; 1456-1478 correctPoses).  The graph is a
// handful of Pose3 variables per lane, so like in the reference it lives on the host; point clouds never do.
//
// Semantics (GTSAM Pose3 conventions): tangent vector = [omega, v] with the rotation first, retraction
// T <- T * [Exp(omega), v], factor error = [Log(dR), dt] of d = measured^-1 * (Ti^-1 * Tj) (prior: measured^-1 * Tj),
// whitened by 1/sqrt(variance).  While the graph is a pure chain every factor was computed from the inserted values,
// the residuals are zero and the estimate IS the inserted values, so no arithmetic happens (pass-through).  From the
// first loop factor on, every update() is one Gauss-Newton step over the whole graph relinearised at the current
// estimate, solved by an envelope (skyline) Cholesky in reverse Cuthill-McKee order of the poses (rcmOrder).  Jacobians are closed-form (right perturbations):
//   d e / d xi_j = [ Jr^-1(phi)  0 ; 0  Rd ]                 d e / d xi_i = [ -Jr^-1(phi) Rh^T  0 ; Rz^T [th]x  -Rz^T ]
// with h = Ti^-1 Tj = (Rh, th), z = (Rz, tz) the measurement, d = (Rd, td) = z^-1 h, phi = Log(Rd).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace lego {
namespace pg {

struct Pose { double R[9]; double t[3]; };

inline void mat_mul(const double A[9], const double B[9], double C[9]) {
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
inline void mat_tmul(const double A[9], const double B[9], double C[9]) {   // A^T B
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C[3 * r + c] = A[r] * B[c] + A[3 + r] * B[3 + c] + A[6 + r] * B[6 + c];
}
inline void skew(const double w[3], double K[9]) { K[0] = 0; K[1] = -w[2]; K[2] = w[1]; K[3] = w[2]; K[4] = 0; K[5] = -w[0]; K[6] = -w[1]; K[7] = w[0]; K[8] = 0; }

// gtsam::Rot3::RzRyRx(x, y, z) = Rz(z) Ry(y) Rx(x)
inline void rzryrx(double x, double y, double z, double R[9]) {
    const double cx = std::cos(x), sx = std::sin(x), cy = std::cos(y), sy = std::sin(y), cz = std::cos(z), sz = std::sin(z);
    R[0] = cz * cy; R[1] = cz * sy * sx - sz * cx; R[2] = cz * sy * cx + sz * sx;
    R[3] = sz * cy; R[4] = sz * sy * sx + cz * cx; R[5] = sz * sy * cx - cz * sx;
    R[6] = -sy;     R[7] = cy * sx;                R[8] = cy * cx;
}
// gtsam::Rot3::roll() / pitch() / yaw()
inline void rpy(const double R[9], double* roll, double* pitch, double* yaw) {
    *roll = std::atan2(R[7], R[8]);
    *pitch = std::atan2(-R[6], std::sqrt(R[7] * R[7] + R[8] * R[8]));
    *yaw = std::atan2(R[3], R[0]);
}
inline Pose make_pose(double rx, double ry, double rz, double tx, double ty, double tz) {
    Pose p; rzryrx(rx, ry, rz, p.R); p.t[0] = tx; p.t[1] = ty; p.t[2] = tz; return p;
}
inline void so3_exp(const double w[3], double R[9]) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
    const double a = th < 1e-8 ? 1.0 - th2 / 6.0 : std::sin(th) / th;
    const double b = th < 1e-8 ? 0.5 - th2 / 24.0 : (1.0 - std::cos(th)) / th2;
    double K[9], K2[9]; skew(w, K); mat_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0 ? 1.0 : 0.0) + a * K[i] + b * K2[i];
}
inline void so3_log(const double R[9], double w[3]) {
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0); c = c > 1 ? 1 : (c < -1 ? -1 : c);
    const double th = std::acos(c);
    const double v[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    if (3.141592653589793 - th < 1e-6) {   // half turn: axis from the diagonal; never reached by key-frame graphs
        for (int a = 0; a < 3; ++a) { double ax = std::sqrt(std::max(0.0, (R[4 * a] + 1.0) * 0.5)); w[a] = th * (v[a] < 0 ? -ax : ax); }
        return;
    }
    const double k = th < 1e-8 ? 0.5 + th * th / 12.0 : th / (2.0 * std::sin(th));
    for (int a = 0; a < 3; ++a) w[a] = k * v[a];
}
// inverse right Jacobian of SO(3)
inline void so3_jr_inv(const double phi[3], double J[9]) {
    const double th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2], th = std::sqrt(th2);
    const double c = th < 1e-5 ? 1.0 / 12.0 + th2 / 720.0 : 1.0 / th2 - (1.0 + std::cos(th)) / (2.0 * th * std::sin(th));
    double K[9], K2[9]; skew(phi, K); mat_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) J[i] = (i % 4 == 0 ? 1.0 : 0.0) + 0.5 * K[i] + c * K2[i];
}
inline Pose between(const Pose& A, const Pose& B) {   // A^-1 * B
    Pose o; mat_tmul(A.R, B.R, o.R);
    const double d[3] = {B.t[0] - A.t[0], B.t[1] - A.t[1], B.t[2] - A.t[2]};
    for (int r = 0; r < 3; ++r) o.t[r] = A.R[r] * d[0] + A.R[3 + r] * d[1] + A.R[6 + r] * d[2];
    return o;
}
inline Pose retract(const Pose& T, const double xi[6]) {
    Pose o; double E[9]; so3_exp(xi, E); mat_mul(T.R, E, o.R);
    for (int r = 0; r < 3; ++r) o.t[r] = T.t[r] + T.R[3 * r] * xi[3] + T.R[3 * r + 1] * xi[4] + T.R[3 * r + 2] * xi[5];
    return o;
}

struct Factor { int i, j; Pose z; double sigma[6]; };   // i < 0: prior on j

class Graph {
public:
    std::vector<Pose> x;
    std::vector<Factor> factors;
    bool hasLoop = false;
    long steps = 0;

    void insert(int key, const Pose& p) { if ((int)x.size() <= key) x.resize(key + 1); x[key] = p; }
    void addPrior(int key, const Pose& z, const double variances[6]) { Factor f; f.i = -1; f.j = key; f.z = z; for (int a = 0; a < 6; ++a) f.sigma[a] = std::sqrt(variances[a]); factors.push_back(f); }
    void addBetween(int k1, int k2, const Pose& z, const double variances[6]) {
        Factor f; f.i = k1; f.j = k2; f.z = z; for (int a = 0; a < 6; ++a) f.sigma[a] = std::sqrt(variances[a]);
        factors.push_back(f);
        if (std::abs(k1 - k2) != 1) hasLoop = true;
    }
    // ISAM2::update(): one Gauss-Newton step once a loop factor exists, nothing before
    void update() { if (hasLoop) { gaussNewtonStep(); ++steps; } }

private:
    // e (whitened), Ji (6x6, unused for priors), Jj (6x6)
    static void linearize(const Factor& F, const Pose* Ti, const Pose& Tj, double e[6], double Ji[36], double Jj[36]) {
        const Pose h = Ti ? between(*Ti, Tj) : Tj;
        const Pose d = between(F.z, h);
        double phi[3]; so3_log(d.R, phi);
        double Jr[9]; so3_jr_inv(phi, Jr);
        for (int a = 0; a < 3; ++a) { e[a] = phi[a]; e[3 + a] = d.t[a]; }
        std::memset(Jj, 0, sizeof(double) * 36);
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { Jj[6 * r + c] = Jr[3 * r + c]; Jj[6 * (3 + r) + 3 + c] = d.R[3 * r + c]; }
        if (Ti) {
            std::memset(Ji, 0, sizeof(double) * 36);
            double K[9]; skew(h.t, K);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) {
                    double a = 0, b = 0;
                    for (int k = 0; k < 3; ++k) { a += Jr[3 * r + k] * h.R[3 * c + k]; b += F.z.R[3 * k + r] * K[3 * k + c]; }
                    Ji[6 * r + c] = -a;                       // -Jr^-1 Rh^T
                    Ji[6 * (3 + r) + c] = b;                  // Rz^T [th]x
                    Ji[6 * (3 + r) + 3 + c] = -F.z.R[3 * c + r];   // -Rz^T
                }
        }
        for (int r = 0; r < 6; ++r) {
            const double w = 1.0 / F.sigma[r];
            e[r] *= w;
            for (int c = 0; c < 6; ++c) { Jj[6 * r + c] *= w; if (Ti) Ji[6 * r + c] *= w; }
        }
    }

    // Reverse Cuthill-McKee order of the key frames (block graph: one node per pose, one edge per between factor).  In key-
    // frame order a loop factor stretches the envelope of its newer pose back to the older one and the Cholesky fills it
    // (a 2000-scan circle: ~500 poses, rows 2000+ wide, ~1 s per update); ordered by breadth-first levels from the newest pose
    // the profile is a few poses per lap wide.  pos[k] = position of pose k in the elimination order.
    void rcmOrder(std::vector<int>& pos) const {
        const int n = (int)x.size();
        std::vector<std::vector<int> > adj(n);
        for (const Factor& F : factors) if (F.i >= 0 && F.i != F.j) { adj[F.i].push_back(F.j); adj[F.j].push_back(F.i); }
        for (auto& a : adj) { std::sort(a.begin(), a.end()); a.erase(std::unique(a.begin(), a.end()), a.end()); }
        std::vector<int> order; order.reserve(n);
        std::vector<char> seen(n, 0);
        for (int root = n - 1; root >= 0; --root) {          // (the graph is connected through the chain; the loop covers any other case)
            if (seen[root]) continue;
            size_t head = order.size();
            order.push_back(root); seen[root] = 1;
            while (head < order.size()) {
                const int u = order[head++];
                std::vector<int> nb;
                for (int v : adj[u]) if (!seen[v]) { nb.push_back(v); seen[v] = 1; }
                std::sort(nb.begin(), nb.end(), [&](int a, int b) { return adj[a].size() != adj[b].size() ? adj[a].size() < adj[b].size() : a < b; });
                order.insert(order.end(), nb.begin(), nb.end());
            }
        }
        pos.assign(n, 0);
        for (int k = 0; k < n; ++k) pos[order[n - 1 - k]] = k;
    }

    void gaussNewtonStep() {
        const int n = (int)x.size(), D = 6 * n;
        if (n == 0) return;
        std::vector<int> pos; rcmOrder(pos);
        std::vector<int> first(D);
        for (int r = 0; r < D; ++r) first[r] = r - r % 6;
        for (const Factor& F : factors)
            if (F.i >= 0) { const int lo = std::min(pos[F.i], pos[F.j]), hi = std::max(pos[F.i], pos[F.j]); for (int a = 0; a < 6; ++a) first[6 * hi + a] = std::min(first[6 * hi + a], 6 * lo); }
        std::vector<size_t> off(D + 1, 0);
        for (int r = 0; r < D; ++r) off[r + 1] = off[r] + (size_t)(r - first[r] + 1);
        std::vector<double> H(off[D], 0.0), g(D, 0.0);
        auto at = [&](int r, int c) -> double& { return H[off[r] + (size_t)(c - first[r])]; };
        for (const Factor& F : factors) {
            double e[6], Ji[36], Jj[36];
            linearize(F, F.i >= 0 ? &x[F.i] : nullptr, x[F.j], e, Ji, Jj);
            const double* J[2] = {F.i >= 0 ? Ji : Jj, Jj};
            const int var[2] = {pos[F.i >= 0 ? F.i : F.j], pos[F.j]};    // rows / columns in elimination order
            const int nv = F.i >= 0 ? 2 : 1;
            for (int v = 0; v < nv; ++v)
                for (int a = 0; a < 6; ++a) {
                    double s = 0; for (int r = 0; r < 6; ++r) s += J[v][6 * r + a] * e[r];
                    g[6 * var[v] + a] -= s;
                    for (int w = 0; w < nv; ++w)
                        for (int b = 0; b < 6; ++b) {
                            const int rr = 6 * var[v] + a, cc = 6 * var[w] + b;
                            if (cc > rr) continue;       // lower triangle; the mirrored block comes with (w, v)
                            double q = 0; for (int r = 0; r < 6; ++r) q += J[v][6 * r + a] * J[w][6 * r + b];
                            at(rr, cc) += q;
                        }
                }
        }
        for (int r = 0; r < D; ++r)
            for (int c = first[r]; c <= r; ++c) {
                double s = at(r, c);
                for (int k = std::max(first[r], first[c]); k < c; ++k) s -= at(r, k) * at(c, k);
                if (c < r) at(r, c) = s / at(c, c); else at(r, r) = std::sqrt(s > 1e-300 ? s : 1e-300);
            }
        for (int r = 0; r < D; ++r) { double s = g[r]; for (int k = first[r]; k < r; ++k) s -= at(r, k) * g[k]; g[r] = s / at(r, r); }
        for (int r = D - 1; r >= 0; --r) { g[r] /= at(r, r); for (int k = first[r]; k < r; ++k) g[k] -= at(r, k) * g[r]; }
        for (int i = 0; i < n; ++i) x[i] = retract(x[i], &g[6 * pos[i]]);
    }
};

}  // namespace pg
}  // namespace lego
