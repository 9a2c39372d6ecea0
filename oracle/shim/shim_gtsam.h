// TEST INFRASTRUCTURE ONLY (oracle).  GTSAM 4.0.0-alpha2 is not vendored in /root/reference.  With
// loopClosureEnableFlag == false (utility.h:104) the factor graph built in mapOptmization.cpp:1375-1409 is a
// prior plus a chain of BetweenFactors whose measurements are computed from the very values inserted, so the
// iSAM2 estimate of the newest pose equals its inserted initial value (to ~1e-7, far inside the 1e-4 pose
// tolerance).  While the graph is such a chain this shim keeps the inserted Pose3 values and hands them back
// (pass-through); from the first loop BetweenFactor (mapOptmization.cpp:939) on, every ISAM2::update() is one
// Gauss-Newton step over the whole graph (oracle/thirdparty/tp_pose_graph.h, PARITY UNPINNED).  Rot3 keeps a
// double rotation matrix; roll/pitch/yaw follow GTSAM's documented RQ-style extraction (x about X first, then y,
// then z for R = Rz*Ry*Rx).
#pragma once
#include <cmath>
#include <map>
#include <memory>
#include <vector>
#include "../thirdparty/tp_pose_graph.h"

namespace gtsam {

struct Point3 {
    double v[3];
    Point3(double x = 0, double y = 0, double z = 0) { v[0] = x; v[1] = y; v[2] = z; }
    double x() const { return v[0]; } double y() const { return v[1]; } double z() const { return v[2]; }
};

struct Rot3 {
    double R[3][3];
    Rot3() { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[i][j] = (i == j); }
    static Rot3 RzRyRx(double x, double y, double z) {
        double cx = std::cos(x), sx = std::sin(x), cy = std::cos(y), sy = std::sin(y), cz = std::cos(z), sz = std::sin(z);
        Rot3 r;
        r.R[0][0] = cz * cy; r.R[0][1] = cz * sy * sx - sz * cx; r.R[0][2] = cz * sy * cx + sz * sx;
        r.R[1][0] = sz * cy; r.R[1][1] = sz * sy * sx + cz * cx; r.R[1][2] = sz * sy * cx - cz * sx;
        r.R[2][0] = -sy;     r.R[2][1] = cy * sx;                r.R[2][2] = cy * cx;
        return r;
    }
    Rot3 operator*(const Rot3& o) const { Rot3 r; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double s = 0; for (int k = 0; k < 3; ++k) s += R[i][k] * o.R[k][j]; r.R[i][j] = s; } return r; }
    Rot3 inverse() const { Rot3 r; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.R[i][j] = R[j][i]; return r; }
    void xyz(double& x, double& y, double& z) const {
        // R = Rz(z)*Ry(y)*Rx(x): third row is [-sy, cy*sx, cy*cx], first column is [cz*cy, sz*cy, -sy].
        x = std::atan2(R[2][1], R[2][2]);
        y = std::atan2(-R[2][0], std::sqrt(R[2][1] * R[2][1] + R[2][2] * R[2][2]));
        z = std::atan2(R[1][0], R[0][0]);
    }
    double roll() const { double x, y, z; xyz(x, y, z); return x; }
    double pitch() const { double x, y, z; xyz(x, y, z); return y; }
    double yaw() const { double x, y, z; xyz(x, y, z); return z; }
};

struct Pose3 {
    Rot3 r; Point3 t;
    Pose3() {}
    Pose3(const Rot3& rr, const Point3& tt) : r(rr), t(tt) {}
    const Rot3& rotation() const { return r; }
    const Point3& translation() const { return t; }
    Pose3 between(const Pose3& o) const {   // this^-1 * o
        Rot3 ri = r.inverse();
        double d[3] = {o.t.v[0] - t.v[0], o.t.v[1] - t.v[1], o.t.v[2] - t.v[2]};
        Point3 tt(ri.R[0][0] * d[0] + ri.R[0][1] * d[1] + ri.R[0][2] * d[2],
                  ri.R[1][0] * d[0] + ri.R[1][1] * d[1] + ri.R[1][2] * d[2],
                  ri.R[2][0] * d[0] + ri.R[2][1] * d[1] + ri.R[2][2] * d[2]);
        return Pose3(ri * o.r, tt);
    }
};

struct Vector {
    std::vector<double> d; size_t fill = 0;
    explicit Vector(int n = 0) : d(n, 0.0) {}
    struct Comma { Vector* v; Comma operator,(double x) { if (v->fill < v->d.size()) v->d[v->fill++] = x; return *this; } };
    Comma operator<<(double x) { fill = 0; if (!d.empty()) d[fill++] = x; return Comma{this}; }
};

namespace noiseModel {
struct Diagonal {
    typedef std::shared_ptr<Diagonal> shared_ptr;
    double sigma[6];
    static shared_ptr Variances(const Vector& v) {
        shared_ptr d = std::make_shared<Diagonal>();
        for (int i = 0; i < 6; ++i) d->sigma[i] = std::sqrt(i < (int)v.d.size() ? v.d[i] : 1.0);
        return d;
    }
};
}

inline tp::PgPose shim_pg(const Pose3& p) {
    tp::PgPose o;
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) o.R[3 * i + j] = p.r.R[i][j]; o.t[i] = p.t.v[i]; }
    return o;
}
inline Pose3 shim_pose(const tp::PgPose& p) {
    Pose3 o;
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) o.r.R[i][j] = p.R[3 * i + j]; o.t.v[i] = p.t[i]; }
    return o;
}

template <class T> struct PriorFactor {
    tp::PgFactor f;
    template <class N> PriorFactor(size_t k, const T& z, const N& noise) { f.i = -1; f.j = (int)k; f.z = shim_pg(z); for (int a = 0; a < 6; ++a) f.sigma[a] = noise->sigma[a]; }
};
template <class T> struct BetweenFactor {
    tp::PgFactor f;
    template <class N> BetweenFactor(size_t k1, size_t k2, const T& z, const N& noise) { f.i = (int)k1; f.j = (int)k2; f.z = shim_pg(z); for (int a = 0; a < 6; ++a) f.sigma[a] = noise->sigma[a]; }
};

struct NonlinearFactorGraph {
    std::vector<tp::PgFactor> f;
    template <class F> void add(const F& x) { f.push_back(x.f); }
    void resize(size_t n) { f.resize(n); }
};

struct Values {
    std::map<size_t, Pose3> m;
    void insert(size_t k, const Pose3& p) { m[k] = p; }
    void clear() { m.clear(); }
    size_t size() const { return m.size(); }
    template <class T> const T& at(size_t k) const { return m.at(k); }
};

struct ISAM2Params { double relinearizeThreshold = 0.1; int relinearizeSkip = 10; };

struct ISAM2 {
    tp::PoseGraph g;
    long gnSteps = 0;
    explicit ISAM2(const ISAM2Params& = ISAM2Params()) {}
    void step() { if (g.hasLoop) { g.gaussNewtonStep(); ++gnSteps; } }
    void update() { step(); }
    void update(const NonlinearFactorGraph& graph) { update(graph, Values()); }
    void update(const NonlinearFactorGraph& graph, const Values& init) {
        for (auto& kv : init.m) { if (g.x.size() <= kv.first) g.x.resize(kv.first + 1); g.x[kv.first] = shim_pg(kv.second); }
        for (const tp::PgFactor& F : graph.f) { g.f.push_back(F); if (F.i >= 0 && std::abs(F.i - F.j) != 1) g.hasLoop = true; }
        step();
    }
    Values calculateEstimate() const { Values v; for (size_t k = 0; k < g.x.size(); ++k) v.m[k] = shim_pose(g.x[k]); return v; }
};

}  // namespace gtsam
