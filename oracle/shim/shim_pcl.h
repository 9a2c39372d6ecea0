// TEST INFRASTRUCTURE ONLY (oracle).  Minimal stand-in for the PCL / Eigen names the reference sources use.
// Containers are plain std::vector wrappers; VoxelGrid and KdTreeFLANN delegate to the restatements in
// oracle/thirdparty/.  IterativeClosestPoint delegates to oracle/thirdparty/tp_icp.h; Eigen::Affine3f,
// pcl::getTransformation and pcl::getTranslationAndEulerAngles restate PCL's common/eigen.hpp (float); they only
// run in the configs generated with loopClosureEnableFlag = true (utility.h:104).  pcl::io::savePCDFileASCII writes no
// file: it keeps the cloud and its PCD text (oracle/thirdparty/tp_pcd.h) in memory under the path it was given.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include "shim_ros.h"
#include "../thirdparty/tp_voxel_grid.h"
#include "../thirdparty/tp_kdtree.h"
#include "../thirdparty/tp_icp.h"
#include "../thirdparty/tp_pcd.h"
#include <map>

#define PCL_ADD_POINT4D union { float data[4]; struct { float x; float y; float z; }; };
#define PCL_ADD_INTENSITY union { struct { float intensity; }; float data_c[4]; }
#define EIGEN_ALIGN16 __attribute__((aligned(16)))
#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW
#define POINT_CLOUD_REGISTER_POINT_STRUCT(name, fields)

namespace Eigen {
struct Matrix4f {
    float m[4][4];
    Matrix4f() { std::memset(m, 0, sizeof(m)); for (int i = 0; i < 4; ++i) m[i][i] = 1.f; }
};
struct Affine3f {
    Matrix4f mat;
    Affine3f() {}
    Affine3f(const Matrix4f& o) : mat(o) {}
    Affine3f& operator=(const Matrix4f& o) { mat = o; return *this; }
    Affine3f operator*(const Affine3f& o) const {
        Affine3f r;
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { float s = 0; for (int k = 0; k < 4; ++k) s += mat.m[i][k] * o.mat.m[k][j]; r.mat.m[i][j] = s; }
        return r;
    }
};
}  // namespace Eigen

namespace pcl {

struct EIGEN_ALIGN16 PointXYZI {
    PCL_ADD_POINT4D
    PCL_ADD_INTENSITY;
    PointXYZI() { x = y = z = 0.f; data[3] = 1.f; intensity = 0.f; }
};

template <class T> struct PointCloud {
    typedef std::shared_ptr<PointCloud<T> > Ptr;
    typedef std::shared_ptr<const PointCloud<T> > ConstPtr;
    std::vector<T> points;
    std_msgs::Header header;
    uint32_t width = 0, height = 1;
    bool is_dense = true;
    void clear() { points.clear(); width = 0; }
    void push_back(const T& p) { points.push_back(p); width = static_cast<uint32_t>(points.size()); }
    void resize(size_t n) { points.resize(n); width = static_cast<uint32_t>(n); }
    size_t size() const { return points.size(); }
    bool empty() const { return points.empty(); }
    T& back() { return points.back(); }
    PointCloud& operator+=(const PointCloud& o) { points.insert(points.end(), o.points.begin(), o.points.end()); width = static_cast<uint32_t>(points.size()); return *this; }
};

// ---- message conversion (xyz + intensity, optional ring picked up by structs that have a `ring` member) ----
template <class T> auto shim_set_ring(T& p, uint16_t r, int) -> decltype(p.ring, void()) { p.ring = r; }
template <class T> void shim_set_ring(T&, uint16_t, long) {}

template <class T> void fromROSMsg(const sensor_msgs::PointCloud2& msg, PointCloud<T>& cloud) {
    size_t n = msg.xyzi.size() / 4;
    cloud.points.resize(n);
    for (size_t i = 0; i < n; ++i) {
        T& p = cloud.points[i];
        p.x = msg.xyzi[4 * i]; p.y = msg.xyzi[4 * i + 1]; p.z = msg.xyzi[4 * i + 2]; p.intensity = msg.xyzi[4 * i + 3];
        if (!msg.ring.empty()) shim_set_ring(p, msg.ring[i], 0);
    }
    cloud.header = msg.header; cloud.width = static_cast<uint32_t>(n); cloud.is_dense = msg.is_dense;
}
template <class T> void toROSMsg(const PointCloud<T>& cloud, sensor_msgs::PointCloud2& msg) {
    size_t n = cloud.points.size();
    msg.xyzi.resize(4 * n); msg.ring.clear(); msg.is_dense = cloud.is_dense;
    for (size_t i = 0; i < n; ++i) {
        const T& p = cloud.points[i];
        msg.xyzi[4 * i] = p.x; msg.xyzi[4 * i + 1] = p.y; msg.xyzi[4 * i + 2] = p.z; msg.xyzi[4 * i + 3] = p.intensity;
    }
}
// Dense clouds are passed through untouched; otherwise points with a non-finite coordinate are dropped in order.
template <class T> void removeNaNFromPointCloud(const PointCloud<T>& in, PointCloud<T>& out, std::vector<int>& index) {
    if (in.is_dense) {
        if (&in != &out) out = in;
        index.resize(in.points.size());
        for (size_t i = 0; i < index.size(); ++i) index[i] = static_cast<int>(i);
        return;
    }
    std::vector<T> kept; index.clear();
    for (size_t i = 0; i < in.points.size(); ++i) {
        const T& p = in.points[i];
        if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
        kept.push_back(p); index.push_back(static_cast<int>(i));
    }
    out.header = in.header; out.points.swap(kept); out.width = static_cast<uint32_t>(out.points.size()); out.is_dense = true;
}
template <class T> void copyPointCloud(const PointCloud<T>& in, PointCloud<T>& out) { out = in; }

inline float rad2deg(float a) { return a * 57.29578f; }
inline double rad2deg(double a) { return a * 57.29577951308232; }

// ---- VoxelGrid (delegates to oracle/thirdparty/tp_voxel_grid.h) ----
template <class T> class VoxelGrid {
public:
    void setLeafSize(float lx, float, float) { leaf_ = lx; }
    void setInputCloud(const typename PointCloud<T>::Ptr& c) { in_ = c; }
    void filter(PointCloud<T>& out) {
        std::vector<float> flat, res;
        size_t n = in_ ? in_->points.size() : 0;
        flat.resize(4 * n);
        for (size_t i = 0; i < n; ++i) { const T& p = in_->points[i]; flat[4 * i] = p.x; flat[4 * i + 1] = p.y; flat[4 * i + 2] = p.z; flat[4 * i + 3] = p.intensity; }
        size_t m = tp::voxel_grid_filter(flat.data(), n, leaf_, res);
        out.points.resize(m);
        for (size_t i = 0; i < m; ++i) { T& p = out.points[i]; p.x = res[4 * i]; p.y = res[4 * i + 1]; p.z = res[4 * i + 2]; p.intensity = res[4 * i + 3]; }
        out.width = static_cast<uint32_t>(m); out.height = 1; out.is_dense = true;
    }
private:
    float leaf_ = 1.f;
    typename PointCloud<T>::Ptr in_;
};

// ---- KdTreeFLANN (delegates to oracle/thirdparty/tp_kdtree.h) ----
template <class T> class KdTreeFLANN {
public:
    typedef std::shared_ptr<KdTreeFLANN<T> > Ptr;
    void setInputCloud(const typename PointCloud<T>::Ptr& c) {
        cloud_ = c; size_t n = c->points.size();
        flat_.resize(3 * n);
        for (size_t i = 0; i < n; ++i) { flat_[3 * i] = c->points[i].x; flat_[3 * i + 1] = c->points[i].y; flat_[3 * i + 2] = c->points[i].z; }
        tree_.build(flat_.data(), n, 3);
    }
    int nearestKSearch(const T& q, int k, std::vector<int>& idx, std::vector<float>& d2) const {
        float qq[3] = {q.x, q.y, q.z};
        return tree_.knn(qq, k, idx, d2);
    }
    int radiusSearch(const T& q, double radius, std::vector<int>& idx, std::vector<float>& d2, unsigned = 0) const {
        float qq[3] = {q.x, q.y, q.z};
        return tree_.radius(qq, static_cast<float>(radius), idx, d2);
    }
private:
    typename PointCloud<T>::Ptr cloud_;
    std::vector<float> flat_;
    tp::KdTree3 tree_;
};

// ---- loop closure (mapOptmization.cpp:875-945) ----
struct ShimIcpLog { long runs = 0; bool converged = false; int iterations = 0; double fitness = 0; float T[16]; long ns = 0, nt = 0; };
inline ShimIcpLog& shim_icp_log() { static ShimIcpLog l; return l; }

template <class S, class D> class IterativeClosestPoint {
public:
    void setMaxCorrespondenceDistance(double d) { maxDist_ = d; }
    void setMaximumIterations(int n) { maxIter_ = n; }
    void setTransformationEpsilon(double e) { epsT_ = e; }
    void setEuclideanFitnessEpsilon(double e) { epsF_ = e; }
    void setRANSACIterations(int) {}
    void setInputSource(const typename PointCloud<S>::Ptr& c) { src_ = c; }
    void setInputTarget(const typename PointCloud<D>::Ptr& c) { tgt_ = c; }
    void align(PointCloud<S>& out) {
        std::vector<float> a, b;
        flatten(*src_, a); flatten(*tgt_, b);
        res_ = tp::icp_align(a.data(), src_->points.size(), b.data(), tgt_->points.size(), maxDist_, maxIter_, epsT_, epsF_);
        out = *src_;
        for (auto& p : out.points) { float q[3] = {p.x, p.y, p.z}, o[3]; tp::icp_apply(res_.T, q, o); p.x = o[0]; p.y = o[1]; p.z = o[2]; }
        ShimIcpLog& l = shim_icp_log();
        l.runs += 1; l.converged = res_.converged; l.iterations = res_.iterations; l.fitness = res_.fitness;
        std::memcpy(l.T, res_.T, sizeof(l.T)); l.ns = (long)src_->points.size(); l.nt = (long)tgt_->points.size();
    }
    bool hasConverged() const { return res_.converged; }
    double getFitnessScore() const { return res_.fitness; }
    Eigen::Matrix4f getFinalTransformation() const { Eigen::Matrix4f m; for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m.m[r][c] = res_.T[4 * r + c]; return m; }
private:
    template <class C> static void flatten(const C& c, std::vector<float>& o) {
        o.resize(4 * c.points.size());
        for (size_t i = 0; i < c.points.size(); ++i) { o[4 * i] = c.points[i].x; o[4 * i + 1] = c.points[i].y; o[4 * i + 2] = c.points[i].z; o[4 * i + 3] = c.points[i].intensity; }
    }
    double maxDist_ = 1e30, epsT_ = 0, epsF_ = 0; int maxIter_ = 10;
    typename PointCloud<S>::Ptr src_; typename PointCloud<D>::Ptr tgt_;
    tp::IcpResult res_;
};
template <class T> void transformPointCloud(const PointCloud<T>& in, PointCloud<T>& out, const Eigen::Matrix4f& m) {
    out = in;
    for (auto& p : out.points) { float q[3] = {p.x, p.y, p.z}, o[3]; tp::icp_apply(&m.m[0][0], q, o); p.x = o[0]; p.y = o[1]; p.z = o[2]; }
}
// pcl::getTransformation (common/impl/eigen.hpp): R = Rz(yaw) Ry(pitch) Rx(roll), float
inline Eigen::Affine3f getTransformation(float x, float y, float z, float roll, float pitch, float yaw) {
    const float A = std::cos(yaw), B = std::sin(yaw), C = std::cos(pitch), D = std::sin(pitch), E = std::cos(roll), F = std::sin(roll);
    const float DE = D * E, DF = D * F;
    Eigen::Affine3f t;
    t.mat.m[0][0] = A * C; t.mat.m[0][1] = A * DF - B * E; t.mat.m[0][2] = B * F + A * DE; t.mat.m[0][3] = x;
    t.mat.m[1][0] = B * C; t.mat.m[1][1] = A * E + B * DF; t.mat.m[1][2] = B * DE - A * F; t.mat.m[1][3] = y;
    t.mat.m[2][0] = -D;    t.mat.m[2][1] = C * F;          t.mat.m[2][2] = C * E;          t.mat.m[2][3] = z;
    t.mat.m[3][0] = 0.f; t.mat.m[3][1] = 0.f; t.mat.m[3][2] = 0.f; t.mat.m[3][3] = 1.f;
    return t;
}
inline void getTranslationAndEulerAngles(const Eigen::Affine3f& t, float& x, float& y, float& z, float& roll, float& pitch, float& yaw) {
    x = t.mat.m[0][3]; y = t.mat.m[1][3]; z = t.mat.m[2][3];
    roll = std::atan2(t.mat.m[2][1], t.mat.m[2][2]);
    pitch = std::asin(-t.mat.m[2][0]);
    yaw = std::atan2(t.mat.m[1][0], t.mat.m[0][0]);
}
// what the reference "saved": path -> (flat xyzi, PCD text); nothing touches the file system
struct ShimPcdFile { std::vector<float> xyzi; std::string text; };
inline std::map<std::string, ShimPcdFile>& shim_pcd_store() { static std::map<std::string, ShimPcdFile> m; return m; }
namespace io {
template <class T> int savePCDFileASCII(const std::string& path, const PointCloud<T>& cloud) {
    ShimPcdFile f;
    f.xyzi.resize(4 * cloud.points.size());
    for (size_t k = 0; k < cloud.points.size(); ++k) {
        f.xyzi[4 * k] = cloud.points[k].x; f.xyzi[4 * k + 1] = cloud.points[k].y; f.xyzi[4 * k + 2] = cloud.points[k].z;
        f.xyzi[4 * k + 3] = cloud.points[k].intensity;
    }
    f.text = tp::pcd_ascii_xyzi(f.xyzi);
    shim_pcd_store()[path] = f;
    return 0;
}
}  // namespace io

}  // namespace pcl
