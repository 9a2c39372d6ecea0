// TEST INFRASTRUCTURE ONLY (oracle). Never linked into the product library.
//
// Minimal stand-in for roscpp / std_msgs / sensor_msgs / nav_msgs / geometry_msgs / tf so that the
// reference's node sources (LeGO-LOAM/src/{imageProjection,featureAssociation,mapOptmization}.cpp)
// compile UNMODIFIED into a ROS-free harness.  Nothing here is copied from ROS:
// only the handful of names the reference touches exist, and transport is replaced by "remember the last
// message per topic" so the harness can hand messages from one stage object to the next synchronously.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <functional>
#include <map>
#include <set>
#include <memory>
#include <string>
#include <vector>

namespace ros {

struct Duration { double sec_ = 0.0; double toSec() const { return sec_; } };
struct Time {
    double sec_ = 0.0;
    Time() {}
    explicit Time(double s) : sec_(s) {}
    double toSec() const { return sec_; }
    Time& fromSec(double s) { sec_ = s; return *this; }
    static Time now() { return Time(0.0); }
    Duration operator-(const Time& o) const { Duration d; d.sec_ = sec_ - o.sec_; return d; }
};

struct Rate {
    explicit Rate(double) {}
    void sleep() {}
};

// Global harness state: last message per topic, publish counters, which topics count as "subscribed".
struct ShimState {
    std::map<std::string, std::shared_ptr<void> > last;
    std::map<std::string, long> count;
    std::set<std::string> subscribed;   // topics for which getNumSubscribers() != 0
    bool all_subscribed = false;
    bool shutdown_requested = false;
    int errors = 0;
    // callbacks registered through NodeHandle::subscribe, per topic; shim_deliver() hands them a message synchronously (the
    // harness of the B200 node shells drives its nodes that way; the reference harness calls the handlers itself)
    std::map<std::string, std::vector<std::function<void(const std::shared_ptr<void>&)> > > callbacks;
    static ShimState& get() { static ShimState s; return s; }
};

struct Publisher {
    std::string topic;
    template <class M> void publish(const M& m) const {
        ShimState& s = ShimState::get();
        s.last[topic] = std::make_shared<M>(m);
        s.count[topic] += 1;
    }
    int getNumSubscribers() const {
        ShimState& s = ShimState::get();
        return (s.all_subscribed || s.subscribed.count(topic)) ? 1 : 0;
    }
};

struct Subscriber {};

struct NodeHandle {
    explicit NodeHandle(const std::string& = "") {}
    template <class M> Publisher advertise(const std::string& topic, int) { Publisher p; p.topic = topic; return p; }
    template <class M, class T, class A>
    Subscriber subscribe(const std::string& topic, int, void (T::*fn)(A), T* obj) {
        ShimState::get().callbacks[topic].push_back([fn, obj](const std::shared_ptr<void>& p) {
            std::shared_ptr<const M> m = std::static_pointer_cast<const M>(p);
            (obj->*fn)(m);
        });
        return Subscriber();
    }
};

inline void init(int, char**, const std::string&) {}
inline void spin() {}
inline void spinOnce() {}
inline bool ok() { return false; }
inline void shutdown() { ShimState::get().shutdown_requested = true; }

template <class M> std::shared_ptr<const M> shim_last(const std::string& topic) {
    ShimState& s = ShimState::get();
    auto it = s.last.find(topic);
    if (it == s.last.end()) return std::shared_ptr<const M>();
    return std::static_pointer_cast<const M>(it->second);
}
inline long shim_count(const std::string& topic) { return ShimState::get().count[topic]; }
// hands `msg` to every callback subscribed to `topic`; returns how many there were
inline int shim_deliver(const std::string& topic, const std::shared_ptr<void>& msg) {
    ShimState& s = ShimState::get();
    auto it = s.callbacks.find(topic);
    if (it == s.callbacks.end()) return 0;
    for (auto& cb : it->second) cb(msg);
    return (int)it->second.size();
}
// the last message published on `topic`, type-erased (for shim_deliver)
inline std::shared_ptr<void> shim_last_raw(const std::string& topic) {
    ShimState& s = ShimState::get();
    auto it = s.last.find(topic);
    return it == s.last.end() ? std::shared_ptr<void>() : it->second;
}

}  // namespace ros

#define ROS_INFO(...)  do { } while (0)
#define ROS_ERROR(...) do { ros::ShimState::get().errors += 1; } while (0)
#define ROS_FATAL(...) do { ros::ShimState::get().errors += 1; } while (0)

namespace std_msgs {
struct Header { ros::Time stamp; std::string frame_id; unsigned seq = 0; };
}

namespace geometry_msgs {
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
struct Point { double x = 0, y = 0, z = 0; };
struct Vector3 { double x = 0, y = 0, z = 0; };
struct Pose { Point position; Quaternion orientation; };
struct PoseWithCovariance { Pose pose; };
struct Twist { Vector3 linear, angular; };
struct TwistWithCovariance { Twist twist; };
}

namespace sensor_msgs {
// Payload is the decoded field arrays rather than a byte blob: x,y,z,intensity per point and an optional
// ring channel.  pcl::fromROSMsg / toROSMsg in shim_pcl.h convert to and from pcl::PointCloud.
struct PointCloud2 {
    std_msgs::Header header;
    std::vector<float> xyzi;        // 4 floats per point
    std::vector<uint16_t> ring;     // empty or one per point
    bool is_dense = true;
    typedef std::shared_ptr<const PointCloud2> ConstPtr;
};
typedef std::shared_ptr<const PointCloud2> PointCloud2ConstPtr;

struct Imu {
    std_msgs::Header header;
    geometry_msgs::Quaternion orientation;
    geometry_msgs::Vector3 angular_velocity, linear_acceleration;
    typedef std::shared_ptr<const Imu> ConstPtr;
};
}

namespace nav_msgs {
struct Odometry {
    std_msgs::Header header;
    std::string child_frame_id;
    geometry_msgs::PoseWithCovariance pose;
    geometry_msgs::TwistWithCovariance twist;
    typedef std::shared_ptr<const Odometry> ConstPtr;
};
}

namespace cloud_msgs {
// std::vector stand-in with a guard band in front of element 0.  The reference reads
// segmentedCloudColInd[ind + l] with ind + l == -1 when the stale smoothness slot points at cloud index 0
// (featureAssociation.cpp:727-732,760-767; quirk Q11): undefined behaviour that in a real run reads allocator
// metadata.  The guard value makes that read deterministic and makes the column-gap test fail, i.e. the oracle
// DEFINES the behaviour as "stop suppressing when the index leaves the array", which the product mirrors.
template <class T> struct GuardedVec {
    enum { GUARD = 8 };
    std::vector<T> v;
    GuardedVec() : v(GUARD, static_cast<T>(0x40000000)) {}
    void assign(size_t n, T x) { v.assign(GUARD + n, x); for (int g = 0; g < GUARD; ++g) v[g] = static_cast<T>(0x40000000); }
    size_t size() const { return v.size() - GUARD; }
    typename std::vector<T>::iterator begin() { return v.begin() + GUARD; }
    typename std::vector<T>::iterator end() { return v.end(); }
    typename std::vector<T>::const_iterator begin() const { return v.begin() + GUARD; }
    typename std::vector<T>::const_iterator end() const { return v.end(); }
    T& operator[](long i) { return v[GUARD + i]; }
    const T& operator[](long i) const { return v[GUARD + i]; }
};

// Field list follows cloud_msgs/msg/cloud_info.msg:1-11 (bool[] becomes a byte vector).
struct cloud_info {
    std_msgs::Header header;
    std::vector<int32_t> startRingIndex, endRingIndex;
    float startOrientation = 0, endOrientation = 0, orientationDiff = 0;
    std::vector<uint8_t> segmentedCloudGroundFlag;
    GuardedVec<uint32_t> segmentedCloudColInd;
    std::vector<float> segmentedCloudRange;
};
typedef std::shared_ptr<const cloud_info> cloud_infoConstPtr;
}

namespace tf {

struct Vector3 { double x_, y_, z_; Vector3(double x = 0, double y = 0, double z = 0) : x_(x), y_(y), z_(z) {} };

struct Quaternion {
    double x_, y_, z_, w_;
    Quaternion() : x_(0), y_(0), z_(0), w_(1) {}
    Quaternion(double x, double y, double z, double w) : x_(x), y_(y), z_(z), w_(w) {}
};

// Rotation matrix from a quaternion and fixed-axis roll/pitch/yaw extraction (the usual ZYX convention).
struct Matrix3x3 {
    double m[3][3];
    explicit Matrix3x3(const Quaternion& q) {
        double d = q.x_ * q.x_ + q.y_ * q.y_ + q.z_ * q.z_ + q.w_ * q.w_;
        double s = 2.0 / d;
        double xs = q.x_ * s, ys = q.y_ * s, zs = q.z_ * s;
        double wx = q.w_ * xs, wy = q.w_ * ys, wz = q.w_ * zs;
        double xx = q.x_ * xs, xy = q.x_ * ys, xz = q.x_ * zs;
        double yy = q.y_ * ys, yz = q.y_ * zs, zz = q.z_ * zs;
        m[0][0] = 1.0 - (yy + zz); m[0][1] = xy - wz;         m[0][2] = xz + wy;
        m[1][0] = xy + wz;         m[1][1] = 1.0 - (xx + zz); m[1][2] = yz - wx;
        m[2][0] = xz - wy;         m[2][1] = yz + wx;         m[2][2] = 1.0 - (xx + yy);
    }
    void getRPY(double& roll, double& pitch, double& yaw) const {
        if (std::fabs(m[2][0]) >= 1.0) {
            yaw = 0.0;
            double delta = std::atan2(m[2][1], m[2][2]);
            if (m[2][0] < 0) { pitch = M_PI / 2.0; roll = delta; }
            else             { pitch = -M_PI / 2.0; roll = delta; }
        } else {
            pitch = -std::asin(m[2][0]);
            double cp = std::cos(pitch);
            roll = std::atan2(m[2][1] / cp, m[2][2] / cp);
            yaw = std::atan2(m[1][0] / cp, m[0][0] / cp);
        }
    }
};

inline void quaternionMsgToTF(const geometry_msgs::Quaternion& g, Quaternion& q) { q = Quaternion(g.x, g.y, g.z, g.w); }

inline geometry_msgs::Quaternion createQuaternionMsgFromRollPitchYaw(double roll, double pitch, double yaw) {
    double hr = roll * 0.5, hp = pitch * 0.5, hy = yaw * 0.5;
    double cr = std::cos(hr), sr = std::sin(hr), cp = std::cos(hp), sp = std::sin(hp), cy = std::cos(hy), sy = std::sin(hy);
    geometry_msgs::Quaternion q;
    q.x = sr * cp * cy - cr * sp * sy;
    q.y = cr * sp * cy + sr * cp * sy;
    q.z = cr * cp * sy - sr * sp * cy;
    q.w = cr * cp * cy + sr * sp * sy;
    return q;
}

struct StampedTransform {
    ros::Time stamp_;
    std::string frame_id_, child_frame_id_;
    void setRotation(const Quaternion&) {}
    void setOrigin(const Vector3&) {}
};

struct TransformBroadcaster { void sendTransform(const StampedTransform&) {} };

}  // namespace tf
