// TEST INFRASTRUCTURE ONLY (oracle).  Flattens the last message published on a topic of THIS library's shim state into
// doubles, so that the topics of the reference's nodes (ref_harness.cpp) and of the B200 node shells (shell_harness.cpp) can
// be compared value for value from Python:
//   sensor_msgs/PointCloud2   x y z intensity per point
//   nav_msgs/Odometry         orientation x y z w | position x y z | twist angular x y z | twist linear x y z   (13 values)
//   cloud_msgs/cloud_info     startOrientation endOrientation orientationDiff | startRingIndex[N_SCAN] | endRingIndex[N_SCAN] |
//                             then per valid point (the first `nvalid` entries): ground flag, column, range
// Returns the number of values (copies min(n, cap)), or -1 when nothing was published on the topic.  *stamp, frame (header
// stamp and frame_id), *child (child_frame_id of an Odometry) and *count (messages published so far) are filled when non-null.
#pragma once
#include <cstring>
#include <string>
#include <vector>

namespace topic_dump {

inline bool is_odometry(const std::string& t) { return t == "/laser_odom_to_init" || t == "/aft_mapped_to_init" || t == "/integrated_to_init"; }
inline bool is_cloud_info(const std::string& t) { return t == "/segmented_cloud_info"; }

inline void put_str(char* dst, int cap, const std::string& s) {
    if (!dst || cap <= 0) return;
    std::strncpy(dst, s.c_str(), cap - 1); dst[cap - 1] = 0;
}

inline long dump(const char* topic_c, double* out, long cap, long nvalid, double* stamp, char* frame, char* child, int str_cap, long* count) {
    const std::string topic(topic_c);
    if (count) *count = ros::shim_count(topic);
    std::vector<double> v;
    if (is_odometry(topic)) {
        auto m = ros::shim_last<nav_msgs::Odometry>(topic);
        if (!m) return -1;
        const auto& p = m->pose.pose; const auto& t = m->twist.twist;
        v = {p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w, p.position.x, p.position.y, p.position.z,
             t.angular.x, t.angular.y, t.angular.z, t.linear.x, t.linear.y, t.linear.z};
        if (stamp) *stamp = m->header.stamp.toSec();
        put_str(frame, str_cap, m->header.frame_id); put_str(child, str_cap, m->child_frame_id);
    } else if (is_cloud_info(topic)) {
        auto m = ros::shim_last<cloud_msgs::cloud_info>(topic);
        if (!m) return -1;
        v = {m->startOrientation, m->endOrientation, m->orientationDiff};
        for (int x : m->startRingIndex) v.push_back(x);
        for (int x : m->endRingIndex) v.push_back(x);
        for (long i = 0; i < nvalid && i < (long)m->segmentedCloudRange.size(); ++i) {
            v.push_back(m->segmentedCloudGroundFlag[i] ? 1.0 : 0.0); v.push_back((double)m->segmentedCloudColInd[i]); v.push_back(m->segmentedCloudRange[i]);
        }
        if (stamp) *stamp = m->header.stamp.toSec();
        put_str(frame, str_cap, m->header.frame_id); put_str(child, str_cap, "");
    } else {
        auto m = ros::shim_last<sensor_msgs::PointCloud2>(topic);
        if (!m) return -1;
        v.assign(m->xyzi.begin(), m->xyzi.end());
        if (stamp) *stamp = m->header.stamp.toSec();
        put_str(frame, str_cap, m->header.frame_id); put_str(child, str_cap, "");
    }
    const long n = (long)v.size();
    if (out) for (long i = 0; i < n && i < cap; ++i) out[i] = v[i];
    return n;
}

}  // namespace topic_dump
