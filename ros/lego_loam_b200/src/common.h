// Shared helpers of the four ROS1 node shells.  The shells keep the reference's topic contract (names,
// types, queue sizes, frame ids: SURVEY.md 8b) and hand the numerical work to liblego_b200.so through include/lego_b200.h.
// ROS1 / catkin / PCL are absent from this repository's build image, so a catkin build has never run.  The sources ARE compiled
// here, unmodified, against the stand-in ROS / PCL headers the reference's own nodes are compiled with (oracle/shim,
// oracle/shell_harness.cpp -> oracle/_ref/liblego_shells_<cfg>.so) and driven message by message next to the reference's nodes:
// tests/test_gpu_ros_shells.py compares every topic (messages published, frame ids, stamps, payloads).
#pragma once
#include <ros/ros.h>
#include <nav_msgs/Odometry.h>
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include <pcl_conversions/pcl_conversions.h>
#include <sensor_msgs/Imu.h>
#include <sensor_msgs/PointCloud2.h>
#include <tf/transform_broadcaster.h>
#include <tf/transform_datatypes.h>
#include <vector>
#include "cloud_msgs/cloud_info.h"
#include "lego_b200.h"
#include "utility.h"   // the reference's own header: sensor block + algorithm constants (LeGO-LOAM/include/utility.h)

namespace lego_ros {

inline lego_params params_from_utility_h() {   // field names = utility.h names
    lego_params p{};
    p.N_SCAN = N_SCAN; p.Horizon_SCAN = Horizon_SCAN; p.ang_res_x = ang_res_x; p.ang_res_y = ang_res_y; p.ang_bottom = ang_bottom;
    p.groundScanInd = groundScanInd; p.useCloudRing = useCloudRing; p.sensorMinimumRange = sensorMinimumRange;
    p.sensorMountAngle = sensorMountAngle; p.segmentTheta = segmentTheta; p.segmentValidPointNum = segmentValidPointNum;
    p.segmentValidLineNum = segmentValidLineNum; p.segmentAlphaX = segmentAlphaX; p.segmentAlphaY = segmentAlphaY;
    p.scanPeriod = scanPeriod; p.edgeFeatureNum = edgeFeatureNum; p.surfFeatureNum = surfFeatureNum; p.sectionsTotal = sectionsTotal;
    p.edgeThreshold = edgeThreshold; p.surfThreshold = surfThreshold; p.nearestFeatureSearchSqDist = nearestFeatureSearchSqDist;
    p.surroundingKeyframeSearchRadius = surroundingKeyframeSearchRadius; p.mappingProcessInterval = mappingProcessInterval;
    p.loopClosureEnableFlag = loopClosureEnableFlag; p.surroundingKeyframeSearchNum = surroundingKeyframeSearchNum;
    p.historyKeyframeSearchRadius = historyKeyframeSearchRadius; p.historyKeyframeSearchNum = historyKeyframeSearchNum;
    p.historyKeyframeFitnessScore = historyKeyframeFitnessScore; p.device = 0;
    return p;
}

// pcl::PointXYZI is 32 bytes; the ABI takes 16-byte points
template <class Cloud> inline void pack(const Cloud& c, std::vector<float>& xyzi) {
    xyzi.resize(4 * c.points.size());
    for (size_t i = 0; i < c.points.size(); ++i) {
        xyzi[4 * i] = c.points[i].x; xyzi[4 * i + 1] = c.points[i].y; xyzi[4 * i + 2] = c.points[i].z; xyzi[4 * i + 3] = c.points[i].intensity;
    }
}
inline void unpack(const lego_cloud_view& v, pcl::PointCloud<pcl::PointXYZI>& c) {
    c.points.resize(v.count); c.width = v.count; c.height = 1; c.is_dense = true;
    for (int i = 0; i < v.count; ++i) {
        c.points[i].x = v.host[4 * i]; c.points[i].y = v.host[4 * i + 1]; c.points[i].z = v.host[4 * i + 2]; c.points[i].intensity = v.host[4 * i + 3];
    }
}
inline lego_cloud_view view_of(const std::vector<float>& xyzi) {
    lego_cloud_view v; v.dev = nullptr; v.host = xyzi.data(); v.count = static_cast<int>(xyzi.size() / 4);
    return v;
}
inline void publish_cloud(ros::Publisher& pub, const pcl::PointCloud<pcl::PointXYZI>& c, const ros::Time& stamp, const std::string& frame) {
    sensor_msgs::PointCloud2 m;
    pcl::toROSMsg(c, m);
    m.header.stamp = stamp; m.header.frame_id = frame;
    pub.publish(m);
}
inline void imu_arrays(const sensor_msgs::Imu& m, double q[4], double w[3], double a[3]) {
    q[0] = m.orientation.x; q[1] = m.orientation.y; q[2] = m.orientation.z; q[3] = m.orientation.w;
    w[0] = m.angular_velocity.x; w[1] = m.angular_velocity.y; w[2] = m.angular_velocity.z;
    a[0] = m.linear_acceleration.x; a[1] = m.linear_acceleration.y; a[2] = m.linear_acceleration.z;
}

}  // namespace lego_ros
