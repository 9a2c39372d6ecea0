// mapOptmization node shell (reference mapOptmization.cpp:226-266 topic wiring, 608-652 handlers, 1487-1522 run, 654-673
// publishTF, 724-800 global map).  Loop closure, when utility.h:104 enables it, runs inside lego_mo_process (no separate
// thread).  The global map the reference serves from visualizeGlobalMapThread is served from the node's own loop (a
// lego_mo handle belongs to one thread): /laser_cloud_surround every 5 s while someone listens, and the four PCD files
// of mapOptmization.cpp:729-755 at shutdown, all through lego_mo_global_map / lego_write_pcd_ascii.
// Compiled and driven through oracle/shim here, see common.h.
#include "common.h"

class MapOptimizationNode {
public:
    MapOptimizationNode() {
        const lego_params p = lego_ros::params_from_utility_h();
        if (lego_mo_create(&p, &mo_) != LEGO_OK) { ROS_FATAL("lego_mo_create: %s", lego_last_error()); ros::shutdown(); return; }
        subCorner_ = nh_.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_corner_last", 2, &MapOptimizationNode::cornerHandler, this);
        subSurf_ = nh_.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_surf_last", 2, &MapOptimizationNode::surfHandler, this);
        subOutlier_ = nh_.subscribe<sensor_msgs::PointCloud2>("/outlier_cloud_last", 2, &MapOptimizationNode::outlierHandler, this);
        subOdometry_ = nh_.subscribe<nav_msgs::Odometry>("/laser_odom_to_init", 5, &MapOptimizationNode::odometryHandler, this);
        subImu_ = nh_.subscribe<sensor_msgs::Imu>(imuTopic, 50, &MapOptimizationNode::imuHandler, this);
        pubAftMapped_ = nh_.advertise<nav_msgs::Odometry>("/aft_mapped_to_init", 5);
        pubSurround_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_surround", 2);
    }
    ~MapOptimizationNode() { if (mo_) lego_mo_destroy(mo_); }

    void cornerHandler(const sensor_msgs::PointCloud2ConstPtr& m) { tCorner_ = m->header.stamp.toSec(); take(m, corner_); newCorner_ = true; }
    void surfHandler(const sensor_msgs::PointCloud2ConstPtr& m) { tSurf_ = m->header.stamp.toSec(); take(m, surf_); newSurf_ = true; }
    void outlierHandler(const sensor_msgs::PointCloud2ConstPtr& m) { tOutlier_ = m->header.stamp.toSec(); take(m, outlier_); newOutlier_ = true; }
    void odometryHandler(const nav_msgs::Odometry::ConstPtr& m) {
        tOdometry_ = m->header.stamp.toSec();
        double roll, pitch, yaw;
        const geometry_msgs::Quaternion& g = m->pose.pose.orientation;
        tf::Matrix3x3(tf::Quaternion(g.z, -g.x, -g.y, g.w)).getRPY(roll, pitch, yaw);
        // lego_mo_process repeats this message round trip on its input; on values that already went through it the
        // repetition changes nothing beyond float rounding
        sum_[0] = -pitch; sum_[1] = -yaw; sum_[2] = roll;
        sum_[3] = m->pose.pose.position.x; sum_[4] = m->pose.pose.position.y; sum_[5] = m->pose.pose.position.z;
        newOdometry_ = true;
    }
    void imuHandler(const sensor_msgs::Imu::ConstPtr& m) {
        double q[4], w[3], a[3]; lego_ros::imu_arrays(*m, q, w, a);
        lego_mo_imu(mo_, m->header.stamp.toSec(), q, w, a);
    }

    void run() {
        // first gate of mapOptmization.cpp:1489-1493; the mappingProcessInterval gate (1499) is inside the library
        if (!(newCorner_ && std::abs(tCorner_ - tOdometry_) < 0.005 && newSurf_ && std::abs(tSurf_ - tOdometry_) < 0.005 &&
              newOutlier_ && std::abs(tOutlier_ - tOdometry_) < 0.005 && newOdometry_)) return;
        newCorner_ = newSurf_ = newOutlier_ = newOdometry_ = false;
        const lego_cloud_view c = lego_ros::view_of(corner_), s = lego_ros::view_of(surf_), o = lego_ros::view_of(outlier_);
        float aft[6], bef[6]; int processed = 0;
        if (lego_mo_process(mo_, &c, &s, &o, sum_, tOdometry_, aft, bef, &processed) != LEGO_OK) { ROS_ERROR("lego_mo_process: %s", lego_last_error()); return; }
        if (!processed) return;
        // publishTF (mapOptmization.cpp:654-673): pose = transformAftMapped, twist = transformBefMapped
        const geometry_msgs::Quaternion g = tf::createQuaternionMsgFromRollPitchYaw(aft[2], -aft[0], -aft[1]);
        nav_msgs::Odometry m;
        m.header.frame_id = "/camera_init"; m.child_frame_id = "/aft_mapped"; m.header.stamp = ros::Time().fromSec(tOdometry_);
        m.pose.pose.orientation.x = -g.y; m.pose.pose.orientation.y = -g.z; m.pose.pose.orientation.z = g.x; m.pose.pose.orientation.w = g.w;
        m.pose.pose.position.x = aft[3]; m.pose.pose.position.y = aft[4]; m.pose.pose.position.z = aft[5];
        m.twist.twist.angular.x = bef[0]; m.twist.twist.angular.y = bef[1]; m.twist.twist.angular.z = bef[2];
        m.twist.twist.linear.x = bef[3]; m.twist.twist.linear.y = bef[4]; m.twist.twist.linear.z = bef[5];
        pubAftMapped_.publish(m);
        tf::StampedTransform t; t.frame_id_ = "/camera_init"; t.child_frame_id_ = "/aft_mapped"; t.stamp_ = m.header.stamp;
        t.setRotation(tf::Quaternion(-g.y, -g.z, g.x, g.w)); t.setOrigin(tf::Vector3(aft[3], aft[4], aft[5]));
        broadcaster_.sendTransform(t);
    }

    // publishGlobalMap (mapOptmization.cpp:757-800)
    void publishGlobalMap() {
        if (pubSurround_.getNumSubscribers() == 0) return;
        std::vector<float> xyzi;
        if (!fetch(LEGO_MAP_GLOBAL, xyzi) || xyzi.empty()) return;
        pcl::PointCloud<pcl::PointXYZI> c; lego_ros::unpack(lego_ros::view_of(xyzi), c);
        lego_ros::publish_cloud(pubSurround_, c, ros::Time().fromSec(tOdometry_), "/camera_init");
    }
    // the tail of visualizeGlobalMapThread (mapOptmization.cpp:729-755)
    void saveMaps() {
        static const struct { int which; const char* file; } kFiles[4] = {{LEGO_MAP_GLOBAL, "finalCloud.pcd"}, {LEGO_MAP_CORNER, "cornerMap.pcd"},
                                                                          {LEGO_MAP_SURFACE, "surfaceMap.pcd"}, {LEGO_MAP_TRAJECTORY, "trajectory.pcd"}};
        for (const auto& f : kFiles) {
            std::vector<float> xyzi;
            if (!fetch(f.which, xyzi)) continue;
            const std::string path = fileDirectory + f.file;
            if (lego_write_pcd_ascii(path.c_str(), xyzi.data(), (long)(xyzi.size() / 4)) != LEGO_OK) ROS_ERROR("%s", lego_last_error());
        }
    }

private:
    bool fetch(int which, std::vector<float>& xyzi) {
        long n = 0;
        if (lego_mo_global_map(mo_, which, nullptr, 0, &n) != LEGO_OK) { ROS_ERROR("lego_mo_global_map: %s", lego_last_error()); return false; }
        xyzi.resize((size_t)n * 4);
        return n == 0 || lego_mo_global_map(mo_, which, xyzi.data(), n, &n) == LEGO_OK;
    }
    static void take(const sensor_msgs::PointCloud2ConstPtr& m, std::vector<float>& xyzi) {
        pcl::PointCloud<pcl::PointXYZI> c; pcl::fromROSMsg(*m, c); lego_ros::pack(c, xyzi);
    }
    ros::NodeHandle nh_;
    ros::Subscriber subCorner_, subSurf_, subOutlier_, subOdometry_, subImu_;
    ros::Publisher pubAftMapped_, pubSurround_;
    tf::TransformBroadcaster broadcaster_;
    lego_mo* mo_ = nullptr;
    std::vector<float> corner_, surf_, outlier_;
    float sum_[6] = {0, 0, 0, 0, 0, 0};
    double tCorner_ = 0, tSurf_ = 0, tOutlier_ = 0, tOdometry_ = 0;
    bool newCorner_ = false, newSurf_ = false, newOutlier_ = false, newOdometry_ = false;
};

int main(int argc, char** argv) {
    ros::init(argc, argv, "lego_loam");
    MapOptimizationNode node;
    ROS_INFO("\033[1;32m---->\033[0m Map Optimization (B200) Started.");
    ros::Rate rate(200);   // mapOptmization.cpp:1537
    ros::Time lastMap = ros::Time::now();
    while (ros::ok()) {
        ros::spinOnce(); node.run();
        if ((ros::Time::now() - lastMap).toSec() >= 5.0) { lastMap = ros::Time::now(); node.publishGlobalMap(); }   // ros::Rate(0.2), mapOptmization.cpp:725
        rate.sleep();
    }
    node.saveMaps();
    return 0;
}
