// featureAssociation node shell (reference featureAssociation.cpp:186-206 topic wiring, 461-489 handlers, 1817-1860
// runFeatureAssociation, 1727-1744 publishOdometry, 1796-1813 cloud publication).  Compiled and driven through oracle/shim here, see common.h.
#include "common.h"

class FeatureAssociationNode {
public:
    FeatureAssociationNode() {
        const lego_params p = lego_ros::params_from_utility_h();
        if (lego_fa_create(&p, &fa_) != LEGO_OK) { ROS_FATAL("lego_fa_create: %s", lego_last_error()); ros::shutdown(); return; }
        subCloud_ = nh_.subscribe<sensor_msgs::PointCloud2>("/segmented_cloud", 1, &FeatureAssociationNode::cloudHandler, this);
        subInfo_ = nh_.subscribe<cloud_msgs::cloud_info>("/segmented_cloud_info", 1, &FeatureAssociationNode::infoHandler, this);
        subOutlier_ = nh_.subscribe<sensor_msgs::PointCloud2>("/outlier_cloud", 1, &FeatureAssociationNode::outlierHandler, this);
        subImu_ = nh_.subscribe<sensor_msgs::Imu>(imuTopic, 50, &FeatureAssociationNode::imuHandler, this);
        pubSharp_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_sharp", 1);
        pubLessSharp_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_less_sharp", 1);
        pubFlat_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_flat", 1);
        pubLessFlat_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_less_flat", 1);
        pubCornerLast_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_corner_last", 2);
        pubSurfLast_ = nh_.advertise<sensor_msgs::PointCloud2>("/laser_cloud_surf_last", 2);
        pubOutlierLast_ = nh_.advertise<sensor_msgs::PointCloud2>("/outlier_cloud_last", 2);
        pubOdometry_ = nh_.advertise<nav_msgs::Odometry>("/laser_odom_to_init", 5);
    }
    ~FeatureAssociationNode() { if (fa_) lego_fa_destroy(fa_); }

    void cloudHandler(const sensor_msgs::PointCloud2ConstPtr& m) {
        header_ = m->header; tCloud_ = m->header.stamp.toSec();
        pcl::PointCloud<pcl::PointXYZI> c; pcl::fromROSMsg(*m, c); lego_ros::pack(c, seg_); newCloud_ = true;
    }
    void outlierHandler(const sensor_msgs::PointCloud2ConstPtr& m) {
        tOutlier_ = m->header.stamp.toSec();
        pcl::PointCloud<pcl::PointXYZI> c; pcl::fromROSMsg(*m, c); lego_ros::pack(c, outlier_); newOutlier_ = true;
    }
    void infoHandler(const cloud_msgs::cloud_infoConstPtr& m) { tInfo_ = m->header.stamp.toSec(); info_ = *m; newInfo_ = true; }
    void imuHandler(const sensor_msgs::Imu::ConstPtr& m) {
        double q[4], w[3], a[3]; lego_ros::imu_arrays(*m, q, w, a);
        lego_fa_imu(fa_, m->header.stamp.toSec(), q, w, a);
    }

    void run() {
        // stamp gate of featureAssociation.cpp:1820-1829
        if (!(newCloud_ && newInfo_ && newOutlier_ && std::abs(tInfo_ - tCloud_) < 0.05 && std::abs(tOutlier_ - tCloud_) < 0.05)) return;
        newCloud_ = newInfo_ = newOutlier_ = false;
        const int S = static_cast<int>(seg_.size() / 4);
        ground_.resize(S); col_.assign(info_.segmentedCloudColInd.begin(), info_.segmentedCloudColInd.begin() + S);
        for (int i = 0; i < S; ++i) ground_[i] = info_.segmentedCloudGroundFlag[i] ? 1 : 0;
        lego_cloud_info_view iv{};
        iv.startRingIndex = info_.startRingIndex.data(); iv.endRingIndex = info_.endRingIndex.data();
        iv.startOrientation = info_.startOrientation; iv.endOrientation = info_.endOrientation; iv.orientationDiff = info_.orientationDiff;
        iv.segmentedCloudGroundFlag = ground_.data(); iv.segmentedCloudColInd = col_.data(); iv.segmentedCloudRange = info_.segmentedCloudRange.data();
        iv.count = S; iv.on_device = 0;
        const lego_cloud_view seg = lego_ros::view_of(seg_), outl = lego_ros::view_of(outlier_);
        lego_fa_outputs out; float sum[6];
        if (lego_fa_process(fa_, &iv, &seg, &outl, tCloud_, sum, &out) != LEGO_OK) { ROS_ERROR("lego_fa_process: %s", lego_last_error()); return; }
        pcl::PointCloud<pcl::PointXYZI> c;
        if (pubSharp_.getNumSubscribers()) { lego_ros::unpack(out.cornerPointsSharp, c); lego_ros::publish_cloud(pubSharp_, c, header_.stamp, "/camera"); }
        if (pubLessSharp_.getNumSubscribers()) { lego_ros::unpack(out.cornerPointsLessSharp, c); lego_ros::publish_cloud(pubLessSharp_, c, header_.stamp, "/camera"); }
        if (pubFlat_.getNumSubscribers()) { lego_ros::unpack(out.surfPointsFlat, c); lego_ros::publish_cloud(pubFlat_, c, header_.stamp, "/camera"); }
        if (pubLessFlat_.getNumSubscribers()) { lego_ros::unpack(out.surfPointsLessFlat, c); lego_ros::publish_cloud(pubLessFlat_, c, header_.stamp, "/camera"); }
        if (out.odometry_valid) {   // publishOdometry: roll/pitch/yaw = (rz, -rx, -ry), then the axis shuffle of the quaternion
            const geometry_msgs::Quaternion g = tf::createQuaternionMsgFromRollPitchYaw(sum[2], -sum[0], -sum[1]);
            nav_msgs::Odometry o;
            o.header.frame_id = "/camera_init"; o.child_frame_id = "/laser_odom"; o.header.stamp = header_.stamp;
            o.pose.pose.orientation.x = -g.y; o.pose.pose.orientation.y = -g.z; o.pose.pose.orientation.z = g.x; o.pose.pose.orientation.w = g.w;
            o.pose.pose.position.x = sum[3]; o.pose.pose.position.y = sum[4]; o.pose.pose.position.z = sum[5];
            pubOdometry_.publish(o);
            tf::StampedTransform t; t.frame_id_ = "/camera_init"; t.child_frame_id_ = "/laser_odom"; t.stamp_ = header_.stamp;
            t.setRotation(tf::Quaternion(-g.y, -g.z, g.x, g.w)); t.setOrigin(tf::Vector3(sum[3], sum[4], sum[5]));
            broadcaster_.sendTransform(t);
        }
        // publishCloudsLast forwards the three clouds every second scan (featureAssociation.cpp:1790-1813); the initialisation scan
        // publishes corner and surf once from checkSystemInitialization (featureAssociation.cpp:1620-1630), no outliers, no odometry
        if (out.published) { lego_ros::unpack(out.outlierCloudLast, c); lego_ros::publish_cloud(pubOutlierLast_, c, header_.stamp, "/camera"); }
        if (out.published || !out.odometry_valid) {
            lego_ros::unpack(out.laserCloudCornerLast, c); lego_ros::publish_cloud(pubCornerLast_, c, header_.stamp, "/camera");
            lego_ros::unpack(out.laserCloudSurfLast, c); lego_ros::publish_cloud(pubSurfLast_, c, header_.stamp, "/camera");
        }
    }

private:
    ros::NodeHandle nh_;
    ros::Subscriber subCloud_, subInfo_, subOutlier_, subImu_;
    ros::Publisher pubSharp_, pubLessSharp_, pubFlat_, pubLessFlat_, pubCornerLast_, pubSurfLast_, pubOutlierLast_, pubOdometry_;
    tf::TransformBroadcaster broadcaster_;
    lego_fa* fa_ = nullptr;
    std_msgs::Header header_;
    cloud_msgs::cloud_info info_;
    std::vector<float> seg_, outlier_;
    std::vector<uint8_t> ground_;
    std::vector<uint32_t> col_;
    double tCloud_ = 0, tInfo_ = 0, tOutlier_ = 0;
    bool newCloud_ = false, newInfo_ = false, newOutlier_ = false;
};

int main(int argc, char** argv) {
    ros::init(argc, argv, "lego_loam");
    FeatureAssociationNode node;
    ROS_INFO("\033[1;32m---->\033[0m Feature Association (B200) Started.");
    ros::Rate rate(200);   // featureAssociation.cpp:1874
    while (ros::ok()) { ros::spinOnce(); node.run(); rate.sleep(); }
    return 0;
}
