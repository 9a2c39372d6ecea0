// imageProjection node shell: /velodyne_points -> /segmented_cloud, /segmented_cloud_info, /outlier_cloud
// (reference imageProjection.cpp:86-107 topic wiring, 181-197 cloudHandler, 463-478 publishCloud).  Compiled and driven through oracle/shim here, see common.h.
#include "common.h"

// PointXYZIR (x y z intensity + uint16 ring) comes from the reference's utility.h:153-165, which common.h includes.

class ImageProjectionNode {
public:
    ImageProjectionNode() {
        const lego_params p = lego_ros::params_from_utility_h();
        if (lego_ip_create(&p, &ip_) != LEGO_OK) { ROS_FATAL("lego_ip_create: %s", lego_last_error()); ros::shutdown(); return; }
        sub_ = nh_.subscribe<sensor_msgs::PointCloud2>(pointCloudTopic, 1, &ImageProjectionNode::cloudHandler, this);
        pubSegmented_ = nh_.advertise<sensor_msgs::PointCloud2>("/segmented_cloud", 1);
        pubInfo_ = nh_.advertise<cloud_msgs::cloud_info>("/segmented_cloud_info", 1);
        pubOutlier_ = nh_.advertise<sensor_msgs::PointCloud2>("/outlier_cloud", 1);
        info_.startRingIndex.assign(N_SCAN, 0); info_.endRingIndex.assign(N_SCAN, 0);
        info_.segmentedCloudGroundFlag.assign(N_SCAN * Horizon_SCAN, false);
        info_.segmentedCloudColInd.assign(N_SCAN * Horizon_SCAN, 0);
        info_.segmentedCloudRange.assign(N_SCAN * Horizon_SCAN, 0);
    }
    ~ImageProjectionNode() { if (ip_) lego_ip_destroy(ip_); }

    void cloudHandler(const sensor_msgs::PointCloud2ConstPtr& msg) {
        pcl::PointCloud<PointXYZIR> in;
        pcl::fromROSMsg(*msg, in);
        xyzi_.resize(4 * in.points.size()); ring_.resize(in.points.size());
        for (size_t i = 0; i < in.points.size(); ++i) {
            xyzi_[4 * i] = in.points[i].x; xyzi_[4 * i + 1] = in.points[i].y; xyzi_[4 * i + 2] = in.points[i].z; xyzi_[4 * i + 3] = in.points[i].intensity;
            ring_[i] = in.points[i].ring;
        }
        lego_cloud_info_view info; lego_cloud_view seg, outl;
        const int rc = lego_ip_process(ip_, xyzi_.data(), useCloudRing ? ring_.data() : nullptr, static_cast<int>(in.points.size()),
                                       msg->is_dense, msg->header.stamp.toSec(), &info, &seg, &outl);
        if (rc == LEGO_ERR_NOT_DENSE) { ROS_ERROR("Point cloud is not in dense format, please remove NaN points first!"); ros::shutdown(); return; }
        if (rc != LEGO_OK) { ROS_ERROR("lego_ip_process: %s", lego_last_error()); return; }
        info_.header = msg->header;
        std::copy(info.startRingIndex, info.startRingIndex + N_SCAN, info_.startRingIndex.begin());
        std::copy(info.endRingIndex, info.endRingIndex + N_SCAN, info_.endRingIndex.begin());
        info_.startOrientation = info.startOrientation; info_.endOrientation = info.endOrientation; info_.orientationDiff = info.orientationDiff;
        for (int i = 0; i < info.count; ++i) {
            info_.segmentedCloudGroundFlag[i] = info.segmentedCloudGroundFlag[i] != 0;
            info_.segmentedCloudColInd[i] = info.segmentedCloudColInd[i];
            info_.segmentedCloudRange[i] = info.segmentedCloudRange[i];
        }
        pcl::PointCloud<pcl::PointXYZI> c;
        lego_ros::unpack(seg, c); lego_ros::publish_cloud(pubSegmented_, c, msg->header.stamp, "base_link");
        lego_ros::unpack(outl, c); lego_ros::publish_cloud(pubOutlier_, c, msg->header.stamp, "base_link");
        pubInfo_.publish(info_);
    }

private:
    ros::NodeHandle nh_;
    ros::Subscriber sub_;
    ros::Publisher pubSegmented_, pubInfo_, pubOutlier_;
    lego_ip* ip_ = nullptr;
    cloud_msgs::cloud_info info_;
    std::vector<float> xyzi_;
    std::vector<uint16_t> ring_;
};

int main(int argc, char** argv) {
    ros::init(argc, argv, "lego_loam");
    ImageProjectionNode node;
    ROS_INFO("\033[1;32m---->\033[0m Image Projection (B200) Started.");
    ros::spin();
    return 0;
}
