// transformFusion node shell (reference transformFusion.cpp:66-76 topic wiring, 181-241 handlers).  Compiled and driven through oracle/shim here, see common.h.
#include "common.h"

class TransformFusionNode {
public:
    TransformFusionNode() {
        const lego_params p = lego_ros::params_from_utility_h();
        if (lego_tf_create(&p, &tf_) != LEGO_OK) { ROS_FATAL("lego_tf_create: %s", lego_last_error()); ros::shutdown(); return; }
        pub_ = nh_.advertise<nav_msgs::Odometry>("/integrated_to_init", 5);
        subOdometry_ = nh_.subscribe<nav_msgs::Odometry>("/laser_odom_to_init", 5, &TransformFusionNode::laserOdometryHandler, this);
        subAftMapped_ = nh_.subscribe<nav_msgs::Odometry>("/aft_mapped_to_init", 5, &TransformFusionNode::odomAftMappedHandler, this);
    }
    ~TransformFusionNode() { if (tf_) lego_tf_destroy(tf_); }

    static void pose_of(const nav_msgs::Odometry& m, float* t) {   // the RPY decoding both handlers share (transformFusion.cpp:185-197, 220-230)
        double roll, pitch, yaw;
        const geometry_msgs::Quaternion& g = m.pose.pose.orientation;
        tf::Matrix3x3(tf::Quaternion(g.z, -g.x, -g.y, g.w)).getRPY(roll, pitch, yaw);
        t[0] = -pitch; t[1] = -yaw; t[2] = roll;
        t[3] = m.pose.pose.position.x; t[4] = m.pose.pose.position.y; t[5] = m.pose.pose.position.z;
    }
    void laserOdometryHandler(const nav_msgs::Odometry::ConstPtr& m) {
        float sum[6], mapped[6], pose[7];
        pose_of(*m, sum);
        if (lego_tf_laser_odometry(tf_, sum, mapped, pose) != LEGO_OK) { ROS_ERROR("lego_tf_laser_odometry: %s", lego_last_error()); return; }
        nav_msgs::Odometry o;
        o.header.frame_id = "/camera_init"; o.child_frame_id = "/camera"; o.header.stamp = m->header.stamp;
        o.pose.pose.orientation.x = pose[0]; o.pose.pose.orientation.y = pose[1]; o.pose.pose.orientation.z = pose[2]; o.pose.pose.orientation.w = pose[3];
        o.pose.pose.position.x = pose[4]; o.pose.pose.position.y = pose[5]; o.pose.pose.position.z = pose[6];
        pub_.publish(o);
        tf::StampedTransform t; t.frame_id_ = "/camera_init"; t.child_frame_id_ = "/camera"; t.stamp_ = m->header.stamp;
        t.setRotation(tf::Quaternion(pose[0], pose[1], pose[2], pose[3])); t.setOrigin(tf::Vector3(pose[4], pose[5], pose[6]));
        broadcaster_.sendTransform(t);
    }
    void odomAftMappedHandler(const nav_msgs::Odometry::ConstPtr& m) {
        float aft[6];
        pose_of(*m, aft);
        const float bef[6] = {(float)m->twist.twist.angular.x, (float)m->twist.twist.angular.y, (float)m->twist.twist.angular.z,
                              (float)m->twist.twist.linear.x, (float)m->twist.twist.linear.y, (float)m->twist.twist.linear.z};
        lego_tf_aft_mapped(tf_, aft, bef);
    }

private:
    ros::NodeHandle nh_;
    ros::Subscriber subOdometry_, subAftMapped_;
    ros::Publisher pub_;
    tf::TransformBroadcaster broadcaster_;
    lego_tf* tf_ = nullptr;
};

int main(int argc, char** argv) {
    ros::init(argc, argv, "lego_loam");
    TransformFusionNode node;
    ROS_INFO("\033[1;32m---->\033[0m Transform Fusion (B200) Started.");
    ros::spin();
    return 0;
}
