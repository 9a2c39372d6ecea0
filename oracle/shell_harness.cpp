// TEST INFRASTRUCTURE ONLY (oracle/_ref/liblego_shells_<cfg>.so).  Compiles the four ROS1 node shells of the B200 port
// (ros/lego_loam_b200/src/*.cpp, the code a maintainer would build with catkin) against the same stand-in ROS / PCL headers the
// reference's own nodes are compiled with (oracle/shim), links them to liblego_b200.so and drives them synchronously: one
// /velodyne_points message in, every message the nodes publish handed to the callbacks subscribed to its topic, in the order
// oracle/ref_harness.cpp drives the reference's nodes (ImageProjection -> FeatureAssociation -> TransformFusion's odometry
// handler -> mapOptimization -> TransformFusion's mapping handler).  tests/test_gpu_ros_shells.py compares the topics with the
// reference nodes' (name, frame_id, stamp, payload).  ROS itself is absent from this image, so this is as close to "the
// shells run" as the environment allows; the shells' sources are compiled unmodified.
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "shim/shim_ros.h"
#include "shim/shim_pcl.h"
#include "shim/shim_cv.h"
#include "shim/shim_gtsam.h"

#define main shell_ip_main
#include SHELL_IP_CPP
#undef main
#define main shell_fa_main
#include SHELL_FA_CPP
#undef main
#define main shell_mo_main
#include SHELL_MO_CPP
#undef main
#define main shell_tf_main
#include SHELL_TF_CPP
#undef main

#include "topic_dump.h"

namespace {

struct Shells {
    ImageProjectionNode ip;
    FeatureAssociationNode fa;
    MapOptimizationNode mo;
    TransformFusionNode tf;
    std::map<std::string, long> delivered;   // messages of a topic already handed to its subscribers
};

// hands the newest message of every listed topic to its subscribers when it is newer than the last one delivered
// (queue sizes 1-5 with a consumer that keeps up: one message per topic and scan)
void pump(Shells& S, const char* const* topics, int n) {
    for (int i = 0; i < n; ++i) {
        const std::string t(topics[i]);
        const long c = ros::shim_count(t);
        if (c == S.delivered[t]) continue;
        S.delivered[t] = c;
        ros::shim_deliver(t, ros::shim_last_raw(t));
    }
}

}  // namespace

// a crash inside the shells or the library under them prints the native frames (module + offset, for addr2line) before the
// interpreter's own handler reports the Python ones
void shell_segv_handler(int sig) {
    void* frames[48];
    const int n = backtrace(frames, 48);
    static const char msg[] = "shell_harness: SIGSEGV, native frames:\n";
    if (write(2, msg, sizeof(msg) - 1) < 0) {}
    backtrace_symbols_fd(frames, n, 2);
    signal(sig, SIG_DFL);
    raise(sig);
}

extern "C" {

void* shell_create(int subscribe_all) {
    if (std::getenv("LEGO_SHELL_BACKTRACE")) signal(SIGSEGV, shell_segv_handler);
    // one set of shells per process at a time: the subscriber table of the stand-in ROS is global to this library, and the
    // callbacks of an earlier, destroyed set would dangle
    ros::ShimState::get() = ros::ShimState();
    ros::ShimState::get().all_subscribed = subscribe_all != 0;
    ros::ShimState::get().errors = 0;
    Shells* s = new Shells();
    if (ros::ShimState::get().shutdown_requested || ros::ShimState::get().errors) { delete s; return nullptr; }
    return s;
}
void shell_destroy(void* h) { delete static_cast<Shells*>(h); ros::ShimState::get().callbacks.clear(); }
int shell_errors() { return ros::ShimState::get().errors; }

// One lidar scan as the driver's /velodyne_points message.  Returns bit0 = delivered, bit3 = /aft_mapped_to_init was published
// during this scan (the mapping body ran), -1 when a node asked for shutdown.
int shell_feed_scan(void* h, const float* xyzi, const uint16_t* ring, int P, int is_dense, double stamp) {
    Shells& S = *static_cast<Shells*>(h);
    std::shared_ptr<sensor_msgs::PointCloud2> msg = std::make_shared<sensor_msgs::PointCloud2>();
    msg->header.stamp = ros::Time(stamp);
    msg->xyzi.assign(xyzi, xyzi + 4 * static_cast<size_t>(P));
    if (ring) msg->ring.assign(ring, ring + P);
    msg->is_dense = is_dense != 0;
    ros::ShimState::get().shutdown_requested = false;
    const long aft_before = ros::shim_count("/aft_mapped_to_init");
    ros::shim_deliver(pointCloudTopic, msg);                        // imageProjection's subscriber (utility.h:53)
    if (ros::ShimState::get().shutdown_requested) return -1;
    static const char* const toFa[] = {"/segmented_cloud", "/segmented_cloud_info", "/outlier_cloud"};
    pump(S, toFa, 3);
    S.fa.run();                                                    // the 200 Hz loop of the node's main()
    static const char* const odom[] = {"/laser_odom_to_init"};      // subscribed by mapOptimization AND transformFusion
    pump(S, odom, 1);
    static const char* const toMo[] = {"/laser_cloud_corner_last", "/laser_cloud_surf_last", "/outlier_cloud_last"};
    pump(S, toMo, 3);
    S.mo.run();
    static const char* const aft[] = {"/aft_mapped_to_init"};
    pump(S, aft, 1);
    return 1 | (ros::shim_count("/aft_mapped_to_init") != aft_before ? 8 : 0);
}

void shell_feed_imu(void* h, double stamp, const double* q_xyzw, const double* w, const double* a) {
    (void)h;
    std::shared_ptr<sensor_msgs::Imu> m = std::make_shared<sensor_msgs::Imu>();
    m->header.stamp = ros::Time(stamp);
    m->orientation.x = q_xyzw[0]; m->orientation.y = q_xyzw[1]; m->orientation.z = q_xyzw[2]; m->orientation.w = q_xyzw[3];
    m->angular_velocity.x = w[0]; m->angular_velocity.y = w[1]; m->angular_velocity.z = w[2];
    m->linear_acceleration.x = a[0]; m->linear_acceleration.y = a[1]; m->linear_acceleration.z = a[2];
    ros::shim_deliver(imuTopic, m);
}

long shell_topic_dump(const char* topic, double* out, long cap, long nvalid, double* stamp, char* frame, char* child, int str_cap, long* count) {
    return topic_dump::dump(topic, out, cap, nvalid, stamp, frame, child, str_cap, count);
}

// the global map as the node serves it (/laser_cloud_surround) -- publishes when "subscribed"
void shell_publish_global_map(void* h) { static_cast<Shells*>(h)->mo.publishGlobalMap(); }

}  // extern "C"
