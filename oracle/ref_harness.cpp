// TEST INFRASTRUCTURE ONLY (oracle/_ref): builds the reference's OWN node sources, byte-for-byte unmodified and
// read where they lie under /root/reference, into a ROS-free shared library.  Only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference arm may load the result; the product never does.
//
// The four sources (the three hot nodes + transformFusion.cpp) are pulled in with #include (paths come from -DREF_IP_CPP=... etc, see oracle/Makefile)
// against the stand-in headers in oracle/shim/.  `private` is made public so the harness can drive the stage
// methods in the order the reference's own cloudHandler (imageProjection.cpp:181-197), runFeatureAssociation
// (featureAssociation.cpp:1817-1860) and run (mapOptmization.cpp:1487-1522) do, and snapshot intermediates
// that those drivers destroy (resetParameters, clearCloud).  ref_set_verbatim_drivers(1) switches to calling
// the reference's driver methods themselves; tests check both give identical poses.
#include <algorithm>
#include <array>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <deque>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <iterator>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <queue>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

// The reference reads four `new[]` work arrays before writing them (featureAssociation.cpp:210-221, quirk
// Q11).  Zero-filled array allocation makes the oracle deterministic; the product mirrors "zero at create".
void* operator new[](std::size_t n) {
    void* p = std::calloc(1, n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void operator delete[](void* p) noexcept { std::free(p); }
void operator delete[](void* p, std::size_t) noexcept { std::free(p); }

#include "shim/shim_ros.h"
#include "shim/shim_pcl.h"
#include "shim/shim_cv.h"
#include "shim/shim_gtsam.h"

#define private public
#define main ref_ip_main
#include REF_IP_CPP
#undef main
#define main ref_fa_main
#include REF_FA_CPP
#undef main
#define main ref_mo_main
#include REF_MO_CPP
#undef main
#define main ref_tf_main
#include REF_TF_CPP
#undef main
#undef private

#include "topic_dump.h"

namespace {

typedef std::chrono::steady_clock Clock;
inline double ms_since(Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count(); }

enum Stage { T_IP_COPY, T_IP_PROJECT, T_IP_GROUND, T_IP_SEGMENT, T_IP_PUBLISH, T_IP_RESET,
             T_FA_DISTORT, T_FA_SMOOTH, T_FA_OCCLUDE, T_FA_EXTRACT, T_FA_LM, T_FA_INTEGRATE, T_FA_CLOUDSLAST,
             T_MO_ASSOC, T_MO_SURROUND, T_MO_DOWNSAMPLE, T_MO_SCAN2MAP, T_MO_SAVE, T_MO_REST, T_COUNT };

struct Snap {   // named float / int snapshots of one scan
    std::map<std::string, std::vector<float> > f;
    std::map<std::string, std::vector<int32_t> > i;
};

template <class C> std::vector<float> flat(const C& cloud) {
    std::vector<float> o(4 * cloud.points.size());
    for (size_t k = 0; k < cloud.points.size(); ++k) {
        o[4 * k] = cloud.points[k].x; o[4 * k + 1] = cloud.points[k].y; o[4 * k + 2] = cloud.points[k].z; o[4 * k + 3] = cloud.points[k].intensity;
    }
    return o;
}
std::vector<float> flat_msg(const std::string& topic) {
    auto m = ros::shim_last<sensor_msgs::PointCloud2>(topic);
    return m ? m->xyzi : std::vector<float>();
}

struct Ref {
    ImageProjection ip;
    FeatureAssociation fa;
    mapOptimization mo;
    TransformFusion tf;
    Snap snap;
    bool verbatim = false;
    bool want_snap = true;
    double t[T_COUNT];
    long scans = 0, mo_calls = 0;
    // loop closure (mapOptmization.cpp:802-812 runs performLoopClosure in a 1 Hz side thread; the harness calls it
    // synchronously after a mapping call once >= 1.0 s of scan time passed since its last call)
    double timeLastLoop = -1e9;
    long loopCalls = 0, loopsClosed = 0;
    Ref() { std::memset(t, 0, sizeof(t)); }
};

}  // namespace

extern "C" {

int ref_n_scan() { return N_SCAN; }
int ref_horizon_scan() { return Horizon_SCAN; }
int ref_ground_scan_ind() { return groundScanInd; }
int ref_use_cloud_ring() { return useCloudRing ? 1 : 0; }
void ref_params(float* out8) {
    out8[0] = ang_res_x; out8[1] = ang_res_y; out8[2] = ang_bottom; out8[3] = segmentAlphaX; out8[4] = segmentAlphaY;
    out8[5] = segmentTheta; out8[6] = sensorMinimumRange; out8[7] = sensorMountAngle;
}

void* ref_create(int viz_subscribed) {
    ros::ShimState& s = ros::ShimState::get();
    s.last.clear(); s.count.clear(); s.subscribed.clear(); s.callbacks.clear(); s.errors = 0; s.shutdown_requested = false;
    s.all_subscribed = false;
    if (viz_subscribed) {   // makes FeatureAssociation::publishCloud (featureAssociation.cpp:786-817) emit the 4 feature clouds
        s.subscribed.insert("/laser_cloud_sharp"); s.subscribed.insert("/laser_cloud_less_sharp");
        s.subscribed.insert("/laser_cloud_flat"); s.subscribed.insert("/laser_cloud_less_flat");
    }
    Ref* r = new Ref();
    r->want_snap = viz_subscribed != 0;
    return r;
}
void ref_destroy(void* h) { delete static_cast<Ref*>(h); }
void ref_set_verbatim_drivers(void* h, int on) { static_cast<Ref*>(h)->verbatim = on != 0; }

}  // extern "C"

// ---------------------------------------------------------------- ImageProjection
static void run_ip(Ref& R, const sensor_msgs::PointCloud2ConstPtr& msg) {
    ImageProjection& ip = R.ip;
    if (R.verbatim) { ip.cloudHandler(msg); return; }
    Clock::time_point t0 = Clock::now();
    ip.copyPointCloud(msg);              R.t[T_IP_COPY] += ms_since(t0); t0 = Clock::now();
    ip.findStartEndAngle();
    ip.projectPointCloud();              R.t[T_IP_PROJECT] += ms_since(t0); t0 = Clock::now();
    ip.groundRemoval();                  R.t[T_IP_GROUND] += ms_since(t0); t0 = Clock::now();
    ip.cloudSegmentation();              R.t[T_IP_SEGMENT] += ms_since(t0); t0 = Clock::now();
    ip.publishCloud();                   R.t[T_IP_PUBLISH] += ms_since(t0);
    if (R.want_snap) {
        const int N = N_SCAN * Horizon_SCAN;
        Snap& s = R.snap;
        s.f["rangeMat"].assign(ip.rangeMat.f(), ip.rangeMat.f() + N);
        std::vector<int32_t>& g = s.i["groundMat"]; g.resize(N);
        for (int k = 0; k < N; ++k) g[k] = reinterpret_cast<const int8_t*>(ip.groundMat.buf.data())[k];
        const int32_t* lab = reinterpret_cast<const int32_t*>(ip.labelMat.buf.data());
        s.i["labelMat"].assign(lab, lab + N);
        s.f["fullCloud"] = flat(*ip.fullCloud);
        s.f["fullInfoCloud"] = flat(*ip.fullInfoCloud);
        s.f["segmentedCloud"] = flat(*ip.segmentedCloud);
        s.f["outlierCloud"] = flat(*ip.outlierCloud);
        s.i["startRingIndex"].assign(ip.segMsg.startRingIndex.begin(), ip.segMsg.startRingIndex.end());
        s.i["endRingIndex"].assign(ip.segMsg.endRingIndex.begin(), ip.segMsg.endRingIndex.end());
        const size_t S = ip.segmentedCloud->points.size();
        std::vector<int32_t>& gf = s.i["segGroundFlag"]; gf.resize(S);
        std::vector<int32_t>& ci = s.i["segColInd"]; ci.resize(S);
        for (size_t k = 0; k < S; ++k) { gf[k] = ip.segMsg.segmentedCloudGroundFlag[k]; ci[k] = static_cast<int32_t>(ip.segMsg.segmentedCloudColInd[k]); }
        s.f["segRange"].assign(ip.segMsg.segmentedCloudRange.begin(), ip.segMsg.segmentedCloudRange.begin() + S);
        s.f["orientation"] = {ip.segMsg.startOrientation, ip.segMsg.endOrientation, ip.segMsg.orientationDiff};
    }
    t0 = Clock::now();
    ip.resetParameters();                R.t[T_IP_RESET] += ms_since(t0);
}

// ---------------------------------------------------------------- FeatureAssociation
// returns bit0 = stage ran, bit1 = odometry LM ran (not the init scan), bit2 = "last" clouds published this scan
static int run_fa(Ref& R) {
    FeatureAssociation& fa = R.fa;
    const long before = ros::shim_count("/laser_cloud_corner_last");
    fa.laserCloudHandler(ros::shim_last<sensor_msgs::PointCloud2>("/segmented_cloud"));
    fa.laserCloudInfoHandler(ros::shim_last<cloud_msgs::cloud_info>("/segmented_cloud_info"));
    fa.outlierCloudHandler(ros::shim_last<sensor_msgs::PointCloud2>("/outlier_cloud"));
    int flags = 1;
    if (R.verbatim) {
        bool inited = fa.systemInitedLM;
        fa.runFeatureAssociation();
        if (inited) flags |= 2;
    } else {
        // gate of featureAssociation.cpp:1820-1829
        if (fa.newSegmentedCloud && fa.newSegmentedCloudInfo && fa.newOutlierCloud &&
            std::abs(fa.timeNewSegmentedCloudInfo - fa.timeNewSegmentedCloud) < 0.05 &&
            std::abs(fa.timeNewOutlierCloud - fa.timeNewSegmentedCloud) < 0.05) {
            fa.newSegmentedCloud = false; fa.newSegmentedCloudInfo = false; fa.newOutlierCloud = false;
        } else {
            return 0;
        }
        Clock::time_point t0 = Clock::now();
        fa.adjustDistortion();      R.t[T_FA_DISTORT] += ms_since(t0); t0 = Clock::now();
        fa.calculateSmoothness();   R.t[T_FA_SMOOTH] += ms_since(t0); t0 = Clock::now();
        fa.markOccludedPoints();    R.t[T_FA_OCCLUDE] += ms_since(t0); t0 = Clock::now();
        fa.extractFeatures();       R.t[T_FA_EXTRACT] += ms_since(t0);
        fa.publishCloud();
        if (R.want_snap) {
            Snap& s = R.snap;
            const size_t S = fa.segmentedCloud->points.size();
            s.f["segmentedCloudAdj"] = flat(*fa.segmentedCloud);
            s.f["cloudCurvature"].assign(fa.cloudCurvature, fa.cloudCurvature + S);
            s.i["cloudNeighborPicked"].assign(fa.cloudNeighborPicked, fa.cloudNeighborPicked + S);
            s.i["cloudLabel"].assign(fa.cloudLabel, fa.cloudLabel + S);
            std::vector<int32_t>& si = s.i["smoothnessInd"]; si.resize(S);
            std::vector<float>& sv = s.f["smoothnessVal"]; sv.resize(S);
            for (size_t k = 0; k < S; ++k) { si[k] = static_cast<int32_t>(fa.cloudSmoothness[k].ind); sv[k] = fa.cloudSmoothness[k].value; }
            s.f["cornerPointsSharp"] = flat(*fa.cornerPointsSharp);
            s.f["cornerPointsLessSharp"] = flat(*fa.cornerPointsLessSharp);
            s.f["surfPointsFlat"] = flat(*fa.surfPointsFlat);
            s.f["surfPointsLessFlat"] = flat(*fa.surfPointsLessFlat);
        }
        if (!fa.systemInitedLM) {
            fa.checkSystemInitialization();
        } else {
            flags |= 2;
            fa.updateInitialGuess();
            t0 = Clock::now();
            fa.updateTransformation();     R.t[T_FA_LM] += ms_since(t0); t0 = Clock::now();
            fa.integrateTransformation();
            fa.publishOdometry();          R.t[T_FA_INTEGRATE] += ms_since(t0); t0 = Clock::now();
            fa.publishCloudsLast();        R.t[T_FA_CLOUDSLAST] += ms_since(t0);
        }
    }
    if (R.want_snap) {
        Snap& s = R.snap;
        s.f["transformCur"].assign(fa.transformCur, fa.transformCur + 6);
        s.f["transformSum"].assign(fa.transformSum, fa.transformSum + 6);
        s.f["laserCloudCornerLast"] = flat(*fa.laserCloudCornerLast);
        s.f["laserCloudSurfLast"] = flat(*fa.laserCloudSurfLast);
        s.i["faDegenerate"] = {fa.isDegenerate ? 1 : 0};
        s.f["imuStart"] = {fa.imuRollStart, fa.imuPitchStart, fa.imuYawStart, fa.imuVeloXStart, fa.imuVeloYStart, fa.imuVeloZStart};
        s.f["imuLast"] = {fa.imuRollLast, fa.imuPitchLast, fa.imuYawLast, fa.imuVeloFromStartX, fa.imuVeloFromStartY, fa.imuVeloFromStartZ,
                          fa.imuAngularFromStartX, fa.imuAngularFromStartY, fa.imuAngularFromStartZ};
    }
    if (ros::shim_count("/laser_cloud_corner_last") != before) flags |= 4;
    return flags;
}

// ---------------------------------------------------------------- mapOptimization
// returns 1 when the mapping body ran (both gates of mapOptmization.cpp:1489-1499 passed)
static int run_mo(Ref& R, int fa_flags, long odo_before, long outl_before) {
    mapOptimization& mo = R.mo;
    if (fa_flags & 4) {
        mo.laserCloudCornerLastHandler(ros::shim_last<sensor_msgs::PointCloud2>("/laser_cloud_corner_last"));
        mo.laserCloudSurfLastHandler(ros::shim_last<sensor_msgs::PointCloud2>("/laser_cloud_surf_last"));
    }
    if (ros::shim_count("/outlier_cloud_last") != outl_before)
        mo.laserCloudOutlierLastHandler(ros::shim_last<sensor_msgs::PointCloud2>("/outlier_cloud_last"));
    if (ros::shim_count("/laser_odom_to_init") != odo_before)
        mo.laserOdometryHandler(ros::shim_last<nav_msgs::Odometry>("/laser_odom_to_init"));
    const long aft_before = ros::shim_count("/aft_mapped_to_init");
    if (R.verbatim) {
        mo.run();
    } else if (mo.newLaserCloudCornerLast && std::abs(mo.timeLaserCloudCornerLast - mo.timeLaserOdometry) < 0.005 &&
               mo.newLaserCloudSurfLast && std::abs(mo.timeLaserCloudSurfLast - mo.timeLaserOdometry) < 0.005 &&
               mo.newLaserCloudOutlierLast && std::abs(mo.timeLaserCloudOutlierLast - mo.timeLaserOdometry) < 0.005 &&
               mo.newLaserOdometry) {
        mo.newLaserCloudCornerLast = false; mo.newLaserCloudSurfLast = false; mo.newLaserCloudOutlierLast = false; mo.newLaserOdometry = false;
        if (mo.timeLaserOdometry - mo.timeLastProcessing >= mappingProcessInterval) {
            mo.timeLastProcessing = mo.timeLaserOdometry;
            Clock::time_point t0 = Clock::now();
            mo.transformAssociateToMap();       R.t[T_MO_ASSOC] += ms_since(t0); t0 = Clock::now();
            mo.extractSurroundingKeyFrames();   R.t[T_MO_SURROUND] += ms_since(t0); t0 = Clock::now();
            mo.downsampleCurrentScan();         R.t[T_MO_DOWNSAMPLE] += ms_since(t0);
            if (R.want_snap) {
                Snap& s = R.snap;
                s.f["moTransformSum"].assign(mo.transformSum, mo.transformSum + 6);
                s.f["moTransformInit"].assign(mo.transformTobeMapped, mo.transformTobeMapped + 6);
                s.f["laserCloudCornerFromMapDS"] = flat(*mo.laserCloudCornerFromMapDS);
                s.f["laserCloudSurfFromMapDS"] = flat(*mo.laserCloudSurfFromMapDS);
                s.f["laserCloudCornerLastDS"] = flat(*mo.laserCloudCornerLastDS);
                s.f["laserCloudSurfLastDS"] = flat(*mo.laserCloudSurfLastDS);
                s.f["laserCloudOutlierLastDS"] = flat(*mo.laserCloudOutlierLastDS);
                s.f["laserCloudSurfTotalLastDS"] = flat(*mo.laserCloudSurfTotalLastDS);
                s.i["surroundingExistingKeyPosesID"].assign(mo.surroundingExistingKeyPosesID.begin(), mo.surroundingExistingKeyPosesID.end());
            }
            t0 = Clock::now();
            if (R.want_snap) {
                // scan2MapOptimization (mapOptmization.cpp:1329-1350) spelled out with the reference's own methods so that the
                // pose after every LM iteration and before transformUpdate's IMU blend can be recorded
                std::vector<float>& tr = R.snap.f["moIterations"]; tr.clear();
                if (mo.laserCloudCornerFromMapDSNum > 10 && mo.laserCloudSurfFromMapDSNum > 100) {
                    mo.kdtreeCornerFromMap->setInputCloud(mo.laserCloudCornerFromMapDS);
                    mo.kdtreeSurfFromMap->setInputCloud(mo.laserCloudSurfFromMapDS);
                    for (int iterCount = 0; iterCount < 10; iterCount++) {
                        mo.laserCloudOri->clear(); mo.coeffSel->clear();
                        mo.cornerOptimization(iterCount);
                        mo.surfOptimization(iterCount);
                        const bool done = mo.LMOptimization(iterCount);
                        tr.insert(tr.end(), mo.transformTobeMapped, mo.transformTobeMapped + 6);
                        tr.push_back((float)mo.laserCloudOri->points.size());
                        if (done) break;
                    }
                    mo.transformUpdate();
                }
            } else mo.scan2MapOptimization();
            R.t[T_MO_SCAN2MAP] += ms_since(t0); t0 = Clock::now();
            if (R.want_snap) R.snap.f["moTransformOpt"].assign(mo.transformTobeMapped, mo.transformTobeMapped + 6);
            mo.saveKeyFramesAndFactor();        R.t[T_MO_SAVE] += ms_since(t0); t0 = Clock::now();
            mo.correctPoses();
            mo.publishTF();
            mo.publishKeyPosesAndFrames();
            mo.clearCloud();                    R.t[T_MO_REST] += ms_since(t0);
        }
    }
    int ran = ros::shim_count("/aft_mapped_to_init") != aft_before;
    if (ran && loopClosureEnableFlag && mo.timeLaserOdometry - R.timeLastLoop >= 1.0) {
        R.timeLastLoop = mo.timeLaserOdometry;
        const long icpBefore = pcl::shim_icp_log().runs;
        const bool closedBefore = mo.aLoopIsClosed;
        mo.aLoopIsClosed = false;
        mo.performLoopClosure();
        R.loopCalls += 1;
        const bool closedNow = mo.aLoopIsClosed;
        if (closedNow) R.loopsClosed += 1;
        mo.aLoopIsClosed = closedBefore || closedNow;
        if (R.want_snap) {
            Snap& s = R.snap;
            const pcl::ShimIcpLog& l = pcl::shim_icp_log();
            const int icpRan = l.runs != icpBefore;
            s.i["loopClosure"] = {1, icpRan, closedNow ? 1 : 0, icpRan ? mo.closestHistoryFrameID : -1, icpRan ? mo.latestFrameIDLoopCloure : -1,
                                  icpRan ? (l.converged ? 1 : 0) : 0, icpRan ? l.iterations : 0, icpRan ? (int32_t)l.ns : 0, icpRan ? (int32_t)l.nt : 0};
            if (icpRan) {
                s.f["icpFitness"] = {(float)l.fitness};
                s.f["icpTransform"].assign(l.T, l.T + 16);
                s.f["latestSurfKeyFrameCloud"] = flat(*mo.latestSurfKeyFrameCloud);
                s.f["nearHistorySurfKeyFrameCloudDS"] = flat(*mo.nearHistorySurfKeyFrameCloudDS);
            }
        }
    }
    if (R.want_snap) {
        Snap& s = R.snap;
        s.f["transformAftMapped"].assign(mo.transformAftMapped, mo.transformAftMapped + 6);
        s.f["transformBefMapped"].assign(mo.transformBefMapped, mo.transformBefMapped + 6);
        s.f["transformTobeMapped"].assign(mo.transformTobeMapped, mo.transformTobeMapped + 6);
        s.i["numKeyFrames"] = {static_cast<int32_t>(mo.cloudKeyPoses3D->points.size())};
        s.i["moDegenerate"] = {mo.isDegenerate ? 1 : 0};
        s.i["loopStats"] = {(int32_t)R.loopCalls, (int32_t)R.loopsClosed};
        std::vector<float>& kp = s.f["cloudKeyPoses6D"];
        kp.clear();
        for (const auto& p : mo.cloudKeyPoses6D->points) { kp.push_back(p.x); kp.push_back(p.y); kp.push_back(p.z); kp.push_back(p.roll); kp.push_back(p.pitch); kp.push_back(p.yaw); }
    }
    if (ran) R.mo_calls += 1;
    return ran;
}

extern "C" {

// One lidar scan through IP -> FA -> MO (-> TransformFusion).  stages: bit0 IP, bit1 FA, bit2 MO, bit3 TransformFusion.
// Returns bit0 = FA ran, bit1 = odometry LM ran, bit2 = clouds forwarded to mapping, bit3 = mapping body ran, bit4 = fused pose published,
// or -1 on the reference's hard failure (non-dense cloud with ring channel, imageProjection.cpp:172-177).
int ref_feed_scan(void* h, const float* xyzi, const uint16_t* ring, int P, int is_dense, double stamp, int stages) {
    Ref& R = *static_cast<Ref*>(h);
    std::shared_ptr<sensor_msgs::PointCloud2> msg = std::make_shared<sensor_msgs::PointCloud2>();
    msg->header.stamp = ros::Time(stamp);
    msg->xyzi.assign(xyzi, xyzi + 4 * static_cast<size_t>(P));
    if (ring) msg->ring.assign(ring, ring + P);
    msg->is_dense = is_dense != 0;
    R.snap.f.clear(); R.snap.i.clear();
    ros::ShimState::get().shutdown_requested = false;
    run_ip(R, msg);
    R.scans += 1;
    if (ros::ShimState::get().shutdown_requested) return -1;
    int out = 0;
    if (stages & 2) {
        const long odo_before = ros::shim_count("/laser_odom_to_init"), outl_before = ros::shim_count("/outlier_cloud_last");
        int f = run_fa(R);
        out |= f & 7;
        // TransformFusion (transformFusion.cpp:181-216): every /laser_odom_to_init message is fused with the last
        // /aft_mapped_to_init it has seen, i.e. the mapping result of an EARLIER scan
        const bool fuse = (stages & 8) && ros::shim_count("/laser_odom_to_init") != odo_before;
        if (fuse) R.tf.laserOdometryHandler(ros::shim_last<nav_msgs::Odometry>("/laser_odom_to_init"));
        if ((stages & 4) && run_mo(R, f, odo_before, outl_before)) {
            out |= 8;
            if (stages & 8) R.tf.odomAftMappedHandler(ros::shim_last<nav_msgs::Odometry>("/aft_mapped_to_init"));   // transformFusion.cpp:218-241
        }
        if (fuse) {
            out |= 16;
            R.snap.f["transformMapped"].assign(R.tf.transformMapped, R.tf.transformMapped + 6);
            auto m = ros::shim_last<nav_msgs::Odometry>("/integrated_to_init");
            if (m) R.snap.f["integratedToInit"] = {(float)m->pose.pose.orientation.x, (float)m->pose.pose.orientation.y, (float)m->pose.pose.orientation.z,
                                                   (float)m->pose.pose.orientation.w, (float)m->pose.pose.position.x, (float)m->pose.pose.position.y,
                                                   (float)m->pose.pose.position.z};
        }
    }
    return out;
}

// The last message the reference's nodes published on `topic`, flattened (oracle/topic_dump.h).  viz topics only exist when
// ref_subscribe_all(1) made getNumSubscribers() non-zero before the scan.
long ref_topic_dump(const char* topic, double* out, long cap, long nvalid, double* stamp, char* frame, char* child, int str_cap, long* count) {
    return topic_dump::dump(topic, out, cap, nvalid, stamp, frame, child, str_cap, count);
}
void ref_subscribe_all(int on) { ros::ShimState::get().all_subscribed = on != 0; }

// One sensor_msgs/Imu message, delivered to the two subscribers the reference has for /imu/data
// (featureAssociation.cpp:193, mapOptmization.cpp:242).
void ref_feed_imu(void* h, double stamp, const double* q_xyzw, const double* w, const double* a) {
    Ref& R = *static_cast<Ref*>(h);
    std::shared_ptr<sensor_msgs::Imu> m = std::make_shared<sensor_msgs::Imu>();
    m->header.stamp = ros::Time(stamp);
    m->orientation.x = q_xyzw[0]; m->orientation.y = q_xyzw[1]; m->orientation.z = q_xyzw[2]; m->orientation.w = q_xyzw[3];
    m->angular_velocity.x = w[0]; m->angular_velocity.y = w[1]; m->angular_velocity.z = w[2];
    m->linear_acceleration.x = a[0]; m->linear_acceleration.y = a[1]; m->linear_acceleration.z = a[2];
    R.fa.imuHandler(m);
    R.mo.imuHandler(m);
}

// mapOptimization::publishGlobalMap (mapOptmization.cpp:757-800) with a subscriber on /laser_cloud_surround, then the
// shutdown half of visualizeGlobalMapThread (729-755; ros::ok() is false in the shim, so its 0.2 Hz loop is skipped).
// Snapshots: "globalMapKeyFramesDS", and per saved file "pcd:finalCloud", "pcd:cornerMap", "pcd:surfaceMap",
// "pcd:trajectory" (flat xyzi); the PCD text of each is read with ref_get_pcd_text.
int ref_global_map(void* h) {
    Ref& R = *static_cast<Ref*>(h);
    ros::ShimState::get().subscribed.insert("/laser_cloud_surround");
    R.mo.publishGlobalMap();
    R.snap.f["globalMapKeyFramesDS"] = flat(*R.mo.globalMapKeyFramesDS);
    pcl::shim_pcd_store().clear();
    R.mo.visualizeGlobalMapThread();
    int n = 0;
    for (auto& kv : pcl::shim_pcd_store()) {
        std::string name = kv.first;
        const size_t slash = name.find_last_of('/');
        if (slash != std::string::npos) name = name.substr(slash + 1);
        const size_t dot = name.find_last_of('.');
        if (dot != std::string::npos) name = name.substr(0, dot);
        R.snap.f["pcd:" + name] = kv.second.xyzi;
        ++n;
    }
    return n;
}
long ref_get_pcd_text(void* h, const char* name, char* buf, long cap) {
    (void)h;
    for (auto& kv : pcl::shim_pcd_store()) {
        std::string base = kv.first;
        const size_t slash = base.find_last_of('/');
        if (slash != std::string::npos) base = base.substr(slash + 1);
        const size_t dot = base.find_last_of('.');
        if (dot != std::string::npos) base = base.substr(0, dot);
        if (base != name) continue;
        const long n = static_cast<long>(kv.second.text.size());
        if (buf) std::memcpy(buf, kv.second.text.data(), std::min(n, cap));
        return n;
    }
    return -1;
}

// Copy a named snapshot (float or int32) of the last scan. Returns the element count, or -1 if unknown.
long ref_get_f32(void* h, const char* name, float* buf, long cap) {
    Ref& R = *static_cast<Ref*>(h);
    auto it = R.snap.f.find(name);
    if (it == R.snap.f.end()) return -1;
    long n = static_cast<long>(it->second.size());
    if (buf) std::memcpy(buf, it->second.data(), sizeof(float) * std::min(n, cap));
    return n;
}
long ref_get_i32(void* h, const char* name, int32_t* buf, long cap) {
    Ref& R = *static_cast<Ref*>(h);
    auto it = R.snap.i.find(name);
    if (it == R.snap.i.end()) return -1;
    long n = static_cast<long>(it->second.size());
    if (buf) std::memcpy(buf, it->second.data(), sizeof(int32_t) * std::min(n, cap));
    return n;
}
// Current poses without needing snapshots: out18 = transformSum(FA) | transformAftMapped | transformTobeMapped
void ref_get_poses(void* h, float* out18) {
    Ref& R = *static_cast<Ref*>(h);
    std::memcpy(out18, R.fa.transformSum, 24); std::memcpy(out18 + 6, R.mo.transformAftMapped, 24); std::memcpy(out18 + 12, R.mo.transformTobeMapped, 24);
}
int ref_num_timers() { return T_COUNT; }
void ref_get_timers(void* h, double* ms_out, long* scans, long* mo_calls) {
    Ref& R = *static_cast<Ref*>(h);
    std::memcpy(ms_out, R.t, sizeof(R.t)); *scans = R.scans; *mo_calls = R.mo_calls;
}
const char* ref_timer_name(int i) {
    static const char* n[T_COUNT] = {"ip_copy", "ip_project", "ip_ground", "ip_segment", "ip_publish", "ip_reset",
        "fa_distort", "fa_smooth", "fa_occlude", "fa_extract", "fa_lm", "fa_integrate", "fa_cloudslast",
        "mo_assoc", "mo_surround", "mo_downsample", "mo_scan2map", "mo_save", "mo_rest"};
    return (i >= 0 && i < T_COUNT) ? n[i] : "";
}

}  // extern "C"
