/* lego_b200.h -- C ABI of the B200-native LeGO-LOAM per-scan hot path.
 *
 * The reference (vigorbird/LeGO-LOAM_comment) exposes no library / plugin / FFI interface: each stage is a
 * file-local C++ class inside a ROS node, and the boundary a maintainer can cut at is the BODY of each node's
 * per-scan method plus the payload of the topics it consumes and produces.  Each
 * create/process/destroy triple below replaces one such body:
 *
 *   lego_ip_*   <- ImageProjection::cloudHandler            LeGO-LOAM/src/imageProjection.cpp:181-197
 *   lego_fa_*   <- FeatureAssociation::runFeatureAssociation LeGO-LOAM/src/featureAssociation.cpp:1817-1860
 *   lego_mo_*   <- mapOptimization::run                      LeGO-LOAM/src/mapOptmization.cpp:1487-1522
 *   lego_pipeline_* : the three stages chained on the device for `lanes` independent sequences at once (the
 *                     throughput path; no equivalent in the reference, which pipelines the nodes over ROS topics).
 *
 * Conventions
 *   - plain C types only; clouds are arrays of 4 floats per point (x, y, z, intensity) like pcl::PointXYZI's
 *     payload (utility.h:51); the ring channel is a parallel uint16 array (PointXYZIR, utility.h:153-165).
 *   - every function returns LEGO_OK (0) or a negative error code; nothing throws, nothing exits.
 *   - a handle owns all its device memory; one host thread per handle; handles are independent.
 *   - there is NO CPU fallback: every entry point runs CUDA kernels built for sm_100a and fails with
 *     LEGO_ERR_CUDA when no usable device is present.
 */
#ifndef LEGO_B200_H_
#define LEGO_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LEGO_OK 0
#define LEGO_ERR_ARG (-1)        /* bad argument (null pointer, size out of range)                      */
#define LEGO_ERR_CUDA (-2)       /* CUDA runtime failure, or no device: lego_last_error() has the text  */
#define LEGO_ERR_NOT_DENSE (-3)  /* useCloudRing with is_dense == 0: the reference logs ROS_ERROR and shuts
                                    down (imageProjection.cpp:172-177); here it is an error code         */
#define LEGO_ERR_CAPACITY (-4)   /* more points / keyframes than the handle was created for             */
#define LEGO_ERR_NAME (-5)       /* unknown array name in lego_*_download / upload                      */

/* Field names and meaning = the reference's compile-time constants, LeGO-LOAM/include/utility.h:60-136. */
typedef struct lego_params {
    int N_SCAN, Horizon_SCAN;                /* utility.h:63-64                                          */
    float ang_res_x, ang_res_y, ang_bottom;  /* utility.h:65-67                                          */
    int groundScanInd;                       /* utility.h:68                                             */
    int useCloudRing;                        /* utility.h:60                                             */
    float sensorMinimumRange;                /* utility.h:111                                            */
    float sensorMountAngle;                  /* utility.h:112                                            */
    float segmentTheta;                      /* utility.h:113                                            */
    int segmentValidPointNum;                /* utility.h:114                                            */
    int segmentValidLineNum;                 /* utility.h:115                                            */
    float segmentAlphaX, segmentAlphaY;      /* utility.h:116-117                                        */
    float scanPeriod;                        /* utility.h:107                                            */
    int edgeFeatureNum, surfFeatureNum, sectionsTotal;   /* utility.h:120-122                            */
    float edgeThreshold, surfThreshold;      /* utility.h:123-124                                        */
    float nearestFeatureSearchSqDist;        /* utility.h:125                                            */
    float surroundingKeyframeSearchRadius;   /* utility.h:129                                            */
    double mappingProcessInterval;           /* utility.h:105                                            */
    int loopClosureEnableFlag;               /* utility.h:104                                            */
    int device;                              /* CUDA device ordinal                                      */
    /* loop closure (only read when loopClosureEnableFlag != 0) */
    int surroundingKeyframeSearchNum;        /* utility.h:130 (50)                                       */
    float historyKeyframeSearchRadius;       /* utility.h:132 (7.0)                                      */
    int historyKeyframeSearchNum;            /* utility.h:133 (25)                                       */
    float historyKeyframeFitnessScore;       /* utility.h:134 (0.3)                                      */
    /* capacity of the handle (the reference's std::vector / deque of key frames grows without bound,
     * mapOptmization.cpp:1443-1453): key frames a sequence may save before LEGO_ERR_CAPACITY; 0 = 640 */
    int maxKeyFrames;
} lego_params;

/* A cloud living on the device (dev) with an optional host mirror (host, may be NULL). Valid until the next
 * process() call on the handle that produced it. */
typedef struct lego_cloud_view {
    const float* dev;    /* device pointer, count*4 floats */
    const float* host;   /* host pointer or NULL           */
    int count;
} lego_cloud_view;

/* cloud_msgs/cloud_info (cloud_msgs/msg/cloud_info.msg:1-11).  Per-ring arrays have N_SCAN entries; per-point
 * arrays have `count` valid entries (the reference allocates N_SCAN*Horizon_SCAN and never clears them,
 * imageProjection.cpp:125-130; only the first `count` are meaningful). Pointers are HOST pointers when the view
 * is an input of lego_fa_process and device pointers in dev_* when handed over on the device. */
typedef struct lego_cloud_info_view {
    const int32_t* startRingIndex;
    const int32_t* endRingIndex;
    float startOrientation, endOrientation, orientationDiff;
    const uint8_t* segmentedCloudGroundFlag;
    const uint32_t* segmentedCloudColInd;
    const float* segmentedCloudRange;
    int count;
    int on_device;       /* 0: the pointers above are host memory, 1: device memory */
} lego_cloud_info_view;

const char* lego_last_error(void);
int lego_device_count(void);

/* ---------------------------------------------------------------- ImageProjection */
typedef struct lego_ip lego_ip;
int lego_ip_create(const lego_params* p, lego_ip** out);
int lego_ip_destroy(lego_ip* h);
/* One lidar scan: xyzi = P*4 floats in sensor firing order, ring = P uint16 or NULL (required when
 * useCloudRing), host memory.  Outputs: segmented cloud + cloud_info + outlier cloud, device-resident with host
 * mirrors filled (payload of /segmented_cloud, /segmented_cloud_info, /outlier_cloud, imageProjection.cpp:463-478). */
int lego_ip_process(lego_ip* h, const float* xyzi, const uint16_t* ring, int P, int is_dense, double stamp,
                    lego_cloud_info_view* info, lego_cloud_view* segmented, lego_cloud_view* outlier);

/* ---------------------------------------------------------------- FeatureAssociation */
typedef struct lego_fa lego_fa;
typedef struct lego_fa_outputs {
    lego_cloud_view cornerPointsSharp, cornerPointsLessSharp, surfPointsFlat, surfPointsLessFlat; /* featureAssociation.cpp:786-817 */
    lego_cloud_view laserCloudCornerLast, laserCloudSurfLast, outlierCloudLast;                   /* featureAssociation.cpp:1796-1813 */
    int odometry_valid;   /* 0 on the initialisation scan (featureAssociation.cpp:1846-1849), else 1      */
    int published;        /* 1 when the three "last" clouds were forwarded to mapping this scan           */
    float transformCur[6];
} lego_fa_outputs;
int lego_fa_create(const lego_params* p, lego_fa** out);
int lego_fa_destroy(lego_fa* h);
/* Inputs = the three ImageProjection outputs of one scan (host or device views). transformSum_out = accumulated
 * odometry [rx,ry,rz,tx,ty,tz] in the LOAM camera frame (featureAssociation.cpp:1697-1725). */
int lego_fa_process(lego_fa* h, const lego_cloud_info_view* info, const lego_cloud_view* segmented,
                    const lego_cloud_view* outlier, double stamp, float transformSum_out[6], lego_fa_outputs* out);

/* FeatureAssociation::imuHandler (featureAssociation.cpp:431-459): one sensor_msgs/Imu message (orientation quaternion
 * x y z w, angular velocity, linear acceleration incl. gravity, all in the IMU's own axes as the reference expects them).
 * Optional: without any call the stage runs the reference's no-IMU path (imuPointerLast == -1). */
int lego_fa_imu(lego_fa* h, double stamp, const double orientation_xyzw[4], const double angular_velocity[3],
                const double linear_acceleration[3]);

/* ---------------------------------------------------------------- mapOptimization (scan-to-map) */
typedef struct lego_mo lego_mo;
int lego_mo_create(const lego_params* p, lego_mo** out);
int lego_mo_destroy(lego_mo* h);
/* Inputs = /laser_cloud_corner_last, /laser_cloud_surf_last, /outlier_cloud_last and the odometry pose of the
 * same stamp.  *processed = 1 when the mapping body ran (time gate mapOptmization.cpp:1499).  Outputs follow
 * publishTF (mapOptmization.cpp:654-673): transformAftMapped and transformBefMapped. */
int lego_mo_process(lego_mo* h, const lego_cloud_view* corner_last, const lego_cloud_view* surf_last,
                    const lego_cloud_view* outlier_last, const float transformSum[6], double stamp,
                    float transformAftMapped_out[6], float transformBefMapped_out[6], int* processed);

/* mapOptimization::imuHandler (mapOptmization.cpp:643-652): roll / pitch for the 0.998 / 0.002 blend of transformUpdate */
int lego_mo_imu(lego_mo* h, double stamp, const double orientation_xyzw[4], const double angular_velocity[3],
                const double linear_acceleration[3]);

/* ---------------------------------------------------------------- TransformFusion node (transformFusion.cpp)
 * The fourth node: re-bases every odometry pose on the latest mapping correction (/integrated_to_init).
 * lego_tf_laser_odometry replaces the body of TransformFusion::laserOdometryHandler (transformFusion.cpp:181-216):
 * transformSum is the pose FeatureAssociation integrated (lego_fa_process's transformSum_out; the quaternion round
 * trip of the message is applied inside), transformMapped_out the fused pose [rx ry rz tx ty tz],
 * integrated_out the published pose: orientation x y z w, position x y z (transformFusion.cpp:201-208).
 * lego_tf_aft_mapped replaces TransformFusion::odomAftMappedHandler (transformFusion.cpp:218-241): the two poses
 * mapOptimization publishes in /aft_mapped_to_init (lego_mo_process's outputs when *processed = 1).
 * Pointers may be host or device memory. */
typedef struct lego_tf lego_tf;
int lego_tf_create(const lego_params* p, lego_tf** out);
int lego_tf_destroy(lego_tf* h);
int lego_tf_laser_odometry(lego_tf* h, const float transformSum[6], float transformMapped_out[6], float integrated_out[7]);
int lego_tf_aft_mapped(lego_tf* h, const float transformAftMapped[6], const float transformBefMapped[6]);

/* ---------------------------------------------------------------- chained, batched pipeline */
typedef struct lego_pipeline lego_pipeline;
/* lanes independent sequences processed in lock step; max_points = capacity per scan (0 = N_SCAN*Horizon_SCAN+slack).
 * Capacity errors: a scan with more than max_points points, or npts > stride_points, fails the step with
 * LEGO_ERR_CAPACITY before anything is enqueued.  A lane that runs out of key-frame pool (lego_params.maxKeyFrames) or
 * local-map capacity is reported by the first lego_pipeline_step / step_lanes that learns of it (the device reports
 * one mapping call late): that call and every later one returns LEGO_ERR_CAPACITY with the lane in lego_last_error();
 * the step was still enqueued, but from then on that lane maps against a truncated map and no longer follows the
 * reference. */
int lego_pipeline_create(const lego_params* p, int lanes, int max_points, lego_pipeline** out);
int lego_pipeline_destroy(lego_pipeline* h);
/* One scan for every lane.  xyzi: lanes * stride_points * 4 floats, ring: lanes * stride_points, npts: lanes
 * ints, stamps: lanes doubles.  src_on_device = 1 when xyzi/ring are device pointers (npts/stamps stay host).
 * stage_mask: bit0 ImageProjection, bit1 FeatureAssociation, bit2 mapOptimization, bit3 TransformFusion (its poses are
 * read with lego_pipeline_download: "transformMapped" 6 floats, "integratedToInit" 7 floats, per lane). */
int lego_pipeline_step(lego_pipeline* h, const float* xyzi, const uint16_t* ring, const int* npts,
                       int stride_points, const double* stamps, int src_on_device, int stage_mask);
/* Same step with ONE POINTER PER LANE and no padding on the wire: xyz[l] points at npts[l] points of lane l,
 * floats_per_point = 4 (x y z intensity, like lego_pipeline_step) or 3 (x y z packed -- ImageProjection never reads the
 * input intensity, imageProjection.cpp:248-255 overwrites it); ring[l] at npts[l] uint16 (ring may be NULL without
 * useCloudRing).  The pointer TABLES are host memory.  src_on_device: 0 = the pointers in them are ordinary host memory
 * (one cudaMemcpyAsync per lane and channel), 1 = device memory (read in place), 2 = PAGE-LOCKED host memory that the
 * device can address (cudaMallocHost / cudaHostAlloc / cudaHostRegister, e.g. torch pin_memory): one packing kernel pulls
 * every lane's points over PCIe, which is what keeps the link busy with hundreds of lanes.  Passing pageable memory with
 * 2 is an error the call cannot detect (the kernel faults).
 * Host sources: exactly npts[l] * (4 * floats_per_point + 2) bytes per lane cross PCIe.  With src_on_device = 0, a lane whose
 * ring channel starts exactly where its points end (ring[l] == (char*)xyz[l] + npts[l] * 4 * floats_per_point) is moved
 * with ONE copy instead of two, which matters at hundreds of lanes (the copy engine spends 2-3 us per copy), and lanes
 * whose blocks touch exactly (xyz[l+1] starts where ring[l] ends) share one copy.  3 = host memory like 0, and the caller
 * states that all lane blocks are slots of ONE allocation (a receive arena): consecutive lanes that lie less than 96 KB
 * apart are then moved by one copy TOGETHER WITH THE BYTES BETWEEN THEM (those bytes must be readable, which only the
 * caller can know; they are never used).  With slots of a fixed size this turns hundreds of copies per scan into a few.
 *
 * Input lifetime (both step forms): host buffers are copied asynchronously on the handle's copy stream -- they must
 * stay valid and unmodified until lego_pipeline_input_wait() returns (or any later synchronising call on the handle:
 * lego_pipeline_sync, get_poses, the poses_wait of this step).  Device buffers are read in place by kernels on the
 * handle's own stream, which is NOT ordered against the caller's streams: the data must be complete before the call
 * (synchronise the producing stream first) and must not be rewritten before lego_pipeline_input_wait() returns. */
int lego_pipeline_step_lanes(lego_pipeline* h, const float* const* xyz, const uint16_t* const* ring, const int* npts,
                             int floats_per_point, const double* stamps, int src_on_device, int stage_mask);
/* Blocks until the input buffers handed to the last step may be refilled (host: copied; device: consumed). */
int lego_pipeline_input_wait(lego_pipeline* h);
/* One /imu/data message for one lane: delivered to that lane's FeatureAssociation and mapOptimization imuHandler
 * (featureAssociation.cpp:431-459, mapOptmization.cpp:643-652).  Call it before the lego_pipeline_step of the scan during
 * which the message would have arrived (the reference's subscribers fill the ring buffer between two scans). */
int lego_pipeline_imu(lego_pipeline* h, int lane, double stamp, const double orientation_xyzw[4], const double angular_velocity[3],
                      const double linear_acceleration[3]);
/* poses_out: lanes * 18 floats = transformSum | transformAftMapped | transformBefMapped; flags_out: lanes ints,
 * bit1 odometry ran, bit2 clouds forwarded to mapping, bit3 mapping body ran.  Synchronises the handle's stream. */
int lego_pipeline_get_poses(lego_pipeline* h, float* poses_out, int* flags_out);
/* Streaming form: poses_async enqueues the read-back of the poses as they stand after the last lego_pipeline_step
 * into pinned slot 0..7 and returns at once; poses_wait blocks until that read-back has landed and unpacks it like
 * lego_pipeline_get_poses.  Lets a caller submit scan k+1 (its host-to-device copy runs on the copy stream) before it
 * collects the result of scan k -- the role the ROS message queues between the reference's nodes play. */
int lego_pipeline_poses_async(lego_pipeline* h, int slot);
int lego_pipeline_poses_wait(lego_pipeline* h, int slot, float* poses_out, int* flags_out);
int lego_pipeline_sync(lego_pipeline* h);
/* number of kernel launches issued by the handle so far (bench.py's gpu_launches) */
long lego_pipeline_launch_count(lego_pipeline* h);
/* number of cudaGraphLaunch submissions so far.  The per-scan launch sequences are replayed from CUDA graphs when the
 * handle has fewer than 148 lanes (a scan is then launch-latency bound; LEGO_GRAPH=0 / 1 in the environment forces
 * either), except while profiling, with IMU data and in loop-closure mode.  Kernels inside a graph still count in
 * lego_pipeline_launch_count.  No counterpart in the reference (its nodes are plain CPU callbacks). */
long lego_pipeline_graph_launch_count(lego_pipeline* h);
/* Timing on the handle's own stream (CUDA events; a caller's torch.cuda.Event cannot see this stream).
 * timer_record(slot) records event `slot` (0..7); timer_elapsed waits for slot_b and returns the device time between
 * the two.  profile(mode): 0 off, 1 bracket every kernel launch with events, 2 only kernels whose name contains
 * `filter`; profile_read sums the brackets per kernel name (names newline-separated) and returns the number of names. */
int lego_pipeline_timer_record(lego_pipeline* h, int slot);
int lego_pipeline_timer_elapsed(lego_pipeline* h, int slot_a, int slot_b, float* ms_out);
/* device time from event slot_a of handle ha to event slot_b of handle hb (handles of one device; several handles =
 * several streams whose kernels overlap on the GPU) */
int lego_pipeline_timer_elapsed_between(lego_pipeline* ha, int slot_a, lego_pipeline* hb, int slot_b, float* ms_out);
int lego_pipeline_profile(lego_pipeline* h, int mode, const char* filter);
long lego_pipeline_profile_read(lego_pipeline* h, char* names, long names_cap, double* ms, long* counts, long max_kernels);
/* Parity / debug access to named device arrays of one lane (see DESIGN.md for the name list).  Returns the
 * element count (and copies min(count, cap) elements) or a negative error code.  elem_size_out gets 1/2/4/16. */
long lego_pipeline_download(lego_pipeline* h, int lane, const char* name, void* dst, long cap, int* elem_size_out);
long lego_pipeline_upload(lego_pipeline* h, int lane, const char* name, const void* src, long count);

/* ---------------------------------------------------------------- loop closure building blocks
 * With loopClosureEnableFlag set, lego_mo_process / lego_pipeline_step run the reference's loop-closure branch:
 * the local map is the surroundingKeyframeSearchNum most recent key frames (mapOptmization.cpp:960-1000) and
 * performLoopClosure (mapOptmization.cpp:875-945; a 1 Hz side thread in the reference) is called after a mapping call
 * once >= 1.0 s of scan time passed since its last call.  Its two third-party pieces are exported on their own:
 *
 * lego_icp_align replaces pcl::IterativeClosestPoint<PointXYZI,PointXYZI>::align + hasConverged + getFitnessScore
 * + getFinalTransformation as configured at mapOptmization.cpp:892-905.  source / target: n*4 floats, host memory.
 * T_out: row-major 4x4.  Runs on the device (voxel-hash exact nearest neighbour, one CTA). */
int lego_icp_align(int device, const float* source, int n_source, const float* target, int n_target,
                   double max_correspondence_distance, int max_iterations, double transformation_epsilon,
                   double euclidean_fitness_epsilon, float T_out[16], double* fitness_out, int* converged_out, int* iterations_out);
/* lego_pg_* replaces the gtsam calls of mapOptmization.cpp:1375-1409 and 929-942 (ISAM2 on a Pose3 graph with diagonal
 * noise): poses are (rx, ry, rz, tx, ty, tz) doubles with R = gtsam::Rot3::RzRyRx(rx, ry, rz); add_between takes the two
 * poses whose relative transform from.between(to) is the measurement.  update = ISAM2::update() (host code, like GTSAM). */
typedef struct lego_pg lego_pg;
int lego_pg_create(lego_pg** out);
int lego_pg_destroy(lego_pg* h);
int lego_pg_add_prior(lego_pg* h, int key, const double pose[6], const double variances[6]);
int lego_pg_add_between(lego_pg* h, int key1, int key2, const double from[6], const double to[6], const double variances[6]);
int lego_pg_insert(lego_pg* h, int key, const double pose[6]);
int lego_pg_update(lego_pg* h);
int lego_pg_size(lego_pg* h);
/* estimate of one key: translation x y z and Rot3::roll() pitch() yaw() */
int lego_pg_estimate(lego_pg* h, int key, double xyz_rpy_out[6]);

/* ---------------------------------------------------------------- global map and PCD output
 * The clouds mapOptimization::visualizeGlobalMapThread publishes every 5 s and saves at shutdown
 * (mapOptmization.cpp:724-800).  which:
 *   LEGO_MAP_GLOBAL      publishGlobalMap's globalMapKeyFramesDS (757-800) = /laser_cloud_surround and finalCloud.pcd (731)
 *   LEGO_MAP_CORNER      cornerMapCloudDS  -> cornerMap.pcd  (734-749, 753)
 *   LEGO_MAP_SURFACE     surfaceMapCloudDS -> surfaceMap.pcd (734-751, 753)
 *   LEGO_MAP_TRAJECTORY  cloudKeyPoses3D   -> trajectory.pcd (754)
 * xyzi_out: cap * 4 floats of host memory (may be NULL with cap 0 to ask for the size); *n_out = points of the cloud,
 * min(*n_out, cap) of them copied.  Key-frame clouds are assembled and filtered on the device.
 * lego_write_pcd_ascii replaces pcl::io::savePCDFileASCII for a PointXYZI cloud (host file output). */
enum { LEGO_MAP_GLOBAL = 0, LEGO_MAP_CORNER = 1, LEGO_MAP_SURFACE = 2, LEGO_MAP_TRAJECTORY = 3 };
int lego_pipeline_global_map(lego_pipeline* h, int lane, int which, float* xyzi_out, long cap, long* n_out);
int lego_mo_global_map(lego_mo* h, int which, float* xyzi_out, long cap, long* n_out);
int lego_write_pcd_ascii(const char* path, const float* xyzi, long n);

/* ---------------------------------------------------------------- allocation-free VoxelGrid
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  pcl::VoxelGrid<PointXYZI>::filter with
 * setLeafSize(leaf, leaf, leaf) on ONE cloud, the same kernels as every other VoxelGrid of this library, but with the
 * sort scratch in a file-scope static __device__ array instead of allocated memory (csrc/voxel_grid_noalloc.cu):
 *   - d_xyzi: n * 4 floats of DEVICE memory; d_out: out_cap * 4 floats of DEVICE memory (out_cap >= n is always enough);
 *     d_n_out: one int of DEVICE memory that receives the size of the filtered cloud; cuda_stream: a cudaStream_t
 *     (NULL = the default stream) of the current device.
 *   - n <= LEGO_VG_NOALLOC_MAX_POINTS, otherwise LEGO_ERR_CAPACITY (lego_last_error() names the limit).  The scratch
 *     (4.3 MB) is held by every GPU that loads the library until the process exits.
 *   - the call only ENQUEUES kernels on cuda_stream and returns; it calls no cudaMalloc / cudaFree / allocating routine
 *     and does not synchronise.  Results are complete when the stream has run.
 *   - all calls on one GPU share the scratch array (it keeps the last call's contents; every call re-initialises what
 *     it reads): calls that OVERLAP on the same GPU -- different streams or host threads -- race.  Serialise them. */
#define LEGO_VG_NOALLOC_MAX_POINTS 262144
int lego_voxel_grid_noalloc(const float* d_xyzi, int n, float leaf, float* d_out, int out_cap, int* d_n_out, void* cuda_stream);
int lego_voxel_grid_noalloc_max_points(void);

/* ---------------------------------------------------------------- device self-tests (parity suite) */
/* out[i] = the device implementation of glibc's atan2f(y[i], x[i]) (csrc/parity_math.cuh); host pointers. */
int lego_selftest_atan2f(const float* y, const float* x, float* out, long n);
/* sin_out[i], cos_out[i] = the device statement of glibc's sinf / cosf (csrc/parity_math.cuh); host pointers. */
int lego_selftest_sincosf(const float* x, float* sin_out, float* cos_out, long n);
/* Sorts nseg independent segments (offsets has nseg+1 entries) on the device with the replayed libstdc++
 * std::sort (csrc/std_sort_replay.cuh); inds_out receives, per segment, the original positions in sorted order. */
int lego_selftest_std_sort(const float* values, const int* offsets, int nseg, int* inds_out);
/* Same contract, sorted by the warp-parallel statement of std::sort that fa_sort_sectors_kernel uses (csrc/warp_sort.cuh:
 * radix sort when all values are distinct, introsort partition replay + stable radix sort otherwise); every segment
 * must hold at most 416 values (LEGO_ERR_ARG otherwise). */
int lego_selftest_warp_sort(const float* values, const int* offsets, int nseg, int* inds_out);

/* Batched pcl::VoxelGrid (csrc/voxel_grid.cu) on nseg independent clouds: offsets has nseg+1 entries (in points);
 * out receives the filtered clouds back to back in segment order, out_counts[s] their sizes.  Host pointers. */
int lego_selftest_voxel_grid(const float* xyzi, const int* offsets, int nseg, float leaf, float* out, int* out_counts);

#ifdef __cplusplus
}
#endif
#endif /* LEGO_B200_H_ */
