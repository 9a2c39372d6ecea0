// Host-side engine: owns the device buffers of `lanes` independent sequences and mirrors the reference's three
// stage classes (ImageProjection / FeatureAssociation / mapOptimization) method for method.  Every method only
// enqueues kernels on the engine's stream; nothing here computes on the CPU except the keyframe bookkeeping of
// mapOptimization, which is host glue in the reference too (mapOptmization.cpp:956-1065, 1353-1454).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <deque>
#include <string>
#include <vector>
#include <unordered_map>
#include <functional>
#include "../../include/lego_b200.h"
#include "pose_graph.h"
#include "voxel_grid.cuh"
#include "voxel_hash.cuh"

namespace lego {

constexpr int MO_RES_THREADS = 256;   // one query per thread in mo_residual_kernel
#ifndef LEGO_MO_RES_INFLIGHT
#define LEGO_MO_RES_INFLIGHT 16
#endif
constexpr int MO_RES_INFLIGHT = LEGO_MO_RES_INFLIGHT;   // 256-query blocks of one lane in flight per launch; more queries loop

// ---- constants derived once from lego_params (host side, with the reference's float/double semantics) ----
struct Consts {
    int R, C, N, G, L, Pcap, Ocap;
    int useRing;
    float ang_res_x, ang_res_y, ang_bottom;
    float sensorMinimumRange, sensorMountAngle, segmentTheta;
    float segTanLo, segTanHi;  // tan(segmentTheta) * (1 -+ 1e-4): bounds of the band in which seg_edge evaluates atan2f (0 = always)
    float sinAlphaX, cosAlphaX, sinAlphaY, cosAlphaY;   // sinf/cosf of segmentAlphaX/Y computed on the host (glibc)
    int segmentValidPointNum, segmentValidLineNum;
    float scanPeriod, edgeThreshold, surfThreshold, nearestFeatureSearchSqDist;
    int rowWords;   // ceil(R/32) words of row bitmask per component
};

// ---- ImageProjection device arrays, all [lane][...] with the strides implied by Consts ----
struct IpDev {
    const float4* const* in_xyzi;     // [L] per-lane pointer to the scan's points (staging buffer or caller's device memory)
    const uint16_t* const* in_ring;   // [L] per-lane pointer to the ring channel
    const int* npts;           // [L]
    int in_floats;             // floats per input point: 4 = x y z intensity, 3 = x y z packed
    int* owner;                // [L][N]   index of the last input point that fell into the cell, -1 = empty
    float* rangeMat;           // [L][N]   FLT_MAX = empty           (imageProjection.cpp:152)
    float4* fullCloud;         // [L][N]   NaN / intensity -1 = empty (imageProjection.cpp:157)
    int8_t* groundMat;         // [L][N]   -1 / 0 / 1                (imageProjection.cpp:263-266)
    int* labelMat;             // [L][N]   -1, 0 (transient), 1..K, 999999
    int* parent;               // [L][N]   union-find forest, root = smallest cell index of the component
    int* compSize;             // [L][N]   at roots: number of non-root members; later: the root's final label
    int* runLen;               // [L][N]   at the last cell of a horizontal run: its length (0 elsewhere)
    unsigned char* cellClass;  // [L][N]   classify_cell bits (valid-segment head | to segmentedCloud | to outlierCloud)
    uint32_t* compRows;        // [L][N][rowWords] at roots: bitmask of rows that hold a non-root member
    int* rowCounts;            // [L][R][4] per row: valid roots, kept cells, outlier cells, unused
    float4* segCloud;          // [L][N]
    uint8_t* segGround;        // [L][N]   persistent, never cleared (imageProjection.cpp:125-130)
    int* segCol;               // [L][N]
    float* segRange;           // [L][N]
    int* startRing; int* endRing;   // [L][R]
    float* orient;             // [L][4]   startOrientation, endOrientation, orientationDiff
    float4* outlier;           // [L][Ocap]
    int* counts;               // [L][8]   0: S (segmented size) 1: O (outliers) 2: number of valid segments
};

// ---- IMU (featureAssociation.cpp:317-459, 525-614, 1639-1664; optional: active once a lane received a message) ----
constexpr int IMU_QUE = 200;          // imuQueLength (utility.h:109)
constexpr int IMU_CH = 12;            // roll pitch yaw | velo xyz | shift xyz | angular rotation xyz
enum ImuState {                       // FaImuDev::state[lane][IMU_STATE]
    IS_ROLL_START = 0, IS_PITCH_START, IS_YAW_START, IS_VELO_START, IS_SHIFT_START = 6,
    IS_COS_ROLL = 9, IS_COS_PITCH, IS_COS_YAW, IS_SIN_ROLL, IS_SIN_PITCH, IS_SIN_YAW,
    IS_ANG_LAST = 15, IS_ANG_FROM_START = 18, IS_RPY_CUR = 21, IS_VELO_FROM_START_CUR = 24, IS_RPY_LAST = 27, IS_VELO_FROM_START = 30,
    IMU_STATE = 36
};
struct FaImuDev {
    double* time;              // [L][IMU_QUE] imuTime
    float* ring;               // [L][IMU_QUE][IMU_CH]
    int* ptr;                  // [L][2] imuPointerLast (-1 = no IMU), imuPointerLastIteration
    double* scanTime;          // [L] timeScanCur
    float* state;              // [L][IMU_STATE]
};
// host side of one lane's imuHandler (featureAssociation.cpp:431-459, mapOptmization.cpp:643-652)
struct ImuLaneHost {
    // FeatureAssociation's ring buffers
    int pointerLast = -1, pointerLastIteration = 0;
    double time[IMU_QUE] = {};
    float roll[IMU_QUE] = {}, pitch[IMU_QUE] = {}, yaw[IMU_QUE] = {};
    float acc[3][IMU_QUE] = {}, velo[3][IMU_QUE] = {}, shift[3][IMU_QUE] = {}, angVelo[3][IMU_QUE] = {}, angRot[3][IMU_QUE] = {};
    bool dirty = false;
    // mapOptimization's own buffers
    int moPointerLast = -1, moPointerFront = 0;
    double moTime[IMU_QUE] = {};
    float moRoll[IMU_QUE] = {}, moPitch[IMU_QUE] = {};
};

// ---- FeatureAssociation device arrays ----
struct FaDev {
    float4* segAdj;            // [L][N]   segmented cloud after adjustDistortion (camera axes)
    float* curv;               // [L][N]   cloudCurvature, persistent
    int* picked;               // [L][N]   cloudNeighborPicked, persistent
    int* label;                // [L][N]   cloudLabel, persistent
    float* slot4val; int* slot4ind;   // [L] stale smoothness slot 4 (quirk Q11)
    int* halfFirst;            // [L]      first index whose branch-A orientation passes pi (quirk Q9)
    // ring-local feature staging, then compacted lists
    float4* ringSharp; float4* ringLessSharp; float4* ringFlat; float4* ringLessFlat;   // [L][R][12|120|24|C]
    int* ringCounts;           // [L][R][4]
    int* cand;                 // [L][N]   per sector: edge candidates from slot sp upward, flat candidates from slot ep downward (index | reach << 24)
    int* candCounts;           // [L][R][6][2]
    void* sortScratch;         // [L][N] {float, int}, only when a sector can exceed what one warp sorts (Horizon_SCAN > 2496)
    float4* sharp; float4* lessSharp; float4* flat; float4* lessFlat;   // [L][12R | 120R | 24R | N]
    int* featCounts;           // [L][4]
    // odometry state
    float4* cornerLast[2]; float4* surfLast[2];   // [L][120R], [L][N] double buffered
    int* lastCounts;           // [L][4]   cornerLastNum, surfLastNum (current "last" buffer), buffer index, frameCount
    float* transformCur;       // [L][6]
    float* transformSum;       // [L][6]
    float* matP;               // [L][9]
    int* lmState;              // [L][8]   0: systemInitedLM 1: isDegenerate 2: surf phase done 3: corner phase done 4: skip 5: rows
    int* cornerInd;            // [L][12R][2]
    int* surfInd;              // [L][24R][3]
    double* partials;          // [L][maxBlocks][16]
    unsigned int* tickets;     // [L]
    float4* outlierLast;       // [L][Ocap]
    int* flags;                // [L]      bit1 odometry ran, bit2 clouds forwarded
    VoxelHash hashCorner, hashSurf;   // 1 m voxel hash over the current "last" clouds (replaces the two kd-trees)
    FaImuDev imu;
};

// ---- TransformFusion (transformFusion.cpp) ----
struct TfDev {
    float* aft; float* bef;    // [L][6]  the node's own transformAftMapped / transformBefMapped (last /aft_mapped_to_init message)
    float* mapped;             // [L][6]  transformMapped
    float* integrated;         // [L][7]  /integrated_to_init pose: orientation x y z w, position x y z
    int* published;            // [L]     1 when the last step published a fused pose
};

// ---- mapOptimization ----
// One surrounding key frame of this mapping call, prepared by the host bookkeeping (extractSurroundingKeyFrames,
// mapOptmization.cpp:1001-1056): where its three down-sampled clouds live in the key-frame pool, where they go in
// the local map, and the sin/cos/translation of its key pose (updateTransformPointCloudSinCos, mapOptmization.cpp:529-543).
struct MoKfEntry {
    int srcC, nC, dstC;       // corner cloud: pool offset, size, offset in laserCloudCornerFromMap
    int srcS, nS, dstS;       // surf cloud -> laserCloudSurfFromMap
    int srcO, nO, dstO;       // outlier cloud -> laserCloudSurfFromMap (right behind the surf cloud of the same key frame)
    float ctRoll, stRoll, ctPitch, stPitch, ctYaw, stYaw, tInX, tInY, tInZ;
    int id;                   // key-frame index (cloudKeyPoses3D position); the incremental local map tags its points with it
};

// ---- incremental local map (incremental_map.cu) ----
// extractSurroundingKeyFrames (mapOptmization.cpp:1013-1063) re-filters the whole local map in every call although its
// key-frame list only loses the frames that left the 50 m neighbourhood and gains the new ones at the END (1013-1049 is a
// cache diff).  The device keeps, per lane and per cloud (corner / surf+outlier), the transformed key-frame points SORTED
// by (voxel of the map's VoxelGrid, position in the list, point index) -- the order pcl::VoxelGrid sums them in -- and per
// call deletes, merges and re-reduces instead of sorting everything again.
constexpr int IM_MAX_DEPARTED = 16;   // more frames leaving in one call -> the lane rebuilds from scratch
constexpr int IM_TILE = 1024;
struct ImCtl {                // per lane, written by the host bookkeeping for every mapping call
    int changed;              // 0: same key-frame list as in the last call, the down-sampled map on the device is still valid
    int rebuild;              // 1: forget the sorted state, every listed frame arrives as new
    int nDeparted;
    int departed[IM_MAX_DEPARTED];   // key-frame ids that left the list
};
struct ImSeg {                // one sorted cloud: segment 2*lane = corner map (leaf 0.2), 2*lane+1 = surf map (leaf 0.4)
    float4* pts[2];           // [cap] ping-pong: the sorted points
    int* frame[2];            // [cap] key-frame id of every point
    int* rank;                // [cap / 32] number of surviving points in front of point 32 k (written by the merge)
    float4* newPts; int* newFrame; unsigned long long* newKey;   // [newCap] the arriving points, sorted
    const float4* raw;        // the arriving points as mo_assemble_map_kernel wrote them (list order)
    const int* nRaw;          // their number
    const VgSeg* sorter;      // radix-sort scratch of this segment: vg_sorted_vals() = order of `raw`
    float4* out; int* nOut;   // down-sampled map + its size (laserCloud*FromMapDS)
    int* nTotal;              // size of the sorted cloud, for the records (laserCloud*FromMap)
    int* state;               // [8] 0: current buffer, 1: points in it, 2: survivors of this call, 3: arrivals of this call
    int* tileKept; int* tileHeads;   // [cap / IM_TILE + 1]
    int* errors;              // moState[lane][5]
    float leaf;
    int cap, newCap, outCap, type;
};
struct ImBatch { ImSeg* segs; int nSeg; int maxTiles; int maxNewTiles; };
// What saveKeyFramesAndFactor (mapOptmization.cpp:1353-1454) decided on the device, read back asynchronously by the
// host bookkeeping before the next mapping call.
struct MoRecord {
    int ran;                  // the mapping body ran for this lane in this step
    int saved;                // a key frame was appended
    int poolOff, nC, nS, nO;  // its clouds in the pool
    int error;                // capacity overflow bits
    int iters;                // LM iterations of scan2MapOptimization
    float robotPos[3];        // currentRobotPosPoint
    float pose6D[6];          // x, y, z, roll, pitch, yaw of cloudKeyPoses6D
    float aftRaw[6];          // the pose the factor was built from: transformAftMapped before the estimate overwrote it
                              // (transformTobeMapped for the first key frame, mapOptmization.cpp:1375-1392)
};
// saveKeyFramesAndFactor decided on the HOST: once a lane's pose graph holds a loop factor the newest estimate is no
// longer the inserted value, so the host runs the graph update between scan2MapOptimization and the key-frame save
// and hands the kernel its decision (mapOptmization.cpp:1395-1440).
struct MoOverride { int use, save; float pose6D[6]; float aft[6]; };

// ---- loop closure (mapOptmization.cpp:814-945): history submap + ICP, a few "slots" shared by all lanes ----
struct LcEntry {              // one key-frame cloud to transform into the ICP source (which = 0) or target (which = 1)
    int src, n, dst, which;
    float ctRoll, stRoll, ctPitch, stPitch, ctYaw, stYaw, tInX, tInY, tInZ;
};
struct LcResult { int converged, iterations, pairs, nSrc; double fitness; float T[16]; };
struct LcIcpParams { float maxDist2; int maxIter; double epsTransform, epsFitness; };
struct LcDev {
    int slots, srcCap, tgtRawCap, tgtDSCap, maxEntries;
    int* slotLane;             // [slots] lane whose key-frame pool the slot reads, -1 = idle
    int* slotActive;           // [slots]
    int* counts;               // [slots][4] nSrc, nTgtRaw, nTgtDS, nEntries
    LcEntry* entries;          // [slots][maxEntries]
    float4* src; float4* work; // [slots][srcCap] latestSurfKeyFrameCloud and the copy ICP transforms in place
    float4* tgtRaw; float4* tgtDS;   // [slots][tgtRawCap], [slots][tgtDSCap] nearHistorySurfKeyFrameCloud(DS)
    int* miss;                 // [slots][srcCap] points whose nearest neighbour lies outside the shell search
    int* nnIdx; float* nnD2;   // [slots][srcCap]
    LcResult* results;         // [slots]
    VoxelHash hash;            // over tgtDS
};

struct MoDev {
    float* transformSum;       // [L][6]  odometry pose after the message round trip (laserOdometryHandler, mapOptmization.cpp:608-621)
    float* transformIncre; float* transformTobeMapped; float* transformBefMapped; float* transformAftMapped; float* transformLast;   // [L][6]
    float* matP;               // [L][36]
    int* state;                // [L][16] 0 isDegenerate, 1 number of key frames, 2 pool points used, 3 LM iterations, 4 rows of the last iteration, 5 error bits
    float* prevRobotPos;       // [L][4]  previousRobotPosPoint
    int* active;               // [L]     host gate decision of this step (mapOptmization.cpp:1489-1499)
    float4* pool;              // [L][poolCap] corner | surf | outlier cloud of every key frame, sensor frame
    MoKfEntry* entries;        // [L][maxSel]
    int* nEntries;             // [L]
    float4* cornerFromMap; float4* surfFromMap;        // [L][cornerRawCap], [L][surfRawCap]
    float4* cornerFromMapDS; float4* surfFromMapDS;    // [L][cornerDSCap], [L][surfDSCap]
    int* mapCounts;            // [L][4]  cornerFromMap, surfFromMap, cornerFromMapDS, surfFromMapDS sizes
    float4* cornerLastDS; float4* surfLastDS; float4* outlierLastDS; float4* surfTotalLast; float4* surfTotalLastDS;
    int* scanCounts;           // [L][8]  cornerLastDS, surfLastDS, outlierLastDS, surfTotalLast, surfTotalLastDS sizes
    VoxelHash hashCornerMap, hashSurfMap;
    MoRecord* records;         // [L]
    MoOverride* overrides;     // [L]
    float* imuBlend;           // [L][4] transformUpdate's IMU blend (mapOptmization.cpp:465-490): active, imuRollLast, imuPitchLast
    double* partials;          // [L][resBlocks][8][28] per-warp partial sums of AtA / AtB / row count
    int4* knnCand;             // [L][resBlocks*256][VH_CACHE_K/4] candidate cache of the scan-to-map 5-NN (map indices per query)
    float4* knnMeta;           // [L][resBlocks*256] query position when the cache was filled, w = squared radius it covers
    int4* knnRes;              // [L][resBlocks*256][2] idx0..3 | idx4, 5th squared distance, neighbours found
    int* knnMissList;          // [L][resBlocks*256] queries to search again in this iteration
    int* knnMissCount;         // [L]
    ImCtl* imCtl;              // [L] incremental local map: this call's list diff
    int* imNewCounts;          // [L][4] points of the arriving key frames (corner, surf+outlier)
    int resBlocks;             // CTAs per lane in mo_residual_kernel
    int poolCap, maxSel, cornerRawCap, surfRawCap, cornerDSCap, surfDSCap, scanCornerCap, scanSurfCap, scanOutlierCap, maxKeyFrames;
};

// host side of one lane's mapOptimization: the key-frame list and the message gates
struct MoKeyFrame { float x, y, z, roll, pitch, yaw; int poolOff, nC, nS, nO; double time; };
struct MoLaneHost {
    std::vector<MoKeyFrame> keyFrames;                // cloudKeyPoses3D / 6D + where the clouds live
    std::vector<int> surroundingExistingKeyPosesID;   // mapOptmization.cpp:90
    float currentRobotPos[3] = {0.f, 0.f, 0.f};       // PointType default = 0 (mapOptmization.cpp:96)
    double timeLastProcessing = -1.0;                 // mapOptmization.cpp:331
    bool newCorner = false, newSurf = false, newOutlier = false, newOdometry = false;
    double timeCorner = 0, timeSurf = 0, timeOutlier = 0, timeOdometry = 0;
    // FeatureAssociation's publish cadence mirrored on the host (featureAssociation.cpp:284, 314, 1790-1794)
    bool faInited = false; int faFrameCount = 1;
    int lastError = 0;
    std::vector<int> imList;                          // key frames in the device's sorted local map, in list order
    bool imValid = false;
    // ---- loop closure state (only used when loopClosureEnableFlag is set) ----
    std::deque<int> recent;                           // recentCorner/Surf/OutlierCloudKeyFrames as key-frame ids (mapOptmization.cpp:960-1000)
    int latestFrameID = 0;                            // mapOptmization.cpp:373
    pg::Graph graph;                                  // gtSAMgraph + isam
    std::vector<pg::Pose> estimate;                   // isamCurrentEstimate: only refreshed when a key frame is saved (mapOptmization.cpp:1408)
    float transformLast[6] = {0, 0, 0, 0, 0, 0};
    float prevRobotPos[3] = {0, 0, 0};
    bool aLoopIsClosed = false;
    double timeLastLoop = -1e9;                       // performLoopClosure cadence (1 Hz thread of mapOptmization.cpp:802-812)
    int loopCalls = 0, loopsClosed = 0;
    int lastLoop[9] = {0, 0, 0, -1, -1, 0, 0, 0, 0};  // called, icp ran, closed, closestHistoryFrameID, latestFrameIDLoopCloure, converged, iterations, nSrc, nTgtDS
    float lastIcpT[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    double lastIcpFitness = 0;
};

void cuda_check(cudaError_t e, const char* what);
void set_last_error(const std::string& s);

class Engine {
public:
    Engine(const lego_params& p, int lanes, int max_points);
    ~Engine();

    // ---- ImageProjection (names follow imageProjection.cpp) ----
    void uploadScan(const float* xyzi, const uint16_t* ring, const int* npts, int stride_points, bool src_on_device);
    void uploadScanLanes(const float* const* xyz, const uint16_t* const* ring, const int* npts, int floats_per_point, bool src_on_device,
                         bool host_is_pinned = false, bool host_is_arena = false);
    void inputWait();             // input buffers of the last step may be refilled once this returns
    void inputConsumed();         // marks the point after which the current input buffers may be overwritten
    void projectPointCloud();     // + findStartEndAngle
    void projectPointCloudLaunch();   // the launches alone (no event record: capturable into a CUDA graph)
    void groundRemoval();
    void cloudSegmentation();
    void finalizeLabelMat();      // numeric labels for every cell (only needed for parity / visualisation)
    // ---- FeatureAssociation (names follow featureAssociation.cpp) ----
    void adjustDistortionCalculateSmoothnessMarkOccluded(const double* stamps = nullptr);   // stamps: timeScanCur per lane (IMU path)
    void extractFeatures();
    void updateTransformation();
    void integrateTransformationPublishCloudsLast();
    void integrateTransformationPublishCloudsLastLaunch();   // the launches alone; the caller flips faBuf
    void frontEnd(const double* stamps);   // ImageProjection + FeatureAssociation of one scan, replayed from a CUDA graph when allowed

    // ---- mapOptimization (names follow mapOptmization.cpp) ----
    void allocMapOptimization();
    VgBatch makeVgBatch(std::vector<VgSeg>& segs, int maxCap);
    VgSeg makeVgSeg(const float4* in, const int* nIn, float4* out, int* nOut, const int* active, float leaf, int cap, int outCap);
    void mapOptimizationRun(const double* stamps);   // gates + the body of mapOptimization::run for every lane that passes them
    bool mapOptimizationGates(const double* stamps, std::vector<int>& active);
    void mapOptimizationBody(const std::vector<int>& active);
    bool lookupMapOptimization(int lane, const std::string& name, void** ptr, int* elem, long* count);
    int firstCapacityError() const {                 // first lane whose mapOptimization ran out of capacity, -1 = none
        for (int l = 0; l < c.L; ++l) if (moHost[l].lastError) return l;
        return -1;
    }
    void consumeMapRecords();                        // host bookkeeping of saveKeyFramesAndFactor from the device records
    void selectSurroundingKeyFrames(int lane, std::vector<MoKfEntry>& out, int* nCorner, int* nSurf);
    // ---- IMU ----
    void imuHandler(int lane, double stamp, const double q_xyzw[4], const double angular_velocity[3], const double linear_acceleration[3]);
    void uploadImu(const double* stamps);            // ring buffers of the lanes that received messages since the last scan
    void mapOptimizationImuBlend(const std::vector<int>& active);
    std::vector<ImuLaneHost> imuHost;
    bool anyImu = false;
    // ---- loop closure (names follow mapOptmization.cpp) ----
    bool loopMode() const { return params.loopClosureEnableFlag != 0; }
    void allocLoopClosure();
    void selectRecentKeyFrames(int lane, std::vector<MoKfEntry>& out, int* nCorner, int* nSurf);   // extractSurroundingKeyFrames, loop-closure branch
    void saveKeyFramesHostGraph(const std::vector<int>& active);   // saveKeyFramesAndFactor for lanes whose graph holds a loop factor
    void correctPoses(int lane);
    void performLoopClosure(const std::vector<int>& active);
    bool detectLoopClosure(int lane, int* closest, int* latest);
    void runIcpBatch(const std::vector<int>& lanes, const std::vector<int>& closest, const std::vector<int>& latest);
    void addLoopConstraint(int lane, int closest, int latest, const LcResult& r);
    // ---- global map (publishGlobalMap / the clouds visualizeGlobalMapThread saves, mapOptmization.cpp:724-800) ----
    void globalMap(int lane, int which, std::vector<float4>& out);

    // ---- TransformFusion (names follow transformFusion.cpp) ----
    void allocTransformFusion();
    void transformFusionLaserOdometry();             // laserOdometryHandler for every lane that published odometry this scan
    void transformFusionAftMapped();                 // odomAftMappedHandler for every lane whose mapping body ran this scan
    bool lookupTransformFusion(int lane, const std::string& name, void** ptr, int* elem, long* count);

    void sync();
    // ---- timing: CUDA events on the engine's own stream (bench.py cannot see this stream through torch) ----
    int profMode = 0;                 // 0 off, 1 every kernel, 2 only kernels whose name contains profFilter
    std::string profFilter;
    struct ProfEntry { const char* name; cudaEvent_t a, b; };
    std::vector<ProfEntry> profLog;
    std::vector<cudaEvent_t> evPool;
    cudaEvent_t timerEv[8] = {};
    static constexpr int POSE_SLOTS = 8;
    float* posePin[POSE_SLOTS] = {};             // pinned: [sum 6L | aft 6L | bef 6L | flags L]
    cudaEvent_t poseEv[POSE_SLOTS] = {};
    std::vector<int> poseActive[POSE_SLOTS];
    cudaEvent_t takeEvent();
    void profBegin(const char* name);
    void profEnd();
    void profReset();
    long profRead(char* names, long cap, double* ms, long* counts, long maxk);
    void timerRecord(int slot);
    float timerElapsed(int a, int b);
    void getPoses(float* poses18, int* flags);   // per lane: transformSum | transformAftMapped | transformBefMapped
    void posesAsync(int slot);                   // enqueue the read-back of the current poses into pinned slot 0..7
    void posesWait(int slot, float* poses18, int* flags);
    long download(int lane, const char* name, void* dst, long cap, int* elem_size);
    long upload(int lane, const char* name, const void* src, long count);
    bool lookup(int lane, const std::string& name, void** ptr, int* elem, long* count);
    bool lookupFeatureAssociation(int lane, const std::string& name, void** ptr, int* elem, long* count);
    void allocFeatureAssociation();
    void allocOdometry();
    VoxelHash allocHash(int cap, float cell, const int* n, int nStride, int lanes = -1);
    bool lookupOdometry(int lane, const std::string& name, void** ptr, int* elem, long* count);

    Consts c;
    lego_params params;
    cudaStream_t stream = nullptr;
    long launches = 0;
    // ---- CUDA graphs (engine.cu): the per-scan launch sequences never change shape (all sizes live in device counters), so
    // the front end (27 launches) and the mapping body (<= 70) are captured once per value of the few host-side indices
    // that select buffers and replayed with one cudaGraphLaunch.  LEGO_GRAPH = 0 / 1 forces either; default: below 148 lanes,
    // where a scan is launch-latency bound.  Off while profiling, with IMU data (host uploads inside the sequence) and in
    // loop-closure mode (host pose graph inside the mapping body).
    struct GraphEntry { cudaGraphExec_t exec = nullptr; long launches = 0; int seen = 0; };
    std::unordered_map<unsigned long long, GraphEntry> graphs;
    int graphMode = 0;
    bool graphBroken = false;          // a capture failed once: stay eager
    long graphLaunches = 0;            // cudaGraphLaunch calls so far
    bool graphsUsable() const { return graphMode != 0 && !graphBroken && profMode == 0 && !anyImu && !loopMode(); }
    void runGraphed(unsigned long long key, const std::function<void()>& body);
    int faBuf = 0;             // which of the double-buffered "last" clouds is current (same for every lane)
    IpDev ip;
    FaDev fa;
    MoDev mo;
    TfDev tf;
    std::vector<MoLaneHost> moHost;
    VgBatch vgMap;             // 2 segments per lane: local corner / surf map
    VgBatch vgMapNew;          // the same scratch, sorting only the arriving key frames (incremental local map)
    ImBatch im = {};
    bool imEnabled = false;    // LEGO_INCR_MAP=0 (or loop closure, which rewrites key poses) keeps the full filter
    void allocIncrementalMap(const std::vector<VgSeg>& mapSegs);
    void diffSurroundingKeyFrames(int lane, const std::vector<MoKfEntry>& all, std::vector<MoKfEntry>& arriving, ImCtl* ctl, int* nCorner, int* nSurf);
    int incrementalLocalMap();   // enqueues the merge / reduce launches, returns their number
    VgBatch vgScan[2];         // per value of faBuf, 3 segments per lane: corner / surf / outlier cloud of the current scan
    VgBatch vgSurfTotal;       // 1 segment per lane
    LcDev lc = {};             // loop-closure slots (allocated only when loopClosureEnableFlag is set)
    VgBatch vgLoop;            // 1 segment per slot: downSizeFilterHistoryKeyFrames (leaf 0.4)
    char* lcPinned = nullptr;  // pinned staging: entries | counts | slotActive | slotLane | results
    MoOverride* moOverridesHost = nullptr;   // pinned [L]
    float* moAftHost = nullptr; int* moScanCountsHost = nullptr;   // pinned [L][6], [L][8]
    char* moPinned = nullptr;  // pinned staging: entries | nEntries | mapCounts | active, double buffered
    size_t moPinnedStride = 0;
    int moPinnedSlot = 0;
    cudaEvent_t moStaged[2] = {nullptr, nullptr};
    MoRecord* moRecordsHost = nullptr;   // pinned [L]
    cudaEvent_t moRecordsReady = nullptr;
    int cclTileW = 0;                // tile width of the shared-memory connected components (image_projection.cu), 0 = three-pass form
    int odoCluster = 1;              // CTAs (one thread-block cluster) per lane in fa_odometry_cluster_kernel; 1 = fa_odometry_kernel
    bool moPersistentLm = true;      // LEGO_LM_PERSISTENT=0: ten launches of {re-rank, search, fit, solve} over all lanes instead of one persistent CTA per lane
    int moLmThreads = 512;
    bool moKnnCache = true;          // LEGO_KNN_CACHE=0 searches the voxel hash in every LM iteration (parity check of the cache)
    bool moRecordsPending = false;
    bool moLastAny = false;
    std::vector<int> moActiveLast;   // per lane: the mapping body ran in the last step
    std::vector<void*> allocations;
    // host mirrors of small per-lane data
    std::vector<int> h_counts;

    template <class T> T* alloc(size_t n, bool zero = true) {
        void* p = nullptr;
        if (n == 0) n = 1;
        cuda_check(cudaMalloc(&p, n * sizeof(T)), "cudaMalloc");
        allocations.push_back(p);
        if (zero) cuda_check(cudaMemsetAsync(p, 0, n * sizeof(T), stream), "cudaMemset");
        return static_cast<T*>(p);
    }
private:
    // double-buffered input staging: the H2D copy of scan k+1 runs on copyStream while scan k is still being processed
    static constexpr int IN_SLOTS = 5;   // staging slots: the uploads of scans k+1 .. k+4 may run while scan k is processed (a mapping scan lasts ~3 uploads)
    float4* d_in_xyzi[IN_SLOTS] = {}; uint16_t* d_in_ring[IN_SLOTS] = {}; int* d_npts[IN_SLOTS] = {};
    char* d_in_blob[IN_SLOTS] = {};
    const float4** d_ptr_xyzi[IN_SLOTS] = {}; const uint16_t** d_ptr_ring[IN_SLOTS] = {};
    const float** d_src_xyz[IN_SLOTS] = {}; const uint16_t** d_src_ring[IN_SLOTS] = {};   // page-locked input mode: the callers' host pointers
    cudaStream_t copyStream = nullptr;
    cudaEvent_t evCopied[IN_SLOTS] = {}, evConsumed[IN_SLOTS] = {};
    int inBuf = 0, curIn = 0;
    bool curInOnDevice = false;
    int beginUpload(const int* npts, bool src_on_device);
    void finishUpload(int b, const int* npts, int floats_per_point);
    char* h_pin[IN_SLOTS] = {};   // pinned: per-lane pointer tables + npts of the step in flight
};

const std::string& last_error();

// pcl::VoxelGrid over a handful of key poses on the host (map_host.cu)
void host_voxel_grid(const std::vector<float4>& in, float leaf, std::vector<float4>& out);

// pcl::io::savePCDFileASCII for a PointXYZI cloud (global_map.cu); 0 on success
int write_pcd_ascii(const char* path, const float* xyzi, long n);

// device memory that lives for one call (lego_icp_align, the global-map queries)
struct DevArena {
    std::vector<void*> ptrs;
    cudaStream_t stream = nullptr;
    template <class T> T* alloc(size_t n, bool zero = true) {
        void* p = nullptr;
        if (n == 0) n = 1;
        cuda_check(cudaMalloc(&p, n * sizeof(T)), "cudaMalloc");
        ptrs.push_back(p);
        if (zero) cuda_check(cudaMemsetAsync(p, 0, n * sizeof(T), stream), "cudaMemset");
        return static_cast<T*>(p);
    }
    ~DevArena() { for (void* p : ptrs) cudaFree(p); }
};

// times the launches enqueued during its lifetime when profiling is on
struct ProfScope {
    Engine* e;
    ProfScope(Engine* eng, const char* name) : e(eng) { if (e->profMode) e->profBegin(name); }
    ~ProfScope() { if (e->profMode) e->profEnd(); }
};

}  // namespace lego
