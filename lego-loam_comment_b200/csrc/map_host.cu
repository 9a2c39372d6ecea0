// Host side of mapOptimization: device memory, the message gates of mapOptimization::run (mapOptmization.cpp:
// 1487-1499), the key-frame bookkeeping of extractSurroundingKeyFrames (mapOptmization.cpp:1001-1049) -- which is host
// code over a few hundred key POSES in the reference too -- and the launch sequence of one mapping call.  All work on
// point CLOUDS happens in kernels; the host only ever sees the small per-call MoRecord the device writes.
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include "engine.h"

namespace lego {

// kernels of map_optimization.cu
__global__ void mo_begin_kernel(FaDev f, MoDev m, int L);
__global__ void mo_assemble_map_kernel(MoDev m);
__global__ void mo_concat_kernel(MoDev m, int surfCap, int outlierCap, int totalCap);
__global__ void mo_lm_begin_kernel(MoDev m, int L);
__global__ void mo_rerank_kernel(MoDev m);
__global__ void mo_search_kernel(MoDev m, int iter);
__global__ void mo_residual_kernel(MoDev m);
__global__ void mo_solve_kernel(MoDev m, int iter);
__global__ void mo_save_keyframe_kernel(MoDev m);
__global__ void mo_lm_persistent_kernel(MoDev m, int useCache);

int mo_lm_threads();                 // map_optimization.cu: block size the persistent LM kernel was compiled for
cudaError_t mo_lm_set_smem(size_t bytes);

static int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v && *v ? std::atoi(v) : dflt;
}

VgBatch Engine::makeVgBatch(std::vector<VgSeg>& segs, int maxCap) {
    VgBatch b; b.nSeg = (int)segs.size(); b.maxBlocks = (maxCap + VG_TILE - 1) / VG_TILE;
    for (auto& g : segs) {
        g.keys[0] = alloc<unsigned>(g.cap, false); g.keys[1] = alloc<unsigned>(g.cap, false);
        g.vals[0] = alloc<unsigned>(g.cap, false); g.vals[1] = alloc<unsigned>(g.cap, false);
        g.blockHist = alloc<int>((size_t)256 * b.maxBlocks); g.blockHeads = alloc<int>(b.maxBlocks); g.state = alloc<int>(16);
    }
    b.segs = alloc<VgSeg>(segs.size());
    cuda_check(cudaMemcpyAsync(b.segs, segs.data(), sizeof(VgSeg) * segs.size(), cudaMemcpyHostToDevice, stream), "vg table");
    cuda_check(cudaStreamSynchronize(stream), "vg table sync");
    return b;
}
VgSeg Engine::makeVgSeg(const float4* in, const int* nIn, float4* out, int* nOut, const int* active, float leaf, int cap, int outCap) {
    VgSeg g; std::memset(&g, 0, sizeof(g));
    g.in = in; g.nIn = nIn; g.out = out; g.nOut = nOut; g.active = active; g.leaf = leaf; g.cap = cap; g.outCap = outCap;
    return g;
}

void Engine::allocMapOptimization() {
    const size_t L = c.L, N = c.N, R = c.R;
    mo.maxKeyFrames = params.maxKeyFrames > 0 ? params.maxKeyFrames : env_int("LEGO_MAX_KEYFRAMES", 640);
    moKnnCache = env_int("LEGO_KNN_CACHE", 1) != 0;
    // One persistent block per lane needs lanes to fill the GPU: measured equal to the launch loop at 592 lanes (with 39 launches
    // fewer per mapping call and no empty tail iterations), 1.7x slower at one lane, where the loop spreads a lane's queries
    // over 16 blocks.  LEGO_LM_PERSISTENT = 0 / 1 forces either.
    const int persist = env_int("LEGO_LM_PERSISTENT", -1);
    moPersistentLm = persist < 0 ? c.L >= 148 : persist != 0;
    if (moPersistentLm) {
        moLmThreads = mo_lm_threads();
        cuda_check(mo_lm_set_smem(sizeof(int) * (size_t)VH_LIST_CELLS * moLmThreads), "smem attr");
    }
    const size_t perKf = std::min<size_t>(N / 4, 16384);
    mo.poolCap = (int)std::min<size_t>((size_t)mo.maxKeyFrames * perKf, (size_t)INT_MAX / 2);
    mo.maxSel = 256;
    mo.cornerRawCap = (int)(4 * N); mo.surfRawCap = (int)(16 * N);
    mo.cornerDSCap = (int)(2 * N); mo.surfDSCap = (int)(4 * N);
    mo.scanCornerCap = (int)(R * 120); mo.scanSurfCap = (int)N; mo.scanOutlierCap = c.Ocap;
    mo.transformSum = alloc<float>(L * 6); mo.transformIncre = alloc<float>(L * 6); mo.transformTobeMapped = alloc<float>(L * 6);
    mo.transformBefMapped = alloc<float>(L * 6); mo.transformAftMapped = alloc<float>(L * 6); mo.transformLast = alloc<float>(L * 6);
    mo.matP = alloc<float>(L * 36); mo.state = alloc<int>(L * 16); mo.prevRobotPos = alloc<float>(L * 4); mo.active = alloc<int>(L);
    mo.pool = alloc<float4>(L * mo.poolCap, false);
    mo.entries = alloc<MoKfEntry>(L * mo.maxSel); mo.nEntries = alloc<int>(L);
    mo.cornerFromMap = alloc<float4>(L * mo.cornerRawCap, false); mo.surfFromMap = alloc<float4>(L * mo.surfRawCap, false);
    mo.cornerFromMapDS = alloc<float4>(L * mo.cornerDSCap, false); mo.surfFromMapDS = alloc<float4>(L * mo.surfDSCap, false);
    mo.mapCounts = alloc<int>(L * 4);
    mo.cornerLastDS = alloc<float4>(L * mo.scanCornerCap); mo.surfLastDS = alloc<float4>(L * mo.scanSurfCap);
    mo.outlierLastDS = alloc<float4>(L * mo.scanOutlierCap);
    const size_t totCap = (size_t)mo.scanSurfCap + mo.scanOutlierCap;
    mo.surfTotalLast = alloc<float4>(L * totCap); mo.surfTotalLastDS = alloc<float4>(L * totCap);
    mo.scanCounts = alloc<int>(L * 8);
    mo.hashCornerMap = allocHash(mo.cornerDSCap, 1.0f, mo.mapCounts + 2, 4); mo.hashSurfMap = allocHash(mo.surfDSCap, 1.0f, mo.mapCounts + 3, 4);
    mo.records = alloc<MoRecord>(L);
    mo.overrides = alloc<MoOverride>(L);
    mo.imuBlend = alloc<float>(L * 4);
    mo.resBlocks = (mo.scanCornerCap + mo.scanSurfCap + mo.scanOutlierCap + 255) / 256;
    mo.partials = alloc<double>(L * mo.resBlocks * 8 * 28);   // per 32-query warp slice
    const size_t nSlots = L * mo.resBlocks * MO_RES_THREADS;           // one per query of every lane
    mo.knnCand = alloc<int4>(nSlots * (VH_CACHE_K / 4), false);    // cached neighbour candidates per query (voxel_hash.cuh)
    mo.knnMeta = alloc<float4>(nSlots, false);      // where the query was when its cache was filled, w = squared radius the candidates cover
    mo.knnRes = alloc<int4>(nSlots * 2, false);     // this iteration's 5 neighbours, 5th distance, count
    mo.knnMissList = alloc<int>(nSlots, false);     // queries whose cache could not prove the answer
    mo.knnMissCount = alloc<int>(L);
    moHost.assign(c.L, MoLaneHost());

    // ---- VoxelGrid segment tables ----
    auto make_batch = [&](std::vector<VgSeg>& segs, int maxCap) { return makeVgBatch(segs, maxCap); };
    auto seg = [&](const float4* in, const int* nIn, float4* out, int* nOut, const int* active, float leaf, int cap, int outCap) {
        return makeVgSeg(in, nIn, out, nOut, active, leaf, cap, outCap);
    };
    // local map: leaf 0.2 for corners, 0.4 for surfaces (mapOptmization.cpp:249-250); the scratch of these segments is
    // shared between the two faBuf variants because only one of them is ever in flight
    std::vector<VgSeg> mapSegs;
    for (size_t l = 0; l < L; ++l) {
        mapSegs.push_back(seg(mo.cornerFromMap + l * mo.cornerRawCap, mo.mapCounts + l * 4 + 0, mo.cornerFromMapDS + l * mo.cornerDSCap,
                              mo.mapCounts + l * 4 + 2, mo.active + l, 0.2f, mo.cornerRawCap, mo.cornerDSCap));
        mapSegs.push_back(seg(mo.surfFromMap + l * mo.surfRawCap, mo.mapCounts + l * 4 + 1, mo.surfFromMapDS + l * mo.surfDSCap,
                              mo.mapCounts + l * 4 + 3, mo.active + l, 0.4f, mo.surfRawCap, mo.surfDSCap));
    }
    vgMap = make_batch(mapSegs, mo.surfRawCap);
    allocIncrementalMap(mapSegs);   // (make_batch filled in the scratch pointers of mapSegs)
    // current scan (mapOptmization.cpp:1067-1081): the inputs are FeatureAssociation's "last" clouds, double buffered
    for (int b = 0; b < 2; ++b) {
        std::vector<VgSeg> s;
        for (size_t l = 0; l < L; ++l) {
            s.push_back(seg(fa.cornerLast[b] + l * R * 120, fa.lastCounts + l * 4 + 0, mo.cornerLastDS + l * mo.scanCornerCap,
                            mo.scanCounts + l * 8 + 0, mo.active + l, 0.2f, (int)(R * 120), mo.scanCornerCap));
            s.push_back(seg(fa.surfLast[b] + l * N, fa.lastCounts + l * 4 + 1, mo.surfLastDS + l * mo.scanSurfCap,
                            mo.scanCounts + l * 8 + 1, mo.active + l, 0.4f, (int)N, mo.scanSurfCap));
            s.push_back(seg(fa.outlierLast + l * c.Ocap, ip.counts + l * 8 + 1, mo.outlierLastDS + l * mo.scanOutlierCap,
                            mo.scanCounts + l * 8 + 2, mo.active + l, 0.4f, c.Ocap, mo.scanOutlierCap));
        }
        vgScan[b] = make_batch(s, (int)N);
    }
    std::vector<VgSeg> tot;
    for (size_t l = 0; l < L; ++l)
        tot.push_back(seg(mo.surfTotalLast + l * totCap, mo.scanCounts + l * 8 + 3, mo.surfTotalLastDS + l * totCap, mo.scanCounts + l * 8 + 4,
                          mo.active + l, 0.4f, (int)totCap, (int)totCap));
    vgSurfTotal = make_batch(tot, (int)totCap);

    // ---- pinned staging for the per-call selection and the records ----
    moPinnedStride = sizeof(MoKfEntry) * L * mo.maxSel + sizeof(int) * L * 6 + sizeof(ImCtl) * L;
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&moPinned), 2 * moPinnedStride), "cudaMallocHost");
    cuda_check(cudaMallocHost(reinterpret_cast<void**>(&moRecordsHost), sizeof(MoRecord) * L), "cudaMallocHost");
    for (int b = 0; b < 2; ++b) cuda_check(cudaEventCreateWithFlags(&moStaged[b], cudaEventDisableTiming), "event");
    cuda_check(cudaEventCreateWithFlags(&moRecordsReady, cudaEventDisableTiming), "event");
    allocLoopClosure();
}

bool Engine::lookupMapOptimization(int lane, const std::string& n, void** ptr, int* elem, long* count) {
    const size_t l = lane;
    auto set = [&](void* p, int e, long cnt) { *ptr = p; *elem = e; *count = cnt; return true; };
    auto cnt_of = [&](int* base, int stride, int which) {
        int v = 0;
        cuda_check(cudaMemcpyAsync(&v, base + l * stride + which, sizeof(int), cudaMemcpyDeviceToHost, stream), "count");
        cuda_check(cudaStreamSynchronize(stream), "count sync");
        return (long)v;
    };
    if (n == "moTransformSum") return set(mo.transformSum + l * 6, 4, 6);
    if (n == "transformTobeMapped") return set(mo.transformTobeMapped + l * 6, 4, 6);
    if (n == "transformAftMapped") return set(mo.transformAftMapped + l * 6, 4, 6);
    if (n == "transformBefMapped") return set(mo.transformBefMapped + l * 6, 4, 6);
    if (n == "moState") return set(mo.state + l * 16, 4, 16);
    if (n == "mapCounts") return set(mo.mapCounts + l * 4, 4, 4);
    if (n == "scanCounts") return set(mo.scanCounts + l * 8, 4, 8);
    if (n == "laserCloudCornerFromMap") return set(mo.cornerFromMap + l * mo.cornerRawCap, 16, cnt_of(mo.mapCounts, 4, 0));
    if (n == "laserCloudSurfFromMap") return set(mo.surfFromMap + l * mo.surfRawCap, 16, cnt_of(mo.mapCounts, 4, 1));
    if (n == "laserCloudCornerFromMapDS") return set(mo.cornerFromMapDS + l * mo.cornerDSCap, 16, cnt_of(mo.mapCounts, 4, 2));
    if (n == "laserCloudSurfFromMapDS") return set(mo.surfFromMapDS + l * mo.surfDSCap, 16, cnt_of(mo.mapCounts, 4, 3));
    if (n == "laserCloudCornerLastDS") return set(mo.cornerLastDS + l * mo.scanCornerCap, 16, cnt_of(mo.scanCounts, 8, 0));
    if (n == "laserCloudSurfLastDS") return set(mo.surfLastDS + l * mo.scanSurfCap, 16, cnt_of(mo.scanCounts, 8, 1));
    if (n == "laserCloudOutlierLastDS") return set(mo.outlierLastDS + l * mo.scanOutlierCap, 16, cnt_of(mo.scanCounts, 8, 2));
    if (n == "laserCloudSurfTotalLastDS")
        return set(mo.surfTotalLastDS + l * ((size_t)mo.scanSurfCap + mo.scanOutlierCap), 16, cnt_of(mo.scanCounts, 8, 4));
    return lookupTransformFusion(lane, n, ptr, elem, count);
}

// ------------------------------------------------------------------------------------------------------------
// pcl::VoxelGrid(leaf) over a handful of key poses, on the host (downSizeFilterSurroundingKeyPoses,
// mapOptmization.cpp:1009-1010).  Same arithmetic as the device filter (voxel_grid.cu): note that the intensity
// channel -- the key-frame index -- is averaged with the coordinates (quirk Q16).
void host_voxel_grid(const std::vector<float4>& in, float leaf, std::vector<float4>& out) {
    out.clear();
    if (in.empty()) return;
    const float inv = 1.0f / leaf;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (const float4& p : in) {
        const float v[3] = {p.x, p.y, p.z};
        for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], v[a]); mx[a] = std::max(mx[a], v[a]); }
    }
    long long dd[3]; int minb[3], divb[3];
    for (int a = 0; a < 3; ++a) {
        dd[a] = (long long)((mx[a] - mn[a]) * inv) + 1;
        minb[a] = (int)std::floor(mn[a] * inv);
        divb[a] = (int)std::floor(mx[a] * inv) - minb[a] + 1;
    }
    if (dd[0] * dd[1] * dd[2] > (long long)INT_MAX) { out = in; return; }
    std::vector<std::pair<int, int> > keys(in.size());
    for (size_t i = 0; i < in.size(); ++i) {
        const int i0 = (int)(std::floor(in[i].x * inv) - (float)minb[0]);
        const int i1 = (int)(std::floor(in[i].y * inv) - (float)minb[1]);
        const int i2 = (int)(std::floor(in[i].z * inv) - (float)minb[2]);
        keys[i] = std::make_pair(i0 + i1 * divb[0] + i2 * divb[0] * divb[1], (int)i);
    }
    std::sort(keys.begin(), keys.end());   // (voxel, input position): members keep their input order
    for (size_t a = 0; a < keys.size();) {
        size_t b = a;
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        for (; b < keys.size() && keys[b].first == keys[a].first; ++b) {
            const float4& p = in[keys[b].second];
            s[0] += p.x; s[1] += p.y; s[2] += p.z; s[3] += p.w;
        }
        const float cnt = (float)(b - a);
        out.push_back(make_float4(s[0] / cnt, s[1] / cnt, s[2] / cnt, s[3] / cnt));
        a = b;
    }
}

// extractSurroundingKeyFrames, loop closure off (mapOptmization.cpp:1001-1056): which key frames make up the local map.
void Engine::selectSurroundingKeyFrames(int lane, std::vector<MoKfEntry>& out, int* nCorner, int* nSurf) {
    MoLaneHost& H = moHost[lane];
    out.clear(); *nCorner = 0; *nSurf = 0;
    if (H.keyFrames.empty()) return;
    // kdtreeSurroundingKeyPoses->radiusSearch(currentRobotPosPoint, 50.0, ...): ascending squared distance
    const float r = (float)(double)params.surroundingKeyframeSearchRadius, r2 = r * r;
    std::vector<std::pair<float, int> > hits;
    for (size_t i = 0; i < H.keyFrames.size(); ++i) {
        const float dx = H.keyFrames[i].x - H.currentRobotPos[0], dy = H.keyFrames[i].y - H.currentRobotPos[1], dz = H.keyFrames[i].z - H.currentRobotPos[2];
        const float d = dx * dx + dy * dy + dz * dz;
        if (d <= r2) hits.push_back(std::make_pair(d, (int)i));
    }
    std::sort(hits.begin(), hits.end());
    std::vector<float4> surrounding, ds;
    for (auto& h : hits) { const MoKeyFrame& k = H.keyFrames[h.second]; surrounding.push_back(make_float4(k.x, k.y, k.z, (float)h.second)); }
    host_voxel_grid(surrounding, 1.0f, ds);
    // drop cached key frames that left the neighbourhood, then append the new ones (mapOptmization.cpp:1013-1049)
    std::vector<int>& ex = H.surroundingExistingKeyPosesID;
    for (size_t i = 0; i < ex.size();) {
        bool keep = false;
        for (const float4& p : ds) if (ex[i] == (int)p.w) { keep = true; break; }
        if (keep) ++i; else ex.erase(ex.begin() + i);
    }
    for (const float4& p : ds) {
        const int id = (int)p.w;
        if (std::find(ex.begin(), ex.end(), id) == ex.end()) ex.push_back(id);
    }
    if ((int)ex.size() > mo.maxSel) { H.lastError |= 2; ex.resize(mo.maxSel); }
    int offC = 0, offS = 0;
    for (int id : ex) {
        const MoKeyFrame& k = H.keyFrames[id];
        MoKfEntry e;
        e.id = id;
        e.srcC = k.poolOff; e.nC = k.nC; e.dstC = offC; offC += k.nC;
        e.srcS = k.poolOff + k.nC; e.nS = k.nS; e.dstS = offS; offS += k.nS;
        e.srcO = k.poolOff + k.nC + k.nS; e.nO = k.nO; e.dstO = offS; offS += k.nO;
        // updateTransformPointCloudSinCos (mapOptmization.cpp:529-543): float sin/cos of the stored key pose
        e.ctRoll = cosf(k.roll); e.stRoll = sinf(k.roll); e.ctPitch = cosf(k.pitch); e.stPitch = sinf(k.pitch);
        e.ctYaw = cosf(k.yaw); e.stYaw = sinf(k.yaw); e.tInX = k.x; e.tInY = k.y; e.tInZ = k.z;
        if (offC > mo.cornerRawCap || offS > mo.surfRawCap) { H.lastError |= 4; offC -= k.nC; offS -= k.nS + k.nO; break; }
        out.push_back(e);
    }
    *nCorner = offC; *nSurf = offS;
}

// The device wrote one MoRecord per lane at the end of a mapping call; fold it into the host key-frame lists (and, with
// loop closure on, into the lane's pose graph: the prior / chain factors of mapOptmization.cpp:1375-1392).
void Engine::consumeMapRecords() {
    if (!moRecordsPending) return;
    cuda_check(cudaEventSynchronize(moRecordsReady), "records sync");
    moRecordsPending = false;
    static const double kVar[6] = {1e-6, 1e-6, 1e-6, 1e-8, 1e-8, 1e-6};   // priorNoise = odometryNoise (mapOptmization.cpp:346-350)
    for (int l = 0; l < c.L; ++l) {
        const MoRecord& r = moRecordsHost[l];
        if (!r.ran) continue;
        MoLaneHost& H = moHost[l];
        for (int a = 0; a < 3; ++a) H.currentRobotPos[a] = r.robotPos[a];
        H.lastError |= r.error;
        if (r.saved) {
            MoKeyFrame k;
            k.x = r.pose6D[0]; k.y = r.pose6D[1]; k.z = r.pose6D[2]; k.roll = r.pose6D[3]; k.pitch = r.pose6D[4]; k.yaw = r.pose6D[5];
            k.poolOff = r.poolOff; k.nC = r.nC; k.nS = r.nS; k.nO = r.nO; k.time = H.timeOdometry;
            if (loopMode() && !H.graph.hasLoop) {   // (lanes with a loop factor were handled by saveKeyFramesHostGraph)
                const int n = (int)H.keyFrames.size();
                const float* a = r.aftRaw;
                const pg::Pose to = pg::make_pose((double)a[2], (double)a[0], (double)a[1], (double)a[5], (double)a[3], (double)a[4]);
                if (n == 0) {
                    H.graph.addPrior(0, to, kVar);
                    for (int i = 0; i < 6; ++i) H.transformLast[i] = a[i];
                } else {
                    const float* t = H.transformLast;
                    const pg::Pose from = pg::make_pose((double)t[2], (double)t[0], (double)t[1], (double)t[5], (double)t[3], (double)t[4]);
                    H.graph.addBetween(n - 1, n, pg::between(from, to), kVar);
                    H.transformLast[0] = k.roll; H.transformLast[1] = k.pitch; H.transformLast[2] = k.yaw;
                    H.transformLast[3] = k.x; H.transformLast[4] = k.y; H.transformLast[5] = k.z;
                }
                H.graph.insert(n, to);
                H.estimate = H.graph.x;
                for (int i = 0; i < 3; ++i) H.prevRobotPos[i] = r.robotPos[i];
            }
            H.keyFrames.push_back(k);
        }
        if (loopMode()) correctPoses(l);
    }
}

// Message gates of FeatureAssociation's publishers and of mapOptimization::run (mapOptmization.cpp:1489-1499) for every
// lane.  stamps: one per lane (the scan that FeatureAssociation just finished).  active[l] = 1 where the body must run.
bool Engine::mapOptimizationGates(const double* stamps, std::vector<int>& active) {
    active.assign(c.L, 0);
    bool any = false;
    for (int l = 0; l < c.L; ++l) {
        MoLaneHost& H = moHost[l];
        const double t = stamps[l];
        if (!H.faInited) {            // checkSystemInitialization publishes corner + surf only (featureAssociation.cpp:1605-1637)
            H.faInited = true;
            H.newCorner = H.newSurf = true; H.timeCorner = H.timeSurf = t;
        } else {
            H.newOdometry = true; H.timeOdometry = t;                       // publishOdometry, every scan
            if (++H.faFrameCount >= 2) {                                    // publishCloudsLast, skipFrameNum = 1
                H.faFrameCount = 0;
                H.newCorner = H.newSurf = H.newOutlier = true; H.timeCorner = H.timeSurf = H.timeOutlier = t;
            }
        }
        if (H.newCorner && std::abs(H.timeCorner - H.timeOdometry) < 0.005 && H.newSurf && std::abs(H.timeSurf - H.timeOdometry) < 0.005 &&
            H.newOutlier && std::abs(H.timeOutlier - H.timeOdometry) < 0.005 && H.newOdometry) {
            H.newCorner = H.newSurf = H.newOutlier = H.newOdometry = false;
            if (H.timeOdometry - H.timeLastProcessing >= params.mappingProcessInterval) {
                H.timeLastProcessing = H.timeOdometry;
                active[l] = 1; any = true;
            }
        }
    }
    return any;
}

void Engine::mapOptimizationRun(const double* stamps) {
    std::vector<int> active;
    const bool any = mapOptimizationGates(stamps, active);
    moLastAny = any;
    moActiveLast = active;
    if (any) mapOptimizationBody(active);
}

// The body of mapOptimization::run (mapOptmization.cpp:1500-1519) for the lanes flagged in `active`.
void Engine::mapOptimizationBody(const std::vector<int>& active) {
    consumeMapRecords();

    // ---- stage this call's selection in pinned memory and send it ----
    const int slot = moPinnedSlot; moPinnedSlot ^= 1;
    cuda_check(cudaEventSynchronize(moStaged[slot]), "staging sync");
    char* base = moPinned + (size_t)slot * moPinnedStride;
    MoKfEntry* hEntries = reinterpret_cast<MoKfEntry*>(base);
    int* hInts = reinterpret_cast<int*>(base + sizeof(MoKfEntry) * (size_t)c.L * mo.maxSel);   // nEntries[L] | active[L] | mapCounts[L][4]
    ImCtl* hCtl = reinterpret_cast<ImCtl*>(hInts + 6 * (size_t)c.L);
    int maxEntries = 0;
    std::vector<MoKfEntry> sel, arriving;
    for (int l = 0; l < c.L; ++l) {
        int nC = 0, nS = 0;
        if (!active[l]) sel.clear();
        else if (loopMode()) selectRecentKeyFrames(l, sel, &nC, &nS);
        else selectSurroundingKeyFrames(l, sel, &nC, &nS);
        if (imEnabled) {
            // only the key frames that entered the list are assembled; the device merges them into its sorted local map
            if (active[l]) { diffSurroundingKeyFrames(l, sel, arriving, hCtl + l, &nC, &nS); sel.swap(arriving); }
            else std::memset(hCtl + l, 0, sizeof(ImCtl));
        }
        std::memcpy(hEntries + (size_t)l * mo.maxSel, sel.data(), sizeof(MoKfEntry) * sel.size());
        hInts[l] = (int)sel.size(); hInts[c.L + l] = active[l];
        hInts[2 * c.L + l * 4 + 0] = nC; hInts[2 * c.L + l * 4 + 1] = nS; hInts[2 * c.L + l * 4 + 2] = 0; hInts[2 * c.L + l * 4 + 3] = 0;
        maxEntries = std::max(maxEntries, (int)sel.size());
    }
    // (only the live part of the entry table travels: maxEntries rows per lane would need a strided copy, the whole table is 2.7 MB)
    cuda_check(cudaMemcpyAsync(mo.entries, hEntries, sizeof(MoKfEntry) * (size_t)c.L * mo.maxSel, cudaMemcpyHostToDevice, stream), "entries");
    cuda_check(cudaMemcpyAsync(mo.nEntries, hInts, sizeof(int) * c.L, cudaMemcpyHostToDevice, stream), "nEntries");
    cuda_check(cudaMemcpyAsync(mo.active, hInts + c.L, sizeof(int) * c.L, cudaMemcpyHostToDevice, stream), "active");
    if (imEnabled) {
        cuda_check(cudaMemcpyAsync(mo.imNewCounts, hInts + 2 * c.L, sizeof(int) * 4 * c.L, cudaMemcpyHostToDevice, stream), "imNewCounts");
        cuda_check(cudaMemcpyAsync(mo.imCtl, hCtl, sizeof(ImCtl) * c.L, cudaMemcpyHostToDevice, stream), "imCtl");
    } else {
        cuda_check(cudaMemcpyAsync(mo.mapCounts, hInts + 2 * c.L, sizeof(int) * 4 * c.L, cudaMemcpyHostToDevice, stream), "mapCounts");
    }
    cuda_check(cudaEventRecord(moStaged[slot], stream), "event record");

    mapOptimizationImuBlend(active);
    // ---- the mapping body ----
    // One fixed launch sequence; the host enters it only through the parity of the "last" clouds and the grid of the assemble
    // kernel, so it is replayed from a CUDA graph where graphs are in use (engine.h: runGraphed).
    const std::function<void()> body = [this, maxEntries, &active]() {
    { ProfScope ps(this, "mo_begin_kernel"); mo_begin_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(fa, mo, c.L); }
    if (maxEntries > 0) {
        dim3 g(std::min(maxEntries, 64), c.L);
        ProfScope ps(this, "mo_assemble_map_kernel");
        mo_assemble_map_kernel<<<g, 256, 0, stream>>>(mo);
        launches += 1;
    }
    if (imEnabled) launches += incrementalLocalMap();
    else { ProfScope ps(this, "vg_filter(localMap)"); launches += vg_filter(vgMap, stream); }
    { ProfScope ps(this, "vg_filter(currentScan)"); launches += vg_filter(vgScan[faBuf], stream); }
    {
        dim3 g(8, c.L);
        ProfScope ps(this, "mo_concat_kernel");
        mo_concat_kernel<<<g, 256, 0, stream>>>(mo, mo.scanSurfCap, mo.scanOutlierCap, mo.scanSurfCap + mo.scanOutlierCap);
    }
    { ProfScope ps(this, "vg_filter(surfTotal)"); launches += vg_filter(vgSurfTotal, stream); }
    { ProfScope ps(this, "vh_build(cornerMap)"); launches += vh_build(mo.hashCornerMap, c.L, mo.cornerFromMapDS, (size_t)mo.cornerDSCap, stream); }
    { ProfScope ps(this, "vh_build(surfMap)"); launches += vh_build(mo.hashSurfMap, c.L, mo.surfFromMapDS, (size_t)mo.surfDSCap, stream); }
    {   // scan2MapOptimization: <= 10 x {residuals, solve}; a converged lane makes its share of later launches a no-op
        { ProfScope ps(this, "mo_lm_begin_kernel"); mo_lm_begin_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(mo, c.L); }
        if (moPersistentLm) {
            // one persistent CTA per lane runs all iterations (map_optimization.cu: mo_lm_persistent_kernel)
            const int threads = moLmThreads;
            const size_t smem = sizeof(int) * (size_t)VH_LIST_CELLS * threads;
            ProfScope ps(this, "mo_lm_persistent_kernel");
            mo_lm_persistent_kernel<<<c.L, threads, smem, stream>>>(mo, moKnnCache ? 1 : 0);
            launches += 2;
        } else {
        dim3 g(std::min(mo.resBlocks, MO_RES_INFLIGHT), c.L);   // 16 x 256 queries per lane in flight; larger scans loop
#ifndef LEGO_MO_GRID_LATER
#define LEGO_MO_GRID_LATER 8
#endif
        // later iterations mostly find lanes that are done or few misses: a smaller grid keeps the no-op launches cheap
        dim3 gq(std::min(mo.resBlocks, 32), c.L), gLater(std::min(mo.resBlocks, LEGO_MO_GRID_LATER), c.L);
        for (int iter = 0; iter < 10; ++iter) {
            {   // cornerOptimization + surfOptimization: re-rank the cached neighbours, search the rest, fit (map_optimization.cu)
#ifdef LEGO_KNN_ITER_NAMES
                static const char* kR[10] = {"rerank#0", "rerank#1", "rerank#2", "rerank#3", "rerank#4", "rerank#5", "rerank#6", "rerank#7", "rerank#8", "rerank#9"};
                static const char* kS[10] = {"search#0", "search#1", "search#2", "search#3", "search#4", "search#5", "search#6", "search#7", "search#8", "search#9"};
                static const char* kF[10] = {"fit#0", "fit#1", "fit#2", "fit#3", "fit#4", "fit#5", "fit#6", "fit#7", "fit#8", "fit#9"};
                const char *nR = kR[iter], *nS = kS[iter], *nF = kF[iter];
#else
                const char *nR = "mo_rerank_kernel", *nS = "mo_search_kernel", *nF = "mo_residual_kernel";
#endif
                const bool all = iter == 0 || !moKnnCache;   // search every query / only those the cache could not answer
                if (!all) { ProfScope ps(this, nR); mo_rerank_kernel<<<gLater, 256, 0, stream>>>(mo); }
                { ProfScope ps(this, nS); mo_search_kernel<<<all ? gq : gLater, 256, 0, stream>>>(mo, all ? 0 : iter); }
                { ProfScope ps(this, nF); mo_residual_kernel<<<g, MO_RES_THREADS, 0, stream>>>(mo); }
            }
            { ProfScope ps(this, "mo_solve_kernel"); mo_solve_kernel<<<c.L, 32, 0, stream>>>(mo, iter); }
        }
        launches += 41;
        }
    }
    if (loopMode()) saveKeyFramesHostGraph(active);   // lanes whose graph holds a loop factor: gate + iSAM update on the host
    { ProfScope ps(this, "mo_save_keyframe_kernel"); mo_save_keyframe_kernel<<<c.L, 256, 0, stream>>>(mo); }
    launches += 3;
    };
    if (graphsUsable())
        runGraphed(2ull | ((unsigned long long)faBuf << 8) | ((unsigned long long)std::min(maxEntries, 64) << 16) |
                   ((unsigned long long)(moPersistentLm ? 1 : 0) << 32) | ((unsigned long long)(imEnabled ? 1 : 0) << 33), body);
    else body();
    cuda_check(cudaMemcpyAsync(moRecordsHost, mo.records, sizeof(MoRecord) * c.L, cudaMemcpyDeviceToHost, stream), "records");
    cuda_check(cudaEventRecord(moRecordsReady, stream), "event record");
    moRecordsPending = true;
    if (loopMode()) {
        // the loop-closure search needs this call's key frame: take the records now instead of one call later
        consumeMapRecords();            // + correctPoses
        performLoopClosure(active);     // detect -> submap + ICP on the device -> constraint on the host graph
    }
}

}  // namespace lego
