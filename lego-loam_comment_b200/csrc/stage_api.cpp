// Per-stage entry points of include/lego_b200.h: one create / process / destroy triple per reference node, each the
// drop-in for the body of that node's per-scan method
//   lego_ip_process  <- ImageProjection::cloudHandler             imageProjection.cpp:181-197
//   lego_fa_process  <- FeatureAssociation::runFeatureAssociation  featureAssociation.cpp:1817-1860
//   lego_mo_process  <- mapOptimization::run                       mapOptmization.cpp:1487-1522
// Every handle is a one-lane lego::Engine; inputs may be host or device memory (cudaMemcpyDefault resolves which),
// outputs stay on the device with host mirrors filled, as the views in the header describe.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <stdexcept>
#include <vector>
#include "engine.h"

using lego::Engine;
using lego::cuda_check;

namespace {

struct HostCloud { std::vector<float> v; };

void fetch_cloud(Engine& e, const float4* dev, int n, std::vector<float>& host, lego_cloud_view* view) {
    host.resize((size_t)4 * (n > 0 ? n : 1));
    if (n > 0) cuda_check(cudaMemcpyAsync(host.data(), dev, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, e.stream), "fetch cloud");
    if (view) { view->dev = reinterpret_cast<const float*>(dev); view->host = host.data(); view->count = n; }
}
void put(Engine& e, void* dst, const void* src, size_t bytes) {
    if (bytes) cuda_check(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, e.stream), "stage input copy");
}
int read_int(Engine& e, const int* dev) {
    int v = 0;
    cuda_check(cudaMemcpyAsync(&v, dev, sizeof(int), cudaMemcpyDeviceToHost, e.stream), "count");
    cuda_check(cudaStreamSynchronize(e.stream), "count sync");
    return v;
}
bool params_ok(const lego_params* p) {
    return p && p->N_SCAN >= 2 && p->N_SCAN <= 128 && p->Horizon_SCAN >= 16;
}

}  // namespace

struct lego_ip {
    Engine* e;
    std::vector<float> seg, outl, range, filtered;
    std::vector<uint16_t> filteredRing;
    std::vector<int32_t> start, end;
    std::vector<uint8_t> ground;
    std::vector<uint32_t> col;
};
struct lego_fa {
    Engine* e;
    std::vector<float> sharp, lessSharp, flat, lessFlat, cornerLast, surfLast, outlierLast;
};
struct lego_mo {
    Engine* e;
    double timeLastProcessing;
};

#define LEGO_TRY try {
#define LEGO_CATCH \
    } catch (const std::length_error& ex) { lego::set_last_error(ex.what()); return LEGO_ERR_CAPACITY; } \
      catch (const std::bad_alloc&) { lego::set_last_error("out of host memory"); return LEGO_ERR_CUDA; } \
      catch (const std::exception& ex) { if (lego::last_error().empty()) lego::set_last_error(ex.what()); return LEGO_ERR_CUDA; }

extern "C" {

// ------------------------------------------------------------------------------------------------ ImageProjection
int lego_ip_create(const lego_params* p, lego_ip** out) {
    if (!params_ok(p) || !out) return LEGO_ERR_ARG;
    LEGO_TRY
    lego_ip* h = new lego_ip();
    h->e = new Engine(*p, 1, 0);
    *out = h;
    return LEGO_OK;
    LEGO_CATCH
}
int lego_ip_destroy(lego_ip* h) {
    if (!h) return LEGO_ERR_ARG;
    delete h->e; delete h;
    return LEGO_OK;
}
int lego_ip_process(lego_ip* h, const float* xyzi, const uint16_t* ring, int P, int is_dense, double stamp,
                    lego_cloud_info_view* info, lego_cloud_view* segmented, lego_cloud_view* outlier) {
    (void)stamp;
    if (!h || !xyzi || P < 0 || !info || !segmented || !outlier) return LEGO_ERR_ARG;
    Engine& e = *h->e;
    if (e.c.useRing) {
        if (!ring) return LEGO_ERR_ARG;
        // the reference logs "Point cloud is not in dense format" and shuts the node down (imageProjection.cpp:172-177)
        if (!is_dense) { lego::set_last_error("point cloud is not in dense format, remove NaN points first"); return LEGO_ERR_NOT_DENSE; }
    }
    LEGO_TRY
    if (!is_dense) {
        // pcl::removeNaNFromPointCloud (imageProjection.cpp:168): the message is decoded into a cloud without the
        // non-finite points, order kept.  Message decoding, not part of the per-point pipeline.
        h->filtered.clear(); h->filteredRing.clear();
        for (int i = 0; i < P; ++i) {
            const float* q = xyzi + 4 * (size_t)i;
            if (std::isfinite(q[0]) && std::isfinite(q[1]) && std::isfinite(q[2])) {
                h->filtered.insert(h->filtered.end(), q, q + 4);
                if (ring) h->filteredRing.push_back(ring[i]);
            }
        }
        P = (int)(h->filtered.size() / 4);
        xyzi = h->filtered.data();
        if (ring) ring = h->filteredRing.data();
    }
    if (P > e.c.Pcap) { lego::set_last_error("scan larger than the handle's capacity"); return LEGO_ERR_CAPACITY; }
    e.uploadScan(xyzi, ring, &P, P > 0 ? P : 1, false);
    e.projectPointCloud();
    e.groundRemoval();
    e.cloudSegmentation();
    const int S = read_int(e, e.ip.counts + 0), O = read_int(e, e.ip.counts + 1);
    const int R = e.c.R;
    fetch_cloud(e, e.ip.segCloud, S, h->seg, segmented);
    fetch_cloud(e, e.ip.outlier, O, h->outl, outlier);
    h->range.resize(S + 1); h->ground.resize(S + 1); h->col.resize(S + 1); h->start.resize(R); h->end.resize(R);
    float orient[4];
    if (S > 0) {
        cuda_check(cudaMemcpyAsync(h->range.data(), e.ip.segRange, sizeof(float) * S, cudaMemcpyDeviceToHost, e.stream), "range");
        cuda_check(cudaMemcpyAsync(h->ground.data(), e.ip.segGround, S, cudaMemcpyDeviceToHost, e.stream), "ground");
        cuda_check(cudaMemcpyAsync(h->col.data(), e.ip.segCol, sizeof(int) * S, cudaMemcpyDeviceToHost, e.stream), "col");
    }
    cuda_check(cudaMemcpyAsync(h->start.data(), e.ip.startRing, sizeof(int) * R, cudaMemcpyDeviceToHost, e.stream), "start");
    cuda_check(cudaMemcpyAsync(h->end.data(), e.ip.endRing, sizeof(int) * R, cudaMemcpyDeviceToHost, e.stream), "end");
    cuda_check(cudaMemcpyAsync(orient, e.ip.orient, sizeof(orient), cudaMemcpyDeviceToHost, e.stream), "orient");
    e.sync();
    info->startRingIndex = h->start.data(); info->endRingIndex = h->end.data();
    info->startOrientation = orient[0]; info->endOrientation = orient[1]; info->orientationDiff = orient[2];
    info->segmentedCloudGroundFlag = h->ground.data(); info->segmentedCloudColInd = h->col.data(); info->segmentedCloudRange = h->range.data();
    info->count = S; info->on_device = 0;
    cuda_check(cudaGetLastError(), "ImageProjection kernels");
    return LEGO_OK;
    LEGO_CATCH
}

// ------------------------------------------------------------------------------------------------ FeatureAssociation
int lego_fa_create(const lego_params* p, lego_fa** out) {
    if (!params_ok(p) || !out) return LEGO_ERR_ARG;
    LEGO_TRY
    lego_fa* h = new lego_fa();
    h->e = new Engine(*p, 1, 0);
    *out = h;
    return LEGO_OK;
    LEGO_CATCH
}
int lego_fa_destroy(lego_fa* h) {
    if (!h) return LEGO_ERR_ARG;
    delete h->e; delete h;
    return LEGO_OK;
}
int lego_fa_process(lego_fa* h, const lego_cloud_info_view* info, const lego_cloud_view* segmented, const lego_cloud_view* outlier,
                    double stamp, float transformSum_out[6], lego_fa_outputs* out) {
    // the three inputs of one call belong to one scan by construction (gate of featureAssociation.cpp:1820-1829); the stamp
    // is timeScanCur for the IMU interpolation of adjustDistortion
    if (!h || !info || !segmented || !outlier || !transformSum_out || !out) return LEGO_ERR_ARG;
    Engine& e = *h->e;
    const int S = segmented->count, O = outlier->count, R = e.c.R;
    if (S < 0 || O < 0 || info->count != S) return LEGO_ERR_ARG;
    if (S > e.c.N || O > e.c.Ocap) { lego::set_last_error("cloud larger than the handle's capacity"); return LEGO_ERR_CAPACITY; }
    const float* segSrc = segmented->dev ? segmented->dev : segmented->host;
    const float* outSrc = outlier->dev ? outlier->dev : outlier->host;
    if ((S > 0 && !segSrc) || (O > 0 && !outSrc) || !info->startRingIndex || !info->endRingIndex) return LEGO_ERR_ARG;
    LEGO_TRY
    put(e, e.ip.segCloud, segSrc, sizeof(float4) * (size_t)S);
    put(e, e.ip.outlier, outSrc, sizeof(float4) * (size_t)O);
    put(e, e.ip.segGround, info->segmentedCloudGroundFlag, (size_t)S);
    put(e, e.ip.segCol, info->segmentedCloudColInd, sizeof(int) * (size_t)S);
    put(e, e.ip.segRange, info->segmentedCloudRange, sizeof(float) * (size_t)S);
    put(e, e.ip.startRing, info->startRingIndex, sizeof(int) * (size_t)R);
    put(e, e.ip.endRing, info->endRingIndex, sizeof(int) * (size_t)R);
    const float orient[4] = {info->startOrientation, info->endOrientation, info->orientationDiff, 0.f};
    const int counts[2] = {S, O};
    put(e, e.ip.orient, orient, sizeof(orient));
    put(e, e.ip.counts, counts, sizeof(counts));
    e.sync();   // orient / counts live on this stack frame
    e.adjustDistortionCalculateSmoothnessMarkOccluded(&stamp);
    e.extractFeatures();
    e.updateTransformation();
    e.integrateTransformationPublishCloudsLast();
    int fc[4], lc[4], flags = 0;
    cuda_check(cudaMemcpyAsync(fc, e.fa.featCounts, sizeof(fc), cudaMemcpyDeviceToHost, e.stream), "featCounts");
    cuda_check(cudaMemcpyAsync(lc, e.fa.lastCounts, sizeof(lc), cudaMemcpyDeviceToHost, e.stream), "lastCounts");
    cuda_check(cudaMemcpyAsync(&flags, e.fa.flags, sizeof(int), cudaMemcpyDeviceToHost, e.stream), "flags");
    cuda_check(cudaMemcpyAsync(transformSum_out, e.fa.transformSum, sizeof(float) * 6, cudaMemcpyDeviceToHost, e.stream), "transformSum");
    cuda_check(cudaMemcpyAsync(out->transformCur, e.fa.transformCur, sizeof(float) * 6, cudaMemcpyDeviceToHost, e.stream), "transformCur");
    e.sync();
    fetch_cloud(e, e.fa.sharp, fc[0], h->sharp, &out->cornerPointsSharp);
    fetch_cloud(e, e.fa.lessSharp, fc[1], h->lessSharp, &out->cornerPointsLessSharp);
    fetch_cloud(e, e.fa.flat, fc[2], h->flat, &out->surfPointsFlat);
    fetch_cloud(e, e.fa.lessFlat, fc[3], h->lessFlat, &out->surfPointsLessFlat);
    fetch_cloud(e, e.fa.cornerLast[e.faBuf], lc[0], h->cornerLast, &out->laserCloudCornerLast);
    fetch_cloud(e, e.fa.surfLast[e.faBuf], lc[1], h->surfLast, &out->laserCloudSurfLast);
    fetch_cloud(e, e.fa.outlierLast, O, h->outlierLast, &out->outlierCloudLast);
    e.sync();
    out->odometry_valid = (flags & 2) ? 1 : 0;
    out->published = (flags & 4) ? 1 : 0;
    cuda_check(cudaGetLastError(), "FeatureAssociation kernels");
    return LEGO_OK;
    LEGO_CATCH
}

// ------------------------------------------------------------------------------------------------ mapOptimization
int lego_mo_create(const lego_params* p, lego_mo** out) {
    if (!params_ok(p) || !out) return LEGO_ERR_ARG;
    LEGO_TRY
    lego_mo* h = new lego_mo();
    h->e = new Engine(*p, 1, 0);
    h->timeLastProcessing = -1.0;   // mapOptmization.cpp:331
    *out = h;
    return LEGO_OK;
    LEGO_CATCH
}
int lego_mo_destroy(lego_mo* h) {
    if (!h) return LEGO_ERR_ARG;
    delete h->e; delete h;
    return LEGO_OK;
}
// imuHandler of the two nodes (featureAssociation.cpp:431-459, mapOptmization.cpp:643-652)
int lego_fa_imu(lego_fa* h, double stamp, const double orientation_xyzw[4], const double angular_velocity[3], const double linear_acceleration[3]) {
    if (!h || !orientation_xyzw || !angular_velocity || !linear_acceleration) return LEGO_ERR_ARG;
    LEGO_TRY h->e->imuHandler(0, stamp, orientation_xyzw, angular_velocity, linear_acceleration); return LEGO_OK; LEGO_CATCH
}
int lego_mo_imu(lego_mo* h, double stamp, const double orientation_xyzw[4], const double angular_velocity[3], const double linear_acceleration[3]) {
    if (!h || !orientation_xyzw || !angular_velocity || !linear_acceleration) return LEGO_ERR_ARG;
    LEGO_TRY h->e->imuHandler(0, stamp, orientation_xyzw, angular_velocity, linear_acceleration); return LEGO_OK; LEGO_CATCH
}

// publishGlobalMap and the clouds saved at shutdown (mapOptmization.cpp:724-800)
static int global_map_out(Engine& e, int lane, int which, float* xyzi_out, long cap, long* n_out) {
    std::vector<float4> cloud;
    e.globalMap(lane, which, cloud);
    *n_out = (long)cloud.size();
    const long n = std::min<long>((long)cloud.size(), cap);
    if (xyzi_out && n > 0) std::memcpy(xyzi_out, cloud.data(), sizeof(float4) * (size_t)n);
    return LEGO_OK;
}
int lego_mo_global_map(lego_mo* h, int which, float* xyzi_out, long cap, long* n_out) {
    if (!h || !n_out || which < 0 || which > 3 || cap < 0 || (cap > 0 && !xyzi_out)) return LEGO_ERR_ARG;
    LEGO_TRY return global_map_out(*h->e, 0, which, xyzi_out, cap, n_out); LEGO_CATCH
}
int lego_write_pcd_ascii(const char* path, const float* xyzi, long n) {
    if (!path || n < 0 || (n > 0 && !xyzi)) return LEGO_ERR_ARG;
    LEGO_TRY
    if (lego::write_pcd_ascii(path, xyzi, n) != 0) { lego::set_last_error(std::string("cannot write ") + path); return LEGO_ERR_ARG; }
    return LEGO_OK;
    LEGO_CATCH
}

int lego_mo_process(lego_mo* h, const lego_cloud_view* corner_last, const lego_cloud_view* surf_last, const lego_cloud_view* outlier_last,
                    const float transformSum[6], double stamp, float transformAftMapped_out[6], float transformBefMapped_out[6], int* processed) {
    if (!h || !corner_last || !surf_last || !outlier_last || !transformSum || !transformAftMapped_out || !processed) return LEGO_ERR_ARG;
    Engine& e = *h->e;
    const int nC = corner_last->count, nS = surf_last->count, nO = outlier_last->count;
    if (nC < 0 || nS < 0 || nO < 0) return LEGO_ERR_ARG;
    if (nC > e.c.R * 120 || nS > e.c.N || nO > e.c.Ocap) { lego::set_last_error("cloud larger than the handle's capacity"); return LEGO_ERR_CAPACITY; }
    LEGO_TRY
    *processed = 0;
    // the four messages of one call carry one stamp, so the first gate of mapOptmization.cpp:1489-1493 is open;
    // what remains is the mappingProcessInterval gate of mapOptmization.cpp:1499
    if (stamp - h->timeLastProcessing >= e.params.mappingProcessInterval) {
        h->timeLastProcessing = stamp;
        const float* cs = corner_last->dev ? corner_last->dev : corner_last->host;
        const float* ss = surf_last->dev ? surf_last->dev : surf_last->host;
        const float* os = outlier_last->dev ? outlier_last->dev : outlier_last->host;
        put(e, e.fa.cornerLast[e.faBuf], cs, sizeof(float4) * (size_t)nC);
        put(e, e.fa.surfLast[e.faBuf], ss, sizeof(float4) * (size_t)nS);
        put(e, e.fa.outlierLast, os, sizeof(float4) * (size_t)nO);
        const int lc[2] = {nC, nS};
        put(e, e.fa.lastCounts, lc, sizeof(lc));
        put(e, e.ip.counts + 1, &nO, sizeof(int));
        put(e, e.fa.transformSum, transformSum, sizeof(float) * 6);
        e.sync();
        e.moHost[0].timeOdometry = stamp;   // timeLaserOdometry
        e.mapOptimizationBody(std::vector<int>(1, 1));
        *processed = 1;
    }
    cuda_check(cudaMemcpyAsync(transformAftMapped_out, e.mo.transformAftMapped, sizeof(float) * 6, cudaMemcpyDeviceToHost, e.stream), "aft");
    if (transformBefMapped_out)
        cuda_check(cudaMemcpyAsync(transformBefMapped_out, e.mo.transformBefMapped, sizeof(float) * 6, cudaMemcpyDeviceToHost, e.stream), "bef");
    e.sync();
    if (e.moHost[0].lastError) { lego::set_last_error("key-frame pool or local-map capacity exceeded"); return LEGO_ERR_CAPACITY; }
    cuda_check(cudaGetLastError(), "mapOptimization kernels");
    return LEGO_OK;
    LEGO_CATCH
}

}  // extern "C"
