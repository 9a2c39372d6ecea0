// C ABI (include/lego_b200.h) over lego::Engine.  Exceptions never cross the boundary: every entry point
// converts them into an error code and a message retrievable with lego_last_error().
#include <cstdio>
#include <cstring>
#include <new>
#include <stdexcept>
#include <vector>
#include "engine.h"

using lego::Engine;

struct lego_pipeline { Engine* e; std::vector<double> stamps; };
struct lego_ip { Engine* e; std::vector<float> h_seg, h_outlier, h_range; std::vector<int32_t> h_start, h_end; std::vector<uint8_t> h_ground; std::vector<uint32_t> h_col; };

#define LEGO_TRY try {
#define LEGO_CATCH \
    } catch (const std::length_error& ex) { lego::set_last_error(ex.what()); return LEGO_ERR_CAPACITY; } \
      catch (const std::invalid_argument& ex) { lego::set_last_error(ex.what()); return LEGO_ERR_ARG; } \
      catch (const std::bad_alloc&) { lego::set_last_error("out of host memory"); return LEGO_ERR_CUDA; } \
      catch (const std::exception& ex) { if (lego::last_error().empty()) lego::set_last_error(ex.what()); return LEGO_ERR_CUDA; }

extern "C" {

const char* lego_last_error(void) { return lego::last_error().c_str(); }

int lego_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ------------------------------------------------------------------ pipeline
int lego_pipeline_create(const lego_params* p, int lanes, int max_points, lego_pipeline** out) {
    if (!p || !out || lanes < 1 || p->N_SCAN < 2 || p->N_SCAN > 128 || p->Horizon_SCAN < 16) return LEGO_ERR_ARG;
    LEGO_TRY
    lego_pipeline* h = new lego_pipeline();
    h->e = new Engine(*p, lanes, max_points);
    *out = h;
    return LEGO_OK;
    LEGO_CATCH
}
int lego_pipeline_destroy(lego_pipeline* h) {
    if (!h) return LEGO_ERR_ARG;
    delete h->e; delete h;
    return LEGO_OK;
}
// the stages of one scan for every lane, after the input was handed to the engine
static int run_stages(lego_pipeline* h, int stage_mask) {
    Engine& e = *h->e;
    if ((stage_mask & 3) == 3) e.frontEnd(h->stamps.data());
    else if (stage_mask & 1) {
        e.projectPointCloud();
        e.groundRemoval();
        e.cloudSegmentation();
    } else if (stage_mask & 2) {
        e.adjustDistortionCalculateSmoothnessMarkOccluded(h->stamps.data());
        e.extractFeatures();
        e.updateTransformation();
        e.integrateTransformationPublishCloudsLast();
    }
    // TransformFusion sees the odometry message of this scan before mapOptimization has answered it
    if ((stage_mask & 10) == 10) e.transformFusionLaserOdometry();
    if ((stage_mask & 6) == 6) {
        e.mapOptimizationRun(h->stamps.data());
        if ((stage_mask & 8) && e.moLastAny) e.transformFusionAftMapped();
    } else e.moActiveLast.assign(e.c.L, 0);
    lego::cuda_check(cudaGetLastError(), "kernel launch");
    // capacity overflows of mapOptimization (key-frame pool, selection table, local-map buffers): sticky per lane
    const int bad = e.firstCapacityError();
    if (bad >= 0) {
        char msg[200];
        snprintf(msg, sizeof msg, "lane %d: mapOptimization capacity exceeded (bits 0x%x: 1 key-frame pool / maxKeyFrames, 2 surrounding key frames, 4 local-map points)",
                 bad, e.moHost[bad].lastError);
        lego::set_last_error(msg);
        return LEGO_ERR_CAPACITY;
    }
    return LEGO_OK;
}

int lego_pipeline_step(lego_pipeline* h, const float* xyzi, const uint16_t* ring, const int* npts, int stride_points,
                       const double* stamps, int src_on_device, int stage_mask) {
    if (!h || !xyzi || !npts || !stamps) return LEGO_ERR_ARG;
    Engine& e = *h->e;
    if (e.c.useRing && !ring) return LEGO_ERR_ARG;
    LEGO_TRY
    h->stamps.assign(stamps, stamps + e.c.L);
    if (stage_mask & 1) e.uploadScan(xyzi, ring, npts, stride_points, src_on_device != 0);
    return run_stages(h, stage_mask);
    LEGO_CATCH
}
int lego_pipeline_step_lanes(lego_pipeline* h, const float* const* xyz, const uint16_t* const* ring, const int* npts,
                             int floats_per_point, const double* stamps, int src_on_device, int stage_mask) {
    if (!h || !xyz || !npts || !stamps || (floats_per_point != 3 && floats_per_point != 4)) {
        lego::set_last_error("lego_pipeline_step_lanes: null argument or floats_per_point not 3 / 4");
        return LEGO_ERR_ARG;
    }
    Engine& e = *h->e;
    if (e.c.useRing && !ring) return LEGO_ERR_ARG;
    LEGO_TRY
    h->stamps.assign(stamps, stamps + e.c.L);
    if (src_on_device < 0 || src_on_device > 3) { lego::set_last_error("src_on_device must be 0, 1, 2 or 3"); return LEGO_ERR_ARG; }
    if (stage_mask & 1) e.uploadScanLanes(xyz, ring, npts, floats_per_point, src_on_device == 1, src_on_device == 2, src_on_device == 3);
    return run_stages(h, stage_mask);
    LEGO_CATCH
}
int lego_pipeline_input_wait(lego_pipeline* h) {
    if (!h) return LEGO_ERR_ARG;
    LEGO_TRY h->e->inputWait(); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_imu(lego_pipeline* h, int lane, double stamp, const double orientation_xyzw[4], const double angular_velocity[3],
                      const double linear_acceleration[3]) {
    if (!h || lane < 0 || lane >= h->e->c.L || !orientation_xyzw || !angular_velocity || !linear_acceleration) return LEGO_ERR_ARG;
    LEGO_TRY h->e->imuHandler(lane, stamp, orientation_xyzw, angular_velocity, linear_acceleration); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_sync(lego_pipeline* h) {
    if (!h) return LEGO_ERR_ARG;
    LEGO_TRY h->e->sync(); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_get_poses(lego_pipeline* h, float* poses_out, int* flags_out) {
    if (!h || !poses_out) return LEGO_ERR_ARG;
    LEGO_TRY h->e->getPoses(poses_out, flags_out); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_poses_async(lego_pipeline* h, int slot) {
    if (!h || slot < 0 || slot >= Engine::POSE_SLOTS) return LEGO_ERR_ARG;
    LEGO_TRY h->e->posesAsync(slot); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_poses_wait(lego_pipeline* h, int slot, float* poses_out, int* flags_out) {
    if (!h || !poses_out || slot < 0 || slot >= Engine::POSE_SLOTS) return LEGO_ERR_ARG;
    LEGO_TRY h->e->posesWait(slot, poses_out, flags_out); return LEGO_OK; LEGO_CATCH
}
long lego_pipeline_launch_count(lego_pipeline* h) { return h ? h->e->launches : 0; }
long lego_pipeline_graph_launch_count(lego_pipeline* h) { return h ? h->e->graphLaunches : 0; }
int lego_pipeline_profile(lego_pipeline* h, int mode, const char* filter) {
    if (!h || mode < 0 || mode > 2) return LEGO_ERR_ARG;
    LEGO_TRY
    h->e->sync();
    h->e->profReset();
    h->e->profMode = mode;
    h->e->profFilter = filter ? filter : "";
    return LEGO_OK;
    LEGO_CATCH
}
long lego_pipeline_profile_read(lego_pipeline* h, char* names, long names_cap, double* ms, long* counts, long max_kernels) {
    if (!h || !ms || !counts) return LEGO_ERR_ARG;
    LEGO_TRY return h->e->profRead(names, names_cap, ms, counts, max_kernels); LEGO_CATCH
}
int lego_pipeline_timer_record(lego_pipeline* h, int slot) {
    if (!h || slot < 0 || slot >= 8) return LEGO_ERR_ARG;
    LEGO_TRY h->e->timerRecord(slot); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_timer_elapsed(lego_pipeline* h, int slot_a, int slot_b, float* ms_out) {
    if (!h || !ms_out || slot_a < 0 || slot_a >= 8 || slot_b < 0 || slot_b >= 8) return LEGO_ERR_ARG;
    LEGO_TRY *ms_out = h->e->timerElapsed(slot_a, slot_b); return LEGO_OK; LEGO_CATCH
}
int lego_pipeline_timer_elapsed_between(lego_pipeline* ha, int slot_a, lego_pipeline* hb, int slot_b, float* ms_out) {
    if (!ha || !hb || !ms_out || slot_a < 0 || slot_a >= 8 || slot_b < 0 || slot_b >= 8) return LEGO_ERR_ARG;
    LEGO_TRY
    if (!ha->e->timerEv[slot_a] || !hb->e->timerEv[slot_b]) { lego::set_last_error("timer slot was never recorded"); return LEGO_ERR_ARG; }
    lego::cuda_check(cudaEventSynchronize(hb->e->timerEv[slot_b]), "event sync");
    lego::cuda_check(cudaEventElapsedTime(ms_out, ha->e->timerEv[slot_a], hb->e->timerEv[slot_b]), "event elapsed");
    return LEGO_OK;
    LEGO_CATCH
}
// publishGlobalMap and the clouds saved at shutdown (mapOptmization.cpp:724-800) for one lane
int lego_pipeline_global_map(lego_pipeline* h, int lane, int which, float* xyzi_out, long cap, long* n_out) {
    if (!h || !n_out || which < 0 || which > 3 || cap < 0 || (cap > 0 && !xyzi_out)) return LEGO_ERR_ARG;
    if (lane < 0 || lane >= h->e->c.L) return LEGO_ERR_ARG;
    LEGO_TRY
    std::vector<float4> cloud;
    h->e->globalMap(lane, which, cloud);
    *n_out = (long)cloud.size();
    const long n = (long)cloud.size() < cap ? (long)cloud.size() : cap;
    if (xyzi_out && n > 0) std::memcpy(xyzi_out, cloud.data(), sizeof(float4) * (size_t)n);
    return LEGO_OK;
    LEGO_CATCH
}

long lego_pipeline_download(lego_pipeline* h, int lane, const char* name, void* dst, long cap, int* elem_size_out) {
    if (!h || !name) return LEGO_ERR_ARG;
    LEGO_TRY
    if (!std::strcmp(name, "surroundingExistingKeyPosesID")) {   // host-side list of the key-frame bookkeeping
        if (lane < 0 || lane >= h->e->c.L) return LEGO_ERR_NAME;
        const std::vector<int>& v = h->e->moHost[lane].surroundingExistingKeyPosesID;
        if (elem_size_out) *elem_size_out = 4;
        const long n = (long)v.size();
        if (dst && n > 0) std::memcpy(dst, v.data(), sizeof(int) * (size_t)(n < cap ? n : cap));
        return n;
    }
    if (!std::strcmp(name, "cloudKeyPoses6D") || !std::strcmp(name, "loopClosure") || !std::strcmp(name, "icpTransform") ||
        !std::strcmp(name, "icpFitness") || !std::strcmp(name, "loopStats")) {   // host-side state of the loop-closure bookkeeping
        if (lane < 0 || lane >= h->e->c.L) return LEGO_ERR_NAME;
        const lego::MoLaneHost& H = h->e->moHost[lane];
        std::vector<float> f; std::vector<int> iv;
        if (name[0] == 'c') for (const lego::MoKeyFrame& k : H.keyFrames) { f.push_back(k.x); f.push_back(k.y); f.push_back(k.z); f.push_back(k.roll); f.push_back(k.pitch); f.push_back(k.yaw); }
        else if (!std::strcmp(name, "icpTransform")) f.assign(H.lastIcpT, H.lastIcpT + 16);
        else if (!std::strcmp(name, "icpFitness")) f.push_back((float)H.lastIcpFitness);
        else if (!std::strcmp(name, "loopStats")) { iv.push_back(H.loopCalls); iv.push_back(H.loopsClosed); }
        else iv.assign(H.lastLoop, H.lastLoop + 9);
        if (elem_size_out) *elem_size_out = 4;
        const long n = (long)(f.empty() ? iv.size() : f.size());
        const void* srcp = f.empty() ? (const void*)iv.data() : (const void*)f.data();
        if (dst && n > 0) std::memcpy(dst, srcp, 4 * (size_t)(n < cap ? n : cap));
        return n;
    }
    if (!std::strcmp(name, "labelMat")) h->e->finalizeLabelMat();
    return h->e->download(lane, name, dst, cap, elem_size_out);
    LEGO_CATCH
}
long lego_pipeline_upload(lego_pipeline* h, int lane, const char* name, const void* src, long count) {
    if (!h || !name || !src) return LEGO_ERR_ARG;
    LEGO_TRY return h->e->upload(lane, name, src, count); LEGO_CATCH
}

}  // extern "C"
