// TransformFusion (reference LeGO-LOAM/src/transformFusion.cpp): every odometry pose FeatureAssociation publishes is
// re-based on the latest mapping correction, giving the high-rate /integrated_to_init pose.
//   laserOdometryHandler (transformFusion.cpp:181-216)  transformSum <- message, transformAssociateToMap (94-179),
//                                                        quaternion + position of the fused pose
//   odomAftMappedHandler (transformFusion.cpp:218-241)   transformAftMapped / transformBefMapped <- message
// Both messages carry their rotation as a quaternion built from roll/pitch/yaw in double and read back with
// tf::Matrix3x3::getRPY; map_kernels.cuh restates that round trip (odometry_message_round_trip) and
// transformAssociateToMap, which this node shares term by term with mapOptimization (mapOptmization.cpp:376-461).
// In the pipeline the poses never leave the device: one thread per lane.
#include <new>
#include <stdexcept>
#include "engine.h"
#include "map_kernels.cuh"

namespace lego {

// sum: [L][6] FeatureAssociation's transformSum; faFlags bit1 = an odometry message was published this scan
__global__ void tf_laser_odometry_kernel(TfDev t, const float* sum, const int* faFlags, int L) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= L) return;
    if (faFlags && !(faFlags[lane] & 2)) { t.published[lane] = 0; return; }
    float S[6], incre[6], M[6];
    odometry_message_round_trip(sum + lane * 6, S);
    transform_associate_to_map(S, t.bef + lane * 6, t.aft + lane * 6, incre, M);
    for (int i = 0; i < 6; ++i) t.mapped[lane * 6 + i] = M[i];
    // tf::createQuaternionMsgFromRollPitchYaw(transformMapped[2], -transformMapped[0], -transformMapped[1]) and the
    // axis shuffle of transformFusion.cpp:201-208
    const double hr = (double)M[2] * 0.5, hp = -(double)M[0] * 0.5, hy = -(double)M[1] * 0.5;
    const double cr = cos(hr), sr = sin(hr), cp = cos(hp), sp = sin(hp), cy = cos(hy), sy = sin(hy);
    const double qx = sr * cp * cy - cr * sp * sy, qy = cr * sp * cy + sr * cp * sy;
    const double qz = cr * cp * sy - sr * sp * cy, qw = cr * cp * cy + sr * sp * sy;
    float* o = t.integrated + lane * 7;
    o[0] = (float)(-qy); o[1] = (float)(-qz); o[2] = (float)qx; o[3] = (float)qw;
    o[4] = M[3]; o[5] = M[4]; o[6] = M[5];
    t.published[lane] = 1;
}

// aft / bef: [L][6] mapOptimization's transformAftMapped / transformBefMapped as published by publishTF
// (mapOptmization.cpp:654-673); active: lanes whose mapping body ran this scan (nullptr = all)
__global__ void tf_aft_mapped_kernel(TfDev t, const float* aft, const float* bef, const int* active, int L) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= L) return;
    if (active && !active[lane]) return;
    float A[6];
    odometry_message_round_trip(aft + lane * 6, A);   // same quaternion encoding as the odometry message
    for (int i = 0; i < 6; ++i) { t.aft[lane * 6 + i] = A[i]; t.bef[lane * 6 + i] = bef[lane * 6 + i]; }
}

void Engine::allocTransformFusion() {
    const size_t L = c.L;
    tf.aft = alloc<float>(L * 6); tf.bef = alloc<float>(L * 6); tf.mapped = alloc<float>(L * 6);
    tf.integrated = alloc<float>(L * 7); tf.published = alloc<int>(L);
}

void Engine::transformFusionLaserOdometry() {
    ProfScope ps(this, "tf_laser_odometry_kernel");
    tf_laser_odometry_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(tf, fa.transformSum, fa.flags, c.L);
    launches += 1;
}

void Engine::transformFusionAftMapped() {
    ProfScope ps(this, "tf_aft_mapped_kernel");
    tf_aft_mapped_kernel<<<(c.L + 63) / 64, 64, 0, stream>>>(tf, mo.transformAftMapped, mo.transformBefMapped, mo.active, c.L);
    launches += 1;
}

bool Engine::lookupTransformFusion(int lane, const std::string& n, void** ptr, int* elem, long* count) {
    const size_t l = lane;
    auto set = [&](void* p, int e, long cnt) { *ptr = p; *elem = e; *count = cnt; return true; };
    if (n == "transformMapped") return set(tf.mapped + l * 6, 4, 6);
    if (n == "integratedToInit") return set(tf.integrated + l * 7, 4, 7);
    if (n == "tfPublished") return set(tf.published + l, 4, 1);
    if (n == "tfTransformAftMapped") return set(tf.aft + l * 6, 4, 6);
    if (n == "tfTransformBefMapped") return set(tf.bef + l * 6, 4, 6);
    return false;
}

}  // namespace lego

// ---- per-stage handle (include/lego_b200.h): the drop-in for the two handlers of the TransformFusion node ----------
struct lego_tf {
    lego::TfDev t;
    float* sum; float* aft; float* bef;   // device staging of the message payloads, 6 floats each
    cudaStream_t stream;
    int device;
};

extern "C" {

int lego_tf_create(const lego_params* p, lego_tf** out) {
    if (!p || !out) return LEGO_ERR_ARG;
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { lego::set_last_error("no CUDA device"); return LEGO_ERR_CUDA; }
        lego::cuda_check(cudaSetDevice(p->device), "cudaSetDevice");
        lego_tf* h = new lego_tf();
        h->device = p->device;
        lego::cuda_check(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "stream create");
        float* blk = nullptr;
        // aft 6 | bef 6 | mapped 6 | integrated 7 | published 1 | sum 6 | aft-in 6 | bef-in 6 (all zero: the reference's constructor)
        lego::cuda_check(cudaMalloc(reinterpret_cast<void**>(&blk), sizeof(float) * 44), "cudaMalloc");
        lego::cuda_check(cudaMemsetAsync(blk, 0, sizeof(float) * 44, h->stream), "cudaMemset");
        h->t.aft = blk; h->t.bef = blk + 6; h->t.mapped = blk + 12; h->t.integrated = blk + 18;
        h->t.published = reinterpret_cast<int*>(blk + 25);
        h->sum = blk + 26; h->aft = blk + 32; h->bef = blk + 38;
        *out = h;
        return LEGO_OK;
    } catch (const std::exception& ex) { if (lego::last_error().empty()) lego::set_last_error(ex.what()); return LEGO_ERR_CUDA; }
}

int lego_tf_destroy(lego_tf* h) {
    if (!h) return LEGO_ERR_ARG;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    cudaFree(h->t.aft);
    cudaStreamDestroy(h->stream);
    delete h;
    return LEGO_OK;
}

int lego_tf_laser_odometry(lego_tf* h, const float transformSum[6], float transformMapped_out[6], float integrated_out[7]) {
    if (!h || !transformSum) return LEGO_ERR_ARG;
    try {
        lego::cuda_check(cudaSetDevice(h->device), "cudaSetDevice");
        lego::cuda_check(cudaMemcpyAsync(h->sum, transformSum, sizeof(float) * 6, cudaMemcpyDefault, h->stream), "tf input");
        lego::tf_laser_odometry_kernel<<<1, 32, 0, h->stream>>>(h->t, h->sum, nullptr, 1);
        float host[13];
        lego::cuda_check(cudaMemcpyAsync(host, h->t.mapped, sizeof(float) * 13, cudaMemcpyDeviceToHost, h->stream), "tf output");
        lego::cuda_check(cudaStreamSynchronize(h->stream), "tf sync");
        if (transformMapped_out) for (int i = 0; i < 6; ++i) transformMapped_out[i] = host[i];
        if (integrated_out) for (int i = 0; i < 7; ++i) integrated_out[i] = host[6 + i];
        return LEGO_OK;
    } catch (const std::exception& ex) { if (lego::last_error().empty()) lego::set_last_error(ex.what()); return LEGO_ERR_CUDA; }
}

int lego_tf_aft_mapped(lego_tf* h, const float transformAftMapped[6], const float transformBefMapped[6]) {
    if (!h || !transformAftMapped || !transformBefMapped) return LEGO_ERR_ARG;
    try {
        lego::cuda_check(cudaSetDevice(h->device), "cudaSetDevice");
        lego::cuda_check(cudaMemcpyAsync(h->aft, transformAftMapped, sizeof(float) * 6, cudaMemcpyDefault, h->stream), "tf input");
        lego::cuda_check(cudaMemcpyAsync(h->bef, transformBefMapped, sizeof(float) * 6, cudaMemcpyDefault, h->stream), "tf input");
        lego::tf_aft_mapped_kernel<<<1, 32, 0, h->stream>>>(h->t, h->aft, h->bef, nullptr, 1);
        lego::cuda_check(cudaStreamSynchronize(h->stream), "tf sync");
        return LEGO_OK;
    } catch (const std::exception& ex) { if (lego::last_error().empty()) lego::set_last_error(ex.what()); return LEGO_ERR_CUDA; }
}

}  // extern "C"
