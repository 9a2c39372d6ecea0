// Engine construction, device memory, host<->device plumbing and the named-array access used by parity tests.
#include <cmath>
#include <cstring>
#include <stdexcept>
#include "engine.h"

namespace lego {

static thread_local std::string g_last_error;
void set_last_error(const std::string& s) { g_last_error = s; }
const std::string& last_error() { return g_last_error; }

void cuda_check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) {
        std::string msg = std::string(what) + ": " + cudaGetErrorString(e);
        set_last_error(msg);
        throw std::runtime_error(msg);
    }
}

Engine::Engine(const lego_params& p, int lanes, int max_points) : params(p) {
    int ndev = 0;
    cuda_check(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount");
    if (ndev <= 0) { set_last_error("no CUDA device"); throw std::runtime_error("no CUDA device"); }
    cuda_check(cudaSetDevice(p.device), "cudaSetDevice");
    cuda_check(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking), "cudaStreamCreate");

    c.R = p.N_SCAN; c.C = p.Horizon_SCAN; c.N = c.R * c.C; c.G = p.groundScanInd; c.L = lanes;
    c.Pcap = max_points > 0 ? max_points : c.N + 64 * c.R;
    c.Ocap = c.N / 5 + c.R + 16;
    c.useRing = p.useCloudRing;
    c.ang_res_x = p.ang_res_x; c.ang_res_y = p.ang_res_y; c.ang_bottom = p.ang_bottom;
    c.sensorMinimumRange = p.sensorMinimumRange; c.sensorMountAngle = p.sensorMountAngle; c.segmentTheta = p.segmentTheta;
    // sin(alpha)/cos(alpha) in labelComponents are float overloads on a float constant (imageProjection.cpp:421);
    // computing them here with the host C library keeps them bit-identical to the reference's values.
    c.sinAlphaX = sinf(p.segmentAlphaX); c.cosAlphaX = cosf(p.segmentAlphaX);
    c.sinAlphaY = sinf(p.segmentAlphaY); c.cosAlphaY = cosf(p.segmentAlphaY);
    c.segmentValidPointNum = p.segmentValidPointNum; c.segmentValidLineNum = p.segmentValidLineNum;
    c.scanPeriod = p.scanPeriod; c.edgeThreshold = p.edgeThreshold; c.surfThreshold = p.surfThreshold;
    c.nearestFeatureSearchSqDist = p.nearestFeatureSearchSqDist;
    c.rowWords = (c.R + 31) / 32;

    const size_t L = lanes, N = c.N, R = c.R;
    d_in_xyzi = alloc<float4>(L * c.Pcap); d_in_ring = alloc<uint16_t>(L * c.Pcap); d_npts = alloc<int>(L);
    ip.in_xyzi = d_in_xyzi; ip.in_ring = d_in_ring; ip.npts = d_npts;
    ip.owner = alloc<int>(L * N); ip.rangeMat = alloc<float>(L * N); ip.fullCloud = alloc<float4>(L * N);
    ip.groundMat = alloc<int8_t>(L * N); ip.labelMat = alloc<int>(L * N); ip.parent = alloc<int>(L * N);
    ip.compSize = alloc<int>(L * N); ip.compRows = alloc<uint32_t>(L * N * c.rowWords); ip.rowCounts = alloc<int>(L * R * 4);
    ip.segCloud = alloc<float4>(L * N); ip.segGround = alloc<uint8_t>(L * N); ip.segCol = alloc<int>(L * N);
    ip.segRange = alloc<float>(L * N); ip.startRing = alloc<int>(L * R); ip.endRing = alloc<int>(L * R);
    ip.orient = alloc<float>(L * 4); ip.outlier = alloc<float4>(L * c.Ocap); ip.counts = alloc<int>(L * 8);

    allocFeatureAssociation();
    cuda_check(cudaStreamSynchronize(stream), "init sync");
}

Engine::~Engine() {
    if (stream) cudaStreamSynchronize(stream);
    for (void* p : allocations) cudaFree(p);
    if (h_stage) cudaFreeHost(h_stage);
    if (stream) cudaStreamDestroy(stream);
}

void Engine::sync() { cuda_check(cudaStreamSynchronize(stream), "stream sync"); }

// copyPointCloud (imageProjection.cpp:163-179): the host message becomes device arrays; one copy per array.
void Engine::uploadScan(const float* xyzi, const uint16_t* ring, const int* npts, int stride_points, bool src_on_device) {
    const cudaMemcpyKind kind = src_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    for (int l = 0; l < c.L; ++l) if (npts[l] > c.Pcap || npts[l] < 0) { set_last_error("scan larger than max_points"); throw std::length_error("capacity"); }
    if (stride_points == c.Pcap) {
        cuda_check(cudaMemcpyAsync(d_in_xyzi, xyzi, sizeof(float4) * (size_t)c.L * c.Pcap, kind, stream), "upload xyzi");
        if (ring) cuda_check(cudaMemcpyAsync(d_in_ring, ring, sizeof(uint16_t) * (size_t)c.L * c.Pcap, kind, stream), "upload ring");
    } else {
        for (int l = 0; l < c.L; ++l) {
            cuda_check(cudaMemcpyAsync(d_in_xyzi + (size_t)l * c.Pcap, xyzi + (size_t)l * stride_points * 4, sizeof(float4) * npts[l], kind, stream), "upload xyzi");
            if (ring) cuda_check(cudaMemcpyAsync(d_in_ring + (size_t)l * c.Pcap, ring + (size_t)l * stride_points, sizeof(uint16_t) * npts[l], kind, stream), "upload ring");
        }
    }
    cuda_check(cudaMemcpyAsync(d_npts, npts, sizeof(int) * c.L, cudaMemcpyHostToDevice, stream), "upload npts");
    // npts is pageable host memory that the caller may reuse: make the copy complete before returning
    if (!src_on_device || true) cuda_check(cudaStreamSynchronize(stream), "upload sync");
}

void Engine::getPoses(float* poses18, int* flags) {
    std::vector<float> sum(6 * (size_t)c.L);
    cuda_check(cudaMemcpyAsync(sum.data(), fa.transformSum, sizeof(float) * 6 * c.L, cudaMemcpyDeviceToHost, stream), "poses");
    std::vector<int> fl(c.L);
    cuda_check(cudaMemcpyAsync(fl.data(), fa.flags, sizeof(int) * c.L, cudaMemcpyDeviceToHost, stream), "flags");
    cuda_check(cudaStreamSynchronize(stream), "poses sync");
    for (int l = 0; l < c.L; ++l) {
        for (int k = 0; k < 6; ++k) poses18[l * 18 + k] = sum[l * 6 + k];
        for (int k = 6; k < 18; ++k) poses18[l * 18 + k] = 0.f;
        if (flags) flags[l] = fl[l];
    }
}

// ---- named arrays (parity tests / debugging) ----
struct NamedArray { void* ptr; int elem; long count; };

static int lane_count(Engine& e, int lane, int which) {
    int v = 0;
    cuda_check(cudaMemcpyAsync(&v, e.ip.counts + lane * 8 + which, sizeof(int), cudaMemcpyDeviceToHost, e.stream), "count");
    cuda_check(cudaStreamSynchronize(e.stream), "count sync");
    return v;
}
static int fa_count(Engine& e, int* base, int lane, int stride, int which) {
    int v = 0;
    cuda_check(cudaMemcpyAsync(&v, base + lane * stride + which, sizeof(int), cudaMemcpyDeviceToHost, e.stream), "count");
    cuda_check(cudaStreamSynchronize(e.stream), "count sync");
    return v;
}

bool Engine::lookup(int lane, const std::string& n, void** ptr, int* elem, long* count) {
    const size_t N = c.N, R = c.R, l = lane;
    auto set = [&](void* p, int e, long cnt) { *ptr = p; *elem = e; *count = cnt; return true; };
    if (n == "rangeMat") return set(ip.rangeMat + l * N, 4, N);
    if (n == "fullCloud") return set(ip.fullCloud + l * N, 16, N);
    if (n == "groundMat") return set(ip.groundMat + l * N, 1, N);
    if (n == "labelMat") return set(ip.labelMat + l * N, 4, N);
    if (n == "owner") return set(ip.owner + l * N, 4, N);
    if (n == "segmentedCloud") return set(ip.segCloud + l * N, 16, lane_count(*this, lane, 0));
    if (n == "segGroundFlag") return set(ip.segGround + l * N, 1, lane_count(*this, lane, 0));
    if (n == "segColInd") return set(ip.segCol + l * N, 4, lane_count(*this, lane, 0));
    if (n == "segRange") return set(ip.segRange + l * N, 4, lane_count(*this, lane, 0));
    if (n == "outlierCloud") return set(ip.outlier + l * c.Ocap, 16, lane_count(*this, lane, 1));
    if (n == "startRingIndex") return set(ip.startRing + l * R, 4, R);
    if (n == "endRingIndex") return set(ip.endRing + l * R, 4, R);
    if (n == "orientation") return set(ip.orient + l * 4, 4, 3);
    if (n == "ipCounts") return set(ip.counts + l * 8, 4, 8);
    return lookupFeatureAssociation(lane, n, ptr, elem, count);
}

long Engine::download(int lane, const char* name, void* dst, long cap, int* elem_size) {
    void* p; int e; long cnt;
    if (lane < 0 || lane >= c.L || !lookup(lane, name, &p, &e, &cnt)) return LEGO_ERR_NAME;
    if (elem_size) *elem_size = e;
    const long ncopy = cnt < cap ? cnt : cap;
    if (dst && ncopy > 0) {
        cuda_check(cudaMemcpyAsync(dst, p, (size_t)ncopy * e, cudaMemcpyDeviceToHost, stream), "download");
        cuda_check(cudaStreamSynchronize(stream), "download sync");
    }
    return cnt;
}

long Engine::upload(int lane, const char* name, const void* src, long count) {
    void* p; int e; long cnt;
    if (lane < 0 || lane >= c.L || !lookup(lane, name, &p, &e, &cnt)) return LEGO_ERR_NAME;
    cuda_check(cudaMemcpyAsync(p, src, (size_t)count * e, cudaMemcpyHostToDevice, stream), "upload");
    cuda_check(cudaStreamSynchronize(stream), "upload sync");
    return count;
}

}  // namespace lego
