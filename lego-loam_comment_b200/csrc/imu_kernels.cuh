// Device side of the optional IMU path of FeatureAssociation (reference LeGO-LOAM/src/featureAssociation.cpp):
// per-point interpolation of the IMU ring buffer inside adjustDistortion (525-614), VeloToStartIMU (346-363),
// TransformToStartIMU (365-390), updateInitialGuess (1639-1664) and the IMU terms of TransformToEnd (928-952).
// The reference walks one pointer through the ring buffer while it loops over the points; the walk restarts from
// imuPointerLastIteration for every point (featureAssociation.cpp:527), so every point can do its own walk.
// imuShiftFromStart{X,Y,Z}Cur are never written in the reference (ShiftToStartIMU is dead code, quirk Q15): they stay 0.
#pragma once
#include "engine.h"
#include "parity_math.cuh"

namespace lego {

struct ImuSample { float rpy[3], velo[3], shift[3], ang[3]; };

// values of the ring buffer at time t = timeScanCur + pointTime (featureAssociation.cpp:526-567, 586-600)
__device__ inline ImuSample imu_interpolate(const FaImuDev& im, int lane, double scanTime, float pointTime) {
    const double* tm = im.time + (size_t)lane * IMU_QUE;
    const float* rg = im.ring + (size_t)lane * IMU_QUE * IMU_CH;
    const int last = im.ptr[lane * 2 + 0];
    int front = im.ptr[lane * 2 + 1];
    while (front != last) {
        if (scanTime + (double)pointTime < tm[front]) break;
        front = (front + 1) % IMU_QUE;
    }
    ImuSample s;
    const float* F = rg + front * IMU_CH;
    if (scanTime + (double)pointTime > tm[front]) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { s.rpy[k] = F[k]; s.velo[k] = F[3 + k]; s.shift[k] = F[6 + k]; s.ang[k] = F[9 + k]; }
        return s;
    }
    const int back = (front + IMU_QUE - 1) % IMU_QUE;
    const float* B = rg + back * IMU_CH;
    const float ratioFront = (float)((scanTime + (double)pointTime - tm[back]) / (tm[front] - tm[back]));
    const float ratioBack = (float)((tm[front] - scanTime - (double)pointTime) / (tm[front] - tm[back]));
    s.rpy[0] = F[0] * ratioFront + B[0] * ratioBack;
    s.rpy[1] = F[1] * ratioFront + B[1] * ratioBack;
    if ((double)(F[2] - B[2]) > M_PI) s.rpy[2] = (float)((double)(F[2] * ratioFront) + ((double)B[2] + 2 * M_PI) * (double)ratioBack);
    else if ((double)(F[2] - B[2]) < -M_PI) s.rpy[2] = (float)((double)(F[2] * ratioFront) + ((double)B[2] - 2 * M_PI) * (double)ratioBack);
    else s.rpy[2] = F[2] * ratioFront + B[2] * ratioBack;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        s.velo[k] = F[3 + k] * ratioFront + B[3 + k] * ratioBack;
        s.shift[k] = F[6 + k] * ratioFront + B[6 + k] * ratioBack;
        s.ang[k] = F[9 + k] * ratioFront + B[9 + k] * ratioBack;
    }
    return s;
}

// VeloToStartIMU (featureAssociation.cpp:346-363)
__device__ inline void imu_velo_to_start(const float* st, const float* veloCur, float* out) {
    const float vx = veloCur[0] - st[IS_VELO_START], vy = veloCur[1] - st[IS_VELO_START + 1], vz = veloCur[2] - st[IS_VELO_START + 2];
    const float x1 = st[IS_COS_YAW] * vx - st[IS_SIN_YAW] * vz;
    const float y1 = vy;
    const float z1 = st[IS_SIN_YAW] * vx + st[IS_COS_YAW] * vz;
    const float x2 = x1;
    const float y2 = st[IS_COS_PITCH] * y1 + st[IS_SIN_PITCH] * z1;
    const float z2 = -st[IS_SIN_PITCH] * y1 + st[IS_COS_PITCH] * z1;
    out[0] = st[IS_COS_ROLL] * x2 + st[IS_SIN_ROLL] * y2;
    out[1] = -st[IS_SIN_ROLL] * x2 + st[IS_COS_ROLL] * y2;
    out[2] = z2;
}

// TransformToStartIMU (featureAssociation.cpp:365-390); imuShiftFromStart*Cur = 0
__device__ inline void imu_transform_to_start(const float* st, const float* rpyCur, float& px, float& py, float& pz) {
    const float cr = pm::cosf_glibc(rpyCur[0]), sr = pm::sinf_glibc(rpyCur[0]), cp = pm::cosf_glibc(rpyCur[1]), sp = pm::sinf_glibc(rpyCur[1]);
    const float cy = pm::cosf_glibc(rpyCur[2]), sy = pm::sinf_glibc(rpyCur[2]);
    const float x1 = cr * px - sr * py;
    const float y1 = sr * px + cr * py;
    const float z1 = pz;
    const float x2 = x1;
    const float y2 = cp * y1 - sp * z1;
    const float z2 = sp * y1 + cp * z1;
    const float x3 = cy * x2 + sy * z2;
    const float y3 = y2;
    const float z3 = -sy * x2 + cy * z2;
    const float x4 = st[IS_COS_YAW] * x3 - st[IS_SIN_YAW] * z3;
    const float y4 = y3;
    const float z4 = st[IS_SIN_YAW] * x3 + st[IS_COS_YAW] * z3;
    const float x5 = x4;
    const float y5 = st[IS_COS_PITCH] * y4 + st[IS_SIN_PITCH] * z4;
    const float z5 = -st[IS_SIN_PITCH] * y4 + st[IS_COS_PITCH] * z4;
    px = st[IS_COS_ROLL] * x5 + st[IS_SIN_ROLL] * y5 + 0.f;
    py = -st[IS_SIN_ROLL] * x5 + st[IS_COS_ROLL] * y5 + 0.f;
    pz = z5 + 0.f;
}

// the IMU tail of TransformToEnd (featureAssociation.cpp:928-952) applied to (x6, y6, z6); imuShiftFromStart = 0
__device__ inline float4 imu_transform_to_end_tail(const float* st, float4 p) {
    const float x7 = st[IS_COS_ROLL] * (p.x - 0.f) - st[IS_SIN_ROLL] * (p.y - 0.f);
    const float y7 = st[IS_SIN_ROLL] * (p.x - 0.f) + st[IS_COS_ROLL] * (p.y - 0.f);
    const float z7 = p.z - 0.f;
    const float x8 = x7;
    const float y8 = st[IS_COS_PITCH] * y7 - st[IS_SIN_PITCH] * z7;
    const float z8 = st[IS_SIN_PITCH] * y7 + st[IS_COS_PITCH] * z7;
    const float x9 = st[IS_COS_YAW] * x8 + st[IS_SIN_YAW] * z8;
    const float y9 = y8;
    const float z9 = -st[IS_SIN_YAW] * x8 + st[IS_COS_YAW] * z8;
    const float rl = st[IS_RPY_LAST], pl = st[IS_RPY_LAST + 1], yl = st[IS_RPY_LAST + 2];
    const float cyl = pm::cosf_glibc(yl), syl = pm::sinf_glibc(yl), cpl = pm::cosf_glibc(pl), spl = pm::sinf_glibc(pl);
    const float crl = pm::cosf_glibc(rl), srl = pm::sinf_glibc(rl);
    const float x10 = cyl * x9 - syl * z9;
    const float y10 = y9;
    const float z10 = syl * x9 + cyl * z9;
    const float x11 = x10;
    const float y11 = cpl * y10 + spl * z10;
    const float z11 = -spl * y10 + cpl * z10;
    return make_float4(crl * x11 + srl * y11, -srl * x11 + crl * y11, z11, p.w);
}

}  // namespace lego
