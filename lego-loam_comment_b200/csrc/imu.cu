// Host side of the optional IMU path: imuHandler of FeatureAssociation (featureAssociation.cpp:431-459 with
// AccumulateIMUShiftAndRotation 392-429) and of mapOptimization (mapOptmization.cpp:643-652), per lane.  These run once
// per IMU message on a few floats -- host code in the reference, host code here; what touches point clouds (the
// per-point interpolation and de-skew of adjustDistortion) runs in the kernels of imu_kernels.cuh.
#include <cmath>
#include <cstring>
#include "engine.h"

namespace lego {

// tf::Matrix3x3(tf::Quaternion).getRPY, double
static void quat_to_rpy(const double q[4], double* roll, double* pitch, double* yaw) {
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    const double d = x * x + y * y + z * z + w * w, s = 2.0 / d;
    const double xs = x * s, ys = y * s, zs = z * s;
    const double wx = w * xs, wy = w * ys, wz = w * zs, xx = x * xs, xy = x * ys, xz = x * zs, yy = y * ys, yz = y * zs, zz = z * zs;
    const double m00 = 1.0 - (yy + zz), m10 = xy + wz, m20 = xz - wy, m21 = yz + wx, m22 = 1.0 - (xx + yy);
    if (std::fabs(m20) >= 1.0) {
        *yaw = 0.0; *roll = std::atan2(m21, m22); *pitch = m20 < 0 ? M_PI / 2.0 : -M_PI / 2.0;
    } else {
        *pitch = -std::asin(m20);
        const double cp = std::cos(*pitch);
        *roll = std::atan2(m21 / cp, m22 / cp);
        *yaw = std::atan2(m10 / cp, m00 / cp);
    }
}

void Engine::imuHandler(int lane, double stamp, const double q[4], const double w[3], const double a[3]) {
    if (imuHost.empty()) imuHost.resize(c.L);
    ImuLaneHost& I = imuHost[lane];
    double roll, pitch, yaw;
    quat_to_rpy(q, &roll, &pitch, &yaw);
    // ---- FeatureAssociation::imuHandler (featureAssociation.cpp:431-459): gravity removed, axes remapped to the camera frame ----
    float accX = (float)(a[1] - std::sin(roll) * std::cos(pitch) * 9.81);
    float accY = (float)(a[2] - std::cos(roll) * std::cos(pitch) * 9.81);
    float accZ = (float)(a[0] + std::sin(pitch) * 9.81);
    const int P = I.pointerLast = (I.pointerLast + 1) % IMU_QUE;
    I.time[P] = stamp;
    I.roll[P] = (float)roll; I.pitch[P] = (float)pitch; I.yaw[P] = (float)yaw;
    I.acc[0][P] = accX; I.acc[1][P] = accY; I.acc[2][P] = accZ;
    for (int k = 0; k < 3; ++k) I.angVelo[k][P] = (float)w[k];
    {   // AccumulateIMUShiftAndRotation (featureAssociation.cpp:392-429)
        const float r = I.roll[P], p = I.pitch[P], y = I.yaw[P];
        const float x1 = cosf(r) * accX - sinf(r) * accY;
        const float y1 = sinf(r) * accX + cosf(r) * accY;
        const float z1 = accZ;
        const float x2 = x1;
        const float y2 = cosf(p) * y1 - sinf(p) * z1;
        const float z2 = sinf(p) * y1 + cosf(p) * z1;
        accX = cosf(y) * x2 + sinf(y) * z2;
        accY = y2;
        accZ = -sinf(y) * x2 + cosf(y) * z2;
        const float acc[3] = {accX, accY, accZ};
        const int B = (P + IMU_QUE - 1) % IMU_QUE;
        const double dt = I.time[P] - I.time[B];
        if (dt < (double)params.scanPeriod) {
            for (int k = 0; k < 3; ++k) {
                I.shift[k][P] = (float)((double)I.shift[k][B] + (double)I.velo[k][B] * dt + (double)acc[k] * dt * dt / 2);
                I.velo[k][P] = (float)((double)I.velo[k][B] + (double)acc[k] * dt);
                I.angRot[k][P] = (float)((double)I.angRot[k][B] + (double)I.angVelo[k][B] * dt);
            }
        }
    }
    I.dirty = true;
    anyImu = true;
    // ---- mapOptimization::imuHandler (mapOptmization.cpp:643-652) ----
    const int M = I.moPointerLast = (I.moPointerLast + 1) % IMU_QUE;
    I.moTime[M] = stamp; I.moRoll[M] = (float)roll; I.moPitch[M] = (float)pitch;
}

// Before adjustDistortion: the ring buffers of the lanes that got messages, the two pointers and timeScanCur.
// (Pageable sources: cudaMemcpyAsync stages them before it returns, so the host may keep filling the buffers.)
void Engine::uploadImu(const double* stamps) {
    if (!anyImu) return;
    std::vector<float> ring((size_t)IMU_QUE * IMU_CH);
    for (int l = 0; l < c.L; ++l) {
        ImuLaneHost& I = imuHost[l];
        if (I.pointerLast < 0) continue;
        if (I.dirty) {
            for (int i = 0; i < IMU_QUE; ++i) {
                float* o = ring.data() + (size_t)i * IMU_CH;
                o[0] = I.roll[i]; o[1] = I.pitch[i]; o[2] = I.yaw[i];
                for (int k = 0; k < 3; ++k) { o[3 + k] = I.velo[k][i]; o[6 + k] = I.shift[k][i]; o[9 + k] = I.angRot[k][i]; }
            }
            cuda_check(cudaMemcpyAsync(fa.imu.ring + (size_t)l * IMU_QUE * IMU_CH, ring.data(), sizeof(float) * ring.size(), cudaMemcpyHostToDevice, stream), "imu ring");
            cuda_check(cudaMemcpyAsync(fa.imu.time + (size_t)l * IMU_QUE, I.time, sizeof(double) * IMU_QUE, cudaMemcpyHostToDevice, stream), "imu time");
            I.dirty = false;
        }
        const int ptr[2] = {I.pointerLast, I.pointerLastIteration};
        cuda_check(cudaMemcpyAsync(fa.imu.ptr + l * 2, ptr, sizeof(ptr), cudaMemcpyHostToDevice, stream), "imu ptr");
        cuda_check(cudaMemcpyAsync(fa.imu.scanTime + l, stamps + l, sizeof(double), cudaMemcpyHostToDevice, stream), "imu scan time");
        I.pointerLastIteration = I.pointerLast;   // featureAssociation.cpp:618
    }
}

// transformUpdate's IMU part (mapOptmization.cpp:465-490) for the lanes whose mapping body runs: the roll / pitch the IMU
// reports at the end of the sweep, blended into transformTobeMapped on the device once the LM loop is done.
void Engine::mapOptimizationImuBlend(const std::vector<int>& active) {
    if (!anyImu) return;
    std::vector<float> blend((size_t)c.L * 4, 0.f);
    for (int l = 0; l < c.L; ++l) {
        ImuLaneHost& I = imuHost[l];
        if (!active[l] || I.moPointerLast < 0) continue;
        const double t = moHost[l].timeOdometry + (double)params.scanPeriod;
        while (I.moPointerFront != I.moPointerLast) {
            if (t < I.moTime[I.moPointerFront]) break;
            I.moPointerFront = (I.moPointerFront + 1) % IMU_QUE;
        }
        float rollLast, pitchLast;
        const int F = I.moPointerFront;
        if (t > I.moTime[F]) { rollLast = I.moRoll[F]; pitchLast = I.moPitch[F]; }
        else {
            const int B = (F + IMU_QUE - 1) % IMU_QUE;
            const float ratioFront = (float)((moHost[l].timeOdometry + (double)params.scanPeriod - I.moTime[B]) / (I.moTime[F] - I.moTime[B]));
            const float ratioBack = (float)((I.moTime[F] - moHost[l].timeOdometry - (double)params.scanPeriod) / (I.moTime[F] - I.moTime[B]));
            rollLast = I.moRoll[F] * ratioFront + I.moRoll[B] * ratioBack;
            pitchLast = I.moPitch[F] * ratioFront + I.moPitch[B] * ratioBack;
        }
        blend[l * 4 + 0] = 1.f; blend[l * 4 + 1] = rollLast; blend[l * 4 + 2] = pitchLast;
    }
    cuda_check(cudaMemcpyAsync(mo.imuBlend, blend.data(), sizeof(float) * blend.size(), cudaMemcpyHostToDevice, stream), "imu blend");
}

}  // namespace lego
