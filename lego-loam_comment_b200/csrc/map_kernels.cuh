// Device side of mapOptimization's scan-to-map step (reference LeGO-LOAM/src/mapOptmization.cpp): pose bookkeeping
// (transformAssociateToMap 376-461, transformUpdate 463-496), the per-point line / plane residuals of
// cornerOptimization (1093-1174) and surfOptimization (1176-1227), the 6x6 Gauss-Newton step of LMOptimization
// (1229-1327) and the key-frame decision of saveKeyFramesAndFactor (1353-1454).
#pragma once
#include <cfloat>
#include "engine.h"
#include "kernel_utils.cuh"
#include "parity_math.cuh"
#include "smallmat.cuh"
#include "voxel_hash.cuh"

namespace lego {

constexpr int MO_ACC = 28;   // 21 upper-triangle entries of AtA, 6 of AtB, 1 row count

// The odometry pose reaches mapOptimization through a nav_msgs/Odometry message: publishOdometry builds a quaternion
// from (roll, pitch, yaw) = (rz, -rx, -ry) (featureAssociation.cpp:1727-1738) and laserOdometryHandler turns it back
// with tf::Matrix3x3::getRPY (mapOptmization.cpp:608-621), all in double.  Restated here so the float values that
// come out are the ones the reference's mapping stage sees.
__device__ inline void odometry_message_round_trip(const float* S, float* out) {
    const double roll = (double)S[2], pitch = -(double)S[0], yaw = -(double)S[1];
    const double hr = roll * 0.5, hp = pitch * 0.5, hy = yaw * 0.5;
    const double cr = cos(hr), sr = sin(hr), cp = cos(hp), sp = sin(hp), cy = cos(hy), sy = sin(hy);
    const double qx = sr * cp * cy - cr * sp * sy, qy = cr * sp * cy + sr * cp * sy;
    const double qz = cr * cp * sy - sr * sp * cy, qw = cr * cp * cy + sr * sp * sy;
    const double d = qx * qx + qy * qy + qz * qz + qw * qw, s = 2.0 / d;
    const double xs = qx * s, ys = qy * s, zs = qz * s;
    const double wx = qw * xs, wy = qw * ys, wz = qw * zs, xx = qx * xs, xy = qx * ys, xz = qx * zs;
    const double yy = qy * ys, yz = qy * zs, zz = qz * zs;
    const double m00 = 1.0 - (yy + zz), m10 = xy + wz, m20 = xz - wy, m21 = yz + wx, m22 = 1.0 - (xx + yy);
    double r, p, y;
    if (fabs(m20) >= 1.0) {
        y = 0.0; r = atan2(m21, m22); p = m20 < 0 ? M_PI / 2.0 : -M_PI / 2.0;
    } else {
        p = -asin(m20);
        const double c = cos(p);
        r = atan2(m21 / c, m22 / c);
        y = atan2(m10 / c, m00 / c);
    }
    out[0] = (float)(-p); out[1] = (float)(-y); out[2] = (float)r;
    out[3] = S[3]; out[4] = S[4]; out[5] = S[5];
}

// transformAssociateToMap (mapOptmization.cpp:376-461), term by term.
__device__ inline void transform_associate_to_map(const float* Sum, const float* Bef, const float* Aft, float* Incre, float* T) {
    float x1 = pm::cosf_glibc(Sum[1]) * (Bef[3] - Sum[3]) - pm::sinf_glibc(Sum[1]) * (Bef[5] - Sum[5]);
    float y1 = Bef[4] - Sum[4];
    float z1 = pm::sinf_glibc(Sum[1]) * (Bef[3] - Sum[3]) + pm::cosf_glibc(Sum[1]) * (Bef[5] - Sum[5]);
    float x2 = x1;
    float y2 = pm::cosf_glibc(Sum[0]) * y1 + pm::sinf_glibc(Sum[0]) * z1;
    float z2 = -pm::sinf_glibc(Sum[0]) * y1 + pm::cosf_glibc(Sum[0]) * z1;
    Incre[3] = pm::cosf_glibc(Sum[2]) * x2 + pm::sinf_glibc(Sum[2]) * y2;
    Incre[4] = -pm::sinf_glibc(Sum[2]) * x2 + pm::cosf_glibc(Sum[2]) * y2;
    Incre[5] = z2;
    const float sbcx = pm::sinf_glibc(Sum[0]), cbcx = pm::cosf_glibc(Sum[0]), sbcy = pm::sinf_glibc(Sum[1]), cbcy = pm::cosf_glibc(Sum[1]), sbcz = pm::sinf_glibc(Sum[2]), cbcz = pm::cosf_glibc(Sum[2]);
    const float sblx = pm::sinf_glibc(Bef[0]), cblx = pm::cosf_glibc(Bef[0]), sbly = pm::sinf_glibc(Bef[1]), cbly = pm::cosf_glibc(Bef[1]), sblz = pm::sinf_glibc(Bef[2]), cblz = pm::cosf_glibc(Bef[2]);
    const float salx = pm::sinf_glibc(Aft[0]), calx = pm::cosf_glibc(Aft[0]), saly = pm::sinf_glibc(Aft[1]), caly = pm::cosf_glibc(Aft[1]), salz = pm::sinf_glibc(Aft[2]), calz = pm::cosf_glibc(Aft[2]);
    const float srx = -sbcx * (salx * sblx + calx * cblx * salz * sblz + calx * calz * cblx * cblz)
                      - cbcx * sbcy * (calx * calz * (cbly * sblz - cblz * sblx * sbly) - calx * salz * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sbly)
                      - cbcx * cbcy * (calx * salz * (cblz * sbly - cbly * sblx * sblz) - calx * calz * (sbly * sblz + cbly * cblz * sblx) + cblx * cbly * salx);
    T[0] = -pm::asinf_glibc(srx);
    const float srycrx = sbcx * (cblx * cblz * (caly * salz - calz * salx * saly) - cblx * sblz * (caly * calz + salx * saly * salz) + calx * saly * sblx)
                         - cbcx * cbcy * ((caly * calz + salx * saly * salz) * (cblz * sbly - cbly * sblx * sblz)
                                          + (caly * salz - calz * salx * saly) * (sbly * sblz + cbly * cblz * sblx) - calx * cblx * cbly * saly)
                         + cbcx * sbcy * ((caly * calz + salx * saly * salz) * (cbly * cblz + sblx * sbly * sblz)
                                          + (caly * salz - calz * salx * saly) * (cbly * sblz - cblz * sblx * sbly) + calx * cblx * saly * sbly);
    const float crycrx = sbcx * (cblx * sblz * (calz * saly - caly * salx * salz) - cblx * cblz * (saly * salz + caly * calz * salx) + calx * caly * sblx)
                         + cbcx * cbcy * ((saly * salz + caly * calz * salx) * (sbly * sblz + cbly * cblz * sblx)
                                          + (calz * saly - caly * salx * salz) * (cblz * sbly - cbly * sblx * sblz) + calx * caly * cblx * cbly)
                         - cbcx * sbcy * ((saly * salz + caly * calz * salx) * (cbly * sblz - cblz * sblx * sbly)
                                          + (calz * saly - caly * salx * salz) * (cbly * cblz + sblx * sbly * sblz) - calx * caly * cblx * sbly);
    T[1] = pm::atan2f_glibc(srycrx / pm::cosf_glibc(T[0]), crycrx / pm::cosf_glibc(T[0]));
    const float srzcrx = (cbcz * sbcy - cbcy * sbcx * sbcz) * (calx * salz * (cblz * sbly - cbly * sblx * sblz) - calx * calz * (sbly * sblz + cbly * cblz * sblx) + cblx * cbly * salx)
                         - (cbcy * cbcz + sbcx * sbcy * sbcz) * (calx * calz * (cbly * sblz - cblz * sblx * sbly) - calx * salz * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sbly)
                         + cbcx * sbcz * (salx * sblx + calx * cblx * salz * sblz + calx * calz * cblx * cblz);
    const float crzcrx = (cbcy * sbcz - cbcz * sbcx * sbcy) * (calx * calz * (cbly * sblz - cblz * sblx * sbly) - calx * salz * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sbly)
                         - (sbcy * sbcz + cbcy * cbcz * sbcx) * (calx * salz * (cblz * sbly - cbly * sblx * sblz) - calx * calz * (sbly * sblz + cbly * cblz * sblx) + cblx * cbly * salx)
                         + cbcx * cbcz * (salx * sblx + calx * cblx * salz * sblz + calx * calz * cblx * cblz);
    T[2] = pm::atan2f_glibc(srzcrx / pm::cosf_glibc(T[0]), crzcrx / pm::cosf_glibc(T[0]));
    x1 = pm::cosf_glibc(T[2]) * Incre[3] - pm::sinf_glibc(T[2]) * Incre[4];
    y1 = pm::sinf_glibc(T[2]) * Incre[3] + pm::cosf_glibc(T[2]) * Incre[4];
    z1 = Incre[5];
    x2 = x1;
    y2 = pm::cosf_glibc(T[0]) * y1 - pm::sinf_glibc(T[0]) * z1;
    z2 = pm::sinf_glibc(T[0]) * y1 + pm::cosf_glibc(T[0]) * z1;
    T[3] = Aft[3] - (pm::cosf_glibc(T[1]) * x2 + pm::sinf_glibc(T[1]) * z2);
    T[4] = Aft[4] - y2;
    T[5] = Aft[5] - (-pm::sinf_glibc(T[1]) * x2 + pm::cosf_glibc(T[1]) * z2);
}

struct MapSinCos { float cRoll, sRoll, cPitch, sPitch, cYaw, sYaw, tX, tY, tZ; };
// updatePointAssociateToMapSinCos (mapOptmization.cpp:498-511)
__device__ __forceinline__ MapSinCos map_sincos(const float* T) {
    MapSinCos m;
    m.cRoll = pm::cosf_glibc(T[0]); m.sRoll = pm::sinf_glibc(T[0]); m.cPitch = pm::cosf_glibc(T[1]); m.sPitch = pm::sinf_glibc(T[1]); m.cYaw = pm::cosf_glibc(T[2]); m.sYaw = pm::sinf_glibc(T[2]);
    m.tX = T[3]; m.tY = T[4]; m.tZ = T[5];
    return m;
}
// The same for a whole CTA: six threads evaluate one value each (the double-precision sin / cos restatements are costly
// when every thread of every CTA repeats them), everybody reads the result.  Contains a __syncthreads().
__device__ __forceinline__ MapSinCos map_sincos_cta(const float* T) {
    __shared__ float sTrig[6];
    if (threadIdx.x < 6) sTrig[threadIdx.x] = pm::sincosf_glibc(T[threadIdx.x >> 1], (threadIdx.x & 1) ^ 1);   // cos, sin, cos, sin, ...
    __syncthreads();
    MapSinCos m;
    m.cRoll = sTrig[0]; m.sRoll = sTrig[1]; m.cPitch = sTrig[2]; m.sPitch = sTrig[3]; m.cYaw = sTrig[4]; m.sYaw = sTrig[5];
    m.tX = T[3]; m.tY = T[4]; m.tZ = T[5];
    return m;
}
// pointAssociateToMap (mapOptmization.cpp:513-527) and transformPointCloud (mapOptmization.cpp:545-575): same formula
__device__ __forceinline__ float4 map_transform(const MapSinCos& m, float4 p) {
    const float x1 = m.cYaw * p.x - m.sYaw * p.y;
    const float y1 = m.sYaw * p.x + m.cYaw * p.y;
    const float z1 = p.z;
    const float x2 = x1;
    const float y2 = m.cRoll * y1 - m.sRoll * z1;
    const float z2 = m.sRoll * y1 + m.cRoll * z1;
    return make_float4(m.cPitch * x2 + m.sPitch * z2 + m.tX, y2 + m.tY, -m.sPitch * x2 + m.cPitch * z2 + m.tZ, p.w);
}

}  // namespace lego
