// Closed-form pose helpers of the LOAM "camera" frame (x left, y up, z forward; pose = [rx, ry, rz, tx, ty, tz],
// rotation R = Ry(ry) * Rx(rx) * Rz(rz)).  They restate, as explicit 3x3 products, what the reference spells
// out term by term in TransformToStart / TransformToEnd (featureAssociation.cpp:860-953), AccumulateRotation
// (featureAssociation.cpp:1015-1032), PluginIMURotation (featureAssociation.cpp:955-1013) and, for mapping,
// transformAssociateToMap / pointAssociateToMap (mapOptmization.cpp:376-461, 498-527).
// Results are tolerance-level (1e-4 m / 1e-4 rad on poses), not bit-exact: CUDA sinf/cosf/asinf/atan2f differ
// from glibc in the last ulp.
#pragma once
#include <cuda_runtime.h>

namespace pose {

struct Mat3 { float m[3][3]; };

__host__ __device__ inline Mat3 rot_yxz(float rx, float ry, float rz) {
    const float sx = sinf(rx), cx = cosf(rx), sy = sinf(ry), cy = cosf(ry), sz = sinf(rz), cz = cosf(rz);
    Mat3 R;
    R.m[0][0] = cy * cz + sy * sx * sz; R.m[0][1] = -cy * sz + sy * sx * cz; R.m[0][2] = sy * cx;
    R.m[1][0] = cx * sz;                R.m[1][1] = cx * cz;                 R.m[1][2] = -sx;
    R.m[2][0] = -sy * cz + cy * sx * sz; R.m[2][1] = sy * sz + cy * sx * cz; R.m[2][2] = cy * cx;
    return R;
}
__host__ __device__ inline Mat3 mul(const Mat3& a, const Mat3& b) {
    Mat3 c;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) c.m[i][j] = a.m[i][0] * b.m[0][j] + a.m[i][1] * b.m[1][j] + a.m[i][2] * b.m[2][j];
    return c;
}
__host__ __device__ inline Mat3 transpose(const Mat3& a) {
    Mat3 c;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) c.m[i][j] = a.m[j][i];
    return c;
}
// Euler angles of R = Ry*Rx*Rz, the extraction the reference uses: rx = -asin(R12), ry = atan2(R02/cx, R22/cx), ...
__host__ __device__ inline void euler_yxz(const Mat3& R, float* rx, float* ry, float* rz) {
    const float ox = -asinf(R.m[1][2]);
    const float cox = cosf(ox);
    *rx = ox;
    *ry = atan2f(R.m[0][2] / cox, R.m[2][2] / cox);
    *rz = atan2f(R.m[1][0] / cox, R.m[1][1] / cox);
}

// AccumulateRotation (featureAssociation.cpp:1015-1032): o = angles of R(c) * R(l)
__host__ __device__ inline void accumulate_rotation(float cx, float cy, float cz, float lx, float ly, float lz,
                                                    float* ox, float* oy, float* oz) {
    euler_yxz(mul(rot_yxz(cx, cy, cz), rot_yxz(lx, ly, lz)), ox, oy, oz);
}
// PluginIMURotation (featureAssociation.cpp:955-1013): ac = angles of R(bc) * R(bl)^T * R(al)
__host__ __device__ inline void plugin_imu_rotation(float bcx, float bcy, float bcz, float blx, float bly, float blz,
                                                    float alx, float aly, float alz, float* acx, float* acy, float* acz) {
    const Mat3 X = mul(transpose(rot_yxz(blx, bly, blz)), rot_yxz(alx, aly, alz));
    euler_yxz(mul(rot_yxz(bcx, bcy, bcz), X), acx, acy, acz);
}

struct SinCos6 { float srx, crx, sry, cry, srz, crz; };

// TransformToStart (featureAssociation.cpp:860-883): undo the fraction s = 10 * frac(intensity) of transformCur.
__host__ __device__ inline float4 transform_to_start(float4 p, const float* T) {
    const float s = 10.f * (p.w - (float)(int)p.w);
    const float rx = s * T[0], ry = s * T[1], rz = s * T[2], tx = s * T[3], ty = s * T[4], tz = s * T[5];
    const float crz = cosf(rz), srz = sinf(rz), crx = cosf(rx), srx = sinf(rx), cry = cosf(ry), sry = sinf(ry);
    const float x1 = crz * (p.x - tx) + srz * (p.y - ty);
    const float y1 = -srz * (p.x - tx) + crz * (p.y - ty);
    const float z1 = p.z - tz;
    const float y2 = crx * y1 + srx * z1;
    const float z2 = -srx * y1 + crx * z1;
    return make_float4(cry * x1 - sry * z2, y2, sry * x1 + cry * z2, p.w);
}

// TransformToEnd without IMU terms (featureAssociation.cpp:885-953; the IMU shift / start / last rotations are
// identities when no IMU message ever arrived): to sweep start, then forward by the whole transformCur.
__host__ __device__ inline float4 transform_to_end(float4 p, const float* T) {
    const float4 a = transform_to_start(p, T);
    const float crx = cosf(T[0]), srx = sinf(T[0]), cry = cosf(T[1]), sry = sinf(T[1]), crz = cosf(T[2]), srz = sinf(T[2]);
    const float x4 = cry * a.x + sry * a.z;
    const float z4 = -sry * a.x + cry * a.z;
    const float y5 = crx * a.y - srx * z4;
    const float z5 = srx * a.y + crx * z4;
    const float x6 = crz * x4 - srz * y5 + T[3];
    const float y6 = srz * x4 + crz * y5 + T[4];
    const float z6 = z5 + T[5];
    return make_float4(x6, y6, z6, (float)(int)p.w);
}

// integrateTransformation (featureAssociation.cpp:1697-1725) without IMU: transformSum <- transformSum (+) -transformCur
__host__ __device__ inline void integrate_transformation(float* Sum, const float* Cur) {
    float rx, ry, rz;
    accumulate_rotation(Sum[0], Sum[1], Sum[2], -Cur[0], -Cur[1], -Cur[2], &rx, &ry, &rz);
    const float x1 = cosf(rz) * Cur[3] - sinf(rz) * Cur[4];
    const float y1 = sinf(rz) * Cur[3] + cosf(rz) * Cur[4];
    const float z1 = Cur[5];
    const float y2 = cosf(rx) * y1 - sinf(rx) * z1;
    const float z2 = sinf(rx) * y1 + cosf(rx) * z1;
    const float tx = Sum[3] - (cosf(ry) * x1 + sinf(ry) * z2);
    const float ty = Sum[4] - y2;
    const float tz = Sum[5] - (-sinf(ry) * x1 + cosf(ry) * z2);
    plugin_imu_rotation(rx, ry, rz, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, &rx, &ry, &rz);
    Sum[0] = rx; Sum[1] = ry; Sum[2] = rz; Sum[3] = tx; Sum[4] = ty; Sum[5] = tz;
}

}  // namespace pose
