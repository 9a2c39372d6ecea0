// Closed-form pose helpers of the LOAM "camera" frame (x left, y up, z forward; pose = [rx, ry, rz, tx, ty, tz],
// rotation R = Ry(ry) * Rx(rx) * Rz(rz)).  They restate, as explicit 3x3 products, what the reference spells
// out term by term in TransformToStart / TransformToEnd (featureAssociation.cpp:860-953), AccumulateRotation
// (featureAssociation.cpp:1015-1032), PluginIMURotation (featureAssociation.cpp:955-1013) and, for mapping,
// transformAssociateToMap / pointAssociateToMap (mapOptmization.cpp:376-461, 498-527).
// sin / cos of the point transforms go through pm::sinf_glibc / cosf_glibc (parity_math.cuh), which repeat glibc's
// evaluation bit for bit; the Euler-angle extraction (asinf / atan2f of CUDA) stays tolerance-level (1e-4 rad on poses).
#pragma once
#include <cuda_runtime.h>
#include "parity_math.cuh"

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

// AccumulateRotation (featureAssociation.cpp:1015-1032), term by term in the reference's order of operations: with the
// glibc-exact sin / cos / asin / atan2 of parity_math.cuh the composed angles come out bit for bit like the reference's.
__host__ __device__ inline void accumulate_rotation(float cx, float cy, float cz, float lx, float ly, float lz,
                                                    float* ox, float* oy, float* oz) {
    const float scx = pm::sinf_glibc(cx), ccx = pm::cosf_glibc(cx), scy = pm::sinf_glibc(cy), ccy = pm::cosf_glibc(cy);
    const float scz = pm::sinf_glibc(cz), ccz = pm::cosf_glibc(cz);
    const float slx = pm::sinf_glibc(lx), clx = pm::cosf_glibc(lx), sly = pm::sinf_glibc(ly), cly = pm::cosf_glibc(ly);
    const float slz = pm::sinf_glibc(lz), clz = pm::cosf_glibc(lz);
    const float srx = clx * ccx * sly * scz - ccx * ccz * slx - clx * cly * scx;
    const float rx = -pm::asinf_glibc(srx);
    const float crx = pm::cosf_glibc(rx);
    const float srycrx = slx * (ccy * scz - ccz * scx * scy) + clx * sly * (ccy * ccz + scx * scy * scz) + clx * cly * ccx * scy;
    const float crycrx = clx * cly * ccx * ccy - clx * sly * (ccz * scy - ccy * scx * scz) - slx * (scy * scz + ccy * ccz * scx);
    const float srzcrx = scx * (clz * sly - cly * slx * slz) + ccx * scz * (cly * clz + slx * sly * slz) + clx * ccx * ccz * slz;
    const float crzcrx = clx * clz * ccx * ccz - ccx * scz * (cly * slz - clz * slx * sly) - scx * (sly * slz + cly * clz * slx);
    *ox = rx;
    *oy = pm::atan2f_glibc(srycrx / crx, crycrx / crx);
    *oz = pm::atan2f_glibc(srzcrx / crx, crzcrx / crx);
}
// PluginIMURotation (featureAssociation.cpp:955-1013), term by term: bc = accumulated angles, bl = IMU angles at the
// start of the sweep, al = IMU angles at its end
__host__ __device__ inline void plugin_imu_rotation(float bcx, float bcy, float bcz, float blx, float bly, float blz,
                                                    float alx, float aly, float alz, float* acx, float* acy, float* acz) {
    const float sbcx = pm::sinf_glibc(bcx), cbcx = pm::cosf_glibc(bcx), sbcy = pm::sinf_glibc(bcy), cbcy = pm::cosf_glibc(bcy);
    const float sbcz = pm::sinf_glibc(bcz), cbcz = pm::cosf_glibc(bcz);
    const float sblx = pm::sinf_glibc(blx), cblx = pm::cosf_glibc(blx), sbly = pm::sinf_glibc(bly), cbly = pm::cosf_glibc(bly);
    const float sblz = pm::sinf_glibc(blz), cblz = pm::cosf_glibc(blz);
    const float salx = pm::sinf_glibc(alx), calx = pm::cosf_glibc(alx), saly = pm::sinf_glibc(aly), caly = pm::cosf_glibc(aly);
    const float salz = pm::sinf_glibc(alz), calz = pm::cosf_glibc(alz);
    const float srx = -sbcx * (salx * sblx + calx * caly * cblx * cbly + calx * cblx * saly * sbly)
                      - cbcx * cbcz * (calx * saly * (cbly * sblz - cblz * sblx * sbly) - calx * caly * (sbly * sblz + cbly * cblz * sblx) + cblx * cblz * salx)
                      - cbcx * sbcz * (calx * caly * (cblz * sbly - cbly * sblx * sblz) - calx * saly * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sblz);
    const float rx = -pm::asinf_glibc(srx);
    const float crx = pm::cosf_glibc(rx);
    const float srycrx = (cbcy * sbcz - cbcz * sbcx * sbcy) * (calx * saly * (cbly * sblz - cblz * sblx * sbly) - calx * caly * (sbly * sblz + cbly * cblz * sblx) + cblx * cblz * salx)
                         - (cbcy * cbcz + sbcx * sbcy * sbcz) * (calx * caly * (cblz * sbly - cbly * sblx * sblz) - calx * saly * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sblz)
                         + cbcx * sbcy * (salx * sblx + calx * caly * cblx * cbly + calx * cblx * saly * sbly);
    const float crycrx = (cbcz * sbcy - cbcy * sbcx * sbcz) * (calx * caly * (cblz * sbly - cbly * sblx * sblz) - calx * saly * (cbly * cblz + sblx * sbly * sblz) + cblx * salx * sblz)
                         - (sbcy * sbcz + cbcy * cbcz * sbcx) * (calx * saly * (cbly * sblz - cblz * sblx * sbly) - calx * caly * (sbly * sblz + cbly * cblz * sblx) + cblx * cblz * salx)
                         + cbcx * cbcy * (salx * sblx + calx * caly * cblx * cbly + calx * cblx * saly * sbly);
    const float srzcrx = sbcx * (cblx * cbly * (calz * saly - caly * salx * salz) - cblx * sbly * (caly * calz + salx * saly * salz) + calx * salz * sblx)
                         - cbcx * cbcz * ((caly * calz + salx * saly * salz) * (cbly * sblz - cblz * sblx * sbly) + (calz * saly - caly * salx * salz) * (sbly * sblz + cbly * cblz * sblx) - calx * cblx * cblz * salz)
                         + cbcx * sbcz * ((caly * calz + salx * saly * salz) * (cbly * cblz + sblx * sbly * sblz) + (calz * saly - caly * salx * salz) * (cblz * sbly - cbly * sblx * sblz) + calx * cblx * salz * sblz);
    const float crzcrx = sbcx * (cblx * sbly * (caly * salz - calz * salx * saly) - cblx * cbly * (saly * salz + caly * calz * salx) + calx * calz * sblx)
                         + cbcx * cbcz * ((saly * salz + caly * calz * salx) * (sbly * sblz + cbly * cblz * sblx) + (caly * salz - calz * salx * saly) * (cbly * sblz - cblz * sblx * sbly) + calx * calz * cblx * cblz)
                         - cbcx * sbcz * ((saly * salz + caly * calz * salx) * (cblz * sbly - cbly * sblx * sblz) + (caly * salz - calz * salx * saly) * (cbly * cblz + sblx * sbly * sblz) - calx * calz * cblx * sblz);
    *acx = rx;
    *acy = pm::atan2f_glibc(srycrx / crx, crycrx / crx);
    *acz = pm::atan2f_glibc(srzcrx / crx, crzcrx / crx);
}

struct SinCos6 { float srx, crx, sry, cry, srz, crz; };

// TransformToStart (featureAssociation.cpp:860-883): undo the fraction s = 10 * frac(intensity) of transformCur.
__host__ __device__ inline float4 transform_to_start(float4 p, const float* T) {
    const float s = 10.f * (p.w - (float)(int)p.w);
    const float rx = s * T[0], ry = s * T[1], rz = s * T[2], tx = s * T[3], ty = s * T[4], tz = s * T[5];
    const float crz = pm::cosf_glibc(rz), srz = pm::sinf_glibc(rz), crx = pm::cosf_glibc(rx), srx = pm::sinf_glibc(rx), cry = pm::cosf_glibc(ry), sry = pm::sinf_glibc(ry);
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
    const float crx = pm::cosf_glibc(T[0]), srx = pm::sinf_glibc(T[0]), cry = pm::cosf_glibc(T[1]), sry = pm::sinf_glibc(T[1]), crz = pm::cosf_glibc(T[2]), srz = pm::sinf_glibc(T[2]);
    const float x4 = cry * a.x + sry * a.z;
    const float z4 = -sry * a.x + cry * a.z;
    const float y5 = crx * a.y - srx * z4;
    const float z5 = srx * a.y + crx * z4;
    const float x6 = crz * x4 - srz * y5 + T[3];
    const float y6 = srz * x4 + crz * y5 + T[4];
    const float z6 = z5 + T[5];
    return make_float4(x6, y6, z6, (float)(int)p.w);
}

// integrateTransformation (featureAssociation.cpp:1697-1725): transformSum <- transformSum (+) -transformCur, then
// PluginIMURotation with imuRot = {imuPitchStart, imuYawStart, imuRollStart, imuPitchLast, imuYawLast, imuRollLast} (all 0
// without IMU data; imuShiftFromStart is always 0, quirk Q15)
__host__ __device__ inline void integrate_transformation(float* Sum, const float* Cur, const float* imuRot) {
    float rx, ry, rz;
    accumulate_rotation(Sum[0], Sum[1], Sum[2], -Cur[0], -Cur[1], -Cur[2], &rx, &ry, &rz);
    const float x1 = pm::cosf_glibc(rz) * Cur[3] - pm::sinf_glibc(rz) * Cur[4];
    const float y1 = pm::sinf_glibc(rz) * Cur[3] + pm::cosf_glibc(rz) * Cur[4];
    const float z1 = Cur[5];
    const float y2 = pm::cosf_glibc(rx) * y1 - pm::sinf_glibc(rx) * z1;
    const float z2 = pm::sinf_glibc(rx) * y1 + pm::cosf_glibc(rx) * z1;
    const float tx = Sum[3] - (pm::cosf_glibc(ry) * x1 + pm::sinf_glibc(ry) * z2);
    const float ty = Sum[4] - y2;
    const float tz = Sum[5] - (-pm::sinf_glibc(ry) * x1 + pm::cosf_glibc(ry) * z2);
    plugin_imu_rotation(rx, ry, rz, imuRot[0], imuRot[1], imuRot[2], imuRot[3], imuRot[4], imuRot[5], &rx, &ry, &rz);
    Sum[0] = rx; Sum[1] = ry; Sum[2] = rz; Sum[3] = tx; Sum[4] = ty; Sum[5] = tz;
}

}  // namespace pose
