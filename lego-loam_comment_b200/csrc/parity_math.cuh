// Bit-exact float32 helpers for kernels whose results feed integer / bit-exact decisions (projection, ground
// removal, segmentation, de-skew, curvature).  The reference calls atan2/sqrt/abs on floats through
// `using namespace std` (utility.h:49), i.e. glibc's atan2f; CUDA's atan2f is a different routine, so the device
// needs its own implementation that reproduces glibc 2.39's results bit for bit.  Every operation below is a
// single IEEE-754 binary32 add/sub/mul/div with round-to-nearest-even (explicit __f*_rn on the device so the
// compiler can never contract them into FMAs); the polynomial and reduction constants are the classic
// fdlibm/float ones, given as bit patterns (verified there against glibc on 2e8 inputs
// and re-verified by tests/test_parity_math.py on the host and tests/test_gpu_parity_math.py on the device).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define PM_HD __host__ __device__ __forceinline__
#define PM_ADD(a, b) __fadd_rn((a), (b))
#define PM_SUB(a, b) __fsub_rn((a), (b))
#define PM_MUL(a, b) __fmul_rn((a), (b))
#define PM_DIV(a, b) __fdiv_rn((a), (b))
#define PM_SQRT(a) __fsqrt_rn((a))
#elif defined(__CUDACC__)
#include <math.h>
#define PM_HD __host__ __device__ __forceinline__
#define PM_ADD(a, b) ((a) + (b))
#define PM_SUB(a, b) ((a) - (b))
#define PM_MUL(a, b) ((a) * (b))
#define PM_DIV(a, b) ((a) / (b))
#define PM_SQRT(a) sqrtf((a))
#else   // plain host build for unit tests: compile with -ffp-contract=off
#include <math.h>
#define PM_HD static inline
#define PM_ADD(a, b) ((a) + (b))
#define PM_SUB(a, b) ((a) - (b))
#define PM_MUL(a, b) ((a) * (b))
#define PM_DIV(a, b) ((a) / (b))
#define PM_SQRT(a) sqrtf((a))
#endif

namespace pm {

PM_HD float from_bits(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
PM_HD uint32_t to_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}

PM_HD float atan_poly_even(float w) {   // coefficients 0,2,4,6,8,10
    float r = PM_MUL(w, from_bits(0x3c8569d7u));
    r = PM_MUL(w, PM_ADD(from_bits(0x3d4bda59u), r));
    r = PM_MUL(w, PM_ADD(from_bits(0x3d886b35u), r));
    r = PM_MUL(w, PM_ADD(from_bits(0x3dba2e6eu), r));
    r = PM_MUL(w, PM_ADD(from_bits(0x3e124925u), r));
    return PM_ADD(from_bits(0x3eaaaaabu), r);
}
PM_HD float atan_poly_odd(float w) {    // coefficients 1,3,5,7,9
    float r = PM_MUL(w, from_bits(0xbd15a221u));
    r = PM_MUL(w, PM_ADD(from_bits(0xbd6ef16bu), r));
    r = PM_MUL(w, PM_ADD(from_bits(0xbd9d8795u), r));
    r = PM_MUL(w, PM_ADD(from_bits(0xbde38e38u), r));
    return PM_ADD(from_bits(0xbe4ccccdu), r);
}

// glibc 2.39 atanf, bit for bit.
PM_HD float atanf_glibc(float x) {
    const uint32_t hx = to_bits(x);
    const uint32_t ix = hx & 0x7fffffffu;
    if (ix >= 0x4c000000u) {                                  // |x| >= 2^25 (or NaN)
        if (ix > 0x7f800000u) return PM_ADD(x, x);
        float big = PM_ADD(from_bits(0x3fc90fdau), from_bits(0x33a22168u));
        return (hx >> 31) ? -big : big;
    }
    int id;
    if (ix < 0x3ee00000u) {                                   // |x| < 0.4375
        if (ix < 0x31000000u) return x;                       // tiny: atan(x) = x
        id = -1;
    } else {
        x = from_bits(ix);                                    // fabsf
        if (ix < 0x3f980000u) {
            if (ix < 0x3f300000u) { id = 0; x = PM_DIV(PM_SUB(PM_MUL(2.0f, x), 1.0f), PM_ADD(2.0f, x)); }
            else                  { id = 1; x = PM_DIV(PM_SUB(x, 1.0f), PM_ADD(x, 1.0f)); }
        } else {
            if (ix < 0x401c0000u) { id = 2; x = PM_DIV(PM_SUB(x, 1.5f), PM_ADD(1.0f, PM_MUL(1.5f, x))); }
            else                  { id = 3; x = PM_DIV(-1.0f, x); }
        }
    }
    const float z = PM_MUL(x, x);
    const float w = PM_MUL(z, z);
    const float s1 = PM_MUL(z, atan_poly_even(w));
    const float s2 = PM_MUL(w, atan_poly_odd(w));
    if (id < 0) return PM_SUB(x, PM_MUL(x, PM_ADD(s1, s2)));
    uint32_t hi_b, lo_b;
    switch (id) {
        case 0: hi_b = 0x3eed6338u; lo_b = 0x31ac3769u; break;
        case 1: hi_b = 0x3f490fdau; lo_b = 0x33222168u; break;
        case 2: hi_b = 0x3f7b985eu; lo_b = 0x33140fb4u; break;
        default: hi_b = 0x3fc90fdau; lo_b = 0x33a22168u; break;
    }
    const float r = PM_SUB(from_bits(hi_b), PM_SUB(PM_SUB(PM_MUL(x, PM_ADD(s1, s2)), from_bits(lo_b)), x));
    return (hx >> 31) ? -r : r;
}

// glibc 2.39 atan2f(y, x), bit for bit.
PM_HD float atan2f_glibc(float y, float x) {
    const float tiny = 1.0e-30f;
    const float pi = from_bits(0x40490fdbu), pi_o_2 = from_bits(0x3fc90fdbu), pi_o_4 = from_bits(0x3f490fdbu);
    const float pi_lo = from_bits(0xb3bbbd2eu);
    const uint32_t hx = to_bits(x), hy = to_bits(y);
    const uint32_t ix = hx & 0x7fffffffu, iy = hy & 0x7fffffffu;
    if (ix > 0x7f800000u || iy > 0x7f800000u) return PM_ADD(x, y);   // NaN
    if (hx == 0x3f800000u) return atanf_glibc(y);                    // x == 1
    const int m = (int)((hy >> 31) & 1u) | (int)((hx >> 30) & 2u);
    if (iy == 0) {
        switch (m) { case 0: case 1: return y; case 2: return PM_ADD(pi, tiny); default: return PM_SUB(-pi, tiny); }
    }
    if (ix == 0) return (hy >> 31) ? PM_SUB(-pi_o_2, tiny) : PM_ADD(pi_o_2, tiny);
    if (ix == 0x7f800000u) {
        if (iy == 0x7f800000u) {
            switch (m) {
                case 0: return PM_ADD(pi_o_4, tiny);
                case 1: return PM_SUB(-pi_o_4, tiny);
                case 2: return PM_ADD(PM_MUL(3.0f, pi_o_4), tiny);
                default: return PM_SUB(PM_MUL(-3.0f, pi_o_4), tiny);
            }
        } else {
            switch (m) { case 0: return 0.0f; case 1: return -0.0f; case 2: return PM_ADD(pi, tiny); default: return PM_SUB(-pi, tiny); }
        }
    }
    if (iy == 0x7f800000u) return (hy >> 31) ? PM_SUB(-pi_o_2, tiny) : PM_ADD(pi_o_2, tiny);
    const int k = ((int)iy - (int)ix) >> 23;
    float z;
    if (k > 60) z = PM_ADD(pi_o_2, PM_MUL(0.5f, pi_lo));
    else if ((hx >> 31) && k < -60) z = 0.0f;
    else z = atanf_glibc(from_bits(to_bits(PM_DIV(y, x)) & 0x7fffffffu));
    switch (m) {
        case 0: return z;
        case 1: return from_bits(to_bits(z) ^ 0x80000000u);
        case 2: return PM_SUB(pi, PM_SUB(z, pi_lo));
        default: return PM_SUB(PM_SUB(z, pi_lo), pi);
    }
}

// ---- glibc 2.39 sinf / cosf (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h), x86-64 FMA variant -------------
// glibc evaluates both in DOUBLE: range reduction by pi/2 (|y| < 120: one multiply and an integer round), then a degree-7
// (sine) or degree-8 (cosine) polynomial, rounded once to float.  Every step is an IEEE double operation, so the same
// steps on the device give the same float, bit for bit.  On x86-64 libm's ifunc picks the build compiled with -mfma, whose
// multiply-adds are contracted; the contraction pattern below was checked against the host's libm on 2*10^8 inputs with 0
// mismatches (the uncontracted form differs in ~2*10^-7 of the inputs).  Used where the reference evaluates sin / cos per
// POINT (TransformToStartIMU), where a last-ulp difference would otherwise show up in thousands of coordinates per scan.
// |y| >= 120 (never reached by angles in radians here) falls back to the CUDA library.
struct SinCosTab { double c0, c1, c2, c3, c4, s1, s2, s3; };
PM_HD float sincosf_poly(double x, double x2, bool negCos, int n) {
    const double sg = negCos ? -1.0 : 1.0;   // quadrants 2, 3 use the table with the cosine polynomial negated
    if ((n & 1) == 0) {
        const double x3 = x * x2;
        const double s1 = fma(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7);
        const double x7 = x3 * x2;
        const double s = fma(x3, -0x1.555545995a603p-3, x);
        return (float)fma(x7, s1, s);
    }
    const double x4 = x2 * x2;
    const double c2 = fma(x2, sg * 0x1.99343027bf8c3p-16, sg * -0x1.6c087e89a359dp-10);
    const double c1 = fma(x2, sg * -0x1.ffffffd0c621cp-2, sg * 0x1p0);
    const double x6 = x4 * x2;
    const double c = fma(x4, sg * 0x1.55553e1068f19p-5, c1);
    return (float)fma(x6, c2, c);
}
PM_HD uint32_t abstop12(float x) { return (to_bits(x) >> 20) & 0x7ffu; }
PM_HD float sincosf_glibc(float y, int cosine) {
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return cosine ? 1.0f : y;
        return sincosf_poly(x, x * x, false, cosine);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        const double r = x * 0x1.45F306DC9C883p+23;
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = fma(-(double)n, 0x1.921FB54442D18p0, x);
        const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;   // sign[] = {1, -1, -1, 1}
        return sincosf_poly(x * sign, x * x, (n & 2) != 0, n ^ cosine);
    }
    return cosine ? cosf(y) : sinf(y);
}
PM_HD float sinf_glibc(float y) { return sincosf_glibc(y, 0); }
PM_HD float cosf_glibc(float y) { return sincosf_glibc(y, 1); }

// ---- glibc 2.39 asinf (sysdeps/ieee754/flt-32/e_asinf.c): float polynomial, every step one IEEE binary32 operation ----
// Checked on the host against libm on 10^8 inputs in [-1, 1]: 0 mismatches.
PM_HD float asinf_glibc(float x) {
    const float pio2_hi = 1.57079637050628662109375f, pio2_lo = -4.37113900018624283e-8f, pio4_hi = 0.785398185253143310546875f;
    const float p0 = 1.666675248e-1f, p1 = 7.495297643e-2f, p2 = 4.547037598e-2f, p3 = 2.417951451e-2f, p4 = 4.216630880e-2f;
    const int32_t hx = (int32_t)to_bits(x), ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) return PM_ADD(PM_MUL(x, pio2_hi), PM_MUL(x, pio2_lo));
    if (ix > 0x3f800000) return PM_DIV(PM_SUB(x, x), PM_SUB(x, x));
    if (ix < 0x3f000000) {
        if (ix < 0x32000000) return x;
        const float t = PM_MUL(x, x);
        const float w = PM_MUL(t, PM_ADD(p0, PM_MUL(t, PM_ADD(p1, PM_MUL(t, PM_ADD(p2, PM_MUL(t, PM_ADD(p3, PM_MUL(t, p4)))))))));
        return PM_ADD(x, PM_MUL(x, w));
    }
    float w = PM_SUB(1.0f, from_bits(to_bits(x) & 0x7fffffffu));
    float t = PM_MUL(w, 0.5f);
    float p = PM_MUL(t, PM_ADD(p0, PM_MUL(t, PM_ADD(p1, PM_MUL(t, PM_ADD(p2, PM_MUL(t, PM_ADD(p3, PM_MUL(t, p4)))))))));
    const float s = sqrtf(t);
    if (ix >= 0x3F79999A) {
        t = PM_SUB(pio2_hi, PM_SUB(PM_MUL(2.0f, PM_ADD(s, PM_MUL(s, p))), pio2_lo));
    } else {
        w = from_bits(to_bits(s) & 0xfffff000u);
        const float c = PM_DIV(PM_SUB(t, PM_MUL(w, w)), PM_ADD(s, w));
        const float r = p;
        p = PM_SUB(PM_MUL(PM_MUL(2.0f, s), r), PM_SUB(pio2_lo, PM_MUL(2.0f, c)));
        const float q = PM_SUB(pio4_hi, PM_MUL(2.0f, w));
        t = PM_SUB(pio4_hi, PM_SUB(p, q));
    }
    return hx > 0 ? t : -t;
}

}  // namespace pm
