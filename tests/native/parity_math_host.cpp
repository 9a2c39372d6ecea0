// Host check of csrc/parity_math.cuh against the C library's atanf/atan2f (glibc 2.39 in this image).
// Build: g++ -O2 -ffp-contract=off.  Prints the mismatch count over n random inputs per family; exit 0 iff all zero.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <random>
#include "../../lego-loam_comment_b200/csrc/parity_math.cuh"

int main(int argc, char** argv) {
    long n = argc > 1 ? atol(argv[1]) : 5000000;
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<float> U(-100.f, 100.f), Ur(1.f, 100.f), Ue(-20.f, 59.f);
    long bad_atan2 = 0, bad_seg = 0, bad_atan = 0, bad_bits = 0;
    for (long i = 0; i < n; ++i) {
        float y = U(rng), x = U(rng);
        if (i % 7 == 0) y *= 1e-6f;
        if (i % 11 == 0) x *= 1e-5f;
        if (pm::to_bits(pm::atan2f_glibc(y, x)) != pm::to_bits(atan2f(y, x))) ++bad_atan2;
        // segmentation form: atan2f(d2*sin(a), d1 - d2*cos(a))
        float d1 = Ur(rng), d2 = d1 * (0.5f + 0.5f * (float)(rng() % 1000) / 1000.f), sa = sinf(0.0034906585f), ca = cosf(0.0034906585f);
        float yy = d2 * sa, xx = d1 - d2 * ca;
        if (pm::to_bits(pm::atan2f_glibc(yy, xx)) != pm::to_bits(atan2f(yy, xx))) ++bad_seg;
        float t = ldexpf(1.f + (float)(rng() % 100000) / 100000.f, (int)Ue(rng)) * ((rng() & 1) ? 1.f : -1.f);
        if (pm::to_bits(pm::atanf_glibc(t)) != pm::to_bits(atanf(t))) ++bad_atan;
        uint32_t b1 = (uint32_t)rng(), b2 = (uint32_t)rng();
        float fy = pm::from_bits(b1), fx = pm::from_bits(b2);
        float r1 = pm::atan2f_glibc(fy, fx), r2 = atan2f(fy, fx);
        if (!(std::isnan(r1) && std::isnan(r2)) && pm::to_bits(r1) != pm::to_bits(r2)) ++bad_bits;
    }
    float sp[][2] = {{0.f, 1.f}, {-0.f, 1.f}, {0.f, -1.f}, {-0.f, -1.f}, {1.f, 0.f}, {-1.f, 0.f}, {1.f, -0.f}, {INFINITY, 1.f}, {1.f, INFINITY}, {1.f, -INFINITY}, {-INFINITY, -INFINITY}, {1e30f, 1e-30f}, {1e-30f, -1e30f}, {3.f, 1.f}};
    long bad_sp = 0;
    for (auto& p : sp) if (pm::to_bits(pm::atan2f_glibc(p[0], p[1])) != pm::to_bits(atan2f(p[0], p[1]))) ++bad_sp;
    printf("{\"n\": %ld, \"atan2\": %ld, \"seg\": %ld, \"atan\": %ld, \"rawbits\": %ld, \"special\": %ld}\n", n, bad_atan2, bad_seg, bad_atan, bad_bits, bad_sp);
    return (bad_atan2 | bad_seg | bad_atan | bad_bits | bad_sp) ? 1 : 0;
}
