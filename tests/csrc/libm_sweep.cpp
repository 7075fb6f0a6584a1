// CPU check of lidar_odometry_demo_b200/csrc/lo_libm.cuh against the host's libm (the functions the reference calls).
// Prints "name mismatches total" per function; tests/test_libm_cpu.py asserts zero mismatches.
#define LO_LIBM_HOST_ONLY
#include "../../lidar_odometry_demo_b200/csrc/lo_libm.cuh"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <random>

static bool same(float a, float b) { return lo::libm::float_word(a) == lo::libm::float_word(b) || (std::isnan(a) && std::isnan(b)); }

int main() {
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<float> coord(-120.f, 120.f), unit(-1.f, 1.f), small(-0.7853f, 0.7853f);
    long bad_atan2 = 0, bad_acos = 0, bad_sin = 0, bad_atan = 0;
    const long n = 4000000;
    for (long i = 0; i < n; i++) {
        const float y = coord(rng), x = coord(rng);
        if (!same(lo::libm::atan2f_fdlibm(y, x), atan2f(y, x))) bad_atan2++;
        const float a = unit(rng);
        if (!same(lo::libm::acosf_fdlibm(a), acosf(a))) bad_acos++;
        const float s = (i % 4 == 3) ? coord(rng) : small(rng) * (i % 3 == 0 ? 0.01f : 1.f);
        if (!same(lo::libm::sinf_glibc(s), sinf(s))) bad_sin++;
        const float t = coord(rng) * (i % 2 ? 0.02f : 1.f);
        if (!same(lo::libm::atanf_fdlibm(t), atanf(t))) bad_atan++;
    }
    // slerp-shaped acos arguments: |d| within 1e-3 of 1
    long bad_acos1 = 0;
    for (uint32_t k = 0; k < 4000000; k++) {
        const float a = lo::libm::word_float(0x3f800000 - static_cast<int32_t>(k));
        if (!same(lo::libm::acosf_fdlibm(a), acosf(a))) bad_acos1++;
    }
    // special values
    const float sp[] = {0.f, -0.f, 1.f, -1.f, INFINITY, -INFINITY, 1e-30f, -1e-30f, 1e30f, 3.4e38f, 0.4375f, 0.6875f, 1.1875f, 2.4375f};
    long bad_sp = 0;
    for (float y : sp)
        for (float x : sp)
            if (!same(lo::libm::atan2f_fdlibm(y, x), atan2f(y, x))) bad_sp++;
    printf("atan2f %ld %ld\nacosf %ld %ld\nsinf %ld %ld\natanf %ld %ld\nacosf_near1 %ld 4000000\natan2f_special %ld %zu\n", bad_atan2, n, bad_acos, n,
           bad_sin, n, bad_atan, n, bad_acos1, bad_sp, sizeof(sp) / sizeof(sp[0]) * sizeof(sp) / sizeof(sp[0]));
    return 0;
}
