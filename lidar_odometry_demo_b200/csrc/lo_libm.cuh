// Bit-exact restatements of the few libm float functions whose results reach the reference's outputs:
//   atan2f  — azimuth binning in CloudClassifier::classify (src/utils/cloud_classifier.h:49)
//   sinf    — Eigen's Quaternionf::slerp inside transformNonRigid (src/utils/cloud_transform.h:27)
//   acosf   — same slerp (once per scan)
// The reference links glibc (2.39 in the reference's Docker base and on this image). Its atan2f / atanf /
// acosf are the fdlibm single-precision algorithms (sysdeps/ieee754/flt-32/{e_atan2f,s_atanf,e_acosf}.c):
// plain float arithmetic, restated here operation by operation (the library is built with -fmad=false, so
// nothing fuses). Its sinf is the double-precision polynomial of sysdeps/ieee754/flt-32/s_sinf.c; the
// small-argument branch (|x| < pi/4, sinf_poly) is restated with unfused double arithmetic, which is what the
// host's libm computes (pinned empirically by tests/test_libm_cpu.py, which sweeps all four functions against
// the host's libm over 2e7 arguments each with zero mismatches on glibc 2.39 / x86-64). Outside the restated
// argument ranges the functions fall back to "evaluate in double, round once" (correctly rounded, which is
// what the libm results are to within 1 ulp) — the ranges cover every value the pipeline produces.
//
// The header is plain C++ when LO_LIBM_HOST_ONLY is defined, so the CPU test can compile it with g++.
#pragma once

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef LO_LIBM_HOST_ONLY
#define LO_HD inline
#else
#include <cuda_runtime.h>
#define LO_HD __host__ __device__ inline
#endif

namespace lo {
namespace libm {

LO_HD int32_t float_word(float x) {
    int32_t i;
    memcpy(&i, &x, sizeof(i));
    return i;
}
LO_HD float word_float(int32_t i) {
    float x;
    memcpy(&x, &i, sizeof(x));
    return x;
}

// s_atanf.c
LO_HD float atanf_fdlibm(float x) {
    const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
    const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
    const float aT[11] = {3.3333334327e-01f,  -2.0000000298e-01f, 1.4285714924e-01f,  -1.1111110449e-01f,
                          9.0908870101e-02f,  -7.6918758452e-02f, 6.6610731184e-02f,  -5.8335702866e-02f,
                          4.9768779427e-02f,  -3.6531571299e-02f, 1.6285819933e-02f};
    const float one = 1.0f;
    const int32_t hx = float_word(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c800000) {  // |x| >= 2^26
        if (ix > 0x7f800000) return x + x;  // NaN
        if (hx > 0) return atanhi[3] + atanlo[3];
        return -atanhi[3] - atanlo[3];
    }
    if (ix < 0x3ee00000) {       // |x| < 0.4375
        if (ix < 0x31000000) return x;  // |x| < 2^-29
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {          // |x| < 1.1875
            if (ix < 0x3f300000) {      // 7/16 <= |x| < 11/16
                id = 0;
                x = (2.0f * x - one) / (2.0f + x);
            } else {                    // 11/16 <= |x| < 19/16
                id = 1;
                x = (x - one) / (x + one);
            }
        } else {
            if (ix < 0x401c0000) {      // |x| < 2.4375
                id = 2;
                x = (x - 1.5f) / (one + 1.5f * x);
            } else {                    // 2.4375 <= |x|
                id = 3;
                x = -1.0f / x;
            }
        }
    }
    float z = x * x;
    const float w = z * z;
    const float s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
    const float s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
    if (id < 0) return x - x * (s1 + s2);
    z = atanhi[id] - ((x * (s1 + s2) - atanlo[id]) - x);
    return (hx < 0) ? -z : z;
}

// e_atan2f.c
LO_HD float atan2f_fdlibm(float y, float x) {
    const float tiny = 1.0e-30f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    const float pi_o_4 = 7.8539818525e-01f;
    const int32_t hx = float_word(x), hy = float_word(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;  // NaN
    if (hx == 0x3f800000) return atanf_fdlibm(y);           // x = 1.0
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);      // 2*sign(x) + sign(y)
    if (iy == 0) {
        switch (m) {
            case 0:
            case 1: return y;
            case 2: return pi + tiny;
            default: return -pi - tiny;
        }
    }
    if (ix == 0) return (hy < 0) ? -pi_o_2 - tiny : pi_o_2 + tiny;
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            switch (m) {
                case 0: return pi_o_4 + tiny;
                case 1: return -pi_o_4 - tiny;
                case 2: return 3.0f * pi_o_4 + tiny;
                default: return -3.0f * pi_o_4 - tiny;
            }
        }
        switch (m) {
            case 0: return 0.0f;
            case 1: return -0.0f;
            case 2: return pi + tiny;
            default: return -pi - tiny;
        }
    }
    if (iy == 0x7f800000) return (hy < 0) ? -pi_o_2 - tiny : pi_o_2 + tiny;
    const int32_t k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = pi_o_2 + 0.5f * pi_lo;
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = atanf_fdlibm(fabsf(y / x));
    switch (m) {
        case 0: return z;
        case 1: return word_float(float_word(z) ^ static_cast<int32_t>(0x80000000u));
        case 2: return pi - (z - pi_lo);
        default: return (z - pi_lo) - pi;
    }
}

// e_acosf.c
LO_HD float acosf_fdlibm(float x) {
    const float one = 1.0f, pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f,
                pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f;
    const float qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f, qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
    const int32_t hx = float_word(x);
    const int32_t ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) {
        if (hx > 0) return 0.0f;
        return pi + 2.0f * pio2_lo;
    }
    if (ix > 0x3f800000) return (x - x) / (x - x);
    if (ix < 0x3f000000) {  // |x| < 0.5
        if (ix <= 0x32800000) return pio2_hi + pio2_lo;
        const float z = x * x;
        const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
        const float q = one + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
        const float r = p / q;
        return pio2_hi - (x - (pio2_lo - r * x));
    }
    if (hx < 0) {  // x < -0.5
        const float z = (one + x) * 0.5f;
        const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
        const float q = one + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
        const float s = sqrtf(z);
        const float r = p / q;
        const float w = r * s - pio2_lo;
        return pi - 2.0f * (s + w);
    }
    // x > 0.5
    const float z = (one - x) * 0.5f;
    const float s = sqrtf(z);
    const float df = word_float(float_word(s) & static_cast<int32_t>(0xfffff000u));
    const float c = (z - df * df) / (s + df);
    const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    const float q = one + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    const float r = p / q;
    const float w = r * s + c;
    return 2.0f * (df + w);
}

// s_sinf.c: |y| < pi/4 -> sinf_poly; |y| < 120 -> reduce_fast + sinf_poly; beyond: double sin, rounded once
LO_HD float sinf_glibc(float y) {
    const uint32_t top12 = (static_cast<uint32_t>(float_word(y)) >> 20) & 0x7ff;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    double x = static_cast<double>(y);
    int n = 0;
    double flip = 1.0;  // second table entry negates the cosine coefficients
    if (top12 < 0x3f4) {              // abstop12(pi/4)
        if (top12 < 0x398) return y;  // |y| < 2^-12
    } else if (top12 < 0x42f) {       // abstop12(120.0f)
        const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
        const double r = x * hpi_inv;
        n = (static_cast<int32_t>(r) + 0x800000) >> 24;
        x = fma(-static_cast<double>(n), hpi, x);
        const double sign = ((n & 3) == 0 || (n & 3) == 3) ? 1.0 : -1.0;
        if (n & 2) flip = -1.0;
        const double x2r = x * x;
        x = x * sign;
        if (n & 1) {
            const double x4 = x2r * x2r;
            const double c2 = flip * C3 + x2r * (flip * C4);
            const double c1 = flip * C0 + x2r * (flip * C1);
            const double x6 = x4 * x2r;
            const double c = c1 + x4 * (flip * C2);
            return static_cast<float>(c + x6 * c2);
        }
        const double x3 = x * x2r;
        const double s1 = S2 + x2r * S3;
        const double x7 = x3 * x2r;
        const double s = x + x3 * S1;
        return static_cast<float>(s + x7 * s1);
    } else {
        return static_cast<float>(sin(x));
    }
    const double x2 = x * x;
    const double x3 = x * x2;
    const double s1 = S2 + x2 * S3;
    const double x7 = x3 * x2;
    const double s = x + x3 * S1;
    return static_cast<float>(s + x7 * s1);
}

}  // namespace libm
}  // namespace lo
