// Per-point / per-cell device functions of the pre-registration stages, shared by the stage-by-stage kernels
// (lo_points.cu, behind lo_deskew / lo_classify / lo_range_filter) and the fused per-scan pipeline (lo_pipeline.cu):
//   K1  utils::pointTimeNormalize + CloudTransformer::transformNonRigid   (src/utils/point_time_normalize.h:14-39,
//                                                                          src/utils/cloud_transform.h:15-40)
//   K8  CloudClassifier::classify                                         (src/utils/cloud_classifier.h:19-168)
//       utils::rangeFilter                                                (src/utils/range_filter.h:12-28)
#pragma once

#include "lo_device.cuh"
#include "lo_libm.cuh"

namespace lo {

// order-preserving float <-> uint mapping for atomicMin / atomicMax on `time`
__device__ inline uint32_t float_to_ordered(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ inline float ordered_to_float(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}
// numeric_limits<float>::max() / lowest() when every time was NaN (the reference's loop never updates them)
__device__ inline float time_min_of(const uint32_t* minmax) { return minmax[0] == 0xFFFFFFFFu ? 3.402823466e+38f : ordered_to_float(minmax[0]); }
__device__ inline float time_max_of(const uint32_t* minmax) { return minmax[1] == 0u ? -3.402823466e+38f : ordered_to_float(minmax[1]); }

// One point of transformNonRigid(start = s, end = e) after time normalisation. a = {x,y,z,pad}, b = {intensity,
// ring|pad, time, pad}; returns the deskewed xyz record and rewrites b.z with the normalised time.
__device__ inline float4 deskew_point(const Pose& s, const Pose& e, float t_min, float range, const float4 a, float4& b) {
    const float t = (b.z - t_min) / range;
    // Quaternionf::slerp(t, end) (Eigen 3.4); glibc's acosf / sinf restated bit for bit in lo_libm.cuh
    const float one = 1.f - 1.1920928955078125e-07f;
    const float d = quat_dot_f(s, e);
    const float ad = fabsf(d);
    float s0, s1;
    if (ad >= one) {
        s0 = 1.f - t;
        s1 = t;
    } else {
        const float theta = libm::acosf_fdlibm(ad);
        const float sin_theta = libm::sinf_glibc(theta);
        s0 = libm::sinf_glibc((1.f - t) * theta) / sin_theta;
        s1 = libm::sinf_glibc(t * theta) / sin_theta;
    }
    if (d < 0.f) s1 = -s1;
    const float qw = s0 * s.w + s1 * e.w, qx = s0 * s.x + s1 * e.x, qy = s0 * s.y + s1 * e.y, qz = s0 * s.z + s1 * e.z;
    const float v[3] = {a.x, a.y, a.z};
    float r[3];
    quat_rotate_f(qw, qx, qy, qz, v, r);
    // start.t * time + end.t * (1.0 - time): the reference's weights; (1.0 - time) is a double
    const float w_end = static_cast<float>(1.0 - static_cast<double>(t));
    float4 o = a;
    o.x = (r[0] + s.t[0] * t) + e.t[0] * w_end;
    o.y = (r[1] + s.t[1] * t) + e.t[1] * w_end;
    o.z = (r[2] + s.t[2] * t) + e.t[2] * w_end;
    b.z = t;
    return o;
}

// rangeFilter: plain C++ float expression, evaluated left to right
__device__ inline bool in_range(float x, float y, float z, float lo2, float hi2) {
    const float r2 = x * x + y * y + z * z;
    return r2 >= lo2 && r2 <= hi2;
}

// cloud_classifier.h:49-51  float azimuth = atan2(-y, x) + pi;  index = abs(azimuth * W / (2.0 * pi)); kept iff < W
__device__ inline bool azimuth_column(float x, float y, uint32_t W, uint32_t& col) {
    const float at = libm::atan2f_fdlibm(-y, x);
    const float azimuth = static_cast<float>(static_cast<double>(at) + 3.141592653589793);
    const double scaled = static_cast<double>(azimuth * static_cast<float>(W)) / (2.0 * 3.141592653589793);
    if (!(scaled == scaled)) return false;
    const double mag = fabs(scaled);
    if (!(mag < static_cast<double>(W))) return false;
    col = static_cast<uint32_t>(mag);
    return true;
}

// cloud_classifier.h:81-105: curvature of cell c of the flattened grid. `cell(k)` returns {x,y,z,intensity} of cell k.
template <typename CellFn>
__device__ inline float cell_curvature(uint32_t c, uint32_t total, CellFn cell) {
    const float4 p = cell(c);
    if (c < 4 || c + 4 >= total) return p.w;   // first / last 4 cells keep their raw intensity
    const float range = p.x * p.x + p.y * p.y + p.z * p.z;  // powf(v, 2) == v * v
    if (static_cast<double>(range) < 0.1) return 1000.0f;
    float dx = static_cast<float>(static_cast<double>(-p.x) * 9.0);
    float dy = static_cast<float>(static_cast<double>(-p.y) * 9.0);
    float dz = static_cast<float>(static_cast<double>(-p.z) * 9.0);
#pragma unroll
    for (int w = -4; w <= 4; w++) {
        const float4 q = cell(c + w);
        dx += q.x;
        dy += q.y;
        dz += q.z;
    }
    return sqrtf(dx * dx + dy * dy + dz * dz) / range;
}

// cloud_classifier.h:107-165: is cell c planar, and its normal from two neighbours on the previous ring
__device__ inline bool cell_planar(uint32_t c, uint32_t W, const float4* __restrict__ cellp, const float* __restrict__ curv, float nv[3]) {
    const uint32_t ray = c / W, col = c % W;
    if (!(ray >= 1 && col >= 4 && col + 4 < W && curv[c] < 0.05f)) return false;
    const double gate = static_cast<double>(0.05f) * 10.0;
    const uint32_t prev = (ray - 1) * W;
    int left = -1, right = -1;
    for (uint32_t k = col - 4; k < col; k++)
        if (static_cast<double>(curv[prev + k]) < gate) { left = static_cast<int>(k); break; }
    for (uint32_t k = col + 4; k > col; k--)
        if (static_cast<double>(curv[prev + k]) < gate) { right = static_cast<int>(k); break; }
    if (left < 0 || right < 0) return false;
    const float4 p = cellp[c], l = cellp[prev + left], r = cellp[prev + right];
    const float a[3] = {l.x - p.x, l.y - p.y, l.z - p.z}, b[3] = {r.x - p.x, r.y - p.y, r.z - p.z};
    cross3f(a, b, nv);
    const float n2 = nv[0] * nv[0] + (nv[1] * nv[1] + nv[2] * nv[2]);
    if (n2 > 0.f) {
        const float len = sqrtf(n2);
        nv[0] /= len; nv[1] /= len; nv[2] /= len;
    }
    return true;
}

}  // namespace lo
