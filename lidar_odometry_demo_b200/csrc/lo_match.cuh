// K4 + K5 device functions: 27-neighbour nearest-point search (VoxelGrid::getCorrespondence,
// src/voxel_grid.h:164-204), the point-to-plane residual with its analytic Jacobian
// (PointToPlaneErrorAnalytic::Evaluate, src/cloud_matcher.cpp:48-102), the Huber corrector Ceres applies
// (HuberLoss(0.15), cloud_matcher.cpp:134) and the reductions to the 6x6 normal equations.
#pragma once

#include "lo_device.cuh"

namespace lo {

constexpr int kAcc = 31;  // 21 (upper JtJ) + 6 (Jtr) + cost + valid pairs + candidates visited + occupied neighbours
constexpr unsigned kFullMask = 0xFFFFFFFFu;

struct Found {        // result of one query, identical in every lane of the warp
    bool valid;
    float o[3];       // plane_origin  (stored point)
    float n[3];       // plane_normal  (its normal)
};

// One warp, one query. Lane l < 27 probes neighbour voxel l, in the reference's scan order
// ix -> iy -> iz (l = 9*(dx+1) + 3*(dy+1) + (dz+1)), and scans that voxel's stored points in slot order.
// The winner is the first strict minimum in (probe, slot) scan order, which is the lexicographic minimum
// of (d2, probe, slot): d2 >= 0, so its float bit pattern orders like its value.
__device__ inline Found nn_search_warp(const GridView& g, const float wq[3], float max_d2, int lane, uint32_t* lane_count = nullptr) {
    const long long ox = axis_index(wq[0], g.voxel_size), oy = axis_index(wq[1], g.voxel_size), oz = axis_index(wq[2], g.voxel_size);
    unsigned long long best = ~0ull;
    uint32_t vid = 0;
    uint32_t count = 0;
    if (lane < 27) {
        const long long ix = ox + (lane / 9) - 1, iy = oy + ((lane / 3) % 3) - 1, iz = oz + (lane % 3) - 1;
        if (axis_in_range(ix) && axis_in_range(iy) && axis_in_range(iz)) {
            const unsigned long long key = pack_key(ix, iy, iz);
            uint32_t h = probe_hash(key) & g.mask;
            while (true) {
                const uint4 raw = __ldg(reinterpret_cast<const uint4*>(g.slots + h));
                const unsigned long long k = (static_cast<unsigned long long>(raw.y) << 32) | raw.x;
                if (k == key) {
                    vid = raw.z;
                    count = raw.w;
                    break;
                }
                if (k == kKeyEmpty) break;
                h = (h + 1) & g.mask;
            }
            const float4* slab = g.points + static_cast<size_t>(vid) * g.width;
            for (uint32_t e = 0; e < count; e += 4) {
                float4 p[4];
#pragma unroll
                for (int u = 0; u < 4; u++) p[u] = (e + u < count) ? __ldg(slab + e + u) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const float dx = wq[0] - p[u].x, dy = wq[1] - p[u].y, dz = wq[2] - p[u].z;
                    const float d2 = dx * dx + (dy * dy + dz * dz);  // Vector3f::squaredNorm order
                    if (e + u < count && d2 < max_d2) {
                        const unsigned long long cand = (static_cast<unsigned long long>(__float_as_uint(d2)) << 32) |
                                                        (static_cast<unsigned long long>(lane) << 24) | (e + u);
                        best = cand < best ? cand : best;
                    }
                }
            }
        }
    }
    unsigned long long m = best;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long other = __shfl_xor_sync(kFullMask, m, off);
        m = other < m ? other : m;
    }
    if (lane_count) *lane_count = count;
    Found f;
    f.valid = (m != ~0ull);
    f.o[0] = f.o[1] = f.o[2] = f.n[0] = f.n[1] = f.n[2] = 0.f;
    if (f.valid) {
        const int win_lane = static_cast<int>((m >> 24) & 0x1F);
        const uint32_t e = static_cast<uint32_t>(m & 0xFFFFFF);
        const uint32_t win_vid = __shfl_sync(kFullMask, vid, win_lane);
        const size_t at = static_cast<size_t>(win_vid) * g.width + e;
        const float4 p = __ldg(g.points + at);
        const float4 nv = __ldg(g.normals + at);
        f.o[0] = p.x; f.o[1] = p.y; f.o[2] = p.z;
        f.n[0] = nv.x; f.n[1] = nv.y; f.n[2] = nv.z;
    }
    return f;
}

// The evaluation point of the Ceres program, prepared once per evaluation.
struct EvalPoint {
    double q[4];        // w,x,y,z as Ceres holds them (not normalised)
    double t[3];
    double pj[4][3];    // QuaternionManifold::PlusJacobian at q
};

__host__ __device__ inline EvalPoint make_eval_point(const double x[7]) {
    EvalPoint e;
    for (int i = 0; i < 4; i++) e.q[i] = x[i];
    for (int i = 0; i < 3; i++) e.t[i] = x[4 + i];
    const double w = x[0], a = x[1], b = x[2], c = x[3];
    e.pj[0][0] = -a; e.pj[0][1] = -b; e.pj[0][2] = -c;
    e.pj[1][0] = w;  e.pj[1][1] = c;  e.pj[1][2] = -b;
    e.pj[2][0] = -c; e.pj[2][1] = w;  e.pj[2][2] = a;
    e.pj[3][0] = b;  e.pj[3][1] = -a; e.pj[3][2] = w;
    return e;
}

__device__ inline double dot3d(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

// One residual block: adds its contribution to acc[0..20] (upper triangle of JtJ, row-major),
// acc[21..26] (Jtr), acc[27] (cost), acc[28] (count). When `jacobian` is false only the cost is added.
__device__ inline void eval_pair(const EvalPoint& ep, const float pf[3], const float of[3], const float nf[3], bool jacobian,
                                 double acc[kAcc]) {
    const double p[3] = {pf[0], pf[1], pf[2]}, o[3] = {of[0], of[1], of[2]}, n[3] = {nf[0], nf[1], nf[2]};
    const double w = ep.q[0], x = ep.q[1], y = ep.q[2], z = ep.q[3];
    // Eigen Quaternion * Vector3 (cloud_matcher.cpp:55)
    double uv[3] = {y * p[2] - z * p[1], z * p[0] - x * p[2], x * p[1] - y * p[0]};
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    const double c[3] = {y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]};
    double d[3];
    for (int i = 0; i < 3; i++) d[i] = (((p[i] + w * uv[i]) + c[i]) + ep.t[i]) - o[i];
    const double r = dot3d(d, n);
    const double s = r * r;
    const double a = 0.15, b = a * a;  // HuberLoss(0.15)
    double rho0, rho1;
    if (s > b) {
        const double rr = sqrt(s);
        rho0 = 2.0 * a * rr - b;
        rho1 = fmax(2.2250738585072014e-308, a / rr);
    } else {
        rho0 = s;
        rho1 = 1.0;
    }
    acc[27] += 0.5 * rho0;
    acc[28] += 1.0;
    if (!jacobian) return;
    const double sc = sqrt(rho1);
    // dR/dq_k * p . n for the four matrices of cloud_matcher.cpp:64-91
    const double q0 = w, q1 = x, q2 = y, q3 = z;
    const double mk[4][9] = {{q0, -q3, q2, q3, q0, -q1, -q2, q1, q0},
                             {q1, q2, q3, q2, -q1, -q0, q3, q0, -q1},
                             {-q2, q1, q0, q1, q2, q3, -q0, q3, -q2},
                             {-q3, -q0, q1, q0, -q3, q2, q1, q2, q3}};
    double jq[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        double v[3];
#pragma unroll
        for (int i = 0; i < 3; i++)
            v[i] = ((2.0 * mk[k][3 * i]) * p[0] + (2.0 * mk[k][3 * i + 1]) * p[1]) + (2.0 * mk[k][3 * i + 2]) * p[2];
        jq[k] = dot3d(v, n);
    }
    double j[6];
#pragma unroll
    for (int col = 0; col < 3; col++) {
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 4; k++) v += jq[k] * ep.pj[k][col];
        j[col] = sc * v;
    }
#pragma unroll
    for (int col = 0; col < 3; col++) j[3 + col] = sc * n[col];
    const double rc = sc * r;
    int at = 0;
#pragma unroll
    for (int a0 = 0; a0 < 6; a0++)
#pragma unroll
        for (int b0 = a0; b0 < 6; b0++) acc[at++] += j[a0] * j[b0];
#pragma unroll
    for (int a0 = 0; a0 < 6; a0++) acc[21 + a0] += j[a0] * rc;
}

// Block-wide sum of the per-thread accumulators; result valid in thread 0. `smem` holds
// (blockDim.x / 32) * kAcc doubles.
__device__ inline void block_reduce_acc(double acc[kAcc], double* smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
    for (int k = 0; k < kAcc; k++) {
        double v = acc[k];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(kFullMask, v, off);
        if (lane == 0) smem[warp * kAcc + k] = v;
    }
    __syncthreads();
    if (threadIdx.x < kAcc) {
        double v = 0.0;
        for (int w = 0; w < nwarps; w++) v += smem[w * kAcc + threadIdx.x];
        smem[threadIdx.x] = v;  // column sums live in smem[0..kAcc)
    }
    __syncthreads();
}

}  // namespace lo
