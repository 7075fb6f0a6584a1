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
// ix -> iy -> iz (l = 9*(dx+1) + 3*(dy+1) + (dz+1)). The stored points of the occupied neighbours are then cut
// into work items of 4 consecutive points (64 B) that are dealt out to all 32 lanes through a small per-warp
// list in shared memory, so the distance scan takes ceil(items/32) rounds of independent 16-byte loads whatever
// the distribution of points over the 27 voxels. The winner is the first strict minimum in (probe, slot) scan
// order, i.e. the lexicographic minimum of (d2, probe, slot): d2 >= 0, so its float bit pattern orders like its
// value and one 64-bit min-reduction decides.
constexpr int kWorkItems = 27 * 5;           // per-warp list capacity: grids with max_points <= 20
constexpr int kWorkListStride = kWorkItems + 1;

struct Probe {             // lane l's look-up of neighbour voxel l, split so that the slot load can be issued early
    unsigned long long key;
    uint32_t h;
    uint4 raw;
    bool active;
};

// Lane constants of the 27-neighbour walk: lane l < 27 looks at offset (l/9-1, (l/3)%3-1, l%3-1)
struct LaneOffset {
    int dx, dy, dz;
    bool probing;
};
__device__ inline LaneOffset lane_offset(int lane) {
    LaneOffset o;
    o.probing = lane < 27;
    o.dx = lane / 9 - 1;
    o.dy = (lane / 3) % 3 - 1;
    o.dz = lane % 3 - 1;
    return o;
}

// (ox,oy,oz) = voxel of the query (VoxelGrid::getIndices), already limited to |o| <= 2^21 so that 32-bit
// arithmetic is exact; every lane of the warp passes the same values. The packed 64-bit key
// (X << 42 | Y << 21 | Z with X = ix + 2^20 ...) is assembled from two 32-bit words.
__device__ inline Probe probe_issue_at(const GridView& g, int ox, int oy, int oz, const LaneOffset& lo) {
    Probe p;
    p.active = false;
    p.key = 0;
    p.h = 0;
    p.raw = make_uint4(0u, 0u, 0u, 0u);
    const uint32_t X = static_cast<uint32_t>(ox + lo.dx + static_cast<int>(kAxisBias)), Y = static_cast<uint32_t>(oy + lo.dy + static_cast<int>(kAxisBias)),
                   Z = static_cast<uint32_t>(oz + lo.dz + static_cast<int>(kAxisBias));
    const uint32_t top = static_cast<uint32_t>(2 * kAxisBias);
    if (lo.probing && X < top && Y < top && Z < top) {   // axis_in_range on all three axes
        p.active = true;
        const uint32_t key_lo = Z | (Y << 21), key_hi = (Y >> 11) | (X << 10);
        p.key = (static_cast<unsigned long long>(key_hi) << 32) | key_lo;
        uint32_t h = ((X >> 1) * 0x9E3779B1u) ^ ((Y >> 1) * 0x85EBCA77u) ^ ((Z >> 1) * 0xC2B2AE3Du);   // probe_hash, 32-bit form
        h ^= h >> 16; h *= 0x85EBCA6Bu;
        h ^= h >> 13; h *= 0xC2B2AE35u;
        h ^= h >> 16;
        p.h = ((h << 3) | ((X & 1u) << 2) | ((Y & 1u) << 1) | (Z & 1u)) & g.mask;
        p.raw = __ldg(reinterpret_cast<const uint4*>(g.slots + p.h));
    }
    return p;
}
__device__ inline int clamp_axis(long long v) {  // keeps out-of-range / non-finite queries outside the table's index range
    const long long lim = 2 * kAxisBias;
    return static_cast<int>(v < -lim ? -lim : (v > lim ? lim : v));
}
__device__ inline Probe probe_issue(const GridView& g, const float wq[3], int lane) {
    return probe_issue_at(g, clamp_axis(axis_index(wq[0], g.voxel_size)), clamp_axis(axis_index(wq[1], g.voxel_size)),
                          clamp_axis(axis_index(wq[2], g.voxel_size)), lane_offset(lane));
}

__device__ inline void probe_resolve(const GridView& g, Probe p, uint32_t& vid, uint32_t& count) {
    vid = 0;
    count = 0;
    if (!p.active) return;
    while (true) {
        const unsigned long long k = (static_cast<unsigned long long>(p.raw.y) << 32) | p.raw.x;
        if (k == p.key) {
            vid = p.raw.z;
            count = p.raw.w;
            return;
        }
        if (k == kKeyEmpty) return;
        p.h = (p.h + 1) & g.mask;
        p.raw = __ldg(reinterpret_cast<const uint4*>(g.slots + p.h));
    }
}

struct Hit {          // nearest stored point of one query, identical in every lane
    bool valid;
    float o[3];       // the stored point (plane_origin)
    uint32_t at;      // its index in the point / normal slabs
    unsigned long long key;  // (d2 bits << 32 | probe << 24 | slot), ~0 when there is no correspondence
};

__device__ inline Hit scan_candidates(const GridView& g, const float wq[3], float max_d2, int lane, uint2* wl, uint32_t vid, uint32_t count) {
    unsigned long long best = ~0ull;
    float bx = 0.f, by = 0.f, bz = 0.f;
    uint32_t bat = 0;
    if (g.width <= 20) {
        const uint32_t items = (count + 3) >> 2;
        uint32_t incl = items;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t v = __shfl_up_sync(kFullMask, incl, off);
            if (lane >= off) incl += v;
        }
        const uint32_t total = __shfl_sync(kFullMask, incl, 31);
        uint32_t at = incl - items;
        for (uint32_t e0 = 0; e0 < count; e0 += 4, at++)
            wl[at] = make_uint2(vid * g.width + e0, (static_cast<uint32_t>(lane) << 24) | (min(4u, count - e0) << 16) | e0);
        __syncwarp();
        // A lane sees its work items in increasing (probe, slot) order, so a strict float `<` keeps the first minimum
        // of the lane; the 64-bit (d2, probe, slot) key is only built once per lane for the warp reduction.
        float best_d2 = max_d2;
        uint32_t best_id = 0;
        for (uint32_t base = 0; base < total; base += 32) {
            const uint32_t t = base + lane;
            if (t < total) {
                const uint2 it = wl[t];
                const uint32_t n = (it.y >> 16) & 7u;
                const float4* slab = g.points + it.x;
                float4 p[4];
#pragma unroll
                for (int u = 0; u < 4; u++) p[u] = (static_cast<uint32_t>(u) < n) ? __ldg(slab + u) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const float dx = wq[0] - p[u].x, dy = wq[1] - p[u].y, dz = wq[2] - p[u].z;
                    const float d2 = dx * dx + (dy * dy + dz * dz);  // Vector3f::squaredNorm order
                    if (static_cast<uint32_t>(u) < n && d2 < best_d2) {
                        best_d2 = d2;
                        bx = p[u].x; by = p[u].y; bz = p[u].z;
                        bat = it.x + u;
                        best_id = (it.y & 0xFF000000u) | ((it.y & 0xFFFFu) + u);
                    }
                }
            }
        }
        if (best_d2 < max_d2) best = (static_cast<unsigned long long>(__float_as_uint(best_d2)) << 32) | best_id;
        __syncwarp();  // the list is reused by the warp's next query
    } else {
        // wide voxels: every probing lane scans its own voxel
        const float4* slab = g.points + static_cast<size_t>(vid) * g.width;
        for (uint32_t e = 0; e < count; e++) {
            const float4 p = __ldg(slab + e);
            const float dx = wq[0] - p.x, dy = wq[1] - p.y, dz = wq[2] - p.z;
            const float d2 = dx * dx + (dy * dy + dz * dz);
            if (d2 < max_d2) {
                const unsigned long long cand = (static_cast<unsigned long long>(__float_as_uint(d2)) << 32) |
                                                (static_cast<unsigned long long>(lane) << 24) | e;
                if (cand < best) {
                    best = cand;
                    bx = p.x; by = p.y; bz = p.z;
                    bat = vid * g.width + e;
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
    Hit hit;
    hit.valid = (m != ~0ull);
    hit.o[0] = hit.o[1] = hit.o[2] = 0.f;
    hit.at = 0;
    hit.key = m;
    if (hit.valid) {
        const int src = __ffs(__ballot_sync(kFullMask, best == m)) - 1;   // (probe, slot) is unique: exactly one lane
        hit.o[0] = __shfl_sync(kFullMask, bx, src);
        hit.o[1] = __shfl_sync(kFullMask, by, src);
        hit.o[2] = __shfl_sync(kFullMask, bz, src);
        hit.at = __shfl_sync(kFullMask, bat, src);
    }
    return hit;
}

__device__ inline Found nn_search_warp(const GridView& g, const float wq[3], float max_d2, int lane, uint2* wl, uint32_t* lane_count = nullptr) {
    uint32_t vid, count;
    probe_resolve(g, probe_issue(g, wq, lane), vid, count);
    if (lane_count) *lane_count = count;
    const Hit hit = scan_candidates(g, wq, max_d2, lane, wl, vid, count);
    Found f;
    f.valid = hit.valid;
    f.o[0] = hit.o[0]; f.o[1] = hit.o[1]; f.o[2] = hit.o[2];
    f.n[0] = f.n[1] = f.n[2] = 0.f;
    if (f.valid) {
        const float4 nv = __ldg(g.normals + hit.at);
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
