// CloudMatcher::align on the device: residual / Jacobian terms, the warp / block / grid reductions of the 6x6 normal
// equations and the persistent cooperative loop. Shared by lo_match.cu (align_kernel, nn_residual_kernel, sharded
// phases) and lo_pipeline.cu (the per-scan kernel).
#pragma once

#include <cooperative_groups.h>

#include "lo_host.h"
#include "lo_lm.cuh"
#include "lo_tile_search.cuh"

namespace lo {

namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------------------------------
// shared accumulation helpers. `row` points at this warp's kRow (= 32 >= kAcc) doubles in shared memory.
constexpr int kRow = 32;
struct Terms {
    double j[6];
    double rc;     // corrected residual
    double cost;   // 0.5 * rho(s)
};

__device__ inline Terms pair_terms(const EvalPoint& ep, const float pf[3], const float of[3], const float nf[3]) {
    Terms t;
    const double p[3] = {pf[0], pf[1], pf[2]}, o[3] = {of[0], of[1], of[2]}, n[3] = {nf[0], nf[1], nf[2]};
    const double w = ep.q[0], x = ep.q[1], y = ep.q[2], z = ep.q[3];
    double uv[3] = {y * p[2] - z * p[1], z * p[0] - x * p[2], x * p[1] - y * p[0]};
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    const double c[3] = {y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]};
    double d[3];
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = (((p[i] + w * uv[i]) + c[i]) + ep.t[i]) - o[i];
    const double r = dot3d(d, n);
    const double s = r * r;
    const double a = 0.15, b = a * a;
    double rho0, rho1;
    if (s > b) {
        const double rr = sqrt(s);
        rho0 = 2.0 * a * rr - b;
        rho1 = fmax(2.2250738585072014e-308, a / rr);
    } else {
        rho0 = s;
        rho1 = 1.0;
    }
    t.cost = 0.5 * rho0;
    const double sc = sqrt(rho1);
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
#pragma unroll
    for (int col = 0; col < 3; col++) {
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 4; k++) v += jq[k] * ep.pj[k][col];
        t.j[col] = sc * v;
    }
#pragma unroll
    for (int col = 0; col < 3; col++) t.j[3 + col] = sc * n[col];
    t.rc = sc * r;
    return t;
}

// one lane adds one pair (no shuffles)
__device__ inline void accumulate_lane(double* row, const Terms& t) {
    int at = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int b = a; b < 6; b++) row[at++] += t.j[a] * t.j[b];
#pragma unroll
    for (int a = 0; a < 6; a++) row[21 + a] += t.j[a] * t.rc;
    row[27] += t.cost;
    row[28] += 1.0;
}

// sum over lanes 0 .. 2^kSteps - 1 (the other lanes are ignored), result in lane 0
template <int kSteps = 5>
__device__ inline double warp_sum(double v) {
#pragma unroll
    for (int off = (1 << kSteps) >> 1; off > 0; off >>= 1) v += __shfl_down_sync(kFullMask, v, off);
    return v;
}

// all 32 lanes hold (at most) one pair each; products are reduced by shuffles one value at a time so that
// only j[6], rc and one temporary are live (keeps the kernel under 64 registers). kSteps < 5: only lanes below
// 2^kSteps hold pairs.
template <int kSteps = 5>
__device__ inline void accumulate_warp(double* row, const Terms& t, bool valid, int lane) {
    const double m = valid ? 1.0 : 0.0;
    int at = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int b = a; b < 6; b++) {
            const double v = warp_sum<kSteps>(m * (t.j[a] * t.j[b]));
            if (lane == 0) row[at] += v;
            at++;
        }
#pragma unroll
    for (int a = 0; a < 6; a++) {
        const double v = warp_sum<kSteps>(m * (t.j[a] * t.rc));
        if (lane == 0) row[21 + a] += v;
    }
    const double c = warp_sum<kSteps>(m * t.cost);
    const double k = warp_sum<kSteps>(m);
    if (lane == 0) {
        row[27] += c;
        row[28] += k;
    }
}

// sum the per-warp rows of a block into out[kAcc] (global); rows are zeroed for the next pass
__device__ inline void block_rows_to_global(double* rows, double* out) {
    __syncthreads();
    const int nwarps = blockDim.x >> 5;
    if (threadIdx.x < kAcc) {
        double v = 0.0;
        for (int w = 0; w < nwarps; w++) {
            v += rows[w * kRow + threadIdx.x];
            rows[w * kRow + threadIdx.x] = 0.0;
        }
        out[threadIdx.x] = v;
    }
    __syncthreads();
}

// every block sums the per-block partials [nblocks][32] into total[kAcc] (shared), in a fixed order.
// Thread t reads elements t, t + blockDim, ... of the flat array (coalesced); because blockDim is a multiple
// of 32 a thread always sees column (t & 31), so it keeps one running sum; the per-warp sums are then added
// in warp order. `scratch` holds (blockDim.x / 32) * 32 doubles (the caller's per-warp rows are reused).
__device__ inline void reduce_partials(const double* partials, int nblocks, double* scratch, double* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    double v = 0.0;
    const int n = nblocks * 32;
    // 10 independent loads in flight (148 rows x 32 over 512 threads = 9.25 per thread: one round trip to L2); a thread
    // adds its elements in ascending order whatever the batch size, so the sum does not depend on it
    constexpr int kInFlight = 10;
    for (int e0 = threadIdx.x; e0 < n; e0 += kInFlight * blockDim.x) {
        double t[kInFlight];
#pragma unroll
        for (int u = 0; u < kInFlight; u++) {
            const int e = e0 + u * blockDim.x;
            t[u] = e < n ? __ldcg(partials + e) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < kInFlight; u++) v += t[u];
    }
    scratch[warp * 32 + lane] = v;
    __syncthreads();
    if (threadIdx.x < kAcc) {
        double t = 0.0;
        for (int w = 0; w < nwarps; w++) t += scratch[w * 32 + threadIdx.x];
        total[threadIdx.x] = t;
    }
    __syncthreads();
    scratch[warp * 32 + lane] = 0.0;  // the rows double as accumulators of the next pass
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------
// CloudMatcher::align (src/cloud_matcher.cpp:105-178) as one persistent cooperative kernel.
constexpr int kAlignThreads = 512;

struct AlignArgs {
    GridView g;
    const PointXYZ* queries;
    uint32_t m;
    const uint32_t* m_dev; // nullable: actual query count on the device (m is then an upper bound)
    float4* pair_o;        // [m] plane_origin, w = 1 if the query has a correspondence
    float4* pair_n;        // [m] plane_normal
    double* partials;      // [2][gridDim.x][32]
    const float* guess;    // device pose7
    float* out_pose;       // device pose7 (written by block 0)
    float* out_pose_block = nullptr;  // nullable: 7 floats in the caller's shared memory, written in every block
    lo_align_stats* stats; // device, nullable
    float max_dist;
    long long* trace;      // nullable: (tag, clock64) pairs written by block 0 thread 0 (LO_ALIGN_TRACE=1)
    int trace_start = 0;   // first free trace entry (the per-scan kernel has written some already)
};

#define LO_TRACE(tag)                                                          \
    do {                                                                       \
        if (a.trace && blockIdx.x == 0 && threadIdx.x == 0 && trace_n < 500) { \
            a.trace[2 * trace_n] = (tag);                                      \
            a.trace[2 * trace_n + 1] = clock64();                              \
            trace_n++;                                                         \
        }                                                                      \
    } while (0)

// The body of align_kernel; also the middle part of the per-scan kernel of lo_pipeline.cu. Every thread of a
// cooperative grid of kAlignThreads-wide blocks must call it.
// `tile_scratch`: nullable, (kAlignThreads / 32) TileScratch in shared memory; with it, large query sets (several queries
// per warp) are searched 32 per warp with exact neighbour pruning (lo_tile_search.cuh) instead of one per warp.
constexpr size_t kAlignTileScratchBytes = (kAlignThreads / 32) * sizeof(TileScratch);
__device__ inline void align_device(AlignArgs a, cg::grid_group& grid, TileScratch* tile_scratch = nullptr) {
    if (a.m_dev) a.m = min(a.m, *a.m_dev);
    __shared__ double rows[(kAlignThreads / 32) * kRow];
    __shared__ double total[32];
    __shared__ LmState lm;
    __shared__ EvalPoint ep_s;
    __shared__ float pose_s[7];
    __shared__ int flag_s;
    __shared__ uint2 work[(kAlignThreads / 32) * kWorkListStride];
    uint2* wl = work + (threadIdx.x >> 5) * kWorkListStride;

    const int lane = threadIdx.x & 31;
    // consecutive queries / tiles go to different SMs: warp w of block b is global warp w * gridDim.x + b
    const uint32_t gwarp = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    const int nblocks = gridDim.x;
    const LaneOffset lofs = lane_offset(lane);
    const uint32_t my_queries = gwarp < a.m ? (a.m - gwarp + total_warps - 1) / total_warps : 0u;
    const bool tiled = tile_scratch != nullptr && a.g.width <= kTileMaxWidth && a.m >= 8 * total_warps;
    TileScratch* ts = tile_scratch ? tile_scratch + (threadIdx.x >> 5) : nullptr;
    double* row = rows + (threadIdx.x >> 5) * kRow;
    for (int k = threadIdx.x; k < (kAlignThreads / 32) * kRow; k += blockDim.x) rows[k] = 0.0;
    if (threadIdx.x < 7) pose_s[threadIdx.x] = a.guess[threadIdx.x];
    __syncthreads();
    const double prior_b[3] = {static_cast<double>(a.guess[0]), static_cast<double>(a.guess[1]), static_cast<double>(a.guess[2])};
    const float max_d2 = a.max_dist * a.max_dist;  // voxel_grid.h:215
    int buf = 0;
    int lm_total = 0;
    bool cont = false;  // meaningful in thread 0 only
    int trace_n = a.trace_start;
    LO_TRACE(0);

    for (int outer = 0; outer < 35; outer++) {  // cloud_matcher.cpp:117
        // ---- correspondences at the current (float) pose + evaluation at x0 = double(pose) ----
        Pose pf;
        pf.t[0] = pose_s[0]; pf.t[1] = pose_s[1]; pf.t[2] = pose_s[2];
        pf.w = pose_s[3]; pf.x = pose_s[4]; pf.y = pose_s[5]; pf.z = pose_s[6];
        const QueryTransform tf = make_query_transform(pf);
        double x0[7] = {pf.w, pf.x, pf.y, pf.z, pf.t[0], pf.t[1], pf.t[2]};  // cloud_matcher.cpp:122-131
        const EvalPoint ep0 = make_eval_point(x0);
        // every warp walks its queries (gwarp, gwarp + total_warps, ...: neighbouring queries cost about the same, so they are
        // dealt out to different warps) in tiles of 32: lane l transforms query l (findMatchingPairs, voxel_grid.h:217-223),
        // the 27-lane search runs once per query with the slot loads of query j+1 in flight while query j is scanned, lane j
        // keeps the correspondence of query j; then the tile's normals are fetched in one round and all lanes evaluate
        // their pair at x0 in parallel.
        // large query sets: tile search, query i in lane (i & 31) of the warp that owns tile i / 32
        for (uint32_t base = gwarp * 32; tiled && base < a.m; base += total_warps * 32) {
            const uint32_t i = base + lane;
            const bool have = i < a.m;
            PointXYZ q{0.f, 0.f, 0.f, 0.f};
            if (have) q = a.queries[i];
            float wq[3];
            transform_query(tf, q.x, q.y, q.z, wq);
            const TileHit h = nn_search_tile(a.g, wq, have, max_d2, lane, ts);
            if (have) {
                a.pair_o[i] = make_float4(h.o[0], h.o[1], h.o[2], h.valid ? 1.f : 0.f);
                a.pair_n[i] = make_float4(h.n[0], h.n[1], h.n[2], 0.f);
            }
            const float pq[3] = {q.x, q.y, q.z};
            const Terms t = pair_terms(ep0, pq, h.o, h.n);
            accumulate_warp<5>(row, t, h.valid, lane);
        }
        for (uint32_t base = 0; !tiled && base < my_queries; base += 32) {
            const uint32_t mine = gwarp + (base + lane) * total_warps;
            const bool have = base + lane < my_queries;
            PointXYZ q{0.f, 0.f, 0.f, 0.f};
            if (have) q = a.queries[mine];
            const uint32_t in_tile = min(32u, my_queries - base);
            float own[3];
            transform_query(tf, q.x, q.y, q.z, own);
            const int oi[3] = {clamp_axis(axis_index(own[0], a.g.voxel_size)), clamp_axis(axis_index(own[1], a.g.voxel_size)),
                               clamp_axis(axis_index(own[2], a.g.voxel_size))};
            bool my_valid = false;
            float my_o[3] = {0.f, 0.f, 0.f}, my_n[3] = {0.f, 0.f, 0.f};
            uint32_t my_at = 0;
            Probe next = probe_issue_at(a.g, __shfl_sync(kFullMask, oi[0], 0), __shfl_sync(kFullMask, oi[1], 0), __shfl_sync(kFullMask, oi[2], 0), lofs);
            for (uint32_t j = 0; j < in_tile; j++) {
                uint32_t vid, cnt;
                probe_resolve(a.g, next, vid, cnt);
                const int jj = static_cast<int>(j);
                const float cur[3] = {__shfl_sync(kFullMask, own[0], jj), __shfl_sync(kFullMask, own[1], jj), __shfl_sync(kFullMask, own[2], jj)};
                if (j + 1 < in_tile)
                    next = probe_issue_at(a.g, __shfl_sync(kFullMask, oi[0], jj + 1), __shfl_sync(kFullMask, oi[1], jj + 1),
                                          __shfl_sync(kFullMask, oi[2], jj + 1), lofs);
                const Hit hit = scan_candidates(a.g, cur, max_d2, lane, wl, vid, cnt);
                if (lane == jj) {
                    my_valid = hit.valid;
                    my_at = hit.at;
                    for (int k = 0; k < 3; k++) my_o[k] = hit.o[k];
                }
            }
            if (my_valid) {
                const float4 nv = __ldg(a.g.normals + my_at);
                my_n[0] = nv.x; my_n[1] = nv.y; my_n[2] = nv.z;
            }
            if (have) {
                a.pair_o[mine] = make_float4(my_o[0], my_o[1], my_o[2], my_valid ? 1.f : 0.f);
                a.pair_n[mine] = make_float4(my_n[0], my_n[1], my_n[2], 0.f);
            }
            const float pq[3] = {q.x, q.y, q.z};
            const Terms t = pair_terms(ep0, pq, my_o, my_n);
            if (in_tile <= 4) accumulate_warp<2>(row, t, my_valid, lane);
            else if (in_tile <= 8) accumulate_warp<3>(row, t, my_valid, lane);
            else accumulate_warp<5>(row, t, my_valid, lane);
        }
        LO_TRACE(1);
        block_rows_to_global(rows, a.partials + (static_cast<size_t>(buf) * nblocks + blockIdx.x) * 32);
        grid.sync();
        LO_TRACE(2);
        // this warp's first tile of pairs stays in registers for every LM evaluation of this outer iteration
        const uint32_t i0 = gwarp * 32 + lane;
        bool valid0 = false;
        float pq0[3] = {0.f, 0.f, 0.f}, of0[3] = {0.f, 0.f, 0.f}, nf0[3] = {0.f, 0.f, 0.f};
        if (i0 < a.m) {
            const float4 o = __ldcg(a.pair_o + i0);
            const float4 nv = __ldcg(a.pair_n + i0);
            const PointXYZ q = a.queries[i0];
            valid0 = o.w != 0.f;
            pq0[0] = q.x; pq0[1] = q.y; pq0[2] = q.z;
            of0[0] = o.x; of0[1] = o.y; of0[2] = o.z;
            nf0[0] = nv.x; nf0[1] = nv.y; nf0[2] = nv.z;
        }
        reduce_partials(a.partials + static_cast<size_t>(buf) * nblocks * 32, nblocks, rows, total);
        LO_TRACE(3);
        buf ^= 1;
        if (threadIdx.x == 0) {
            add_prior(total, ep0.t, prior_b);
            cont = lm_begin(lm, x0, total);
            if (a.stats && blockIdx.x == 0) {
                a.stats->pairs_log[outer] = static_cast<int32_t>(total[28] + 0.5);
                a.stats->valid_pairs = static_cast<int32_t>(total[28] + 0.5);
            }
        }
        // ---- ceres::Solve: at most 4 LM iterations on the fixed correspondences ----
        while (true) {
            if (threadIdx.x == 0) {
                const bool need_eval = cont && lm_propose(lm);
                flag_s = need_eval ? 1 : 0;
                if (need_eval) ep_s = make_eval_point(lm.cand);
            }
            LO_TRACE(4);
            __syncthreads();
            if (!flag_s) break;
            if (gwarp * 32 < a.m) {
                Terms t;
                if (valid0) {
                    t = pair_terms(ep_s, pq0, of0, nf0);
                } else {
                    for (int k = 0; k < 6; k++) t.j[k] = 0.0;
                    t.rc = 0.0;
                    t.cost = 0.0;
                }
                accumulate_warp(row, t, valid0, lane);
            }
            for (uint32_t base = (gwarp + total_warps) * 32; base < a.m; base += total_warps * 32) {
                const uint32_t i = base + lane;
                bool valid = false;
                Terms t;
                if (i < a.m) {
                    const float4 o = __ldcg(a.pair_o + i);
                    valid = o.w != 0.f;
                    if (valid) {
                        const float4 nv = __ldcg(a.pair_n + i);
                        const PointXYZ q = a.queries[i];
                        const float pq[3] = {q.x, q.y, q.z}, of[3] = {o.x, o.y, o.z}, nf[3] = {nv.x, nv.y, nv.z};
                        t = pair_terms(ep_s, pq, of, nf);
                    }
                }
                if (!valid) {
                    for (int k = 0; k < 6; k++) t.j[k] = 0.0;
                    t.rc = 0.0;
                    t.cost = 0.0;
                }
                accumulate_warp(row, t, valid, lane);
            }
            LO_TRACE(5);
            block_rows_to_global(rows, a.partials + (static_cast<size_t>(buf) * nblocks + blockIdx.x) * 32);
            grid.sync();
            LO_TRACE(6);
            reduce_partials(a.partials + static_cast<size_t>(buf) * nblocks * 32, nblocks, rows, total);
            LO_TRACE(7);
            buf ^= 1;
            if (threadIdx.x == 0) {
                add_prior(total, ep_s.t, prior_b);
                cont = lm_update(lm, total);
            }
            LO_TRACE(8);
        }
        // ---- float round trip of the pose (cloud_matcher.cpp:161-167) and the outer stop rule (:169) ----
        if (threadIdx.x == 0) {
            pose_s[3] = static_cast<float>(lm.x[0]); pose_s[4] = static_cast<float>(lm.x[1]);
            pose_s[5] = static_cast<float>(lm.x[2]); pose_s[6] = static_cast<float>(lm.x[3]);
            pose_s[0] = static_cast<float>(lm.x[4]); pose_s[1] = static_cast<float>(lm.x[5]); pose_s[2] = static_cast<float>(lm.x[6]);
            lm_total += lm.recorded;
            flag_s = (lm.last_step_norm < 1e-4 && outer > 3) ? 1 : 0;
            if (a.stats && blockIdx.x == 0) {
                for (int k = 0; k < 7; k++) a.stats->pose_log[outer][k] = pose_s[k];
                a.stats->outer_iterations = outer + 1;
                a.stats->lm_iterations = lm_total;
                a.stats->termination = lm.termination;
                a.stats->final_cost = lm.x_cost;
                a.stats->last_step_norm = lm.last_step_norm;
            }
        }
        __syncthreads();
        if (flag_s) break;
        __syncthreads();  // flag_s is rewritten by thread 0 at the top of the next solve
    }
    LO_TRACE(9);
    if (a.trace && blockIdx.x == 0 && threadIdx.x == 0) a.trace[1000] = trace_n;
    if (threadIdx.x == 0) {  // cloud_matcher.cpp:175 Quaternionf::normalize()
        const float w = pose_s[3], x = pose_s[4], y = pose_s[5], z = pose_s[6];
        const float nrm = sqrtf((x * x + z * z) + (y * y + w * w));
        const float out[7] = {pose_s[0], pose_s[1], pose_s[2], w / nrm, x / nrm, y / nrm, z / nrm};
        for (int k = 0; k < 7; k++) {
            if (blockIdx.x == 0) a.out_pose[k] = out[k];
            if (a.out_pose_block) a.out_pose_block[k] = out[k];
        }
    }
    __syncthreads();
}

}  // namespace lo
