// Per-rank phases of CloudMatcher::align against a key-frame map sharded by voxel-hash range, as device
// bodies shared by the stand-alone kernels of lo_match.cu (lo_shard_search_dev / lo_shard_evaluate_dev, pose and
// evaluation point passed by value) and the device-resident sharded matcher of lo_sharded.cu (pose and evaluation
// point read from its control block).
#pragma once

#include "lo_align.cuh"

namespace lo {

constexpr int kShardThreads = 256;
constexpr int kMaxShardWorld = 16;

// Exchange 1 fused into the search (nullable in every body below): query i has a HOME rank, i / chunk; a rank that finds a
// stored point for it folds its packed (d2 bits, probe, slot) key into the home rank's best[] with an atomicMin through
// peer memory (NVLink P2P, or plain device memory for shards that share a GPU). A voxel lives on exactly one rank, so keys
// are unique across ranks and the minimum is the reference's first strict minimum in scan order (voxel_grid.h:175-190).
struct PeerMin {
    unsigned long long* best[kMaxShardWorld];   // best[r] = rank r's array for the buffer in use, as seen from this device
    uint32_t chunk;                              // queries per home rank: ceil(m / world)
};
__device__ __forceinline__ void peer_min_push(const PeerMin* pm, uint32_t i, unsigned long long key) {
    if (pm && key != ~0ull) atomicMin(pm->best[i / pm->chunk] + i, key);
}

__device__ __forceinline__ void shard_search_body(GridView g, const PointXYZ* __restrict__ queries, uint32_t m, const QueryTransform& tf,
                                                                      float max_d2, unsigned long long* __restrict__ packed,
                                                                      float4* __restrict__ hit, const PeerMin* pm = nullptr) {
    __shared__ uint2 work[(kShardThreads / 32) * kWorkListStride];
    const int lane = threadIdx.x & 31;
    uint2* wl = work + (threadIdx.x >> 5) * kWorkListStride;
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    const LaneOffset lofs = lane_offset(lane);
    for (uint32_t qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; qi < m; qi += total_warps) {
        const PointXYZ q = queries[qi];
        float wq[3];
        transform_query(tf, q.x, q.y, q.z, wq);
        uint32_t vid, count;
        probe_resolve(g, probe_issue_at(g, clamp_axis(axis_index(wq[0], g.voxel_size)), clamp_axis(axis_index(wq[1], g.voxel_size)),
                                        clamp_axis(axis_index(wq[2], g.voxel_size)), lofs), vid, count);
        const Hit h = scan_candidates(g, wq, max_d2, lane, wl, vid, count);
        if (lane == 0) {
            packed[qi] = h.key;
            hit[qi] = make_float4(h.o[0], h.o[1], h.o[2], __uint_as_float(h.at));
            peer_min_push(pm, qi, h.key);
        }
    }
}

// the same with the tile search (lo_tile_search.cuh): a warp takes 32 consecutive queries
__device__ __forceinline__ void shard_search_tile_body(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                           const QueryTransform& tf, float max_d2,
                                                                           unsigned long long* __restrict__ packed, float4* __restrict__ hit,
                                                                           const PeerMin* pm = nullptr) {
    __shared__ TileScratch scratch[kShardThreads / 32];
    const int lane = threadIdx.x & 31;
    TileScratch* ts = scratch + (threadIdx.x >> 5);
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t base = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; base < m; base += total_warps * 32) {
        const uint32_t i = base + lane;
        const bool have = i < m;
        PointXYZ q{0.f, 0.f, 0.f, 0.f};
        if (have) q = queries[i];
        float wq[3];
        transform_query(tf, q.x, q.y, q.z, wq);
        const TileBest b = nn_search_tile_best<true>(g, wq, have, max_d2, lane, ts);
        if (have) {
            float4 o = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
            if (b.key != ~0ull) {
                const float4 p = __ldg(g.points + b.at);
                o = make_float4(p.x, p.y, p.z, __uint_as_float(b.at));
            }
            packed[i] = b.key;
            hit[i] = o;
            peer_min_push(pm, i, b.key);
        }
    }
}

// terms of the correspondences this rank owns; per-block partial sums in a fixed order
__device__ __forceinline__ void shard_evaluate_body(GridView g, const PointXYZ* __restrict__ queries, uint32_t m, const EvalPoint& ep,
                                                                        const unsigned long long* __restrict__ local,
                                                                        const unsigned long long* __restrict__ global,
                                                                        const float4* __restrict__ hit, double* __restrict__ partials) {
    __shared__ double rows[(kShardThreads / 32) * kRow];
    const int lane = threadIdx.x & 31;
    for (int k = threadIdx.x; k < (kShardThreads / 32) * kRow; k += blockDim.x) rows[k] = 0.0;
    __syncthreads();
    double* row = rows + (threadIdx.x >> 5) * kRow;
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t base = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; base < m; base += total_warps * 32) {
        const uint32_t i = base + lane;
        bool mine = false;
        Terms t;
        if (i < m) {
            const unsigned long long gl = global[i];
            mine = gl != ~0ull && local[i] == gl;
            if (mine) {
                const float4 o = hit[i];
                const float4 nv = __ldg(g.normals + __float_as_uint(o.w));
                const PointXYZ q = queries[i];
                const float pq[3] = {q.x, q.y, q.z}, of[3] = {o.x, o.y, o.z}, nf[3] = {nv.x, nv.y, nv.z};
                t = pair_terms(ep, pq, of, nf);
            }
        }
        if (!mine) {
            for (int k = 0; k < 6; k++) t.j[k] = 0.0;
            t.rc = 0.0;
            t.cost = 0.0;
        }
        accumulate_warp(row, t, mine, lane);
    }
    block_rows_to_global(rows, partials + blockIdx.x * 32);
}

}  // namespace lo
