// K4, tile form: one warp finds the correspondences of 32 queries at once (VoxelGrid::getCorrespondence,
// src/voxel_grid.h:164-204, for every point of findMatchingPairs' loop, :217-232).
//
// The reference visits all 27 neighbour voxels of a query and keeps the first strict minimum of the squared distance
// in (ix, iy, iz, slot) scan order, i.e. the lexicographic minimum of (d2, probe, slot) over the stored points with
// d2 < max_d2. That minimum does not depend on the order of the visit, so the tile search
//   A. looks at the query's own voxel first (probe 13),
//   B. drops every neighbour voxel whose box is farther away than the best distance found so far (it can only hold
//      points with a strictly larger d2; the box is shrunk by a margin that covers the rounding of the reference's
//      float division index, so the test errs on the side of visiting),
//   C. deals the remaining (query, voxel) look-ups of the whole tile out to the 32 lanes through a list in shared memory,
//      32 per round, and the stored points of the voxels they find through a second list, in pieces of 4 points,
//      each piece folding its best candidate into the query's 64-bit (d2 bits, probe, slot) key with a shared-memory
//      atomicMin.
// Typical key-frame queries have a stored point within a few centimetres in their own voxel, so only the one or two
// neighbours next to the query survive B; queries in unmapped space cost 26 empty look-ups that are spread over
// the warp instead of serialising one lane. The result is bit-identical to the exhaustive visit.
#pragma once

#include "lo_match.cuh"

namespace lo {

constexpr int kTileProbeCap = 26 * 32;     // neighbour look-ups of a tile
constexpr int kTilePieceCap = 32 * 5;      // pieces of one round: 32 voxels of <= 20 points
constexpr uint32_t kTileMaxWidth = 20;     // wider grids use the query-per-warp search of lo_match.cuh

struct alignas(16) TileScratch {           // per warp, shared memory
    unsigned long long best[32];           // per query: d2 bits << 32 | probe << 24 | slot ; ~0 = none
    float wq[3][32];                       // transformed queries
    uint32_t best_at[32];                  // per query: index of the best point in the point / normal slabs
    uint2 pieces[kTilePieceCap];           // x: index of the first point in the slab, y: query << 24 | probe << 16 | n << 8 | slot of x
    uint16_t probes[kTileProbeCap];        // query << 5 | probe
};

#ifdef LO_TILE_DEBUG
static __device__ unsigned long long g_tile_dbg[16];
#define LO_TILE_STAGE(k)                                                   \
    do {                                                                   \
        __syncwarp();                                                      \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                         \
            const long long now = clock64();                               \
            atomicAdd(&g_tile_dbg[k], static_cast<unsigned long long>(now - stage_t)); \
            stage_t = now;                                                 \
        }                                                                  \
    } while (0)
#else
#define LO_TILE_STAGE(k)
#endif

struct TileHit {      // per lane: the correspondence of the lane's own query
    bool valid;
    float o[3];       // plane_origin  (stored point)
    float n[3];       // plane_normal  (its normal)
};

// every lane brings one voxel (vid, count; count = 0: nothing) found for query `q` at neighbour index `probe`
__device__ inline void tile_scan_round(const GridView& g, TileScratch* ts, float max_d2, int lane, uint32_t vid, uint32_t count, uint32_t q,
                                       uint32_t probe) {
    const uint32_t n_pieces = (count + 3) >> 2;
    uint32_t incl = n_pieces;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t v = __shfl_up_sync(kFullMask, incl, off);
        if (lane >= off) incl += v;
    }
    const uint32_t total = __shfl_sync(kFullMask, incl, 31);
    if (total == 0) return;
    uint32_t at = incl - n_pieces;
    for (uint32_t e0 = 0; e0 < count; e0 += 4, at++)
        ts->pieces[at] = make_uint2(vid * g.width + e0, (q << 24) | (probe << 16) | (min(4u, count - e0) << 8) | e0);
    __syncwarp();
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        unsigned long long key = ~0ull;
        uint32_t qi = 0, at_point = 0;
        if (t < total) {
            const uint2 it = ts->pieces[t];
            const uint32_t n = (it.y >> 8) & 7u;
            qi = it.y >> 24;
            const float4* slab = g.points + it.x;
            float4 p[4];
#pragma unroll
            for (int u = 0; u < 4; u++) p[u] = (static_cast<uint32_t>(u) < n) ? __ldg(slab + u) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float qx = ts->wq[0][qi], qy = ts->wq[1][qi], qz = ts->wq[2][qi];
            float best_d2 = max_d2;
            uint32_t best_u = 0;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float dx = qx - p[u].x, dy = qy - p[u].y, dz = qz - p[u].z;
                const float d2 = dx * dx + (dy * dy + dz * dz);  // Vector3f::squaredNorm order
                if (static_cast<uint32_t>(u) < n && d2 < best_d2) {   // strict: the first minimum of the piece in slot order
                    best_d2 = d2;
                    best_u = u;
                }
            }
            if (best_d2 < max_d2) {
                key = (static_cast<unsigned long long>(__float_as_uint(best_d2)) << 32) | (((it.y >> 16) & 0xFFu) << 24) | ((it.y & 0xFFu) + best_u);
                at_point = it.x + best_u;
                if (key < ts->best[qi]) atomicMin(&ts->best[qi], key);
            }
        }
        __syncwarp();
        // keys are unique, so at most one piece of the round now owns its query's minimum
        if (key != ~0ull && ts->best[qi] == key) ts->best_at[qi] = at_point;
    }
    __syncwarp();
}

struct TileBest {     // per lane: the lane's own query
    unsigned long long key;   // d2 bits << 32 | probe << 24 | slot ; ~0 = no stored point within max_dist
    uint32_t at;              // index of that point in the point / normal slabs
};

// `visited` / `occupied` (nullable): this lane's share of the stored points / occupied voxels the tile looked at.
// kSharded: the grid holds one hash range of a sharded key frame (lo_grid_set_shard). Which rank owns a voxel follows from
// its indices alone, so look-ups of other ranks' voxels are dropped before they cost a slot load: the look-ups a rank
// issues shrink with the world size instead of staying at 27 per query.
template <bool kSharded = false>
__device__ inline TileBest nn_search_tile_best(const GridView& g, const float wq[3], bool have, float max_d2, int lane, TileScratch* ts,
                                               uint32_t* visited = nullptr, uint32_t* occupied = nullptr) {
    const float vs = g.voxel_size;
    const int ox = clamp_axis(axis_index(wq[0], vs)), oy = clamp_axis(axis_index(wq[1], vs)), oz = clamp_axis(axis_index(wq[2], vs));
#ifdef LO_TILE_DEBUG
    long long stage_t = clock64();
#endif
    ts->best[lane] = ~0ull;
    ts->wq[0][lane] = wq[0];
    ts->wq[1][lane] = wq[1];
    ts->wq[2][lane] = wq[2];
    __syncwarp();
    // A. the query's own voxel
    LaneOffset centre;
    centre.dx = centre.dy = centre.dz = 0;
    centre.probing = have;
    if constexpr (kSharded) centre.probing = have && (g.shard_world <= 1 || shard_owner(ox, oy, oz, g.shard_world) == g.shard_rank);
    uint32_t vid, count;
    probe_resolve(g, probe_issue_at(g, ox, oy, oz, centre), vid, count);
    if (visited) *visited += count;
    if (occupied) *occupied += count ? 1u : 0u;
    LO_TILE_STAGE(4);
    tile_scan_round(g, ts, max_d2, lane, vid, count, static_cast<uint32_t>(lane), 13u);
    LO_TILE_STAGE(5);
    // B. neighbours that can still hold a nearer (or equally near) point
    uint32_t mask = 0;
    if (have) {
        const unsigned long long b = ts->best[lane];
        const float bound = b == ~0ull ? max_d2 : __uint_as_float(static_cast<uint32_t>(b >> 32));
        float m2[3], p2[3];   // squared distance from the query to the neighbour boxes on the minus / plus side of each axis
        const int o[3] = {ox, oy, oz};
#pragma unroll
        for (int k = 0; k < 3; k++) {
            // index i holds x with trunc(x / vs) == i: (i*vs, (i+1)*vs) for i > 0, ((i-1)*vs, i*vs) for i < 0, (-vs, vs) for 0.
            // The boxes are shrunk by a margin that covers the rounding of x / vs.
            const float lo_edge = static_cast<float>(o[k] > 0 ? o[k] : o[k] - 1) * vs, hi_edge = static_cast<float>(o[k] >= 0 ? o[k] + 1 : o[k]) * vs;
            const float margin = 1e-3f * vs + 1e-6f * fabsf(wq[k]);
            const float dm = fmaxf(0.f, (wq[k] - lo_edge) - margin), dp = fmaxf(0.f, (hi_edge - wq[k]) - margin);
            m2[k] = dm * dm;
            p2[k] = dp * dp;
        }
#pragma unroll
        for (int l = 0; l < 27; l++) {
            if (l == 13) continue;
            const int dx = l / 9 - 1, dy = (l / 3) % 3 - 1, dz = l % 3 - 1;
            const float far2 = (dx < 0 ? m2[0] : (dx > 0 ? p2[0] : 0.f)) + (dy < 0 ? m2[1] : (dy > 0 ? p2[1] : 0.f)) +
                               (dz < 0 ? m2[2] : (dz > 0 ? p2[2] : 0.f));
            bool visit = !(far2 > bound);   // NaN: visit
            if constexpr (kSharded) visit = visit && (g.shard_world <= 1 || shard_owner(ox + dx, oy + dy, oz + dz, g.shard_world) == g.shard_rank);
            if (visit) mask |= 1u << l;
        }
    }
    LO_TILE_STAGE(6);
    // C. the tile's look-ups, 32 per round
    uint32_t n_mine = __popc(mask), incl = n_mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t v = __shfl_up_sync(kFullMask, incl, off);
        if (lane >= off) incl += v;
    }
    const uint32_t total = __shfl_sync(kFullMask, incl, 31);
    {
        uint32_t at = incl - n_mine, m = mask;
        while (m) {
            const uint32_t l = __ffs(m) - 1;
            m &= m - 1;
            ts->probes[at++] = static_cast<uint16_t>((lane << 5) | l);
        }
    }
    __syncwarp();
#ifdef LO_TILE_DEBUG
    {
        const unsigned hits = __popc(__ballot_sync(kFullMask, have && ts->best[lane] != ~0ull)), haves = __popc(__ballot_sync(kFullMask, have));
        if (lane == 0) {
            atomicAdd(&g_tile_dbg[0], static_cast<unsigned long long>(total));
            atomicAdd(&g_tile_dbg[1], 1ull);
            atomicAdd(&g_tile_dbg[2], static_cast<unsigned long long>(hits));
            atomicAdd(&g_tile_dbg[3], static_cast<unsigned long long>(haves));
        }
    }
#endif
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        const uint32_t entry = t < total ? ts->probes[t] : 0u;
        const int q = static_cast<int>(entry >> 5), l = static_cast<int>(entry & 31u);
        const int qx = __shfl_sync(kFullMask, ox, q), qy = __shfl_sync(kFullMask, oy, q), qz = __shfl_sync(kFullMask, oz, q);
        LaneOffset lo;
        lo.dx = l / 9 - 1;
        lo.dy = (l / 3) % 3 - 1;
        lo.dz = l % 3 - 1;
        lo.probing = t < total;
        LO_TILE_STAGE(7);
        probe_resolve(g, probe_issue_at(g, qx, qy, qz, lo), vid, count);
        if (visited) *visited += count;
        if (occupied) *occupied += count ? 1u : 0u;
        LO_TILE_STAGE(8);
        tile_scan_round(g, ts, max_d2, lane, vid, count, static_cast<uint32_t>(q), static_cast<uint32_t>(l));
        LO_TILE_STAGE(9);
    }
    LO_TILE_STAGE(10);
    TileBest b;
    b.key = have ? ts->best[lane] : ~0ull;
    b.at = b.key != ~0ull ? ts->best_at[lane] : 0u;
    __syncwarp();   // the scratch is reused by the warp's next tile
    return b;
}

// Lane l holds query l of the tile (world coordinates wq; `have` = the lane has a query). All 32 lanes must call.
__device__ inline TileHit nn_search_tile(const GridView& g, const float wq[3], bool have, float max_d2, int lane, TileScratch* ts,
                                         uint32_t* visited = nullptr, uint32_t* occupied = nullptr) {
    const TileBest b = nn_search_tile_best(g, wq, have, max_d2, lane, ts, visited, occupied);
    // D. the winner's point and normal
    TileHit hit;
    hit.valid = false;
    hit.o[0] = hit.o[1] = hit.o[2] = 0.f;
    hit.n[0] = hit.n[1] = hit.n[2] = 0.f;
    if (b.key != ~0ull) {
        const float4 p = __ldg(g.points + b.at), nv = __ldg(g.normals + b.at);
        hit.valid = true;
        hit.o[0] = p.x; hit.o[1] = p.y; hit.o[2] = p.z;
        hit.n[0] = nv.x; hit.n[1] = nv.y; hit.n[2] = nv.z;
    }
    return hit;
}

}  // namespace lo
