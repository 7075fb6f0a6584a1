// CloudMatcher::align on the device: residual / Jacobian terms, the warp / block / grid reductions of the 6x6 normal
// equations and the persistent cooperative loop. Shared by lo_match.cu (align_kernel, nn_residual_kernel, sharded
// phases) and lo_pipeline.cu (the per-scan kernel).
#pragma once

#include <cooperative_groups.h>

#include <cstddef>

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
// Grid-wide sum of the blocks' column sums with the barrier folded into the exchange. Every block writes its kAcc column
// sums into its row of `partials` (256 B), publishes the row by storing the epoch into the row's last slot, then waits for
// all rows of this epoch and adds them up in block order — one L2 round trip from "my sums are ready" to "everybody's
// sums are here", instead of a grid barrier (fence + arrival counter + spin) followed by the reads. The sum is the same
// in every block, bit for bit (fixed order). Two row sets alternate: a block can be at most one exchange ahead of the
// slowest one (it needs that block's row to finish the exchange in between), so a row is never overwritten while
// somebody still reads it, and a flag never moves on before everybody has seen it.
//   cols    : the block's kAcc column sums (shared), zeroed on return
//   scratch : (blockDim.x / 32) * 32 doubles of shared memory, zeroed on return
//   total   : kAcc sums, shared memory
//   epoch   : unique per exchange among all launches that share `partials` (never 0), identical in all blocks
__device__ inline void exchange_sum(double* cols, double* scratch, double* partials, int nblocks, int buf, unsigned long long epoch,
                                    double* total) {
    double* mine = partials + (static_cast<size_t>(buf) * nblocks + blockIdx.x) * 32;
    __syncthreads();
    if (threadIdx.x < kAcc) {
        __stcg(mine + threadIdx.x, cols[threadIdx.x]);
        cols[threadIdx.x] = 0.0;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long*>(mine + 31) = epoch;
    const double* all = partials + static_cast<size_t>(buf) * nblocks * 32;
    // one poller per row (measured: one warp sweeping 32 flags per load was 1.5x slower — five dependent sweeps)
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) {
        const volatile unsigned long long* flag = reinterpret_cast<const volatile unsigned long long*>(all + static_cast<size_t>(b) * 32 + 31);
        while (*flag != epoch) {
        }
        __threadfence();
    }
    __syncthreads();
    reduce_partials(all, nblocks, scratch, total);
}

// per-warp rows (accumulate_warp) -> the block's column sums; rows are zeroed for the next pass
__device__ inline void rows_to_cols(double* rows, double* cols) {
    __syncthreads();
    const int nwarps = blockDim.x >> 5;
    if (threadIdx.x < kAcc) {
        double v = cols[threadIdx.x];
        for (int w = 0; w < nwarps; w++) {
            v += rows[w * kRow + threadIdx.x];
            rows[w * kRow + threadIdx.x] = 0.0;
        }
        cols[threadIdx.x] = v;
    }
    __syncthreads();
}

// The block's pairs (one per thread, threads 0 .. kPairs-1) added to its column sums through a transposed staging area in
// shared memory: every thread stores its 29 products (conflict-free), then 8 threads per column add kPairs / 8 values each
// in a fixed order and finish with three shuffles. ~600 cycles for 64 pairs; reducing the 29 values one after the other
// by warp shuffles into a shared accumulator took ~6 000. `stage` holds 29 * kPairs doubles. Every thread of the block
// calls; `t` is only meaningful where `valid`.
template <int kPairs>
__device__ inline void block_accumulate(const Terms& t, bool valid, double* stage, double* cols) {
    static_assert(kPairs % 8 == 0, "8 adders per column");
    const int tid = threadIdx.x;
    if (tid < kPairs) {
        int at = 0;
#pragma unroll
        for (int a = 0; a < 6; a++)
#pragma unroll
            for (int b = a; b < 6; b++) stage[(at++) * kPairs + tid] = valid ? t.j[a] * t.j[b] : 0.0;
#pragma unroll
        for (int a = 0; a < 6; a++) stage[(21 + a) * kPairs + tid] = valid ? t.j[a] * t.rc : 0.0;
        stage[27 * kPairs + tid] = valid ? t.cost : 0.0;
        stage[28 * kPairs + tid] = valid ? 1.0 : 0.0;
    }
    __syncthreads();
    if (tid < 29 * 8) {
        const int c = tid >> 3, part = tid & 7;
        constexpr int kEach = kPairs / 8;
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < kEach; k++) v += stage[c * kPairs + part * kEach + k];
        const unsigned mask = tid < 224 ? kFullMask : 0xFFu;   // 29 * 8 = 232 threads: seven whole warps and eight lanes of the eighth
        v += __shfl_xor_sync(mask, v, 4);
        v += __shfl_xor_sync(mask, v, 2);
        v += __shfl_xor_sync(mask, v, 1);
        if (part == 0) cols[c] += v;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------
// Correspondence search of one block's share of the queries, all threads of the block working on all of them
// (VoxelGrid::findMatchingPairs + getCorrespondence, src/voxel_grid.h:164-234). A VLP16 scan leaves ~40 queries per
// block: with one warp per query a block waited for its slowest warp and every warp walked through five dependent
// memory round trips per query. Here the block goes through the same work in flat steps, each one round of
// independent loads over all threads:
//   A  threads 0..nq-1: transform their query (double) and take its voxel;
//   B  one thread per (query, neighbour) pair — 27 nq look-ups: key, probe hash, slot load, resolve;
//   C  block-wide prefix sum over the pieces (4 stored points = 64 B) of the voxels found;
//   D  one thread per piece: distances, the piece's first strict minimum folded into the query's 64-bit
//      (d2 bits, neighbour, slot) key with a shared-memory atomicMin — the lexicographic minimum is the reference's
//      "first strict minimum in ix -> iy -> iz -> slot scan order", whatever the order of the visit;
//   E  threads 0..nq-1: fetch the winner's point and normal.
constexpr int kAlignThreads = 512;
constexpr int kBlockQueries = 64;                 // queries of one block per pass (two warps hold one each in registers)
constexpr int kBlockProbes = kBlockQueries * 27;

struct BlockSearch {                              // shared memory
    float wq[3][kBlockQueries];
    int oi[3][kBlockQueries];
    unsigned long long best[kBlockQueries];
    uint2 found[kBlockProbes];                    // vid, count of neighbour l of query q at [l * nq + q]
    uint32_t piece_start[kBlockProbes + 1];       // exclusive prefix of ceil(count / 4)
};

struct PairRegs {                                 // one correspondence, kept by the thread that owns the query
    float p[3], o[3], n[3];
    bool valid;
};

// query k of this block in pass `pass` is global query (pass * kBlockQueries + k) * gridDim.x + blockIdx.x: neighbouring
// queries (which cost about the same) go to different blocks
__device__ inline uint32_t block_query_index(uint32_t pass, uint32_t k) { return (pass * kBlockQueries + k) * gridDim.x + blockIdx.x; }

#define LO_BS_TRACE(tag)                                                               \
    do {                                                                               \
        if (trace && blockIdx.x == trace_block && threadIdx.x == 0 && *trace_n < 500) { \
            trace[2 * *trace_n] = (tag);                                               \
            trace[2 * *trace_n + 1] = clock64();                                       \
            (*trace_n)++;                                                              \
        }                                                                              \
    } while (0)

__device__ inline PairRegs block_search(const GridView& g, const PointXYZ* __restrict__ queries, uint32_t m, uint32_t pass, const QueryTransform& tf,
                                        float max_d2, BlockSearch& bs, long long* trace = nullptr, int* trace_n = nullptr, uint32_t trace_block = 0) {
    const uint32_t tid = threadIdx.x;
    const uint32_t first = block_query_index(pass, 0);
    const uint32_t nq = first < m ? min(static_cast<uint32_t>(kBlockQueries), (m - first + gridDim.x - 1) / gridDim.x) : 0u;
    PairRegs pr;
    pr.valid = false;
    for (int k = 0; k < 3; k++) pr.p[k] = pr.o[k] = pr.n[k] = 0.f;
    // A
    if (tid < nq) {
        const PointXYZ q = queries[block_query_index(pass, tid)];
        pr.p[0] = q.x; pr.p[1] = q.y; pr.p[2] = q.z;
        float w[3];
        transform_query(tf, q.x, q.y, q.z, w);
        for (int k = 0; k < 3; k++) {
            bs.wq[k][tid] = w[k];
            bs.oi[k][tid] = clamp_axis(axis_index(w[k], g.voxel_size));
        }
        bs.best[tid] = ~0ull;
    }
    __syncthreads();
    LO_BS_TRACE(10);
    // B: the look-ups, neighbour-major (entry e = l * nq + q): consecutive threads look at different queries, so that the
    // pieces consecutive threads fold in step D mostly belong to different queries
    const uint32_t n_probe = nq * 27;
    for (uint32_t base = 0; base < n_probe; base += 4 * blockDim.x) {   // 4 look-ups per thread in flight
        Probe pb[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t e = base + tid + u * blockDim.x;
            pb[u].active = false;
            if (e < n_probe) {
                const uint32_t l = e / nq, q = e - l * nq;
                LaneOffset lo;
                lo.probing = true;
                lo.dx = static_cast<int>(l / 9) - 1;
                lo.dy = static_cast<int>((l / 3) % 3) - 1;
                lo.dz = static_cast<int>(l % 3) - 1;
                pb[u] = probe_issue_at(g, bs.oi[0][q], bs.oi[1][q], bs.oi[2][q], lo);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t e = base + tid + u * blockDim.x;
            if (e < n_probe) {
                uint32_t vid, cnt;
                probe_resolve(g, pb[u], vid, cnt);
                bs.found[e] = make_uint2(vid, cnt);
            }
        }
    }
    __syncthreads();
    LO_BS_TRACE(11);
    // C: thread t owns look-ups kPer * t .. kPer * t + kPer - 1 of the prefix sum
    uint32_t total_pieces;
    {
        constexpr int kPer = (kBlockProbes + kAlignThreads) / kAlignThreads;
        static_assert(kPer * kAlignThreads > kBlockProbes, "every entry and the total need an owner");
        uint32_t c[kPer], sum = 0;
#pragma unroll
        for (int u = 0; u < kPer; u++) {
            const uint32_t e = tid * kPer + u;
            c[u] = e < n_probe ? (bs.found[e].y + 3) >> 2 : 0u;
            sum += c[u];
        }
        __shared__ uint32_t warp_sums[kAlignThreads / 32];
        const int lane = tid & 31, warp = tid >> 5;
        uint32_t incl = sum;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t v = __shfl_up_sync(kFullMask, incl, off);
            if (lane >= off) incl += v;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        uint32_t before = 0, all = 0;
        for (int w = 0; w < kAlignThreads / 32; w++) {
            const uint32_t ws = warp_sums[w];
            if (w < warp) before += ws;
            all += ws;
        }
        total_pieces = all;
        uint32_t run = before + incl - sum;
#pragma unroll
        for (int u = 0; u < kPer; u++) {
            const uint32_t e = tid * kPer + u;
            if (e <= n_probe) bs.piece_start[e] = run;   // entry n_probe = total
            run += c[u];
        }
    }
    __syncthreads();
    LO_BS_TRACE(12);
    // D
    for (uint32_t pc = tid; pc < total_pieces; pc += blockDim.x) {
        uint32_t lo_e = 0, hi_e = n_probe;               // last e with piece_start[e] <= pc
        while (hi_e - lo_e > 1) {
            const uint32_t mid = (lo_e + hi_e) >> 1;
            if (bs.piece_start[mid] <= pc) lo_e = mid;
            else hi_e = mid;
        }
        const uint32_t e = lo_e, l = e / nq, q = e - l * nq;
        const uint2 f = bs.found[e];
        const uint32_t e0 = (pc - bs.piece_start[e]) * 4, cnt = min(4u, f.y - e0);
        const float4* slab = g.points + static_cast<size_t>(f.x) * g.width + e0;
        float4 p[4];
#pragma unroll
        for (int u = 0; u < 4; u++) p[u] = (static_cast<uint32_t>(u) < cnt) ? __ldg(slab + u) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float qx = bs.wq[0][q], qy = bs.wq[1][q], qz = bs.wq[2][q];
        float best_d2 = max_d2;
        uint32_t best_u = 0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const float dx = qx - p[u].x, dy = qy - p[u].y, dz = qz - p[u].z;
            const float d2 = dx * dx + (dy * dy + dz * dz);   // Vector3f::squaredNorm order
            if (static_cast<uint32_t>(u) < cnt && d2 < best_d2) {   // strict: the piece's first minimum in slot order
                best_d2 = d2;
                best_u = u;
            }
        }
        if (best_d2 < max_d2) {
            const unsigned long long key = (static_cast<unsigned long long>(__float_as_uint(best_d2)) << 32) | (l << 24) | (e0 + best_u);
            if (key < bs.best[q]) atomicMin(&bs.best[q], key);
        }
    }
    __syncthreads();
    LO_BS_TRACE(13);
    // E
    if (tid < nq) {
        const unsigned long long key = bs.best[tid];
        if (key != ~0ull) {
            const uint32_t l = (static_cast<uint32_t>(key) >> 24) & 31u, slot = static_cast<uint32_t>(key) & 0xFFFFFFu;
            const size_t at = static_cast<size_t>(bs.found[l * nq + tid].x) * g.width + slot;
            const float4 o = __ldg(g.points + at), nv = __ldg(g.normals + at);
            pr.valid = true;
            pr.o[0] = o.x; pr.o[1] = o.y; pr.o[2] = o.z;
            pr.n[0] = nv.x; pr.n[1] = nv.y; pr.n[2] = nv.z;
        }
    }
    __syncthreads();   // bs is reused by the next pass
    LO_BS_TRACE(14);
    return pr;
}

// ---------------------------------------------------------------------------------------------------
// CloudMatcher::align (src/cloud_matcher.cpp:105-178) as one persistent kernel of co-resident blocks.
struct AlignArgs {
    GridView g;
    const PointXYZ* queries;
    uint32_t m;
    const uint32_t* m_dev; // nullable: actual query count on the device (m is then an upper bound)
    float4* pair_o;        // [m] plane_origin, w = 1 if the query has a correspondence
    float4* pair_n;        // [m] plane_normal
    double* partials;      // [2][gridDim.x][32]: rows of exchange_sum; zero-filled when allocated, written by nothing else
    unsigned long long epoch_base = 0;   // unique per launch among the launches that share `partials`; low 16 bits zero
    const float* guess;    // device pose7
    float* out_pose;       // device pose7 (written by block 0)
    float* out_pose_block = nullptr;  // nullable: 7 floats in the caller's shared memory, written in every block
    lo_align_stats* stats; // device, nullable
    float max_dist;
    long long* trace;      // nullable: (tag, clock64) pairs written by thread 0 of block trace_block (LO_ALIGN_TRACE / LO_SCAN_TRACE)
    int trace_start = 0;   // first free trace entry (the per-scan kernel has written some already)
    uint32_t trace_block = 0;
};

#define LO_TRACE(tag)                                                                      \
    do {                                                                                   \
        if (a.trace && blockIdx.x == a.trace_block && threadIdx.x == 0 && trace_n < 500) { \
            a.trace[2 * trace_n] = (tag);                                                  \
            a.trace[2 * trace_n + 1] = clock64();                                          \
            trace_n++;                                                                     \
        }                                                                                  \
    } while (0)

// The body of align_kernel; also the middle part of the per-scan kernel of lo_pipeline.cu. Every thread of a grid of
// co-resident kAlignThreads-wide blocks must call it. The blocks meet only in exchange_sum: there is no other
// cross-block data flow (a block evaluates the correspondences it found itself; the LM step logic runs redundantly in
// every block on identical sums).
// `tile_scratch`: nullable, (kAlignThreads / 32) TileScratch in shared memory; with it, large query sets (>= 8 queries
// per warp) are searched 32 per warp with exact neighbour pruning (lo_tile_search.cuh) instead of block-wide.
constexpr size_t kAlignTileScratchBytes = (kAlignThreads / 32) * sizeof(TileScratch);
__device__ inline void align_device(AlignArgs a, TileScratch* tile_scratch = nullptr) {
    if (a.m_dev) a.m = min(a.m, *a.m_dev);
    __shared__ double rows[(kAlignThreads / 32) * kRow];
    __shared__ double cols[32];   // the block's column sums of the evaluation in progress
    __shared__ double total[32];
    __shared__ LmState lm;
    __shared__ EvalPoint ep_s;
    __shared__ float pose_s[7];
    __shared__ int flag_s;
    __shared__ BlockSearch bs;

    const int lane = threadIdx.x & 31;
    // tile search: consecutive tiles go to different SMs: warp w of block b is global warp w * gridDim.x + b
    const uint32_t gwarp = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    const int nblocks = gridDim.x;
    // the tile search handles voxels of up to kTileMaxWidth points; the block search walks a voxel in pieces of 4 whatever its width
    const bool tiled = tile_scratch != nullptr && a.g.width <= kTileMaxWidth && a.m >= 8 * total_warps;
    const uint32_t passes = (a.m + kBlockQueries * gridDim.x - 1) / (kBlockQueries * gridDim.x);   // block search: passes of 64 queries per block
    TileScratch* ts = tile_scratch ? tile_scratch + (threadIdx.x >> 5) : nullptr;
    double* row = rows + (threadIdx.x >> 5) * kRow;
    for (int k = threadIdx.x; k < (kAlignThreads / 32) * kRow; k += blockDim.x) rows[k] = 0.0;
    if (threadIdx.x < 32) cols[threadIdx.x] = 0.0;
    double* stage = reinterpret_cast<double*>(bs.found);   // 29 x 64 doubles over found + piece_start, free between two searches
    static_assert(sizeof(bs.found) + sizeof(bs.piece_start) >= 29 * kBlockQueries * sizeof(double), "staging area of block_accumulate");
    static_assert(offsetof(BlockSearch, piece_start) == offsetof(BlockSearch, found) + sizeof(bs.found), "found and piece_start are contiguous");
    if (threadIdx.x < 7) pose_s[threadIdx.x] = a.guess[threadIdx.x];
    __syncthreads();
    const double prior_b[3] = {static_cast<double>(a.guess[0]), static_cast<double>(a.guess[1]), static_cast<double>(a.guess[2])};
    const float max_d2 = a.max_dist * a.max_dist;  // voxel_grid.h:215
    int buf = 0;
    unsigned long long epoch = a.epoch_base;
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
        PairRegs mine;   // block search: the correspondence of this thread's query of pass 0 stays in registers for the LM evaluations
        mine.valid = false;
        // large query sets: query i in lane (i & 31) of the warp that owns tile i / 32
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
        for (uint32_t pass = 0; !tiled && pass < passes; pass++) {
            const PairRegs pr = block_search(a.g, a.queries, a.m, pass, tf, max_d2, bs, a.trace, &trace_n, a.trace_block);
            if (pass == 0) mine = pr;
            Terms t;
            if (threadIdx.x < kBlockQueries) {   // two warps hold the block's pairs, one per lane
                const uint32_t i = block_query_index(pass, threadIdx.x);
                if (i < a.m && pass > 0) {   // later passes are re-read from here by the evaluations
                    a.pair_o[i] = make_float4(pr.o[0], pr.o[1], pr.o[2], pr.valid ? 1.f : 0.f);
                    a.pair_n[i] = make_float4(pr.n[0], pr.n[1], pr.n[2], 0.f);
                }
                if (pr.valid) t = pair_terms(ep0, pr.p, pr.o, pr.n);
            }
            block_accumulate<kBlockQueries>(t, pr.valid, stage, cols);
        }
        if (tiled) rows_to_cols(rows, cols);
        LO_TRACE(1);
        exchange_sum(cols, rows, a.partials, nblocks, buf, ++epoch, total);
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
            if (tiled) {
                for (uint32_t base = gwarp * 32; base < a.m; base += total_warps * 32) {
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
                rows_to_cols(rows, cols);
            } else {
                for (uint32_t pass = 0; pass < passes; pass++) {
                    PairRegs pr = mine;
                    if (pass > 0) {
                        pr.valid = false;
                        const uint32_t i = block_query_index(pass, threadIdx.x);
                        if (threadIdx.x < kBlockQueries && i < a.m) {
                            const float4 o = __ldcg(a.pair_o + i);
                            pr.valid = o.w != 0.f;
                            if (pr.valid) {
                                const float4 nv = __ldcg(a.pair_n + i);
                                const PointXYZ q = a.queries[i];
                                pr.p[0] = q.x; pr.p[1] = q.y; pr.p[2] = q.z;
                                pr.o[0] = o.x; pr.o[1] = o.y; pr.o[2] = o.z;
                                pr.n[0] = nv.x; pr.n[1] = nv.y; pr.n[2] = nv.z;
                            }
                        }
                    }
                    Terms t;
                    if (pr.valid) t = pair_terms(ep_s, pr.p, pr.o, pr.n);
                    block_accumulate<kBlockQueries>(t, pr.valid, stage, cols);
                }
            }
            LO_TRACE(5);
            exchange_sum(cols, rows, a.partials, nblocks, buf, ++epoch, total);
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
    if (a.trace && blockIdx.x == a.trace_block && threadIdx.x == 0) a.trace[1000] = trace_n;
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
