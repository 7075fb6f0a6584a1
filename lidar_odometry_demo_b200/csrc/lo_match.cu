// K4/K5/K6 kernels and their C-ABI entry points:
//   find_pairs_kernel      VoxelGrid::findMatchingPairs with the query index kept (parity dumps)
//   nn_residual_kernel     the fused metric kernel: 27-neighbour search + residual/Jacobian/Huber weight +
//                          warp-shuffle / shared-memory reduction of the 6x6 normal equations
//   align_kernel           CloudMatcher::align as ONE persistent cooperative kernel: up to 35 outer
//                          iterations x (search + evaluate, <= 4 LM iterations) with grid-wide barriers and
//                          no host round trip until the final pose

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "lo_shard.cuh"

namespace lo {

// ---------------------------------------------------------------------------------------------------
__global__ void find_pairs_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m, QueryTransform tf, float max_d2,
                                  lo_pair* __restrict__ pairs, uint8_t* __restrict__ valid, uint32_t* __restrict__ n_valid) {
    __shared__ uint2 work[(256 / 32) * kWorkListStride];
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= m) return;
    const PointXYZ q = queries[warp];
    float wq[3];
    transform_query(tf, q.x, q.y, q.z, wq);
    const Found f = nn_search_warp(g, wq, max_d2, lane, work + (threadIdx.x >> 5) * kWorkListStride);
    if (lane == 0) {
        lo_pair out;
        memset(&out, 0, sizeof(out));
        if (f.valid) {
            out.source_point_local[0] = q.x; out.source_point_local[1] = q.y; out.source_point_local[2] = q.z;
            for (int k = 0; k < 3; k++) {
                out.plane_origin[k] = f.o[k];
                out.plane_normal[k] = f.n[k];
            }
            atomicAdd(n_valid, 1u);
        }
        pairs[warp] = out;
        valid[warp] = f.valid ? 1 : 0;
    }
}

// Fused NN + residual. Each warp walks tiles of 32 consecutive queries: the 27-lane search runs once per
// query, lane j keeps the correspondence of query j, then all 32 lanes evaluate their residual/Jacobian in
// parallel and the warp reduces the 28 sums by shuffles into its shared-memory row.
constexpr int kNnThreads = 256;
constexpr int kNnBlocksPerSm = 4;   // resident blocks per SM of the tile kernel (64 registers)
__global__ void __launch_bounds__(kNnThreads, 4) nn_residual_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                  QueryTransform tf, EvalPoint ep, float max_d2,
                                                                  double* __restrict__ partials) {
    __shared__ double rows[(kNnThreads / 32) * kRow];
    __shared__ uint2 work[(kNnThreads / 32) * kWorkListStride];
    const int lane = threadIdx.x & 31, warp_in_block = threadIdx.x >> 5;
    uint2* wl = work + warp_in_block * kWorkListStride;
    const LaneOffset lofs = lane_offset(lane);
    for (int k = threadIdx.x; k < (kNnThreads / 32) * kRow; k += blockDim.x) rows[k] = 0.0;
    __syncthreads();
    double* row = rows + warp_in_block * kRow;
    // every warp owns one contiguous range of ceil(m / warps) queries (balanced: no partial last wave) and walks it in
    // tiles of 32
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t per_warp = (m + total_warps - 1) / total_warps;
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t range_begin = min(m, gw * per_warp), range_end = min(m, range_begin + per_warp);
    for (uint32_t base = range_begin; base < range_end; base += 32) {
        const uint32_t mine = base + lane;
        PointXYZ q{0.f, 0.f, 0.f, 0.f};
        if (mine < range_end) q = queries[mine];  // coalesced 16 B loads
        bool my_valid = false;
        float my_o[3] = {0.f, 0.f, 0.f}, my_n[3] = {0.f, 0.f, 0.f};
        uint32_t my_at = 0;
        uint32_t visited = 0, occupied = 0;
        const uint32_t in_tile = min(32u, range_end - base);
        // Lane l transforms query l once (findMatchingPairs, voxel_grid.h:217-223) and takes its voxel index; the loop
        // broadcasts them. Software pipeline: the slot loads of query j+1 are in flight while query j is scanned.
        float own[3];
        transform_query(tf, q.x, q.y, q.z, own);
        const int oi[3] = {clamp_axis(axis_index(own[0], g.voxel_size)), clamp_axis(axis_index(own[1], g.voxel_size)),
                           clamp_axis(axis_index(own[2], g.voxel_size))};
        Probe next = probe_issue_at(g, __shfl_sync(kFullMask, oi[0], 0), __shfl_sync(kFullMask, oi[1], 0), __shfl_sync(kFullMask, oi[2], 0), lofs);
        for (uint32_t j = 0; j < in_tile; j++) {
            uint32_t vid, cnt;
            probe_resolve(g, next, vid, cnt);
            const int jj = static_cast<int>(j);
            const float cur[3] = {__shfl_sync(kFullMask, own[0], jj), __shfl_sync(kFullMask, own[1], jj), __shfl_sync(kFullMask, own[2], jj)};
            if (j + 1 < in_tile)
                next = probe_issue_at(g, __shfl_sync(kFullMask, oi[0], jj + 1), __shfl_sync(kFullMask, oi[1], jj + 1),
                                      __shfl_sync(kFullMask, oi[2], jj + 1), lofs);
            const Hit hit = scan_candidates(g, cur, max_d2, lane, wl, vid, cnt);
            visited += cnt;
            occupied += cnt ? 1u : 0u;
            if (lane == static_cast<int>(j)) {
                my_valid = hit.valid;
                my_at = hit.at;
                for (int k = 0; k < 3; k++) my_o[k] = hit.o[k];
            }
        }
        if (my_valid) {  // the 32 winning normals of the tile in one round of loads
            const float4 nv = __ldg(g.normals + my_at);
            my_n[0] = nv.x; my_n[1] = nv.y; my_n[2] = nv.z;
        }
        const float pf[3] = {q.x, q.y, q.z};
        const Terms t = pair_terms(ep, pf, my_o, my_n);
        accumulate_warp(row, t, my_valid, lane);
        const double vis = warp_sum(static_cast<double>(visited)), occ = warp_sum(static_cast<double>(occupied));
        if (lane == 0) {
            row[29] += vis;
            row[30] += occ;
        }
    }
    block_rows_to_global(rows, partials + blockIdx.x * 32);
}

// The same contract with the tile search of lo_tile_search.cuh (own voxel first, neighbours pruned exactly by the best
// distance so far): a warp takes 32 consecutive queries, lane l owns query l from the transform to the residual.
// rows[29] / rows[30] count the stored points / occupied voxels actually visited.
__global__ void __launch_bounds__(kNnThreads, kNnBlocksPerSm) nn_residual_tile_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                       QueryTransform tf, EvalPoint ep, float max_d2,
                                                                       double* __restrict__ partials) {
    __shared__ double rows[(kNnThreads / 32) * kRow];
    __shared__ TileScratch scratch[kNnThreads / 32];
    const int lane = threadIdx.x & 31, warp_in_block = threadIdx.x >> 5;
    TileScratch* ts = scratch + warp_in_block;
    for (int k = threadIdx.x; k < (kNnThreads / 32) * kRow; k += blockDim.x) rows[k] = 0.0;
    __syncthreads();
    double* row = rows + warp_in_block * kRow;
    const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (uint32_t base = gw * 32; base < m; base += total_warps * 32) {
        const uint32_t i = base + lane;
        const bool have = i < m;
        PointXYZ q{0.f, 0.f, 0.f, 0.f};
        if (have) q = queries[i];
        float wq[3];
        transform_query(tf, q.x, q.y, q.z, wq);
        uint32_t visited = 0, occupied = 0;
        const TileHit h = nn_search_tile(g, wq, have, max_d2, lane, ts, &visited, &occupied);
        const float pf[3] = {q.x, q.y, q.z};
        const Terms t = pair_terms(ep, pf, h.o, h.n);
        accumulate_warp(row, t, h.valid, lane);
        const double vis = warp_sum(static_cast<double>(visited)), occ = warp_sum(static_cast<double>(occupied));
        if (lane == 0) {
            row[29] += vis;
            row[30] += occ;
        }
    }
    block_rows_to_global(rows, partials + blockIdx.x * 32);
}

// final reduction of nn_residual_kernel's partials + the prior, by one block
__global__ void normal_eq_finalize_kernel(const double* __restrict__ partials, int nblocks, EvalPoint ep, double b0, double b1, double b2,
                                          lo_normal_eq* __restrict__ out) {
    __shared__ double total[32];
    __shared__ double scratch[(256 / 32) * kRow];
    reduce_partials(partials, nblocks, scratch, total);
    if (threadIdx.x == 0) {
        double acc[kAcc];
        for (int k = 0; k < kAcc; k++) acc[k] = total[k];
        const double b[3] = {b0, b1, b2};
        add_prior(acc, ep.t, b);
        out->cost = acc[27];
        for (int a = 0; a < 6; a++) {
            out->g[a] = acc[21 + a];
            for (int c = 0; c < 6; c++) out->jtj[a * 6 + c] = sym(acc, a, c);
        }
        out->n_valid = static_cast<uint64_t>(acc[28] + 0.5);
        out->candidates = static_cast<uint64_t>(acc[29] + 0.5);
        out->occupied = static_cast<uint64_t>(acc[30] + 0.5);
    }
}

// ---------------------------------------------------------------------------------------------------
// CloudMatcher::align (src/cloud_matcher.cpp:105-178) as one persistent cooperative kernel (body: lo_align.cuh)
__global__ void __launch_bounds__(kAlignThreads, 1) align_kernel(AlignArgs a) {
    extern __shared__ __align__(16) unsigned char tile_scratch[];
    cg::grid_group grid = cg::this_grid();
    align_device(a, grid, reinterpret_cast<TileScratch*>(tile_scratch));
}

// enqueue align on the context's stream; every pointer is a device pointer
int align_enqueue(lo_grid* kf, const PointXYZ* d_queries, uint32_t m, const uint32_t* d_m, const float* d_guess7, float* d_out7,
                  lo_align_stats* d_stats) {
    lo_ctx* ctx = kf->ctx;
    if (ctx->align_blocks == 0) {
        int per_sm = 0;
        LO_CUDA(cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kAlignTileScratchBytes)));
        LO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, align_kernel, kAlignThreads, kAlignTileScratchBytes));
        if (per_sm < 1) {
            set_error("align kernel does not fit on an SM");
            return LO_E_CUDA;
        }
        ctx->align_blocks = per_sm * ctx->sm_count;
    }
    // do not spin up more blocks than there is work for: one warp per query in the search phase
    int blocks = ctx->align_blocks;
    const int wanted = static_cast<int>((static_cast<uint64_t>(m) * 32 + kAlignThreads - 1) / kAlignThreads);
    blocks = std::max(1, std::min(blocks, wanted));
    LO_TRY(ctx->tmp_a.ensure(std::max<size_t>(m, 1) * sizeof(float4)));
    LO_TRY(ctx->tmp_b.ensure(std::max<size_t>(m, 1) * sizeof(float4)));
    LO_TRY(ctx->tmp_c.ensure(static_cast<size_t>(2) * ctx->align_blocks * 32 * sizeof(double)));
    AlignArgs args;
    args.g = kf->view();
    args.queries = d_queries;
    args.m = m;
    args.m_dev = d_m;
    args.pair_o = ctx->tmp_a.as<float4>();
    args.pair_n = ctx->tmp_b.as<float4>();
    args.partials = ctx->tmp_c.as<double>();
    args.guess = d_guess7;
    args.out_pose = d_out7;
    args.stats = d_stats;
    args.max_dist = 0.3f;  // cloud_matcher.cpp:139
    static const bool want_trace = getenv("LO_ALIGN_TRACE") != nullptr;
    args.trace = nullptr;
    if (want_trace) {
        LO_TRY(ctx->tmp_d.ensure(1001 * sizeof(long long)));
        args.trace = ctx->tmp_d.as<long long>();
    }
    void* kargs[] = {&args};
    LO_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(align_kernel), dim3(blocks), dim3(kAlignThreads), kargs, kAlignTileScratchBytes, ctx->stream));
    ctx->launches++;
    return LO_OK;
}

}  // namespace lo

using namespace lo;

static int stage_queries(lo_ctx* ctx, const float* xyz16, size_t m) {
    LO_TRY(ctx->in_a.ensure(std::max<size_t>(m, 1) * sizeof(PointXYZ)));
    if (m) LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, xyz16, m * sizeof(PointXYZ), cudaMemcpyHostToDevice, ctx->stream));
    return LO_OK;
}

extern "C" int lo_grid_find_matching_pairs(lo_grid* g, const float* xyz16, size_t m, const float pose7[7], float max_dist,
                                           lo_pair* pairs, uint8_t* valid, size_t* n_valid) {
    if (!g || (!xyz16 && m) || !pose7 || !pairs || !valid) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (n_valid) *n_valid = 0;
    if (m == 0) return LO_OK;
    LO_TRY(stage_queries(ctx, xyz16, m));
    LO_TRY(ctx->out_a.ensure(m * sizeof(lo_pair)));
    LO_TRY(ctx->out_b.ensure(m));
    uint32_t* d_n = ctx->small.as<uint32_t>();
    LO_CUDA(cudaMemsetAsync(d_n, 0, sizeof(uint32_t), ctx->stream));
    const QueryTransform tf = make_query_transform(pose_from7(pose7));
    const float max_d2 = max_dist * max_dist;
    find_pairs_kernel<<<blocks_for(m * 32, 256), 256, 0, ctx->stream>>>(g->view(), ctx->in_a.as<PointXYZ>(), static_cast<uint32_t>(m), tf,
                                                                        max_d2, ctx->out_a.as<lo_pair>(), ctx->out_b.as<uint8_t>(), d_n);
    LO_LAUNCHED(ctx);
    uint32_t h_n = 0;
    LO_CUDA(cudaMemcpyAsync(pairs, ctx->out_a.p, m * sizeof(lo_pair), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(valid, ctx->out_b.p, m, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(&h_n, d_n, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (n_valid) *n_valid = h_n;
    return LO_OK;
}

// VoxelGrid::getCorrespondence (src/voxel_grid.h:164-204): one query, already in the grid's frame. The reference compares the
// float squared distance, widened to double, with the double bound: d2 < T. For a float d2 that is d2 < (T rounded UP to
// float), which is what the kernel's float comparison gets.
extern "C" int lo_grid_get_correspondence(lo_grid* g, const float query3[3], double max_correspondence_distance_sq, lo_pair* out, int* valid) {
    if (!g || !query3 || !out || !valid) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    float bound = static_cast<float>(max_correspondence_distance_sq);
    if (static_cast<double>(bound) < max_correspondence_distance_sq) bound = nextafterf(bound, INFINITY);
    if (!(max_correspondence_distance_sq > 0.0)) bound = 0.f;   // nothing is nearer than a non-positive (or NaN) bound
    const float q4[4] = {query3[0], query3[1], query3[2], 1.f};
    LO_TRY(stage_queries(ctx, q4, 1));
    LO_TRY(ctx->out_a.ensure(sizeof(lo_pair)));
    LO_TRY(ctx->out_b.ensure(16));
    uint32_t* d_n = ctx->small.as<uint32_t>();
    LO_CUDA(cudaMemsetAsync(d_n, 0, sizeof(uint32_t), ctx->stream));
    QueryTransform tf;   // identity, exact: 1*x + 0*y + 0*z + 0
    for (int k = 0; k < 9; k++) tf.r[k] = (k % 4 == 0) ? 1.0 : 0.0;
    tf.t[0] = tf.t[1] = tf.t[2] = 0.0;
    find_pairs_kernel<<<1, 32, 0, ctx->stream>>>(g->view(), ctx->in_a.as<PointXYZ>(), 1u, tf, bound, ctx->out_a.as<lo_pair>(),
                                                 ctx->out_b.as<uint8_t>(), d_n);
    LO_LAUNCHED(ctx);
    uint8_t v = 0;
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, sizeof(lo_pair), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(&v, ctx->out_b.p, 1, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    *valid = v ? 1 : 0;
    // Correspondence::source_point_local is left default-constructed by getCorrespondence (findMatchingPairs fills it, :225)
    out->source_point_local[0] = out->source_point_local[1] = out->source_point_local[2] = 0.0;
    return LO_OK;
}

extern "C" int lo_nn_residual_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], const double x7d[7],
                                  const double prior_b[3], float max_dist, lo_normal_eq* out) {
    if (!g || (!d_xyz16 && m) || !pose7 || !x7d || !prior_b) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const QueryTransform tf = make_query_transform(pose_from7(pose7));
    const EvalPoint ep = make_eval_point(x7d);
    const uint32_t tiles = static_cast<uint32_t>((m + 31) / 32);
    // at most one resident wave (4 blocks of 8 warps per SM); each warp takes an equal share of the queries
    const int max_blocks = ctx->sm_count * kNnBlocksPerSm;
    const int blocks = std::max(1, std::min<int>(max_blocks, static_cast<int>((tiles + kNnThreads / 32 - 1) / (kNnThreads / 32))));
    LO_TRY(ctx->tmp_c.ensure(static_cast<size_t>(blocks) * 32 * sizeof(double)));
    lo_normal_eq* d_out = reinterpret_cast<lo_normal_eq*>(ctx->small.as<char>() + 1024);
    // LO_B200_NN_EXHAUSTIVE=1: visit all 27 neighbours of every query (lo_normal_eq::candidates is then the reference's count)
    const bool exhaustive = g->width > kTileMaxWidth || getenv("LO_B200_NN_EXHAUSTIVE") != nullptr;
    if (exhaustive)
        nn_residual_kernel<<<blocks, kNnThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16), static_cast<uint32_t>(m),
                                                                   tf, ep, max_dist * max_dist, ctx->tmp_c.as<double>());
    else
        nn_residual_tile_kernel<<<blocks, kNnThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16),
                                                                        static_cast<uint32_t>(m), tf, ep, max_dist * max_dist,
                                                                        ctx->tmp_c.as<double>());
    LO_LAUNCHED(ctx);
    normal_eq_finalize_kernel<<<1, 256, 0, ctx->stream>>>(ctx->tmp_c.as<double>(), blocks, ep, prior_b[0], prior_b[1], prior_b[2], d_out);
    LO_LAUNCHED(ctx);
    if (out) {
        LO_CUDA(cudaMemcpyAsync(out, d_out, sizeof(lo_normal_eq), cudaMemcpyDeviceToHost, ctx->stream));
        LO_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return LO_OK;
}

// ---- the fused NN + residual pass for callers that must not allocate (CUDA graph capture, serving loops) -------------
// No counterpart in the reference. Same kernels as lo_nn_residual_dev, but
// the per-block partial sums live in a file-scope device array instead of the context's growable scratch, and the result
// goes to a device block the caller owns, so the call is launches only: no cudaMalloc / cudaFree, no copy, no
// synchronisation. The array exists once per device for the life of the process (256 KiB); every call uses it, so two
// calls that overlap on the GPU (different streams / contexts of one device) race. Nothing in it has to start at zero:
// every block overwrites its whole row before the finalize kernel reads it, so a call never sees the previous call's sums.
constexpr int kStaticNnMaxBlocks = 1024;                          // rows of 32 doubles
static __device__ __align__(256) double g_static_nn_partials[kStaticNnMaxBlocks * 32];

// In host code `g_static_nn_partials` is a host-side stand-in, not a device address: the address is asked for once per
// device, with that device current, and handed to the kernels as an ordinary pointer.
static int static_nn_partials(int device, double** out) {
    constexpr int kMaxDevices = 64;
    static double* ptr[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) {
        set_error("lo_nn_residual_static_dev: device %d out of range (at most %d devices)", device, kMaxDevices);
        return LO_E_ARG;
    }
    if (!ptr[device]) {
        void* p = nullptr;
        LO_CUDA(cudaGetSymbolAddress(&p, g_static_nn_partials));   // the caller made `device` current
        ptr[device] = static_cast<double*>(p);
    }
    *out = ptr[device];
    return LO_OK;
}

extern "C" int lo_nn_residual_static_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], const double x7d[7],
                                         const double prior_b[3], float max_dist, lo_normal_eq* d_out) {
    if (!g || (!d_xyz16 && m) || !pose7 || !x7d || !prior_b || !d_out) return LO_E_ARG;
    if (m > LO_NN_STATIC_MAX_QUERIES) {
        set_error("lo_nn_residual_static_dev takes at most %u queries per call (got %zu): its scratch is a fixed device array; "
                  "split the batch or use lo_nn_residual_dev", static_cast<unsigned>(LO_NN_STATIC_MAX_QUERIES), m);
        return LO_E_CAPACITY;
    }
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    double* partials = nullptr;
    LO_TRY(static_nn_partials(ctx->device, &partials));
    const QueryTransform tf = make_query_transform(pose_from7(pose7));
    const EvalPoint ep = make_eval_point(x7d);
    const uint32_t tiles = static_cast<uint32_t>((m + 31) / 32);
    const int max_blocks = std::min(ctx->sm_count * kNnBlocksPerSm, kStaticNnMaxBlocks);   // the same launch shape as lo_nn_residual_dev on any GPU of <= 256 SMs
    const int blocks = std::max(1, std::min<int>(max_blocks, static_cast<int>((tiles + kNnThreads / 32 - 1) / (kNnThreads / 32))));
    if (g->width > kTileMaxWidth)
        nn_residual_kernel<<<blocks, kNnThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16), static_cast<uint32_t>(m),
                                                                   tf, ep, max_dist * max_dist, partials);
    else
        nn_residual_tile_kernel<<<blocks, kNnThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16),
                                                                        static_cast<uint32_t>(m), tf, ep, max_dist * max_dist, partials);
    LO_LAUNCHED(ctx);
    normal_eq_finalize_kernel<<<1, 256, 0, ctx->stream>>>(partials, blocks, ep, prior_b[0], prior_b[1], prior_b[2], d_out);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

extern "C" int lo_nn_residual(lo_grid* g, const float* xyz16, size_t m, const float pose7[7], const double x7d[7],
                              const double prior_b[3], float max_dist, lo_normal_eq* out) {
    if (!g || (!xyz16 && m) || !out) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(stage_queries(ctx, xyz16, m));
    return lo_nn_residual_dev(g, ctx->in_a.p, m, pose7, x7d, prior_b, max_dist, out);
}

extern "C" int lo_align_dev(lo_grid* kf, const void* d_xyz16, size_t m, const float guess7[7], float out7[7], lo_align_stats* stats) {
    if (!kf || (!d_xyz16 && m) || !guess7 || !out7) return LO_E_ARG;
    lo_ctx* ctx = kf->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    float* d_guess = ctx->small.as<float>() + 16;
    float* d_out = ctx->small.as<float>() + 32;
    lo_align_stats* d_stats = reinterpret_cast<lo_align_stats*>(ctx->small.as<char>() + 4096);
    LO_CUDA(cudaMemcpyAsync(d_guess, guess7, 7 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    if (stats) LO_CUDA(cudaMemsetAsync(d_stats, 0, sizeof(lo_align_stats), ctx->stream));
    LO_TRY(align_enqueue(kf, static_cast<const PointXYZ*>(d_xyz16), static_cast<uint32_t>(m), nullptr, d_guess, d_out, stats ? d_stats : nullptr));
    LO_CUDA(cudaMemcpyAsync(out7, d_out, 7 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (stats) LO_CUDA(cudaMemcpyAsync(stats, d_stats, sizeof(lo_align_stats), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (getenv("LO_ALIGN_TRACE")) {  // developer aid: phase timeline of block 0 in SM clocks
        static long long h[1001];
        LO_CUDA(cudaMemcpy(h, ctx->tmp_d.p, sizeof(h), cudaMemcpyDeviceToHost));
        for (long long k = 1; k < h[1000]; k++) fprintf(stderr, "trace %lld %lld\n", h[2 * k], h[2 * k + 1] - h[2 * (k - 1) + 1]);
    }
    return LO_OK;
}

extern "C" int lo_align(lo_grid* kf, const float* xyz16, size_t m, const float guess7[7], float out7[7], lo_align_stats* stats) {
    if (!kf || (!xyz16 && m)) return LO_E_ARG;
    lo_ctx* ctx = kf->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(stage_queries(ctx, xyz16, m));
    return lo_align_dev(kf, ctx->in_a.p, m, guess7, out7, stats);
}

// ---------------------------------------------------------------------------------------------------
// BASELINE.json config 4 / SURVEY.md 8(e): CloudMatcher::align against a key-frame map that is partitioned by
// voxel-hash range over several GPUs. The library stays collective-free: it offers the per-rank phases, and the host
// (lidar_odometry_demo_b200/sharded.py over torch.distributed / NCCL in the tests and the bench; two ncclAllReduce calls
// for a C++ caller) places two all-reduces between them per evaluation:
//
//   lo_shard_search_dev    every rank: 27-neighbour search of the replicated queries in ITS voxels ->
//                          packed[i] = (d2 bits << 32 | probe << 24 | slot), ~0 if none          [exchange 1: MIN over u64]
//   lo_shard_evaluate_dev  the rank whose packed[i] equals the reduced value owns query i's correspondence
//                          (VoxelGrid::getCorrespondence's first strict minimum, voxel_grid.h:175-190) and adds
//                          its residual / Jacobian terms -> 32 doubles                            [exchange 2: SUM over f64]
//   lo_lm_*                ceres::Solve's trust-region step logic (csrc/lo_lm.cuh) on the host, run redundantly and
//                          deterministically by every rank on the identical reduced sums
namespace lo {

__global__ void __launch_bounds__(kShardThreads) shard_search_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m, QueryTransform tf,
                                                                      float max_d2, unsigned long long* __restrict__ packed,
                                                                      float4* __restrict__ hit) {
    shard_search_body(g, queries, m, tf, max_d2, packed, hit);
}
__global__ void __launch_bounds__(kShardThreads) shard_search_tile_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                           QueryTransform tf, float max_d2,
                                                                           unsigned long long* __restrict__ packed, float4* __restrict__ hit) {
    shard_search_tile_body(g, queries, m, tf, max_d2, packed, hit);
}
__global__ void __launch_bounds__(kShardThreads) shard_evaluate_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m, EvalPoint ep,
                                                                        const unsigned long long* __restrict__ local,
                                                                        const unsigned long long* __restrict__ global,
                                                                        const float4* __restrict__ hit, double* __restrict__ partials) {
    shard_evaluate_body(g, queries, m, ep, local, global, hit, partials);
}

__global__ void shard_sum_kernel(const double* __restrict__ partials, int nblocks, double* __restrict__ acc32) {
    __shared__ double total[32];
    __shared__ double scratch[(256 / 32) * kRow];
    reduce_partials(partials, nblocks, scratch, total);
    if (threadIdx.x < 32) acc32[threadIdx.x] = threadIdx.x < kAcc ? total[threadIdx.x] : 0.0;
}

}  // namespace lo

extern "C" int lo_grid_set_shard(lo_grid* g, uint32_t rank, uint32_t world) {
    if (!g || world == 0 || rank >= world) return LO_E_ARG;
    LO_TRY(grid_refresh_counters(g));
    if (g->h_counters[GC_VOXELS] != 0) {
        set_error("lo_grid_set_shard must be called on an empty grid");
        return LO_E_ARG;
    }
    g->shard_rank = rank;
    g->shard_world = world;
    return LO_OK;
}

extern "C" int lo_shard_search_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], float max_dist, uint64_t* d_packed,
                                   void* d_hit16) {
    if (!g || (!d_xyz16 && m) || !pose7 || (!d_packed && m) || (!d_hit16 && m)) return LO_E_ARG;
    if (m == 0) return LO_OK;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const QueryTransform tf = make_query_transform(pose_from7(pose7));
    if (g->width <= kTileMaxWidth) {
        const int blocks = std::max(1, std::min<int>(ctx->sm_count * 4, static_cast<int>((m + kShardThreads - 1) / kShardThreads)));
        shard_search_tile_kernel<<<blocks, kShardThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16),
                                                                            static_cast<uint32_t>(m), tf, max_dist * max_dist,
                                                                            reinterpret_cast<unsigned long long*>(d_packed),
                                                                            static_cast<float4*>(d_hit16));
    } else {
        const int blocks = std::max(1, std::min<int>(ctx->sm_count * 8, static_cast<int>((m * 32 + kShardThreads - 1) / kShardThreads)));
        shard_search_kernel<<<blocks, kShardThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16), static_cast<uint32_t>(m),
                                                                       tf, max_dist * max_dist, reinterpret_cast<unsigned long long*>(d_packed),
                                                                       static_cast<float4*>(d_hit16));
    }
    LO_LAUNCHED(ctx);
    return LO_OK;
}

extern "C" int lo_shard_evaluate_dev(lo_grid* g, const void* d_xyz16, size_t m, const uint64_t* d_packed_local, const uint64_t* d_packed_global,
                                     const void* d_hit16, const double x7d[7], double* d_acc32) {
    if (!g || !x7d || !d_acc32 || (m && (!d_xyz16 || !d_packed_local || !d_packed_global || !d_hit16))) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const EvalPoint ep = make_eval_point(x7d);
    const int blocks = std::max(1, std::min<int>(ctx->sm_count * 2, static_cast<int>((m + kShardThreads - 1) / kShardThreads)));
    LO_TRY(ctx->tmp_c.ensure(static_cast<size_t>(blocks) * 32 * sizeof(double)));
    shard_evaluate_kernel<<<blocks, kShardThreads, 0, ctx->stream>>>(g->view(), static_cast<const PointXYZ*>(d_xyz16), static_cast<uint32_t>(m), ep,
                                                                     reinterpret_cast<const unsigned long long*>(d_packed_local),
                                                                     reinterpret_cast<const unsigned long long*>(d_packed_global),
                                                                     static_cast<const float4*>(d_hit16), ctx->tmp_c.as<double>());
    LO_LAUNCHED(ctx);
    shard_sum_kernel<<<1, 256, 0, ctx->stream>>>(ctx->tmp_c.as<double>(), blocks, d_acc32);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// ---- ceres::Solve's step logic on the host (the same lo_lm.cuh the device kernels use) -------------------------
struct lo_lm {
    lo::LmState s;
    double prior_b[3];
};
extern "C" lo_lm* lo_lm_create(void) { return new lo_lm(); }
extern "C" void lo_lm_destroy(lo_lm* h) { delete h; }
// acc32: the all-reduced sums of lo_shard_evaluate_dev at x0 (21 upper JtJ, 6 Jtr, cost, count); the NormalPrior of
// cloud_matcher.cpp:153-154 is added here, once. Returns 1 if the solve continues.
extern "C" int lo_lm_begin(lo_lm* h, const double x0[7], const double acc32[32], const double prior_b[3]) {
    if (!h || !x0 || !acc32 || !prior_b) return LO_E_ARG;
    double acc[kAcc];
    for (int k = 0; k < kAcc; k++) acc[k] = acc32[k];
    for (int k = 0; k < 3; k++) h->prior_b[k] = prior_b[k];
    add_prior(acc, x0 + 4, prior_b);
    return lm_begin(h->s, x0, acc) ? 1 : 0;
}
// returns 1 and writes the candidate (7 doubles) when a cost/Jacobian evaluation at cand is needed, 0 when the solve ended
extern "C" int lo_lm_propose(lo_lm* h, double cand7[7]) {
    if (!h || !cand7) return LO_E_ARG;
    if (!lm_propose(h->s)) return 0;
    for (int k = 0; k < 7; k++) cand7[k] = h->s.cand[k];
    return 1;
}
extern "C" int lo_lm_update(lo_lm* h, const double acc32[32]) {
    if (!h || !acc32) return LO_E_ARG;
    double acc[kAcc];
    for (int k = 0; k < kAcc; k++) acc[k] = acc32[k];
    add_prior(acc, h->s.cand + 4, h->prior_b);
    return lm_update(h->s, acc) ? 1 : 0;
}
extern "C" int lo_lm_result(const lo_lm* h, double x7[7], double* last_step_norm, double* cost, int32_t* termination, int32_t* recorded) {
    if (!h || !x7) return LO_E_ARG;
    for (int k = 0; k < 7; k++) x7[k] = h->s.x[k];
    if (last_step_norm) *last_step_norm = h->s.last_step_norm;
    if (cost) *cost = h->s.x_cost;
    if (termination) *termination = h->s.termination;
    if (recorded) *recorded = h->s.recorded;
    return LO_OK;
}
