// CloudMatcher::align (src/cloud_matcher.cpp:105-178) against a key-frame map sharded by voxel-hash range over several
// GPUs (BASELINE.json config 4), with the whole solve resident on the devices: one process per GPU, the
// host only enqueues.
//
// Per outer iteration every rank runs, on its context's stream:
//   search      27-neighbour search of the replicated queries in ITS voxels at the current pose -> packed (d2, probe, slot)
//   exchange 1  ncclAllReduce(MIN, u64[m]) over NVLink: the globally nearest stored point of every query, with
//               VoxelGrid::getCorrespondence's first-strict-minimum tie-break (voxel_grid.h:175-190) in the low word
//   evaluate    the owner of each winner adds its residual / Jacobian terms (PointToPlaneErrorAnalytic + Huber corrector)
//   step        ONE single-block kernel: fixed-order sum of the per-block partials -> 32 doubles, exchange 2 (all-to-all
//               of those 32 doubles through peer memory: every rank stores its sums and then a sequence flag into a
//               mailbox on each peer, waits for the peers' flags in its own mailbox and adds the contributions in rank
//               order, so all ranks hold bit-identical normal equations), then ceres::Solve's trust-region step logic
//               (lo_lm.cuh) redundantly on every rank: next candidate, or the end of the solve with the float round trip
//               of the pose and the outer stop rule (cloud_matcher.cpp:161-172).
//   up to 4 x (evaluate at the candidate -> step), predicated on the control block: once the solve or the align has
//   ended the remaining kernels of the fixed schedule return at once, on every rank alike.
// The host enqueues five outer iterations (the stop rule needs outer > 3) and then reads the `done` word once per
// further iteration: one synchronisation per align in the usual case, against 18 with a host-side LM.
// LO_SHARDED_SUM=nccl (or peer memory that cannot be mapped) replaces the mailbox exchange by ncclAllReduce(SUM, f64[32]).
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy already loaded by the process, e.g. torch's, else the
// system's), so liblo_b200.so itself has no NCCL dependency and single-GPU users never load it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "lo_shard.cuh"

namespace lo {

constexpr int kEvalBlocksPerSm = 4;   // evaluate passes are latency-bound gathers: one or two queries per thread
constexpr int kStepThreads = 256;     // (512 threads and half the partials measured no better: the sums are L2-resident)

struct ShardMail {          // one sender's contribution to one exchange
    double v[32];
    uint32_t flag;          // sequence number of the exchange the values belong to
    uint32_t pad[15];
};
struct ShardMailbox {       // lives on every rank; peers write into it through their mapping of this allocation
    ShardMail slot[2][kMaxShardWorld];
};

// One device allocation per rank, which the peers map (CUDA IPC between processes) or simply point at (shards of one
// process): the mailbox of exchange 2, the arrival flags of the cross-rank barrier, and the two best[] buffers of the fused
// exchange 1 (max_q packed keys each; only this rank's home range of each is used).
constexpr size_t kSharedBarOffset = (sizeof(ShardMailbox) + 255) & ~static_cast<size_t>(255);
constexpr size_t kSharedBestOffset = kSharedBarOffset + 256;   // uint32_t bar[2][kMaxShardWorld] = 128 B
static_assert(2 * kMaxShardWorld * sizeof(uint32_t) <= 256, "barrier flags fit their slot");

struct ShardPeers {
    ShardMailbox* box[kMaxShardWorld];          // box[r] = rank r's mailbox as seen from this device (own one included)
    uint32_t* bar[kMaxShardWorld];              // bar[r][(seq & 1) * kMaxShardWorld + sender] = seq
    unsigned long long* best[kMaxShardWorld];   // best[r][buf * max_q + i]
    uint32_t rank, world, max_q, pad;
};
__host__ __device__ inline void shard_peer_pointers(ShardPeers& p, int r, void* base) {
    p.box[r] = static_cast<ShardMailbox*>(base);
    p.bar[r] = reinterpret_cast<uint32_t*>(static_cast<char*>(base) + kSharedBarOffset);
    p.best[r] = reinterpret_cast<unsigned long long*>(static_cast<char*>(base) + kSharedBestOffset);
}

struct ShardCtl {
    LmState lm;
    EvalPoint ep;           // evaluation point of the next evaluate pass
    QueryTransform tf;      // of the current float pose
    double prior_b[3];
    double x0[7];
    float pose[7];
    float out_pose[7];
    uint32_t done, active, outer, lm_total;
    uint32_t seq;           // exchanges performed so far
    uint32_t error;         // 1: a peer did not arrive
    uint32_t evaluations;
    uint32_t bar_seq;       // cross-rank barriers passed so far
    lo_align_stats stats;
};

__device__ inline void shard_prepare_outer(ShardCtl* c) {
    Pose pf;
    pf.t[0] = c->pose[0]; pf.t[1] = c->pose[1]; pf.t[2] = c->pose[2];
    pf.w = c->pose[3]; pf.x = c->pose[4]; pf.y = c->pose[5]; pf.z = c->pose[6];
    c->tf = make_query_transform(pf);
    const double x0[7] = {pf.w, pf.x, pf.y, pf.z, pf.t[0], pf.t[1], pf.t[2]};   // cloud_matcher.cpp:122-131
    for (int k = 0; k < 7; k++) c->x0[k] = x0[k];
    c->ep = make_eval_point(x0);
}

__global__ void shard_ctl_init_kernel(ShardCtl* c, const float* __restrict__ guess7) {
    for (int k = 0; k < 7; k++) c->pose[k] = guess7[k];
    for (int k = 0; k < 3; k++) c->prior_b[k] = static_cast<double>(guess7[k]);   // NormalPrior mean, cloud_matcher.cpp:153
    c->done = 0;
    c->active = 0;
    c->outer = 0;
    c->lm_total = 0;
    c->evaluations = 0;
    memset(&c->stats, 0, sizeof(c->stats));
    shard_prepare_outer(c);
}

__global__ void __launch_bounds__(kShardThreads) shard_search_ctl_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                          const ShardCtl* __restrict__ c, float max_d2, int tiled,
                                                                          unsigned long long* __restrict__ packed, float4* __restrict__ hit) {
    if (c->done) return;
    const QueryTransform tf = c->tf;
    if (tiled) shard_search_tile_body(g, queries, m, tf, max_d2, packed, hit);
    else shard_search_body(g, queries, m, tf, max_d2, packed, hit);
}

// phase 0: evaluation at x0 right after the search; phase 1: evaluation at an LM candidate (only while the solve is active)
__global__ void __launch_bounds__(kShardThreads) shard_evaluate_ctl_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                            const ShardCtl* __restrict__ c, int phase,
                                                                            const unsigned long long* __restrict__ local,
                                                                            const unsigned long long* __restrict__ global,
                                                                            const float4* __restrict__ hit, double* __restrict__ partials) {
    if (c->done || (phase == 1 && !c->active)) return;
    const EvalPoint ep = c->ep;
    shard_evaluate_body(g, queries, m, ep, local, global, hit, partials);
}

// ---- exchange 1 without NCCL: search + atomicMin into the home rank's best[] | barrier | pull ------------------------------
__global__ void __launch_bounds__(kShardThreads) shard_search_peer_kernel(GridView g, const PointXYZ* __restrict__ queries, uint32_t m,
                                                                           const ShardCtl* __restrict__ c, float max_d2, int tiled,
                                                                           unsigned long long* __restrict__ packed, float4* __restrict__ hit, PeerMin pm) {
    if (c->done) return;
    const QueryTransform tf = c->tf;
    if (tiled) shard_search_tile_body(g, queries, m, tf, max_d2, packed, hit, &pm);
    else shard_search_body(g, queries, m, tf, max_d2, packed, hit, &pm);
}

// Cross-rank barrier, one block: my arrival into every peer's flag array, then wait for theirs in mine. Everything this
// rank's earlier kernels wrote through peer memory (the atomicMins of the search) is out before the flags go. Two flag
// sets alternate: a peer can be one barrier ahead of me, never two (it needs my next arrival for that).
__global__ void shard_barrier_kernel(ShardCtl* c, ShardPeers peers, int always, long long timeout_cycles) {
    __shared__ uint32_t seq_s;
    __shared__ int late_s;
    if (!always && c->done) return;   // identical on every rank
    if (threadIdx.x == 0) {
        seq_s = c->bar_seq + 1;
        c->bar_seq = seq_s;
        late_s = 0;
    }
    __syncthreads();
    const uint32_t seq = seq_s, set = (seq & 1u) * kMaxShardWorld, me = peers.rank;
    __threadfence_system();
    if (threadIdx.x < peers.world && threadIdx.x != me) {
        *reinterpret_cast<volatile uint32_t*>(peers.bar[threadIdx.x] + set + me) = seq;
        const volatile uint32_t* flag = peers.bar[me] + set + threadIdx.x;
        const long long t0 = clock64();
        while (*flag != seq) {
            if (clock64() - t0 > timeout_cycles) {
                late_s = 1;
                break;
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (late_s && threadIdx.x == 0) {   // a peer is gone: stop this rank's schedule instead of hanging the stream
        c->error = 1;
        c->done = 1;
    }
}

// After the barrier: the winners of all queries from their home ranks' best[] (coalesced reads over NVLink, m x 8 B in
// all) into this rank's `global` array, which the evaluate passes compare with `local` exactly as after the NCCL MIN;
// and this rank's home range of the OTHER buffer back to "none" for the next outer iteration (nobody touches that buffer
// before the step kernel's exchange of this iteration, which every rank enters after this kernel).
__global__ void shard_pull_kernel(const ShardCtl* __restrict__ c, ShardPeers peers, uint32_t m, uint32_t chunk, uint32_t buf,
                                  unsigned long long* __restrict__ global) {
    if (c->done) return;
    const size_t cur = static_cast<size_t>(buf) * peers.max_q, nxt = static_cast<size_t>(buf ^ 1u) * peers.max_q;
    const uint32_t lo_i = peers.rank * chunk, hi_i = min(m, lo_i + chunk);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        global[i] = *reinterpret_cast<const volatile unsigned long long*>(peers.best[i / chunk] + cur + i);
        if (i >= lo_i && i < hi_i) peers.best[peers.rank][nxt + i] = ~0ull;
    }
}

// start of an align: both buffers of this rank's home range to "none" (followed by a barrier before anybody searches)
__global__ void shard_best_reset_kernel(ShardPeers peers, uint32_t m) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        peers.best[peers.rank][i] = ~0ull;
        peers.best[peers.rank][static_cast<size_t>(peers.max_q) + i] = ~0ull;
    }
}

// end of one ceres::Solve: float round trip of the pose (cloud_matcher.cpp:161-167), outer stop rule (:169), and either
// the final normalisation (:175) or the set-up of the next outer iteration
__device__ inline void shard_finish_outer(ShardCtl* c) {
    const LmState& lm = c->lm;
    c->pose[3] = static_cast<float>(lm.x[0]); c->pose[4] = static_cast<float>(lm.x[1]);
    c->pose[5] = static_cast<float>(lm.x[2]); c->pose[6] = static_cast<float>(lm.x[3]);
    c->pose[0] = static_cast<float>(lm.x[4]); c->pose[1] = static_cast<float>(lm.x[5]); c->pose[2] = static_cast<float>(lm.x[6]);
    c->lm_total += lm.recorded;
    const uint32_t outer = c->outer;
    for (int k = 0; k < 7; k++) c->stats.pose_log[outer][k] = c->pose[k];
    c->stats.outer_iterations = static_cast<int32_t>(outer + 1);
    c->stats.lm_iterations = static_cast<int32_t>(c->lm_total);
    c->stats.termination = lm.termination;
    c->stats.final_cost = lm.x_cost;
    c->stats.last_step_norm = lm.last_step_norm;
    c->outer = outer + 1;
    if ((lm.last_step_norm < 1e-4 && outer > 3) || outer + 1 >= 35) {
        const float w = c->pose[3], x = c->pose[4], y = c->pose[5], z = c->pose[6];
        const float nrm = sqrtf((x * x + z * z) + (y * y + w * w));
        c->out_pose[0] = c->pose[0]; c->out_pose[1] = c->pose[1]; c->out_pose[2] = c->pose[2];
        c->out_pose[3] = w / nrm; c->out_pose[4] = x / nrm; c->out_pose[5] = y / nrm; c->out_pose[6] = z / nrm;
        c->done = 1;
    } else {
        shard_prepare_outer(c);
    }
}

// sum of the partials + exchange 2 + LM step logic, one block. kStepLocal: the sums in `partials` are already global
// (one shard, or ncclAllReduce ran before this kernel).
enum { kStepLocal = 0, kStepExchange = 1, kStepSumOnly = 2 };
__global__ void __launch_bounds__(kStepThreads) shard_step_kernel(const double* __restrict__ partials, int nblocks, ShardCtl* c, ShardPeers peers, int exchange,
                                                         int phase, long long timeout_cycles, double* __restrict__ sums_out) {
    __shared__ double total[32];
    __shared__ double scratch[(kStepThreads / 32) * kRow];
    __shared__ uint32_t seq_s;
    __shared__ int late_s;
    if (c->done || (phase == 1 && !c->active)) return;   // identical on every rank: nobody waits for a rank that skips
    reduce_partials(partials, nblocks, scratch, total);
    if (threadIdx.x >= kAcc && threadIdx.x < 32) total[threadIdx.x] = 0.0;
    __syncthreads();
    if (exchange == kStepSumOnly) {   // first half of the NCCL variant: this rank's sums for ncclAllReduce
        if (threadIdx.x < 32) sums_out[threadIdx.x] = total[threadIdx.x];
        return;
    }
    if (exchange == kStepExchange && peers.world > 1) {
        if (threadIdx.x == 0) {
            seq_s = c->seq + 1;
            c->seq = seq_s;
            late_s = 0;
        }
        __syncthreads();
        const uint32_t seq = seq_s, slot = seq & 1u, me = peers.rank;
        // my 32 sums into every peer's mailbox (one thread per value and peer), then the flags
        for (uint32_t e = threadIdx.x; e < 32 * peers.world; e += blockDim.x) {
            const uint32_t r = e >> 5, k = e & 31u;
            if (r != me) peers.box[r]->slot[slot][me].v[k] = total[k];
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x < peers.world && threadIdx.x != me) {
            *reinterpret_cast<volatile uint32_t*>(&peers.box[threadIdx.x]->slot[slot][me].flag) = seq;
            // wait for that peer's contribution to arrive in MY mailbox
            const volatile uint32_t* flag = &peers.box[me]->slot[slot][threadIdx.x].flag;
            const long long t0 = clock64();
            while (*flag != seq) {
                if (clock64() - t0 > timeout_cycles) {
                    late_s = 1;
                    break;
                }
            }
        }
        __threadfence_system();
        __syncthreads();
        if (late_s) {   // a peer is gone: stop this rank's schedule instead of hanging the stream
            if (threadIdx.x == 0) {
                c->error = 1;
                c->done = 1;
            }
            return;
        }
        if (threadIdx.x < 32) {   // contributions in rank order: the same sum, bit for bit, on every rank
            double v = 0.0;
            for (uint32_t r = 0; r < peers.world; r++)
                v += (r == me) ? total[threadIdx.x] : *reinterpret_cast<const volatile double*>(&peers.box[me]->slot[slot][r].v[threadIdx.x]);
            total[threadIdx.x] = v;
        }
        __syncthreads();
    }
    if (threadIdx.x != 0) return;
    c->evaluations++;
    bool cont;
    if (phase == 0) {
        add_prior(total, c->x0 + 4, c->prior_b);
        cont = lm_begin(c->lm, c->x0, total);
        c->stats.pairs_log[c->outer] = static_cast<int32_t>(total[28] + 0.5);
        c->stats.valid_pairs = static_cast<int32_t>(total[28] + 0.5);
    } else {
        add_prior(total, c->lm.cand + 4, c->prior_b);
        cont = lm_update(c->lm, total);
    }
    const bool need_eval = cont && lm_propose(c->lm);
    if (need_eval) {
        c->ep = make_eval_point(c->lm.cand);
        c->active = 1;
    } else {
        c->active = 0;
        shard_finish_outer(c);
    }
}

// ---- NCCL, bound at run time ---------------------------------------------------------------------------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};

static NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.lib ? &api : nullptr;
    tried = true;
    const char* names[] = {getenv("LO_B200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
    }
    if (!api.lib) {
        set_error("NCCL could not be loaded (libnccl.so.2): %s", dlerror());
        return nullptr;
    }
    bool ok = true;
    auto sym = [&](const char* name) {
        void* p = dlsym(api.lib, name);
        if (!p) ok = false;
        return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
    if (!ok) {
        set_error("libnccl.so.2 lacks an expected symbol");
        dlclose(api.lib);
        api.lib = nullptr;
        return nullptr;
    }
    return &api;
}

#define LO_NCCL(api, expr)                                                                              \
    do {                                                                                                \
        ncclResult_t _r = (expr);                                                                       \
        if (_r != ncclSuccess) {                                                                        \
            lo::set_error("%s failed: %s (%s:%d)", #expr, (api)->GetErrorString(_r), __FILE__, __LINE__); \
            return LO_E_CUDA;                                                                           \
        }                                                                                               \
    } while (0)

}  // namespace lo

using namespace lo;

struct lo_sharded {
    lo_ctx* ctx = nullptr;
    lo_grid* grid = nullptr;
    int rank = 0, world = 1;
    NcclApi* api = nullptr;
    ncclComm_t comm = nullptr;          // null for world = 1 and for the shards of a process-local group
    ShardCtl* d_ctl = nullptr;
    ShardCtl* h_ctl = nullptr;          // pinned read-back copy
    void* d_shared = nullptr;           // this rank's mailbox + barrier flags + best[] buffers (ShardPeers layout)
    uint32_t max_q = 0;                 // queries per align the best[] buffers hold
    ShardPeers peers;
    void* peer_mapped[kMaxShardWorld] = {nullptr};
    bool fused_sum = false;             // exchange 2 through peer memory inside shard_step_kernel
    bool fused_min = false;             // exchange 1 through peer memory inside the search kernel
    bool peers_mapped = false;          // every rank can reach every peer's shared block
    bool local_group = false;           // created by lo_sharded_create_group: the peers are objects of this process
    DevBuf queries, local, global, hit, partials, acc32, guess;
    long long timeout_cycles = 19000000000LL;   // ~10 s of SM clocks (LO_SHARDED_TIMEOUT_MS)
    lo_sharded_info info;
};

static void sharded_release(lo_sharded* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    for (int r = 0; r < s->world && r < kMaxShardWorld; r++)
        if (s->peer_mapped[r]) cudaIpcCloseMemHandle(s->peer_mapped[r]);
    if (s->comm && s->api) s->api->CommDestroy(s->comm);
    if (s->grid) lo_grid_destroy(s->grid);
    if (s->d_ctl) cudaFree(s->d_ctl);
    if (s->h_ctl) cudaFreeHost(s->h_ctl);
    if (s->d_shared) cudaFree(s->d_shared);
    for (DevBuf* b : {&s->queries, &s->local, &s->global, &s->hit, &s->partials, &s->acc32, &s->guess}) b->release();
    delete s;
}

extern "C" int lo_sharded_unique_id(uint8_t id[LO_SHARDED_ID_BYTES]) {
    if (!id) return LO_E_ARG;
    static_assert(sizeof(ncclUniqueId) == LO_SHARDED_ID_BYTES, "ncclUniqueId is 128 bytes");
    NcclApi* api = nccl_api();
    if (!api) return LO_E_CUDA;
    ncclUniqueId u;
    LO_NCCL(api, api->GetUniqueId(&u));
    std::memcpy(id, &u, sizeof(u));
    return LO_OK;
}

static uint32_t sharded_max_queries() {
    const char* e = getenv("LO_SHARDED_MAX_QUERIES");
    return e && atoll(e) > 0 ? static_cast<uint32_t>(std::min<long long>(atoll(e), 1ll << 28)) : (1u << 20);
}

// everything of a shard that does not depend on how its peers are reached
static int sharded_new(lo_ctx* ctx, int rank, int world, float voxel_size, size_t max_points, size_t capacity_hint, lo_sharded** out) {
    LO_CUDA(cudaSetDevice(ctx->device));
    lo_sharded* s = new lo_sharded();
    s->ctx = ctx;
    s->rank = rank;
    s->world = world;
    std::memset(&s->peers, 0, sizeof(s->peers));
    s->peers.rank = static_cast<uint32_t>(rank);
    s->peers.world = static_cast<uint32_t>(world);
    s->max_q = world > 1 ? sharded_max_queries() : 0;
    s->peers.max_q = s->max_q;
    std::memset(&s->info, 0, sizeof(s->info));
    int st = lo_grid_create(ctx, voxel_size, max_points, capacity_hint, &s->grid);
    if (st == LO_OK) st = lo_grid_set_shard(s->grid, static_cast<uint32_t>(rank), static_cast<uint32_t>(world));
    if (st != LO_OK) {
        sharded_release(s);
        return st;
    }
    const size_t shared_bytes = kSharedBestOffset + 2 * static_cast<size_t>(s->max_q) * sizeof(unsigned long long);
    if (cudaMalloc(&s->d_ctl, sizeof(ShardCtl)) != cudaSuccess || cudaMallocHost(&s->h_ctl, sizeof(ShardCtl)) != cudaSuccess ||
        cudaMalloc(&s->d_shared, shared_bytes) != cudaSuccess || cudaMemsetAsync(s->d_ctl, 0, sizeof(ShardCtl), ctx->stream) != cudaSuccess ||
        cudaMemsetAsync(s->d_shared, 0, kSharedBestOffset, ctx->stream) != cudaSuccess ||
        cudaMemsetAsync(static_cast<char*>(s->d_shared) + kSharedBestOffset, 0xFF, shared_bytes - kSharedBestOffset, ctx->stream) != cudaSuccess ||
        s->acc32.ensure(32 * sizeof(double)) != LO_OK || s->guess.ensure(8 * sizeof(float)) != LO_OK ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        set_error("lo_sharded_create: device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        sharded_release(s);
        return LO_E_CUDA;
    }
    std::memset(s->h_ctl, 0, sizeof(ShardCtl));
    shard_peer_pointers(s->peers, rank, s->d_shared);
    const char* to = getenv("LO_SHARDED_TIMEOUT_MS");
    if (to && atof(to) > 0) s->timeout_cycles = static_cast<long long>(atof(to) * 1.9e6);
    s->info.rank = rank;
    s->info.world = world;
    *out = s;
    return LO_OK;
}

// Map every peer's shared block (processes: CUDA IPC). Every rank takes part in the all-gather of the handles whatever
// happened before it on this rank (a rank that could not make its handle sends zeros and an ok flag of 0), so no rank is
// left waiting in the collective; any failure leaves the NCCL exchanges in place on all ranks (agreed right after).
static bool sharded_map_peers(lo_sharded* s) {
    lo_ctx* ctx = s->ctx;
    struct Entry {
        cudaIpcMemHandle_t handle;
        int ok;
        int pad[15];
    };
    static_assert(sizeof(Entry) == 128, "all-gather record");
    Entry mine;
    std::memset(&mine, 0, sizeof(mine));
    mine.ok = cudaIpcGetMemHandle(&mine.handle, s->d_shared) == cudaSuccess ? 1 : 0;
    DevBuf hs;
    const bool have_buf = hs.ensure(sizeof(Entry) * (s->world + 1)) == LO_OK;
    std::vector<Entry> all(s->world);
    bool ok = have_buf;
    if (have_buf) {
        char* d_all = hs.as<char>();
        char* d_mine = d_all + sizeof(Entry) * s->world;
        ok = cudaMemcpyAsync(d_mine, &mine, sizeof(Entry), cudaMemcpyHostToDevice, ctx->stream) == cudaSuccess;
        // the collective itself is entered unconditionally (with whatever d_mine holds)
        ok = (s->api->AllGather(d_mine, d_all, sizeof(Entry), ncclChar, s->comm, ctx->stream) == ncclSuccess) && ok;
        ok = ok && cudaMemcpyAsync(all.data(), d_all, sizeof(Entry) * s->world, cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess &&
             cudaStreamSynchronize(ctx->stream) == cudaSuccess;
        hs.release();
    }
    int mapped_ok = ok ? 1 : 0;
    for (int r = 0; r < s->world && mapped_ok; r++) mapped_ok = all[r].ok;
    for (int r = 0; r < s->world && mapped_ok; r++) {
        if (r == s->rank) continue;
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            mapped_ok = 0;
            break;
        }
        s->peer_mapped[r] = p;
        shard_peer_pointers(s->peers, r, p);
    }
    cudaGetLastError();   // a failed IPC call must not poison later launches
    return mapped_ok == 1;
}

extern "C" int lo_sharded_create(lo_ctx* ctx, const uint8_t id[LO_SHARDED_ID_BYTES], int rank, int world, float voxel_size, size_t max_points,
                                 size_t capacity_hint, lo_sharded** out) {
    if (!ctx || !out || world < 1 || world > kMaxShardWorld || rank < 0 || rank >= world || (world > 1 && !id)) return LO_E_ARG;
    lo_sharded* s = nullptr;
    LO_TRY(sharded_new(ctx, rank, world, voxel_size, max_points, capacity_hint, &s));
    auto fail = [&](int code) {
        sharded_release(s);
        return code;
    };
    if (world > 1) {
        s->api = nccl_api();
        if (!s->api) return fail(LO_E_CUDA);
        ncclUniqueId u;
        std::memcpy(&u, id, sizeof(u));
        ncclResult_t r = s->api->CommInitRank(&s->comm, world, u, rank);
        if (r != ncclSuccess) {
            set_error("ncclCommInitRank failed: %s", s->api->GetErrorString(r));
            s->comm = nullptr;
            return fail(LO_E_CUDA);
        }
        const char* mode = getenv("LO_SHARDED_SUM");
        const bool want_fused_sum = !(mode && mode[0] == 'n');
        const char* min_mode = getenv("LO_SHARDED_MIN");
        const bool want_fused_min = !(min_mode && min_mode[0] == 'n');
        // every rank maps (the all-gather of the handles is a collective), then all must agree, or the ranks would mix
        // the exchange protocols: MIN over "I can"
        const bool mapped = sharded_map_peers(s);
        int both = mapped ? 1 : 0;
        int* d_flag = s->acc32.as<int>();
        if (cudaMemcpyAsync(d_flag, &both, sizeof(int), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess ||
            s->api->AllReduce(d_flag, d_flag, 1, ncclInt, ncclMin, s->comm, ctx->stream) != ncclSuccess ||
            cudaMemcpyAsync(&both, d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
            cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
            set_error("lo_sharded_create: the ranks could not agree on the exchange protocol");
            return fail(LO_E_CUDA);
        }
        s->peers_mapped = both == 1;
        s->fused_sum = s->peers_mapped && want_fused_sum;   // the LO_SHARDED_* switches are read by every rank of the job alike
        s->fused_min = s->peers_mapped && want_fused_min;
        int version = 0;
        s->api->GetVersion(&version);
        s->info.nccl_version = version;
    }
    s->info.fused_sum = s->fused_sum ? 1 : 0;
    s->info.fused_min = s->fused_min ? 1 : 0;
    *out = s;
    return LO_OK;
}

// The shards of ONE process: `world` contexts (each its own stream; on different GPUs of the box, or several on the same
// GPU), no NCCL at all — both exchanges run through peer memory, which for two shards on one device is plain device
// memory. The shards must be driven together (lo_sharded_align_group): a shard's kernels wait on the device for its
// peers' arrivals, so a blocking call on one shard alone would wait for work nobody has queued.
extern "C" int lo_sharded_create_group(lo_ctx* const* ctxs, int world, float voxel_size, size_t max_points, size_t capacity_hint,
                                       lo_sharded** out) {
    if (!ctxs || !out || world < 1 || world > kMaxShardWorld) return LO_E_ARG;
    for (int r = 0; r < world; r++) {
        if (!ctxs[r]) return LO_E_ARG;
        for (int q = 0; q < r; q++)
            if (ctxs[q] == ctxs[r]) {
                set_error("lo_sharded_create_group: every shard needs a context (stream) of its own");
                return LO_E_ARG;
            }
        out[r] = nullptr;
    }
    auto fail = [&](int code) {
        for (int r = 0; r < world; r++) {
            sharded_release(out[r]);
            out[r] = nullptr;
        }
        return code;
    };
    for (int r = 0; r < world; r++) {
        const int st = sharded_new(ctxs[r], r, world, voxel_size, max_points, capacity_hint, &out[r]);
        if (st != LO_OK) return fail(st);
        out[r]->local_group = true;
    }
    for (int r = 0; r < world; r++) {
        if (cudaSetDevice(ctxs[r]->device) != cudaSuccess) return fail(LO_E_CUDA);
        for (int q = 0; q < world; q++) {
            if (ctxs[q]->device != ctxs[r]->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, ctxs[r]->device, ctxs[q]->device);
                const cudaError_t e = can ? cudaDeviceEnablePeerAccess(ctxs[q]->device, 0) : cudaErrorPeerAccessUnsupported;
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                    set_error("lo_sharded_create_group: device %d cannot access device %d: %s", ctxs[r]->device, ctxs[q]->device, cudaGetErrorString(e));
                    cudaGetLastError();
                    return fail(LO_E_CUDA);
                }
                cudaGetLastError();
            }
            shard_peer_pointers(out[r]->peers, q, out[q]->d_shared);
        }
        out[r]->peers_mapped = true;
        out[r]->fused_sum = out[r]->fused_min = world > 1;
        out[r]->info.fused_sum = out[r]->info.fused_min = world > 1 ? 1 : 0;
    }
    return LO_OK;
}

extern "C" void lo_sharded_destroy(lo_sharded* s) { sharded_release(s); }

extern "C" lo_grid* lo_sharded_grid(lo_sharded* s) { return s ? s->grid : nullptr; }

extern "C" int lo_sharded_set_fused_sum(lo_sharded* s, int enable) {
    if (!s) return LO_E_ARG;
    if (!enable && s->local_group && s->world > 1) {
        set_error("the shards of a process-local group have no NCCL communicator: the peer-memory exchanges stay in place");
        return LO_E_ARG;
    }
    if (enable && s->world > 1 && !s->peers_mapped) {
        set_error("the peers' mailboxes could not be mapped on this job (no peer access / IPC): the NCCL sum stays in place");
        return LO_E_ARG;
    }
    s->fused_sum = enable && s->world > 1;
    s->info.fused_sum = s->fused_sum ? 1 : 0;
    return LO_OK;
}

extern "C" int lo_sharded_set_fused_min(lo_sharded* s, int enable) {
    if (!s) return LO_E_ARG;
    if (!enable && s->local_group && s->world > 1) {
        set_error("the shards of a process-local group have no NCCL communicator: the peer-memory exchanges stay in place");
        return LO_E_ARG;
    }
    if (enable && s->world > 1 && !s->peers_mapped) {
        set_error("the peers' memory could not be mapped on this job (no peer access / IPC): the NCCL MIN stays in place");
        return LO_E_ARG;
    }
    s->fused_min = enable && s->world > 1;
    s->info.fused_min = s->fused_min ? 1 : 0;
    return LO_OK;
}

extern "C" int lo_sharded_info_get(const lo_sharded* s, lo_sharded_info* out) {
    if (!s || !out) return LO_E_ARG;
    *out = s->info;
    return LO_OK;
}

// One outer iteration is a fixed schedule of kOuterParts steps per shard. A caller that drives several shards issues it
// BREADTH-FIRST — step p for every shard before step p + 1 for any: a kernel that waits for a peer on the device (barrier,
// step kernel) is then always followed at once by the peer's counterpart, with nothing of the first shard's own stream
// in between. Streams may share a hardware work queue (CUDA_DEVICE_MAX_CONNECTIONS), which dispatches in issue order and
// stalls at an entry whose stream predecessor is still running: depth-first issue would put shard 0's later kernels
// between its spinning barrier and shard 1's arrival, and nothing would ever move.
//   part 0          search (+ exchange 1 fused into it, or ncclAllReduce(MIN) after it)
//   part 1          cross-rank barrier              (peer-memory MIN only)
//   part 2          pull of the winners             (peer-memory MIN only)
//   part 3 + 2 k    evaluate, k = 0: at x0; k = 1..4: at the LM candidates (max_num_iterations = 4, cloud_matcher.cpp:111)
//   part 4 + 2 k    step: sum, exchange 2, LM logic
constexpr int kOuterParts = 3 + 2 * 5;
static int enqueue_outer_part(lo_sharded* s, const PointXYZ* q, uint32_t m, float max_d2, uint32_t outer, int part) {
    lo_ctx* ctx = s->ctx;
    lo_grid* g = s->grid;
    LO_CUDA(cudaSetDevice(ctx->device));
    const GridView v = g->view();
    unsigned long long* local = s->local.as<unsigned long long>();
    unsigned long long* global = s->world > 1 ? s->global.as<unsigned long long>() : local;
    float4* hit = s->hit.as<float4>();
    const int tiled = g->width <= kTileMaxWidth ? 1 : 0;
    const int sblocks = tiled ? std::max(1, std::min<int>(ctx->sm_count * 4, static_cast<int>((m + kShardThreads - 1) / kShardThreads)))
                              : std::max(1, std::min<int>(ctx->sm_count * 8, static_cast<int>((static_cast<size_t>(m) * 32 + kShardThreads - 1) / kShardThreads)));
    const int eblocks = std::max(1, std::min<int>(ctx->sm_count * kEvalBlocksPerSm, static_cast<int>((m + kShardThreads - 1) / kShardThreads)));
    double* partials = s->partials.as<double>();
    double* acc32 = s->acc32.as<double>();
    const bool peer_min = s->world > 1 && m && s->fused_min;
    const uint32_t buf = outer & 1u, chunk = m ? (m + s->world - 1) / static_cast<uint32_t>(s->world) : 1u;
    if (part == 0) {
        if (peer_min) {   // exchange 1 inside the search: atomicMin into the home ranks' best[]
            PeerMin pm;
            for (int r = 0; r < kMaxShardWorld; r++) pm.best[r] = r < s->world ? s->peers.best[r] + static_cast<size_t>(buf) * s->max_q : nullptr;
            pm.chunk = chunk;
            shard_search_peer_kernel<<<sblocks, kShardThreads, 0, ctx->stream>>>(v, q, m, s->d_ctl, max_d2, tiled, local, hit, pm);
            LO_LAUNCHED(ctx);
            s->info.min_exchanges_fused++;
        } else {
            shard_search_ctl_kernel<<<sblocks, kShardThreads, 0, ctx->stream>>>(v, q, m, s->d_ctl, max_d2, tiled, local, hit);
            LO_LAUNCHED(ctx);
            if (s->world > 1 && m) {
                LO_NCCL(s->api, s->api->AllReduce(local, global, m, ncclUint64, ncclMin, s->comm, ctx->stream));   // exchange 1
                s->info.min_allreduces++;
            }
        }
        return LO_OK;
    }
    if (part == 1) {
        if (peer_min) {
            shard_barrier_kernel<<<1, 32, 0, ctx->stream>>>(s->d_ctl, s->peers, 0, s->timeout_cycles);
            LO_LAUNCHED(ctx);
        }
        return LO_OK;
    }
    if (part == 2) {
        if (peer_min) {
            shard_pull_kernel<<<std::max(1, std::min<int>(ctx->sm_count * 2, static_cast<int>((m + 255) / 256))), 256, 0, ctx->stream>>>(
                s->d_ctl, s->peers, m, chunk, buf, global);
            LO_LAUNCHED(ctx);
        }
        return LO_OK;
    }
    const int ph = (part - 3) / 2 ? 1 : 0;   // phase 0: x0; phase 1: an LM candidate
    const int exchange = (s->world > 1 && m) ? kStepExchange : kStepLocal;   // without queries every rank's sums are zero
    if ((part - 3) % 2 == 0) {
        shard_evaluate_ctl_kernel<<<eblocks, kShardThreads, 0, ctx->stream>>>(v, q, m, s->d_ctl, ph, local, global, hit, partials);
        LO_LAUNCHED(ctx);
        return LO_OK;
    }
    if (exchange == kStepExchange && !s->fused_sum) {
        // NCCL variant: local sums -> acc32, all-reduce in place, then the step logic on the global sums
        shard_step_kernel<<<1, kStepThreads, 0, ctx->stream>>>(partials, eblocks, s->d_ctl, s->peers, kStepSumOnly, ph, s->timeout_cycles, acc32);
        LO_LAUNCHED(ctx);
        LO_NCCL(s->api, s->api->AllReduce(acc32, acc32, 32, ncclDouble, ncclSum, s->comm, ctx->stream));   // exchange 2
        s->info.sum_allreduces++;
        shard_step_kernel<<<1, kStepThreads, 0, ctx->stream>>>(acc32, 1, s->d_ctl, s->peers, kStepLocal, ph, s->timeout_cycles, nullptr);
    } else {
        shard_step_kernel<<<1, kStepThreads, 0, ctx->stream>>>(partials, eblocks, s->d_ctl, s->peers, exchange, ph, s->timeout_cycles, nullptr);
    }
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// buffers first, for every shard, before anything is enqueued: a growing buffer frees its old storage, which waits for the
// device — and a device on which a peer's barrier kernel is already spinning would never get there
static int align_reserve(lo_sharded* s, size_t m) {
    lo_ctx* ctx = s->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (s->world > 1 && s->fused_min && m > s->max_q) {
        set_error("sharded align: %zu queries exceed the %u the peer-memory MIN exchange was sized for (LO_SHARDED_MAX_QUERIES)", m, s->max_q);
        return LO_E_CAPACITY;
    }
    const size_t mr = std::max<size_t>(m, 1);
    LO_TRY(s->local.ensure(mr * 8));
    if (s->world > 1) LO_TRY(s->global.ensure(mr * 8));
    LO_TRY(s->hit.ensure(mr * 16));
    LO_TRY(s->partials.ensure(static_cast<size_t>(ctx->sm_count) * kEvalBlocksPerSm * 32 * sizeof(double)));
    return LO_OK;
}

// control block and (fused MIN) clean best[] buffers; part 1: the cross-rank barrier behind them (breadth-first, see above)
static int align_begin(lo_sharded* s, size_t m, const float guess7[7], int part) {
    lo_ctx* ctx = s->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const bool peer_min = s->world > 1 && s->fused_min && m;
    if (part == 0) {
        LO_CUDA(cudaMemcpyAsync(s->guess.p, guess7, 7 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        shard_ctl_init_kernel<<<1, 1, 0, ctx->stream>>>(s->d_ctl, s->guess.as<float>());
        LO_LAUNCHED(ctx);
        if (peer_min) {
            const uint32_t mm = static_cast<uint32_t>(m);
            shard_best_reset_kernel<<<std::max(1, std::min<int>(ctx->sm_count * 2, static_cast<int>((mm + 255) / 256))), 256, 0, ctx->stream>>>(s->peers, mm);
            LO_LAUNCHED(ctx);
        }
    } else if (peer_min) {
        shard_barrier_kernel<<<1, 32, 0, ctx->stream>>>(s->d_ctl, s->peers, 1, s->timeout_cycles);   // nobody searches before everybody has reset
        LO_LAUNCHED(ctx);
    }
    return LO_OK;
}

// CloudMatcher::align over `n` shards driven by this process (n = 1: this rank of a multi-process job, the peers are
// driven by theirs). Every shard's schedule is enqueued before anything is waited for.
static int align_many(lo_sharded* const* ss, int n, const void* const* d_xyz16, size_t m, const float guess7[7], float max_dist, float* out7s,
                      lo_align_stats* stats) {
    const uint32_t mm = static_cast<uint32_t>(m);
    const float max_d2 = max_dist * max_dist;   // voxel_grid.h:215
    for (int k = 0; k < n; k++) LO_TRY(align_reserve(ss[k], m));
    for (int part = 0; part < 2; part++)
        for (int k = 0; k < n; k++) LO_TRY(align_begin(ss[k], m, guess7, part));
    for (uint32_t outer = 0; outer < 35; outer++) {
        for (int part = 0; part < kOuterParts; part++)
            for (int k = 0; k < n; k++) LO_TRY(enqueue_outer_part(ss[k], static_cast<const PointXYZ*>(d_xyz16[k]), mm, max_d2, outer, part));
        if (outer >= 4) {   // the stop rule can first hold after the fifth solve (cloud_matcher.cpp:169)
            for (int k = 0; k < n; k++) {
                LO_CUDA(cudaSetDevice(ss[k]->ctx->device));
                LO_CUDA(cudaMemcpyAsync(ss[k]->h_ctl, ss[k]->d_ctl, sizeof(ShardCtl), cudaMemcpyDeviceToHost, ss[k]->ctx->stream));
            }
            for (int k = 0; k < n; k++) {
                LO_CUDA(cudaSetDevice(ss[k]->ctx->device));
                LO_CUDA(cudaStreamSynchronize(ss[k]->ctx->stream));
                ss[k]->info.host_syncs++;
            }
            if (ss[0]->h_ctl->done) break;   // every shard holds the same control state
        }
    }
    for (int k = 0; k < n; k++) {
        if (ss[k]->h_ctl->error) {
            const ShardCtl* c = ss[k]->h_ctl;
            set_error("sharded align: a peer rank did not arrive within the time limit (rank %d of %d: outer iteration %u, %u barriers and %u sum "
                      "exchanges entered, %u evaluations)", ss[k]->rank, ss[k]->world, c->outer, c->bar_seq, c->seq, c->evaluations);
            return LO_E_CUDA;
        }
        std::memcpy(out7s + 7 * k, ss[k]->h_ctl->out_pose, 7 * sizeof(float));
        ss[k]->info.evaluations = ss[k]->h_ctl->evaluations;
    }
    if (stats) *stats = ss[0]->h_ctl->stats;
    return LO_OK;
}

extern "C" int lo_sharded_align_dev(lo_sharded* s, const void* d_xyz16, size_t m, const float guess7[7], float max_dist, float out7[7],
                                    lo_align_stats* stats) {
    if (!s || !guess7 || !out7 || (!d_xyz16 && m) || m > 0x7FFFFFFFu) return LO_E_ARG;
    if (s->local_group && s->world > 1) {
        set_error("this shard belongs to a process-local group: drive the group with lo_sharded_align_group");
        return LO_E_ARG;
    }
    return align_many(&s, 1, &d_xyz16, m, guess7, max_dist, out7, stats);
}

extern "C" int lo_sharded_align_group(lo_sharded* const* shards, int world, const void* const* d_xyz16, size_t m, const float guess7[7],
                                      float max_dist, float* out7s, lo_align_stats* stats) {
    if (!shards || world < 1 || world > kMaxShardWorld || !guess7 || !out7s || (!d_xyz16 && m) || m > 0x7FFFFFFFu) return LO_E_ARG;
    for (int k = 0; k < world; k++)
        if (!shards[k] || shards[k]->world != world || shards[k]->rank != k || !shards[k]->local_group || (m && !d_xyz16[k])) {
            set_error("lo_sharded_align_group wants the %d shards lo_sharded_create_group made, in rank order, and one query pointer per shard", world);
            return LO_E_ARG;
        }
    return align_many(shards, world, d_xyz16, m, guess7, max_dist, out7s, stats);
}

extern "C" int lo_sharded_align(lo_sharded* s, const float* xyz16, size_t m, const float guess7[7], float max_dist, float out7[7],
                                lo_align_stats* stats) {
    if (!s || (!xyz16 && m)) return LO_E_ARG;
    lo_ctx* ctx = s->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(s->queries.ensure(std::max<size_t>(m, 1) * sizeof(PointXYZ)));
    if (m) LO_CUDA(cudaMemcpyAsync(s->queries.p, xyz16, m * sizeof(PointXYZ), cudaMemcpyHostToDevice, ctx->stream));
    return lo_sharded_align_dev(s, s->queries.p, m, guess7, max_dist, out7, stats);
}

// host-pointer form of lo_sharded_align_group: the queries are uploaded to every shard's device first
extern "C" int lo_sharded_align_group_host(lo_sharded* const* shards, int world, const float* xyz16, size_t m, const float guess7[7],
                                           float max_dist, float* out7s, lo_align_stats* stats) {
    if (!shards || world < 1 || world > kMaxShardWorld || (!xyz16 && m)) return LO_E_ARG;
    const void* d_q[kMaxShardWorld];
    for (int k = 0; k < world; k++) {
        if (!shards[k]) return LO_E_ARG;
        lo_ctx* ctx = shards[k]->ctx;
        LO_CUDA(cudaSetDevice(ctx->device));
        LO_TRY(shards[k]->queries.ensure(std::max<size_t>(m, 1) * sizeof(PointXYZ)));
        if (m) LO_CUDA(cudaMemcpyAsync(shards[k]->queries.p, xyz16, m * sizeof(PointXYZ), cudaMemcpyHostToDevice, ctx->stream));
        d_q[k] = shards[k]->queries.p;
    }
    return lo_sharded_align_group(shards, world, d_q, m, guess7, max_dist, out7s, stats);
}
