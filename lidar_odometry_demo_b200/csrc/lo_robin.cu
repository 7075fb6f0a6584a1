// Emission of a grid's content in the iteration order of the reference's tsl::robin_map
// (third-party/robin-map, robin_hash.h) — the order VoxelGrid::getCloud* walk (src/voxel_grid.h:112-162)
// and therefore the order in which LidarOdometry feeds down-sampled points onward
// (src/lidar_odometry.cpp:37-50,69-70).
//
// What defines that order for a map that only ever saw insertions of n distinct keys:
//   bucket_count C = 2 * pow2ceil(n) (growth x2 from 2, max load 0.5); ideal bucket b = hash22 & (C-1);
//   iteration = ascending bucket position; robin-hood probing keeps elements sorted by ideal bucket, so
//   the order is "by b", ties (same ideal bucket) in a history-dependent order, and a cluster that runs
//   past the last bucket wraps to the front.
// The order inside a tie group depends on the whole insertion history (an insertion into an earlier group of the same
// cluster moves the first element of every following group to that group's end, and every growth re-inserts the
// elements in the old iteration order), so this export REPLAYS the history: the voxels are sorted by the arrival ticket
// of their first point (= the order in which the reference created them) and one thread runs tsl::robin_hash's own
// insertion / rehash procedure (robin_hash.h:1219-1341,1379-1403) over 4-byte bucket records in HBM; the bucket array is
// then walked in order. Exact for any key set, wrap-around and growth included (tests/test_gpu_robin.py compares with
// the real tsl::robin_map on random and adversarial key sets). A sequential pass of ~2 n dependent look-ups: this is the
// VoxelGrid::getCloud* API path (and the stage-by-stage pipeline), not the per-scan kernel.
#include <cub/cub.cuh>

#include "lo_host.h"
#include <cstdio>
#include <cstdlib>

#include "lo_robin_grid.cuh"

namespace lo {

__global__ void robin_keys_kernel(GridView g, uint32_t capacity, const uint32_t* __restrict__ counters,
                                  unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ cursor,
                                  uint32_t* __restrict__ hash_of_slot) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= capacity) return;
    const Slot slot = g.slots[s];
    if (slot.key == kKeyEmpty || slot.key == kKeyTomb || slot.count == 0) return;
    long long ix, iy, iz;
    unpack_key(slot.key, ix, iy, iz);
    hash_of_slot[s] = reference_hash22(ix, iy, iz);
    // the voxel's first point carries the ticket it arrived with: voxels sorted by it are in creation order
    const uint32_t ticket = __float_as_uint(g.points[static_cast<size_t>(slot.vid) * g.width].w);
    const uint32_t at = atomicAdd(cursor, 1u);
    keys[at] = ticket;
    vals[at] = s;
}

// keys are sorted by arrival ticket: element e = e-th voxel the reference created. One thread replays the map.
//   idx[2][cap], dist[2][cap]: two bucket arrays (current and rehash target); dist -1 = empty
constexpr int kRobinDistLimit = 8192;   // robin_hash.h:327 DIST_FROM_IDEAL_BUCKET_LIMIT

struct RobinTable {
    int* idx;
    int* dist;
    uint32_t mask;
};

__device__ inline void robin_displace(RobinTable t, uint32_t ib, int carry_idx, int carry_dist, bool may_flag, bool& grow_flag) {
    // robin_hash.h:1283-1317: the carried element takes the bucket of any element strictly closer to its ideal bucket
    while (true) {
        const int d = t.dist[ib];
        if (d < 0) {
            t.idx[ib] = carry_idx;
            t.dist[ib] = carry_dist;
            return;
        }
        if (carry_dist > d) {
            if (may_flag && carry_dist > kRobinDistLimit) grow_flag = true;
            const int oi = t.idx[ib];
            t.idx[ib] = carry_idx;
            t.dist[ib] = carry_dist;
            carry_idx = oi;
            carry_dist = d;
        }
        ib = (ib + 1) & t.mask;
        carry_dist++;
    }
}

__global__ void robin_replay_kernel(const unsigned long long* __restrict__ keys_sorted, const uint32_t* __restrict__ hash_of_slot,
                                    const uint32_t* __restrict__ vals_sorted, const uint32_t* __restrict__ counters, int* __restrict__ idx2,
                                    int* __restrict__ dist2, uint32_t cap, uint32_t* __restrict__ order) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const uint32_t n = counters[GC_VOXELS];
    int cur = 0;
    RobinTable t{idx2, dist2, 0u};
    uint32_t buckets = 0, count = 0, threshold = 0;
    bool grow_flag = false;
    for (uint32_t e = 0; e < n; e++) {
        const uint32_t h = hash_of_slot[vals_sorted[e]];
        uint32_t ib = 0;
        int dist = 0;
        auto probe = [&]() {   // robin_hash.h:1226-1243 (the key is known to be absent: one voxel per slot)
            ib = h & t.mask;
            dist = 0;
            while (dist <= t.dist[ib]) {
                ib = (ib + 1) & t.mask;
                dist++;
            }
        };
        if (buckets) probe();
        while (grow_flag || dist > kRobinDistLimit || count >= threshold) {   // rehash_on_extreme_load, :1379-1403
            const uint32_t nb = buckets ? buckets * 2 : 2;                   // power_of_two_growth_policy<2>
            RobinTable nt{idx2 + (cur ^ 1) * static_cast<size_t>(cap), dist2 + (cur ^ 1) * static_cast<size_t>(cap), nb - 1};
            for (uint32_t k = 0; k < nb; k++) nt.dist[k] = -1;
            bool unused = false;
            for (uint32_t k = 0; k < buckets; k++) {                         // rehash_impl, :1319-1341: old buckets in order
                if (t.dist[k] < 0) continue;
                const int el = t.idx[k];
                robin_displace(nt, hash_of_slot[vals_sorted[el]] & nt.mask, el, 0, false, unused);
            }
            cur ^= 1;
            t = nt;
            buckets = nb;
            threshold = static_cast<uint32_t>(static_cast<float>(nb) * 0.5f);
            grow_flag = false;
            probe();
        }
        if (t.dist[ib] < 0) {
            t.idx[ib] = static_cast<int>(e);
            t.dist[ib] = dist;
        } else {
            robin_displace(t, ib, static_cast<int>(e), dist, true, grow_flag);
        }
        count++;
    }
    uint32_t k = 0;   // iteration = ascending bucket position (robin_hash.h begin()/++)
    for (uint32_t b = 0; b < buckets; b++)
        if (t.dist[b] >= 0) order[k++] = static_cast<uint32_t>(t.idx[b]);
}

__global__ void robin_gather_kernel(const uint32_t* __restrict__ order, const uint32_t* __restrict__ vals_sorted, GridView g,
                                    const uint32_t* __restrict__ counters, int mode, uint32_t bound, uint32_t* __restrict__ slot_rot,
                                    uint32_t* __restrict__ count_rot) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= bound) return;
    const uint32_t n = counters[GC_VOXELS];
    if (i >= n) {
        count_rot[i] = 0;
        return;
    }
    const uint32_t s = vals_sorted[order[i]];
    slot_rot[i] = s;
    count_rot[i] = (mode == 2) ? 1u : g.slots[s].count;
}

__global__ void robin_emit_kernel(GridView g, const uint32_t* __restrict__ counters, const uint32_t* __restrict__ slot_rot,
                                  const uint32_t* __restrict__ offsets, int mode, void* __restrict__ out, uint32_t* __restrict__ total) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = counters[GC_VOXELS];
    if (i >= n) return;
    const Slot slot = g.slots[slot_rot[i]];
    const uint32_t cnt = (mode == 2) ? 1u : slot.count;
    const uint32_t off = offsets[i];
    if (i == n - 1) *total = off + cnt;
    for (uint32_t e = 0; e < cnt; e++) {
        const float4 p = g.points[static_cast<size_t>(slot.vid) * g.width + e];
        if (mode == 0) {
            const float4 nv = g.normals[static_cast<size_t>(slot.vid) * g.width + e];
            float4* o = reinterpret_cast<float4*>(static_cast<PointNormal*>(out) + off + e);
            o[0] = make_float4(p.x, p.y, p.z, 1.f);
            o[1] = make_float4(nv.x, nv.y, nv.z, 0.f);
            o[2] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            reinterpret_cast<float4*>(out)[off + e] = make_float4(p.x, p.y, p.z, 1.f);
        }
    }
}

// `bound` = host-known upper bound of the number of live voxels (the sort needs a host-side item count).
int robin_export_bounded(lo_grid* g, int mode, void* d_out, uint32_t* d_count, uint32_t bound) {
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), ctx->stream));
    if (bound == 0 || !g->capacity) return LO_OK;
    uint32_t cap = 2;                       // buckets of the map after `bound` insertions: 2 * pow2ceil(bound)
    while (cap < 2u * bound) cap <<= 1;
    // scratch in tmp_c: keys_in | keys_out | vals_in | vals_out | slot_rot | count_rot | offsets | order | hash_of_slot | idx[2] | dist[2] | cursor
    const size_t kb = static_cast<size_t>(bound) * sizeof(unsigned long long), vb = static_cast<size_t>(bound) * sizeof(uint32_t);
    const size_t hb = static_cast<size_t>(g->capacity) * sizeof(uint32_t), tb = static_cast<size_t>(cap) * sizeof(int);
    LO_TRY(ctx->tmp_c.ensure(2 * kb + 6 * vb + hb + 4 * tb + 64));
    char* base = ctx->tmp_c.as<char>();
    auto* keys_in = reinterpret_cast<unsigned long long*>(base);
    auto* keys_out = reinterpret_cast<unsigned long long*>(base + kb);
    auto* vals_in = reinterpret_cast<uint32_t*>(base + 2 * kb);
    auto* vals_out = vals_in + bound;
    auto* slot_rot = vals_out + bound;
    auto* count_rot = slot_rot + bound;
    auto* offsets = count_rot + bound;
    auto* order = offsets + bound;
    auto* hash_of_slot = order + bound;
    int* idx2 = reinterpret_cast<int*>(hash_of_slot + g->capacity);
    int* dist2 = idx2 + 2 * static_cast<size_t>(cap);
    auto* cursor = reinterpret_cast<uint32_t*>(dist2 + 2 * static_cast<size_t>(cap));
    LO_CUDA(cudaMemsetAsync(keys_in, 0xFF, kb, ctx->stream));
    LO_CUDA(cudaMemsetAsync(cursor, 0, sizeof(uint32_t), ctx->stream));
    const GridView v = g->view();
    robin_keys_kernel<<<blocks_for(g->capacity, 256), 256, 0, ctx->stream>>>(v, g->capacity, g->counters, keys_in, vals_in, cursor, hash_of_slot);
    LO_LAUNCHED(ctx);
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(bound), 0, 64, ctx->stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, count_rot, offsets, static_cast<int>(bound), ctx->stream);
    LO_TRY(ctx->tmp_d.ensure(std::max(sort_bytes, scan_bytes)));
    LO_CUDA(cub::DeviceRadixSort::SortPairs(ctx->tmp_d.p, sort_bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(bound), 0, 32,
                                            ctx->stream));
    ctx->launches += 4;  // cub's sort pipeline (histogram + onesweep passes); counted approximately
    robin_replay_kernel<<<1, 32, 0, ctx->stream>>>(keys_out, hash_of_slot, vals_out, g->counters, idx2, dist2, cap, order);
    LO_LAUNCHED(ctx);
    const unsigned blocks = blocks_for(bound, 256);
    robin_gather_kernel<<<blocks, 256, 0, ctx->stream>>>(order, vals_out, v, g->counters, mode, bound, slot_rot, count_rot);
    LO_LAUNCHED(ctx);
    LO_CUDA(cub::DeviceScan::ExclusiveSum(ctx->tmp_d.p, scan_bytes, count_rot, offsets, static_cast<int>(bound), ctx->stream));
    ctx->launches += 2;
    robin_emit_kernel<<<blocks, 256, 0, ctx->stream>>>(v, g->counters, slot_rot, offsets, mode, d_out, d_count);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// the block-local epoch replay of lo_robin_block.cuh (what the per-scan kernel runs) on its own
__global__ void __launch_bounds__(512, 1) robin_order_block_kernel(const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n,
                                                                   uint32_t* __restrict__ out, uint32_t* __restrict__ refused, long long* prof) {
    extern __shared__ __align__(16) unsigned char robin_smem[];
    const bool ok = robin_order_block(uh, ids, n, robin_smem, out, prof);
    if (threadIdx.x == 0) *refused = ok ? 0u : 1u;
}

// the same for any n up to kRobinGridMax: block replay of the first 2^14 elements, the later epochs on the whole cooperative grid
__global__ void __launch_bounds__(512, 1) robin_order_grid_kernel(const uint32_t* __restrict__ uh, const uint32_t* __restrict__ ids, uint32_t n,
                                                                  uint32_t* scratch, uint32_t* __restrict__ out, uint32_t* __restrict__ refused,
                                                                  uint32_t handover) {
    extern __shared__ __align__(16) unsigned char robin_smem[];
    cg::grid_group grid = cg::this_grid();
    const RobinGridBufs B = robin_grid_carve(scratch, n);
    const bool ok = robin_order_any(grid, uh, ids, n, robin_smem, B, out, handover);
    if (blockIdx.x == 0 && threadIdx.x == 0) *refused = ok ? 0u : 1u;
}

int robin_export_dev(lo_grid* g, int mode, void* d_out, uint32_t* d_count) {
    return robin_export_bounded(g, mode, d_out, d_count, g->h_counters[GC_VOXELS]);
}

}  // namespace lo

using namespace lo;

// order_out[d] = index (into hash22) of the key at iteration position d of a tsl::robin_map<..., IndicesHash> into which n
// DISTINCT keys with these hashes were inserted in index order (src/voxel_grid.h:82-87 on an empty map)
extern "C" int lo_robin_iteration_order(lo_ctx* ctx, const uint32_t* hash22, size_t n, uint32_t* order_out) {
    if (!ctx || (!hash22 && n) || (!order_out && n)) return LO_E_ARG;
    if (n > kRobinGridMax) {
        set_error("lo_robin_iteration_order takes at most %u keys (got %zu)", kRobinGridMax, n);
        return LO_E_CAPACITY;
    }
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(ctx->tmp_a.ensure(n * sizeof(uint32_t)));
    LO_TRY(ctx->tmp_b.ensure(n * sizeof(uint32_t)));
    LO_TRY(ctx->tmp_c.ensure(n * sizeof(uint32_t)));
    std::vector<uint32_t> iota(n);
    for (size_t k = 0; k < n; k++) iota[k] = static_cast<uint32_t>(k);
    LO_CUDA(cudaMemcpyAsync(ctx->tmp_a.p, hash22, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(ctx->tmp_b.p, iota.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaFuncSetAttribute(robin_order_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kRobinBlockSmemBytes)));
    uint32_t* d_refused = ctx->small.as<uint32_t>() + 96;
    const char* force = getenv("LO_ROBIN_GRID");   // developer switch: 1 = every epoch on the grid, whatever n
    const char* hand = getenv("LO_ROBIN_HANDOVER");   // developer switch: the block replays the first 2^k elements (k <= 14), the grid the rest
    const int force_grid = force && force[0] == '1' ? 1 : 0;
    uint32_t handover = force_grid ? 0u : kRobinHandover;
    if (hand && !force_grid) handover = static_cast<uint32_t>(std::max(1, std::min(14, atoi(hand))));
    if (n > (1u << handover) || force_grid) {
        uint32_t nn = static_cast<uint32_t>(n);
        LO_TRY(ctx->tmp_d.ensure(robin_grid_words(nn) * sizeof(uint32_t)));
        LO_CUDA(cudaFuncSetAttribute(robin_order_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kRobinBlockSmemBytes)));
        int per_sm = 0;
        LO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, robin_order_grid_kernel, 512, kRobinBlockSmemBytes));
        if (per_sm < 1) {
            set_error("the grid replay kernel does not fit on an SM");
            return LO_E_CUDA;
        }
        const int blocks = std::min<int>(ctx->sm_count, static_cast<int>(kRobinGridMaxBlocks));
        const uint32_t* d_uh = ctx->tmp_a.as<uint32_t>();
        const uint32_t* d_ids = ctx->tmp_b.as<uint32_t>();
        uint32_t* d_scratch = ctx->tmp_d.as<uint32_t>();
        uint32_t* d_out = ctx->tmp_c.as<uint32_t>();
        void* kargs[] = {&d_uh, &d_ids, &nn, &d_scratch, &d_out, &d_refused, &handover};
        LO_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(robin_order_grid_kernel), dim3(blocks), dim3(512), kargs, kRobinBlockSmemBytes,
                                            ctx->stream));
        ctx->launches++;
        uint32_t refused_g = 0;
        LO_CUDA(cudaMemcpyAsync(order_out, ctx->tmp_c.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        LO_CUDA(cudaMemcpyAsync(&refused_g, d_refused, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        LO_CUDA(cudaStreamSynchronize(ctx->stream));
        if (refused_g) {
            set_error("lo_robin_iteration_order: a run of more than %u occupied buckets (degenerate hashes); the epoch replay refuses it — "
                      "VoxelGrid::getCloud* (lo_grid_get_cloud, order 1) replays any key set", kRobinBlockMaxCluster);
            return LO_E_RANGE;
        }
        return LO_OK;
    }
    static const bool profile = getenv("LO_ROBIN_PROF") != nullptr;   // developer switch: cycles per phase of the replay on stderr
    long long* d_prof = nullptr;
    if (profile) {
        LO_CUDA(cudaMalloc(&d_prof, 64 * sizeof(long long)));
        LO_CUDA(cudaMemsetAsync(d_prof, 0, 64 * sizeof(long long), ctx->stream));
    }
    robin_order_block_kernel<<<1, 512, kRobinBlockSmemBytes, ctx->stream>>>(ctx->tmp_a.as<uint32_t>(), ctx->tmp_b.as<uint32_t>(), static_cast<uint32_t>(n),
                                                                             ctx->tmp_c.as<uint32_t>(), d_refused, d_prof);
    LO_LAUNCHED(ctx);
    if (profile) {
        long long h_prof[64];
        LO_CUDA(cudaMemcpyAsync(h_prof, d_prof, sizeof(h_prof), cudaMemcpyDeviceToHost, ctx->stream));
        LO_CUDA(cudaStreamSynchronize(ctx->stream));
        LO_CUDA(cudaFree(d_prof));
        static const char* names[14] = {"", "E1 clear", "E2 histogram", "E3 scan", "E4 wrap", "E5 place", "A labels", "R2 lengths", "R3 rank",
                                        "P park", "W walk", "-", "-", "E7 order"};
        fprintf(stderr, "robin replay n = %zu:", n);
        for (int k = 1; k <= 13; k++) fprintf(stderr, " %s %lld |", names[k], h_prof[k]);
        fprintf(stderr, "\n  per epoch:");
        for (int k = 32; k < 48; k++) fprintf(stderr, " %lld", h_prof[k]);
        fprintf(stderr, "\n");
    }
    uint32_t refused = 0;
    LO_CUDA(cudaMemcpyAsync(order_out, ctx->tmp_c.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(&refused, d_refused, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));   // iota must outlive the copy
    if (refused) {
        set_error("lo_robin_iteration_order: a run of more than %u occupied buckets (degenerate hashes); the block-local replay refuses it — "
                  "VoxelGrid::getCloud* (lo_grid_get_cloud, order 1) replays any key set", kRobinBlockMaxCluster);
        return LO_E_RANGE;
    }
    return LO_OK;
}
