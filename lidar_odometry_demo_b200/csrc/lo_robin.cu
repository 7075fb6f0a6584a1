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
// Round 1 implementation: sort by (b, arrival ticket) and rotate the wrapped tail to the front. Ties are
// emitted in arrival order, which is exact unless a later insertion displaced through the tie group
// (DESIGN.md "robin order" quantifies the residual mismatch against the oracle).
#include <cub/cub.cuh>

#include "lo_host.h"

namespace lo {

__global__ void robin_keys_kernel(GridView g, uint32_t capacity, const uint32_t* __restrict__ counters,
                                  unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ cursor) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= capacity) return;
    const Slot slot = g.slots[s];
    if (slot.key == kKeyEmpty || slot.key == kKeyTomb || slot.count == 0) return;
    const uint32_t n = counters[GC_VOXELS];
    uint32_t c = 2;
    while (c < 2 * n) c <<= 1;  // 2 * pow2ceil(n)
    long long ix, iy, iz;
    unpack_key(slot.key, ix, iy, iz);
    const uint32_t b = reference_hash22(ix, iy, iz) & (c - 1);
    const uint32_t ticket = __float_as_uint(g.points[static_cast<size_t>(slot.vid) * g.width].w);
    const uint32_t at = atomicAdd(cursor, 1u);
    keys[at] = (static_cast<unsigned long long>(b) << 32) | ticket;
    vals[at] = s;
}

// after the sort: how far does the last element sit past the end of the bucket array?
__global__ void robin_wrap_kernel(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ counters,
                                  int* __restrict__ max_shift) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = counters[GC_VOXELS];
    if (i >= n) return;
    const int b = static_cast<int>(keys[i] >> 32);
    atomicMax(max_shift, b - static_cast<int>(i));
}

__global__ void robin_rotate_kernel(const uint32_t* __restrict__ sorted_slots, GridView g, const uint32_t* __restrict__ counters,
                                    const int* __restrict__ max_shift, int mode, uint32_t bound, uint32_t* __restrict__ slot_rot,
                                    uint32_t* __restrict__ count_rot) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= bound) return;
    const uint32_t n = counters[GC_VOXELS];
    if (i >= n) {
        count_rot[i] = 0;
        return;
    }
    uint32_t c = 2;
    while (c < 2 * n) c <<= 1;
    // positions are pos_i = i + max_{j<=i}(b_j - j); the last one is n-1+max_shift; everything at >= C wraps
    const long long over = static_cast<long long>(n) + static_cast<long long>(*max_shift) - static_cast<long long>(c);
    const uint32_t w = over > 0 ? static_cast<uint32_t>(over) : 0u;
    const uint32_t dst = (i + w) % n;
    const uint32_t s = sorted_slots[i];
    slot_rot[dst] = s;
    count_rot[dst] = (mode == 2) ? 1u : g.slots[s].count;
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
    // scratch layout in tmp_c: keys_in | keys_out | vals_in | vals_out | slot_rot | count_rot | offsets | cursor,max_shift
    const size_t kb = static_cast<size_t>(bound) * sizeof(unsigned long long), vb = static_cast<size_t>(bound) * sizeof(uint32_t);
    LO_TRY(ctx->tmp_c.ensure(2 * kb + 5 * vb + 64));
    char* base = ctx->tmp_c.as<char>();
    auto* keys_in = reinterpret_cast<unsigned long long*>(base);
    auto* keys_out = reinterpret_cast<unsigned long long*>(base + kb);
    auto* vals_in = reinterpret_cast<uint32_t*>(base + 2 * kb);
    auto* vals_out = vals_in + bound;
    auto* slot_rot = vals_out + bound;
    auto* count_rot = slot_rot + bound;
    auto* offsets = count_rot + bound;
    auto* cursor = offsets + bound;
    int* max_shift = reinterpret_cast<int*>(cursor + 1);
    LO_CUDA(cudaMemsetAsync(keys_in, 0xFF, kb, ctx->stream));
    LO_CUDA(cudaMemsetAsync(cursor, 0, sizeof(uint32_t), ctx->stream));
    const int neg = INT_MIN;
    LO_CUDA(cudaMemcpyAsync(max_shift, &neg, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    const GridView v = g->view();
    robin_keys_kernel<<<blocks_for(g->capacity, 256), 256, 0, ctx->stream>>>(v, g->capacity, g->counters, keys_in, vals_in, cursor);
    LO_LAUNCHED(ctx);
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(bound), 0, 64, ctx->stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, count_rot, offsets, static_cast<int>(bound), ctx->stream);
    LO_TRY(ctx->tmp_d.ensure(std::max(sort_bytes, scan_bytes)));
    LO_CUDA(cub::DeviceRadixSort::SortPairs(ctx->tmp_d.p, sort_bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(bound), 0, 54,
                                            ctx->stream));
    ctx->launches += 4;  // cub's sort pipeline (histogram + onesweep passes); counted approximately
    const unsigned blocks = blocks_for(bound, 256);
    robin_wrap_kernel<<<blocks, 256, 0, ctx->stream>>>(keys_out, g->counters, max_shift);
    LO_LAUNCHED(ctx);
    robin_rotate_kernel<<<blocks, 256, 0, ctx->stream>>>(vals_out, v, g->counters, max_shift, mode, bound, slot_rot, count_rot);
    LO_LAUNCHED(ctx);
    LO_CUDA(cub::DeviceScan::ExclusiveSum(ctx->tmp_d.p, scan_bytes, count_rot, offsets, static_cast<int>(bound), ctx->stream));
    ctx->launches += 2;
    robin_emit_kernel<<<blocks, 256, 0, ctx->stream>>>(v, g->counters, slot_rot, offsets, mode, d_out, d_count);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

int robin_export_dev(lo_grid* g, int mode, void* d_out, uint32_t* d_count) {
    return robin_export_bounded(g, mode, d_out, d_count, g->h_counters[GC_VOXELS]);
}

}  // namespace lo
