// K2/K3: the GPU voxel grid that replaces VoxelGrid's tsl::robin_map (src/voxel_grid.h:17-257).
//
// Layout in HBM (DESIGN.md "data layout"):
//   slots   [capacity]              16 B {key64, vid, count}; open addressing, linear probing, load <= 0.5
//   points  [payload_cap * width]   float4 xyz + w (w = arrival ticket of the stored point)
//   normals [payload_cap * width]   float4
// A voxel keeps its first `max_points` points in arrival order (voxel_grid.h:82-91). Parallel insertion
// reproduces that order without sorting: every point carries a ticket (its position in the arrival
// sequence) and is pushed through the free entries of its voxel's slab with a chain of atomicMin on the
// w lane; entry k ends up holding the k-th smallest ticket, whatever the thread interleaving.
#include <algorithm>
#include <cstring>

#include "lo_grid.cuh"

namespace lo {

// thin kernels over the per-item device functions of lo_grid.cuh
__global__ void grid_locate_kernel(PointSource src, uint32_t n_bound, const uint32_t* __restrict__ d_n, const Pose* __restrict__ pose, GridView g,
                                   uint32_t* __restrict__ counters, const uint32_t* __restrict__ free_list,
                                   uint32_t* __restrict__ slot_of) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = d_n ? min(n_bound, *d_n) : n_bound;
    if (i < n) grid_locate_item(i, src, pose, g, counters, free_list, slot_of);
}
__global__ void grid_chain_kernel(uint32_t n_bound, const uint32_t* __restrict__ d_n, uint32_t ticket_base, GridView g, const uint32_t* __restrict__ slot_of,
                                  uint32_t* __restrict__ base_of) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = d_n ? min(n_bound, *d_n) : n_bound;
    if (i < n) grid_chain_item(i, ticket_base, g, slot_of, base_of);
}
__global__ void grid_write_kernel(PointSource src, uint32_t n_bound, const uint32_t* __restrict__ d_n, uint32_t ticket_base, const Pose* __restrict__ pose,
                                  GridView g, uint32_t* __restrict__ counters, const uint32_t* __restrict__ slot_of,
                                  const uint32_t* __restrict__ base_of) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = d_n ? min(n_bound, *d_n) : n_bound;
    if (i < n) grid_write_item(i, src, ticket_base, pose, g, counters, slot_of, base_of);
}
__global__ void grid_fix_counters_kernel(uint32_t* counters) { grid_fix_counters(counters); }
__global__ void grid_cleanup_kernel(GridView g, uint32_t capacity, const float* __restrict__ center, float radius,
                                    uint32_t* __restrict__ counters, uint32_t* __restrict__ free_list) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < capacity) grid_cleanup_slot(s, g, center, radius, counters, free_list);
}

// move live slots into a fresh table (growth, or purge of tombstones)
__global__ void grid_rebuild_kernel(const Slot* __restrict__ old_slots, uint32_t old_capacity, Slot* __restrict__ new_slots,
                                    uint32_t new_mask) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= old_capacity) return;
    const Slot slot = old_slots[s];
    if (slot.key == kKeyEmpty || slot.key == kKeyTomb) return;
    uint32_t h = probe_hash(slot.key) & new_mask;
    while (atomicCAS(&new_slots[h].key, kKeyEmpty, slot.key) != kKeyEmpty) h = (h + 1) & new_mask;
    new_slots[h].vid = slot.vid;
    new_slots[h].count = slot.count;
}

// getCloud / getCloudWithoutNormals / getSparseCloudWithoutNormals, unspecified voxel order
__global__ void grid_export_kernel(GridView g, uint32_t capacity, int mode, void* __restrict__ out, uint32_t* __restrict__ cursor) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= capacity) return;
    const Slot slot = g.slots[s];
    if (slot.key == kKeyEmpty || slot.key == kKeyTomb || slot.count == 0) return;
    const uint32_t cnt = (mode == 2) ? 1u : slot.count;
    const uint32_t off = atomicAdd(cursor, cnt);
    if (!out) return;
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

// ---------------------------------------------------------------------------------------------------
static uint32_t next_pow2(uint64_t v) {
    uint64_t p = 1;
    while (p < v) p <<= 1;
    return static_cast<uint32_t>(p);
}

// All (re)allocations of a grid are stream-ordered (cudaMallocAsync / cudaFreeAsync on the context's stream, from the
// device's default pool whose release threshold lo_ctx_create lifts): a table or payload that grows is allocated,
// filled and swapped in behind the work already queued and in front of the work queued next, with no host
// synchronisation and no cudaFree. Kernels already queued hold the old view and run before the old buffers are
// released; callers take g->view() after grid_reserve.
static int alloc_table(lo_grid* g, uint32_t capacity, Slot** out) {
    LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(out), static_cast<size_t>(capacity) * sizeof(Slot), g->ctx->stream));
    LO_CUDA(cudaMemsetAsync(*out, 0xFF, static_cast<size_t>(capacity) * sizeof(Slot), g->ctx->stream));
    return LO_OK;
}

static int grow_payload(lo_grid* g, uint32_t new_cap) {
    lo_ctx* ctx = g->ctx;
    float4 *np = nullptr, *nn = nullptr;
    uint32_t* nf = nullptr;
    const size_t new_elems = static_cast<size_t>(new_cap) * g->width, old_elems = static_cast<size_t>(g->payload_cap) * g->width;
    LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&np), new_elems * sizeof(float4), ctx->stream));
    LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&nn), new_elems * sizeof(float4), ctx->stream));
    LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&nf), static_cast<size_t>(new_cap) * sizeof(uint32_t), ctx->stream));
    if (old_elems) {
        LO_CUDA(cudaMemcpyAsync(np, g->points, old_elems * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
        LO_CUDA(cudaMemcpyAsync(nn, g->normals, old_elems * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
        LO_CUDA(cudaMemcpyAsync(nf, g->free_list, static_cast<size_t>(g->payload_cap) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    LO_CUDA(cudaMemsetAsync(np + old_elems, 0xFF, (new_elems - old_elems) * sizeof(float4), ctx->stream));
    if (g->points) LO_CUDA(cudaFreeAsync(g->points, ctx->stream));
    if (g->normals) LO_CUDA(cudaFreeAsync(g->normals, ctx->stream));
    if (g->free_list) LO_CUDA(cudaFreeAsync(g->free_list, ctx->stream));
    g->points = np; g->normals = nn; g->free_list = nf;
    g->payload_cap = new_cap;
    g->grow_count++;
    return LO_OK;
}

static int rebuild_table(lo_grid* g, uint32_t new_capacity) {
    lo_ctx* ctx = g->ctx;
    Slot* fresh = nullptr;
    LO_TRY(alloc_table(g, new_capacity, &fresh));
    if (g->slots && g->capacity) {
        grid_rebuild_kernel<<<blocks_for(g->capacity, 256), 256, 0, ctx->stream>>>(g->slots, g->capacity, fresh, new_capacity - 1);
        LO_LAUNCHED(ctx);
        LO_CUDA(cudaMemsetAsync(g->counters + GC_TOMBS, 0, sizeof(uint32_t), ctx->stream));
        LO_CUDA(cudaFreeAsync(g->slots, ctx->stream));
        g->rebuild_count++;   // counter blocks read back from work queued before this point still carry the old tombstone count
    }
    g->slots = fresh;
    g->capacity = new_capacity;
    g->h_counters[GC_TOMBS] = 0;
    return LO_OK;
}

int grid_refresh_counters(lo_grid* g) {
    const uint32_t ticket = g->h_counters[GC_TICKET];   // arrival tickets are handed out by the host only
    LO_CUDA(cudaMemcpyAsync(g->h_counters, g->counters, sizeof(g->h_counters), cudaMemcpyDeviceToHost, g->ctx->stream));
    LO_CUDA(cudaStreamSynchronize(g->ctx->stream));
    g->h_counters[GC_TICKET] = ticket;
    return LO_OK;
}

// Tickets only order the points of ONE add_cloud batch inside a voxel (entries below the voxel's count are never
// touched again) and the voxels of a never-cleared grid for the reference-order export; 0xFFFFFFFF marks a free
// entry. Restart the sequence before it could reach that value (after ~4e9 offered points).
uint32_t grid_take_tickets(lo_grid* g, uint32_t n) {
    if (g->h_counters[GC_TICKET] > 0xF0000000u - n) g->h_counters[GC_TICKET] = 0;
    const uint32_t base = g->h_counters[GC_TICKET];
    g->h_counters[GC_TICKET] = base + n;
    return base;
}

// Make room for `incoming` more points (each may open a new voxel). Uses the host mirror of the counters,
// which every mutating entry point refreshes, so no device synchronisation happens here unless we grow.
int grid_reserve(lo_grid* g, size_t incoming) {
    const uint64_t live = g->h_counters[GC_VOXELS], tombs = g->h_counters[GC_TOMBS];
    const uint64_t need_slots = (live + tombs + incoming) * 2;
    if (need_slots > g->capacity) {
        const uint32_t cap = next_pow2(std::max<uint64_t>((live + incoming) * 4, 1024));
        LO_TRY(rebuild_table(g, cap));
    } else if (tombs > g->capacity / 8) {
        LO_TRY(rebuild_table(g, g->capacity));
    }
    const int64_t free_top = static_cast<int32_t>(g->h_counters[GC_FREE_TOP]);
    const uint64_t need_payload = g->h_counters[GC_BUMP] + incoming - static_cast<uint64_t>(std::max<int64_t>(free_top, 0) > static_cast<int64_t>(incoming) ? incoming : std::max<int64_t>(free_top, 0));
    if (need_payload > g->payload_cap) {
        const uint32_t cap = next_pow2(std::max<uint64_t>(need_payload * 2, 1024));
        LO_TRY(grow_payload(g, cap));
    }
    return LO_OK;
}

int grid_add_dev(lo_grid* g, const void* d_points, size_t n, int has_normals, const Pose* d_pose, const uint32_t* d_n) {
    if (n == 0) return LO_OK;
    lo_ctx* ctx = g->ctx;
    LO_TRY(grid_reserve(g, n));
    LO_TRY(ctx->tmp_a.ensure(n * sizeof(uint32_t)));
    LO_TRY(ctx->tmp_b.ensure(n * sizeof(uint32_t)));
    PointSource src{static_cast<const float*>(d_points), has_normals ? 12u : 4u, has_normals};
    const uint32_t nn = static_cast<uint32_t>(n);
    const uint32_t ticket_base = grid_take_tickets(g, nn);
    const GridView v = g->view();
    const unsigned blocks = blocks_for(n, 256);
    grid_locate_kernel<<<blocks, 256, 0, ctx->stream>>>(src, nn, d_n, d_pose, v, g->counters, g->free_list, ctx->tmp_a.as<uint32_t>());
    LO_LAUNCHED(ctx);
    grid_chain_kernel<<<blocks, 256, 0, ctx->stream>>>(nn, d_n, ticket_base, v, ctx->tmp_a.as<uint32_t>(), ctx->tmp_b.as<uint32_t>());
    LO_LAUNCHED(ctx);
    grid_write_kernel<<<blocks, 256, 0, ctx->stream>>>(src, nn, d_n, ticket_base, d_pose, v, g->counters, ctx->tmp_a.as<uint32_t>(),
                                                       ctx->tmp_b.as<uint32_t>());
    LO_LAUNCHED(ctx);
    grid_fix_counters_kernel<<<1, 1, 0, ctx->stream>>>(g->counters);
    LO_LAUNCHED(ctx);
    // conservative host-side bounds until the next refresh
    g->h_counters[GC_VOXELS] += nn;
    g->h_counters[GC_BUMP] += nn;
    return LO_OK;
}

int grid_clear(lo_grid* g) {
    lo_ctx* ctx = g->ctx;
    if (g->capacity) LO_CUDA(cudaMemsetAsync(g->slots, 0xFF, static_cast<size_t>(g->capacity) * sizeof(Slot), ctx->stream));
    // only the slabs that were ever handed out can hold tickets
    const size_t used = static_cast<size_t>(std::min<uint64_t>(g->h_counters[GC_BUMP], g->payload_cap)) * g->width;
    if (used) LO_CUDA(cudaMemsetAsync(g->points, 0xFF, used * sizeof(float4), ctx->stream));
    LO_CUDA(cudaMemsetAsync(g->counters, 0, GC_COUNT * sizeof(uint32_t), ctx->stream));
    for (auto& c : g->h_counters) c = 0;
    g->cleaned = false;
    return LO_OK;
}

int grid_cleanup_dev(lo_grid* g, const float* d_center3, float radius) {
    if (!g->capacity) return LO_OK;
    lo_ctx* ctx = g->ctx;
    grid_cleanup_kernel<<<blocks_for(g->capacity, 256), 256, 0, ctx->stream>>>(g->view(), g->capacity, d_center3, radius, g->counters,
                                                                               g->free_list);
    LO_LAUNCHED(ctx);
    g->cleaned = true;
    return LO_OK;
}

int grid_export_dev(lo_grid* g, int mode, void* d_out, uint32_t* d_count) {
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), ctx->stream));
    if (!g->capacity) return LO_OK;
    grid_export_kernel<<<blocks_for(g->capacity, 256), 256, 0, ctx->stream>>>(g->view(), g->capacity, mode, d_out, d_count);
    LO_LAUNCHED(ctx);
    return LO_OK;
}


// ---- snapshot / restore: the table is dumped as it lies in HBM (slots, the slabs handed out so far, the free list,
// the counters), so a restored grid behaves bit for bit like the original: same probe positions, same slab order.
struct GridSnapHeader {
    uint64_t magic;
    uint32_t header_bytes, width;
    float voxel_size;
    uint32_t capacity;
    uint64_t max_points;
    uint32_t payload_used, free_n;
    uint32_t shard_rank, shard_world;
    uint32_t cleaned, reserved;
    uint32_t counters[GC_COUNT];
};
constexpr uint64_t kGridMagic = 0x0001444952474F4CULL;   // "LOGRID\1\0"

static size_t grid_snapshot_payload(const GridSnapHeader& h) {
    return static_cast<size_t>(h.capacity) * sizeof(Slot) + 2 * static_cast<size_t>(h.payload_used) * h.width * sizeof(float4) +
           static_cast<size_t>(h.free_n) * sizeof(uint32_t);
}

int grid_snapshot(lo_grid* g, void* out, size_t cap, size_t* bytes) {
    lo_ctx* ctx = g->ctx;
    LO_TRY(grid_refresh_counters(g));
    GridSnapHeader h{};
    h.magic = kGridMagic;
    h.header_bytes = sizeof(GridSnapHeader);
    h.width = g->width;
    h.voxel_size = g->voxel_size;
    h.capacity = g->capacity;
    h.max_points = g->max_points;
    h.payload_used = std::min<uint32_t>(g->h_counters[GC_BUMP], g->payload_cap);
    h.free_n = static_cast<uint32_t>(std::max<int32_t>(static_cast<int32_t>(g->h_counters[GC_FREE_TOP]), 0));
    h.shard_rank = g->shard_rank;
    h.shard_world = g->shard_world;
    h.cleaned = g->cleaned ? 1u : 0u;
    for (int k = 0; k < GC_COUNT; k++) h.counters[k] = g->h_counters[k];
    const size_t total = sizeof(h) + grid_snapshot_payload(h);
    *bytes = total;
    if (!out) return LO_OK;
    if (cap < total) return LO_E_CAPACITY;
    char* o = static_cast<char*>(out);
    std::memcpy(o, &h, sizeof(h));
    o += sizeof(h);
    const size_t slab = static_cast<size_t>(h.payload_used) * h.width * sizeof(float4);
    LO_CUDA(cudaMemcpyAsync(o, g->slots, static_cast<size_t>(h.capacity) * sizeof(Slot), cudaMemcpyDeviceToHost, ctx->stream));
    o += static_cast<size_t>(h.capacity) * sizeof(Slot);
    if (slab) {
        LO_CUDA(cudaMemcpyAsync(o, g->points, slab, cudaMemcpyDeviceToHost, ctx->stream));
        LO_CUDA(cudaMemcpyAsync(o + slab, g->normals, slab, cudaMemcpyDeviceToHost, ctx->stream));
    }
    o += 2 * slab;
    if (h.free_n) LO_CUDA(cudaMemcpyAsync(o, g->free_list, static_cast<size_t>(h.free_n) * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

int grid_restore(lo_grid* g, const void* blob, size_t bytes) {
    lo_ctx* ctx = g->ctx;
    GridSnapHeader h;
    if (bytes < sizeof(h)) {
        set_error("grid snapshot: %zu bytes is shorter than the header", bytes);
        return LO_E_IO;
    }
    std::memcpy(&h, blob, sizeof(h));
    if (h.magic != kGridMagic || h.header_bytes != sizeof(h) || h.capacity == 0 || (h.capacity & (h.capacity - 1)) != 0 ||
        h.free_n > h.payload_used || bytes != sizeof(h) + grid_snapshot_payload(h)) {
        set_error("grid snapshot: bad magic, version or size");
        return LO_E_IO;
    }
    if (h.width > g->width) LO_TRY(lo_grid_set_max_points(g, h.width));   // slabs laid out for a larger cap (setMaxPoints history)
    if (h.width != g->width) {
        set_error("grid snapshot was taken with max_points %llu (slab width %u), this grid has %zu (slab width %u)",
                  static_cast<unsigned long long>(h.max_points), h.width, g->max_points, g->width);
        return LO_E_ARG;
    }
    g->max_points = static_cast<size_t>(h.max_points);
    if (h.capacity != g->capacity) {
        Slot* fresh = nullptr;
        LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&fresh), static_cast<size_t>(h.capacity) * sizeof(Slot), ctx->stream));
        if (g->slots) LO_CUDA(cudaFreeAsync(g->slots, ctx->stream));
        g->slots = fresh;
        g->capacity = h.capacity;
    }
    if (h.payload_used > g->payload_cap) LO_TRY(grow_payload(g, next_pow2(std::max<uint64_t>(h.payload_used, 1024))));
    const char* in = static_cast<const char*>(blob) + sizeof(h);
    const size_t slab = static_cast<size_t>(h.payload_used) * h.width * sizeof(float4);
    LO_CUDA(cudaMemcpyAsync(g->slots, in, static_cast<size_t>(h.capacity) * sizeof(Slot), cudaMemcpyHostToDevice, ctx->stream));
    in += static_cast<size_t>(h.capacity) * sizeof(Slot);
    // entries never handed out must read as free (ticket lane 0xFFFFFFFF)
    LO_CUDA(cudaMemsetAsync(g->points, 0xFF, static_cast<size_t>(g->payload_cap) * g->width * sizeof(float4), ctx->stream));
    if (slab) {
        LO_CUDA(cudaMemcpyAsync(g->points, in, slab, cudaMemcpyHostToDevice, ctx->stream));
        LO_CUDA(cudaMemcpyAsync(g->normals, in + slab, slab, cudaMemcpyHostToDevice, ctx->stream));
    }
    in += 2 * slab;
    if (h.free_n) LO_CUDA(cudaMemcpyAsync(g->free_list, in, static_cast<size_t>(h.free_n) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(g->counters, h.counters, sizeof(h.counters), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < GC_COUNT; k++) g->h_counters[k] = h.counters[k];
    g->voxel_size = h.voxel_size;
    g->shard_rank = h.shard_rank;
    g->shard_world = h.shard_world;
    g->cleaned = h.cleaned != 0;
    return LO_OK;
}

}  // namespace lo

using namespace lo;

extern "C" int lo_grid_snapshot(lo_grid* g, void* out, size_t cap, size_t* bytes) {
    if (!g || !bytes) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(g->ctx->device));
    return grid_snapshot(g, out, cap, bytes);
}

extern "C" int lo_grid_restore(lo_grid* g, const void* blob, size_t bytes) {
    if (!g || !blob) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(g->ctx->device));
    return grid_restore(g, blob, bytes);
}

extern "C" int lo_grid_create(lo_ctx* ctx, float voxel_size, size_t max_points, size_t capacity_hint, lo_grid** out) {
    if (!ctx || !out || !(voxel_size > 0.f)) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(ctx->device));
    lo_grid* g = new lo_grid();
    g->ctx = ctx;
    g->voxel_size = voxel_size;
    g->max_points = max_points;
    g->width = static_cast<uint32_t>(max_points ? max_points : 1);
    LO_CUDA(cudaMalloc(&g->counters, GC_COUNT * sizeof(uint32_t)));
    LO_CUDA(cudaMemsetAsync(g->counters, 0, GC_COUNT * sizeof(uint32_t), ctx->stream));
    const uint64_t hint = capacity_hint ? capacity_hint : 4096;
    int st = rebuild_table(g, next_pow2(std::max<uint64_t>(hint * 2, 1024)));
    if (st == LO_OK) st = grow_payload(g, next_pow2(std::max<uint64_t>(hint, 1024)));
    if (st == LO_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) {   // out of memory shows up here, not in a later call
        set_error("allocating a voxel grid for %llu voxels of %u points failed: %s", static_cast<unsigned long long>(hint), g->width,
                  cudaGetErrorString(cudaGetLastError()));
        st = LO_E_CUDA;
    }
    g->rebuild_count = g->grow_count = 0;
    if (st != LO_OK) {
        lo_grid_destroy(g);
        return st;
    }
    *out = g;
    return LO_OK;
}

extern "C" void lo_grid_destroy(lo_grid* g) {
    if (!g) return;
    cudaSetDevice(g->ctx->device);
    cudaStreamSynchronize(g->ctx->stream);
    if (g->slots) cudaFreeAsync(g->slots, g->ctx->stream);
    if (g->points) cudaFreeAsync(g->points, g->ctx->stream);
    if (g->normals) cudaFreeAsync(g->normals, g->ctx->stream);
    if (g->free_list) cudaFreeAsync(g->free_list, g->ctx->stream);
    if (g->counters) cudaFree(g->counters);
    delete g;
}

extern "C" int lo_grid_set_voxel_size(lo_grid* g, float voxel_size) {
    if (!g || !(voxel_size > 0.f)) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(g->ctx->device));
    LO_TRY(grid_clear(g));
    g->voxel_size = voxel_size;
    LO_CUDA(cudaStreamSynchronize(g->ctx->stream));
    return LO_OK;
}

// setMaxPoints (voxel_grid.h:56-59) only stores the new cap: voxels keep what they hold, later points are appended while
// a voxel holds fewer than the cap. A larger cap than the slabs were laid out for re-lays them out (stream-ordered).
__global__ void grid_relayout_kernel(const float4* __restrict__ old_p, const float4* __restrict__ old_n, uint32_t old_w, float4* __restrict__ new_p,
                                     float4* __restrict__ new_n, uint32_t new_w, size_t n_new) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_new; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const size_t vid = i / new_w;
        const uint32_t e = static_cast<uint32_t>(i % new_w);
        const uint32_t free_bits = 0xFFFFFFFFu;
        float4 p = make_float4(__uint_as_float(free_bits), __uint_as_float(free_bits), __uint_as_float(free_bits), __uint_as_float(free_bits));
        float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e < old_w) {
            p = old_p[vid * old_w + e];
            nv = old_n[vid * old_w + e];
        }
        new_p[i] = p;
        new_n[i] = nv;
    }
}

extern "C" int lo_grid_set_max_points(lo_grid* g, size_t max_points) {
    if (!g) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const uint32_t want = static_cast<uint32_t>(std::min<size_t>(max_points ? max_points : 1, 1u << 16));
    if (want > g->width) {
        float4 *np = nullptr, *nn = nullptr;
        const size_t n_new = static_cast<size_t>(g->payload_cap) * want;
        LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&np), n_new * sizeof(float4), ctx->stream));
        LO_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&nn), n_new * sizeof(float4), ctx->stream));
        grid_relayout_kernel<<<std::max(1u, std::min<unsigned>(blocks_for(n_new, 256), ctx->sm_count * 16)), 256, 0, ctx->stream>>>(
            g->points, g->normals, g->width, np, nn, want, n_new);
        LO_LAUNCHED(ctx);
        LO_CUDA(cudaFreeAsync(g->points, ctx->stream));
        LO_CUDA(cudaFreeAsync(g->normals, ctx->stream));
        g->points = np;
        g->normals = nn;
        g->width = want;
        g->grow_count++;
    }
    g->max_points = max_points;
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" int lo_grid_size(lo_grid* g, size_t* n_voxels, size_t* n_points) {
    if (!g) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(g->ctx->device));
    LO_TRY(grid_refresh_counters(g));
    if (n_voxels) *n_voxels = g->h_counters[GC_VOXELS];
    if (n_points) *n_points = g->h_counters[GC_POINTS];
    return LO_OK;
}

static int finish_mutation(lo_grid* g) {
    LO_TRY(grid_refresh_counters(g));
    if (g->h_counters[GC_RANGE_ERR]) {
        set_error("a point was not finite or its voxel index fell outside +-2^20; it was skipped");
        LO_CUDA(cudaMemsetAsync(g->counters + GC_RANGE_ERR, 0, sizeof(uint32_t), g->ctx->stream));
        return LO_E_RANGE;
    }
    return LO_OK;
}

extern "C" int lo_grid_add_cloud_dev(lo_grid* g, const void* d_points, size_t n, int has_normals) {
    if (!g || (!d_points && n)) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(g->ctx->device));
    LO_TRY(grid_add_dev(g, d_points, n, has_normals, nullptr, nullptr));
    return finish_mutation(g);
}

extern "C" int lo_grid_add_cloud(lo_grid* g, const void* points, size_t n, int has_normals) {
    if (!g || (!points && n)) return LO_E_ARG;
    if (n == 0) return LO_OK;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * (has_normals ? sizeof(PointNormal) : sizeof(PointXYZ));
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, points, bytes, cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(grid_add_dev(g, ctx->in_a.p, n, has_normals, nullptr, nullptr));
    return finish_mutation(g);
}

extern "C" int lo_grid_radius_cleanup(lo_grid* g, const float center[3], float radius) {
    if (!g || !center) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    float* d_c = ctx->small.as<float>();
    LO_CUDA(cudaMemcpyAsync(d_c, center, 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(grid_cleanup_dev(g, d_c, radius));
    return finish_mutation(g);
}

extern "C" int lo_grid_get_cloud(lo_grid* g, int mode, int order, void* out, size_t cap, size_t* n) {
    if (!g || mode < 0 || mode > 2 || !n) return LO_E_ARG;
    lo_ctx* ctx = g->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(grid_refresh_counters(g));
    const size_t total = (mode == 2) ? g->h_counters[GC_VOXELS] : g->h_counters[GC_POINTS];
    *n = total;
    if (!out) return LO_OK;
    if (cap < total) return LO_E_CAPACITY;
    if (total == 0) return LO_OK;
    if (order == 1 && g->cleaned) {
        set_error("reference iteration order is only emulated for grids without radius_cleanup");
        return LO_E_ARG;
    }
    const size_t rec = (mode == 0) ? sizeof(PointNormal) : sizeof(PointXYZ);
    LO_TRY(ctx->out_a.ensure(total * rec));
    uint32_t* d_count = ctx->small.as<uint32_t>() + 64;
    if (order == 1) LO_TRY(robin_export_dev(g, mode, ctx->out_a.p, d_count));
    else LO_TRY(grid_export_dev(g, mode, ctx->out_a.p, d_count));
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, total * rec, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}
