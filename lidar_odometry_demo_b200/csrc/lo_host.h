// Host-side plumbing shared by the translation units of liblo_b200: error reporting, the context
// (device, stream, scratch), growable device buffers and the grid object. Plain CUDA runtime; no torch.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/lo_b200.h"
#include "lo_device.cuh"

namespace lo {

void set_error(const char* fmt, ...);

#define LO_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            lo::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return LO_E_CUDA;                                                                      \
        }                                                                                          \
    } while (0)

#define LO_TRY(expr)               \
    do {                           \
        int _s = (expr);           \
        if (_s != LO_OK) return _s; \
    } while (0)

#define LO_LAUNCHED(ctx)                                                                           \
    do {                                                                                           \
        (ctx)->launches++;                                                                         \
        cudaError_t _e = cudaPeekAtLastError();                                                    \
        if (_e != cudaSuccess) {                                                                   \
            lo::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
            return LO_E_CUDA;                                                                      \
        }                                                                                          \
    } while (0)

struct DevBuf {  // growable device scratch buffer
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);
    void release();
    template <typename T>
    T* as() const { return static_cast<T*>(p); }
};

struct PinnedBuf {  // growable pinned host staging buffer
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);
    void release();
    template <typename T>
    T* as() const { return static_cast<T*>(p); }
};

inline unsigned blocks_for(size_t n, unsigned threads) { return static_cast<unsigned>((n + threads - 1) / threads); }

}  // namespace lo

struct lo_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    uint64_t launches = 0;
    lo::DevBuf in_a, in_b;        // staged inputs (host-pointer API)
    lo::DevBuf out_a, out_b;      // staged outputs
    lo::DevBuf tmp_a, tmp_b, tmp_c, tmp_d;  // per-op scratch
    lo::DevBuf small;             // small result block (counters, normal equations, poses)
    lo::PinnedBuf pin_in, pin_out;
    int align_blocks = 0;         // co-resident blocks of the align kernel (computed once)
};

// Device counters of a grid (one uint32 array)
enum GridCounter {
    GC_VOXELS = 0,     // live voxels
    GC_POINTS = 1,     // stored points
    GC_BUMP = 2,       // next never-used payload id
    GC_FREE_TOP = 3,   // entries in the free list (signed, may dip below 0 transiently)
    GC_TOMBS = 4,      // tombstoned slots
    GC_RANGE_ERR = 5,  // set if a voxel index was outside the packable range
    GC_TICKET = 6,     // arrival tickets handed out so far (points ever offered to add_cloud)
    GC_EXPORT = 7,     // export cursor
    GC_COUNT = 8
};

struct lo_grid {
    lo_ctx* ctx = nullptr;
    float voxel_size = 0.5f;
    size_t max_points = 10;
    uint32_t shard_rank = 0, shard_world = 1;  // lo_grid_set_shard
    uint32_t width = 10;         // slab stride: the largest max(max_points, 1) the grid has had
    uint32_t capacity = 0;       // slots (power of two)
    uint32_t payload_cap = 0;    // payload ids allocated
    lo::Slot* slots = nullptr;
    float4* points = nullptr;
    float4* normals = nullptr;
    uint32_t* free_list = nullptr;  // [payload_cap]
    uint32_t* counters = nullptr;   // [GC_COUNT] device
    uint32_t h_counters[GC_COUNT] = {0, 0, 0, 0, 0, 0, 0, 0};  // host mirror (valid after refresh)
    bool cleaned = false;        // radius_cleanup ran: reference iteration order can no longer be emulated
    uint64_t rebuild_count = 0;  // table rebuilds (growth or tombstone purge) since creation
    uint64_t grow_count = 0;     // payload growths since creation

    lo::GridView view() const {
        lo::GridView v;
        v.slots = slots; v.points = points; v.normals = normals;
        v.mask = capacity - 1; v.width = width; v.voxel_size = voxel_size;
        v.cap = static_cast<uint32_t>(max_points ? (max_points < width ? max_points : width) : 1);
        v.shard_rank = shard_rank; v.shard_world = shard_world;
        return v;
    }
};

namespace lo {
// grid internals used across translation units (lo_grid.cu)
int grid_reserve(lo_grid* g, size_t incoming_points);             // make room for `incoming_points` more inserts
// n = host-side upper bound of the point count; d_n (nullable) = device-side actual count
int grid_add_dev(lo_grid* g, const void* d_points, size_t n, int has_normals, const Pose* d_pose_or_null, const uint32_t* d_n);
int grid_refresh_counters(lo_grid* g);                            // D2H of the counter block (synchronises)
uint32_t grid_take_tickets(lo_grid* g, uint32_t n);               // arrival tickets of the next n offered points
int grid_snapshot(lo_grid* g, void* out, size_t cap, size_t* bytes);
int grid_restore(lo_grid* g, const void* blob, size_t bytes);
int grid_clear(lo_grid* g);
int grid_cleanup_dev(lo_grid* g, const float* d_center3, float radius);
int grid_export_dev(lo_grid* g, int mode, void* d_out, uint32_t* d_count);
int robin_export_dev(lo_grid* g, int mode, void* d_out, uint32_t* d_count);   // lo_robin.cu
int robin_export_bounded(lo_grid* g, int mode, void* d_out, uint32_t* d_count, uint32_t bound);
// lo_match.cu: enqueue CloudMatcher::align; d_m (nullable) = device-side query count, m = host upper bound
int align_enqueue(lo_grid* kf, const PointXYZ* d_queries, uint32_t m, const uint32_t* d_m, const float* d_guess7, float* d_out7,
                  lo_align_stats* d_stats);
}  // namespace lo
