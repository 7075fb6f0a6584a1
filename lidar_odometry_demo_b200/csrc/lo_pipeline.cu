// LidarOdometry::processCloud (src/lidar_odometry.cpp:22-76) as ONE persistent cooperative kernel per scan, for scans
// of up to kFusedMaxPoints points (a VLP16 revolution is ~30 k). One block per SM; the stages below are separated by
// grid-wide barriers instead of kernel launches, poses and counts never leave the device (not even the decision whether
// this is the bootstrap scan, so the host can queue a scan behind the one in flight), and the kernel ends by writing
// an 88-byte result block straight into pinned host memory.
//
//   P1 prep        constant-velocity prediction (cpp:27-28,51) | min/max of `time` | clear scratch and the two
//                  down-sampler tables
//   P2 deskew      pointTimeNormalize + transformNonRigid (cpp:25,30) | ring histogram of the classifier
//   P3 cell_fill   classifier grid: ring x azimuth bin, last writer wins (cloud_classifier.h:21-71)
//   P4 curvature   gather cell points + LOAM curvature (cloud_classifier.h:81-105)
//   P5 planar      planar test + normals (cloud_classifier.h:107-165) fused with rangeFilter (cpp:35) -> flags, block counts
//   P6 scatter     ordered compaction into the filtered planar cloud | insert every point into both
//                  VoxelGrid(vs,1) down-samplers (cpp:37-38,46-47): find-or-create voxel, atomicMin of the arrival index
//                  (the voxels created = the sizes of the update set U and the query set M)
//   P7 select      which points are the first of their voxel (winners): U winners get their tsl::robin_map bucket
//                  (histogram), M winners are counted per block
//   P8 emit        query cloud M (arrival order) | start of every robin_map bucket
//   P9 place       the update set in the reference's robin_map iteration order: counting sort by bucket,
//                  tie groups ordered by arrival; the insert below reads U through the wrap rotation
//   align          CloudMatcher::align (lo_align.cuh), cpp:49-51
//   gate           Euler-angle divergence gate (cpp:53-65), evaluated by every block on the identical result
//   clean-up       keyframe_.radiusCleanup (cpp:67)
//   insert         transformWithNormals + keyframe_.addCloud (cpp:69-70): locate, ticket chain, write (lo_grid.cuh)
//   read-back      pose + counters into pinned host memory, sequence word last
//
// Every per-point formula is the shared code of lo_stages.cuh / lo_grid.cuh / lo_align.cuh, so this kernel and the
// stage-by-stage entry points (lo_points.cu, lo_grid.cu, lo_match.cu) produce identical bits; the GPU tests run both
// against the oracle.
#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "lo_align.cuh"
#include "lo_grid.cuh"
#include "lo_odometry.cuh"
#include "lo_robin_grid.cuh"
#include "lo_stages.cuh"

namespace lo {

constexpr int kPipeThreads = kAlignThreads;   // the whole scan runs in align's launch shape
constexpr uint32_t kRobinBins = 256;       // coarse bins over the robin_map bucket array (<= kPipeThreads), see P9
constexpr uint32_t kMaxPipeBlocks = 512;   // upper bound of the chunked grids (block_counts rows)

struct alignas(16) DsSlot {                // slot of a VoxelGrid(vs, 1) down-sampler: the voxel's first point
    unsigned long long key;                // packed voxel key or kKeyEmpty
    uint32_t ticket;                       // smallest arrival index (position in the filtered cloud) seen so far
    uint32_t pad;
};

struct PipeBufs {                          // carved out of lo_odometry::fused
    uint32_t* hist;                        // [256] ring populations
    uint32_t* layout;                      // [4]   W, H, overflow, W*H
    uint32_t* block_counts;                // [3][kMaxPipeBlocks] planar / U winners / M winners per block
    uint32_t* cell_idx;                    // [cell_cap] 1 + index of the point that owns the cell
    float* curv;                           // [cell_cap]
    float4* cellp;                         // [cell_cap] x,y,z,intensity of the cell's point
    float4* cnormal;                       // [cell_cap] normal, w = 1 if the cell is planar and in range
    DsSlot* ds_u;                          // [ds_cap] keyframe_downsampler   (cpp:37)
    DsSlot* ds_m;                          // [ds_cap] matching_downsampler   (cpp:46)
    uint32_t* slot_u;                      // [n] slot of filtered point i in ds_u
    uint32_t* slot_m;                      // [n]
    uint32_t* ubucket;                     // [n] ideal robin_map bucket of filtered point i if it is an update winner, else ~0
    uint32_t* uarr;                        // [n] update winners in arrival order (index into filtered) = the reference's insertion order
    uint32_t* uhash;                       // [n] their 22-bit reference hashes
    uint32_t* robin_grid;                  // [robin_grid_words(n)] scratch of the grid-wide replay (update sets of more than 2^14 points)
    uint32_t* usorted;                     // [n] update winners in robin_map order (index into filtered)
    uint32_t* umember;                     // [n] bucket members before a tie group is ordered
    uint32_t* rb_hist;                     // [rb_cap] winners per ideal bucket, then exclusive start of each bucket
    uint32_t* rb_fill;                     // [rb_cap] placement cursors
    uint32_t* rb_done;                     // [rb_cap] members of a tie group that have been placed
    uint32_t* rb_coarse;                   // [kRobinBins] update winners per coarse bin of buckets
    int* rb_shift;                         // [1] max over sorted elements of (bucket - position)
    uint32_t* robin_flags;                 // [0] a tie group with two members in one key-frame voxel was seen; [2] the rest of the
                                           //     insert is left to scan_replay_kernel (1: with the pose, 2: bootstrap scan)
    uint32_t* kf_slot;                     // [n] key-frame insert: slot of update point i
    uint32_t* kf_base;                     // [n] key-frame insert: first free entry of its voxel
    float4* pair_o;                        // [n] align: plane origin per query (w = valid)
    float4* pair_n;                        // [n] align: plane normal per query
    double* partials;                      // [2][kMaxPipeBlocks][32] align: per-block partial sums
    uint32_t cell_cap, ds_cap, rb_cap;
};

static size_t carve(PipeBufs& b, char* base, uint32_t n) {
    b.cell_cap = std::max<uint32_t>(8u * n, 65536u);
    uint32_t cap = 1024;
    while (cap < 2u * n) cap <<= 1;
    b.ds_cap = cap;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* p = base ? base + off : nullptr;
        off += (bytes + 255) & ~static_cast<size_t>(255);
        return p;
    };
    b.hist = reinterpret_cast<uint32_t*>(take(kRobinBins * 4));
    b.layout = reinterpret_cast<uint32_t*>(take(16));
    b.block_counts = reinterpret_cast<uint32_t*>(take(3 * kMaxPipeBlocks * 4));
    b.cell_idx = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(b.cell_cap) * 4));
    b.curv = reinterpret_cast<float*>(take(static_cast<size_t>(b.cell_cap) * 4));
    b.cellp = reinterpret_cast<float4*>(take(static_cast<size_t>(b.cell_cap) * 16));
    b.cnormal = reinterpret_cast<float4*>(take(static_cast<size_t>(b.cell_cap) * 16));
    b.ds_u = reinterpret_cast<DsSlot*>(take(static_cast<size_t>(b.ds_cap) * 16));
    b.ds_m = reinterpret_cast<DsSlot*>(take(static_cast<size_t>(b.ds_cap) * 16));
    b.slot_u = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.slot_m = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.ubucket = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.uarr = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.uhash = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.robin_grid = reinterpret_cast<uint32_t*>(take(robin_grid_words(n) * 4));
    b.usorted = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.umember = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    uint32_t rb = 2;
    while (rb < 2u * n) rb <<= 1;
    b.rb_cap = rb;
    b.rb_hist = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(rb) * 4 + 16));
    b.rb_fill = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(rb) * 4));
    b.rb_done = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(rb) * 4));
    b.rb_shift = reinterpret_cast<int*>(take(16));
    b.robin_flags = reinterpret_cast<uint32_t*>(take(16));
    b.rb_coarse = reinterpret_cast<uint32_t*>(take(kRobinBins * 4));
    b.kf_slot = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.kf_base = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(n) * 4));
    b.pair_o = reinterpret_cast<float4*>(take(static_cast<size_t>(n) * 16));
    b.pair_n = reinterpret_cast<float4*>(take(static_cast<size_t>(n) * 16));
    b.partials = reinterpret_cast<double*>(take(static_cast<size_t>(2) * kMaxPipeBlocks * 32 * 8));
    return off;
}

// ---------------------------------------------------------------------------------------------------
// P1. One thread does the pose prediction; everybody reduces min/max(time) and clears scratch.
__device__ inline void phase_prep(OdomDev* s, const PointXYZIRT* __restrict__ in, uint32_t n,
                                                                  uint32_t* __restrict__ minmax, PipeBufs b) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    if (tid == 0) {  // lidar_odometry.cpp:27-30,51
        const Pose rel = pose_relative_to(s->previous, s->current);
        s->previous = s->current;
        s->relative = rel;
        s->relative_inv = pose_inverse(rel);
        s->guess = pose_compose(s->current, rel);
        s->unstable = 0;
        s->n_planar = s->n_update = s->n_query = 0;
        s->align.outer_iterations = 0;
    }
    uint32_t lo_v = 0xFFFFFFFFu, hi_v = 0u;
    for (uint32_t i = tid; i < n; i += nthreads) {
        const float t = in[i].time;
        if (t == t) {  // std::min / std::max keep the running value when the operand is NaN
            const uint32_t o = float_to_ordered(t);
            lo_v = min(lo_v, o);
            hi_v = max(hi_v, o);
        }
    }
    for (int off = 16; off > 0; off >>= 1) {
        lo_v = min(lo_v, __shfl_xor_sync(0xFFFFFFFFu, lo_v, off));
        hi_v = max(hi_v, __shfl_xor_sync(0xFFFFFFFFu, hi_v, off));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&minmax[0], lo_v);
        atomicMax(&minmax[1], hi_v);
    }
    if (tid < 256) b.hist[tid] = 0;
    if (tid < kRobinBins) b.rb_coarse[tid] = 0;
    if (tid == 0) {
        *b.rb_shift = INT_MIN;
        b.robin_flags[0] = b.robin_flags[1] = b.robin_flags[2] = 0u;
    }
    for (uint32_t c = tid; c < b.rb_cap; c += nthreads) {
        b.rb_hist[c] = 0;
        b.rb_done[c] = 0;
    }
    for (uint32_t c = tid; c < b.cell_cap; c += nthreads) b.cell_idx[c] = 0;
    const uint4 empty = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
    uint4* u = reinterpret_cast<uint4*>(b.ds_u);
    uint4* m = reinterpret_cast<uint4*>(b.ds_m);
    for (uint32_t c = tid; c < b.ds_cap; c += nthreads) {
        u[c] = empty;
        m[c] = empty;
    }
}

// P2. deskew + ring histogram (the reference keys its ring map by uint8_t, cloud_classifier.h:21-23)
__device__ inline void phase_deskew(const OdomDev* __restrict__ s, const PointXYZIRT* __restrict__ in, uint32_t n,
                                                                    const uint32_t* __restrict__ minmax, PointXYZIRT* __restrict__ out,
                                                                    uint32_t* __restrict__ hist) {
    __shared__ uint32_t local[256];
    if (threadIdx.x < 256) local[threadIdx.x] = 0;
    __syncthreads();
    const Pose st = s->relative_inv, en = s->identity;   // transformNonRigid(cloud, relative.inverse(), Pose3D()) cpp:30
    const float t_min = time_min_of(minmax), t_max = time_max_of(minmax);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(in)[2 * i];
        float4 bb = reinterpret_cast<const float4*>(in)[2 * i + 1];
        const float4 o = deskew_point(st, en, t_min, t_max - t_min, a, bb);
        reinterpret_cast<float4*>(out)[2 * i] = o;
        reinterpret_cast<float4*>(out)[2 * i + 1] = bb;
        atomicAdd(&local[__float_as_uint(bb.y) & 0xFF], 1u);   // ring is the low 16 bits of the second word
    }
    __syncthreads();
    if (threadIdx.x < 256 && local[threadIdx.x]) atomicAdd(&hist[threadIdx.x], local[threadIdx.x]);
}

// Block-wide layout of the classifier grid from the ring histogram: W = largest ring population, row of ring r =
// number of populated rings below r (std::map order), H = populated rings. Needs blockDim.x >= 256.
__device__ inline void block_layout(const uint32_t* __restrict__ hist, uint32_t* row_of_s, uint32_t* wh_s) {
    __shared__ uint32_t warp_cnt[8], warp_max[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned present = 0;
    if (threadIdx.x < 256) {   // whole warps: 8 of them
        const uint32_t c = hist[threadIdx.x];
        present = __ballot_sync(0xFFFFFFFFu, c != 0);
        uint32_t mx = c;
        for (int off = 16; off > 0; off >>= 1) mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, off));
        if (lane == 0) {
            warp_cnt[warp] = __popc(present);
            warp_max[warp] = mx;
        }
    }
    __syncthreads();
    if (threadIdx.x < 256) {
        uint32_t before = 0, total = 0, width = 0;
        for (int w = 0; w < 8; w++) {
            if (w < warp) before += warp_cnt[w];
            total += warp_cnt[w];
            width = max(width, warp_max[w]);
        }
        row_of_s[threadIdx.x] = before + __popc(present & ((1u << lane) - 1u));
        if (threadIdx.x == 0) {
            wh_s[0] = width;
            wh_s[1] = total;
        }
    }
    __syncthreads();
}

// P3. cloud_classifier.h:44-56: cell (ring row, azimuth bin) <- point, last writer (largest input index) wins
__device__ inline void phase_cell_fill(const PointXYZIRT* __restrict__ desk, uint32_t n, PipeBufs b) {
    __shared__ uint32_t row_of[256], wh[2];
    block_layout(b.hist, row_of, wh);
    const uint32_t W = wh[0], H = wh[1];
    const bool overflow = static_cast<unsigned long long>(W) * H > b.cell_cap;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        b.layout[0] = W;
        b.layout[1] = H;
        b.layout[2] = overflow ? 1u : 0u;
        b.layout[3] = overflow ? 0u : W * H;
    }
    if (overflow) return;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(desk)[2 * i];
        const uint32_t ring = desk[i].ring & 0xFF;
        uint32_t col;
        if (azimuth_column(a.x, a.y, W, col)) atomicMax(&b.cell_idx[row_of[ring] * W + col], i + 1);
    }
}

__device__ inline float4 cell_point(const PointXYZIRT* __restrict__ desk, const uint32_t* __restrict__ cell_idx, uint32_t c) {
    const uint32_t idx = cell_idx[c];
    if (!idx) return make_float4(0.f, 0.f, 0.f, 0.f);  // PointType(): value-initialised
    const float4 a = reinterpret_cast<const float4*>(desk)[2 * (idx - 1)];
    const float4 bb = reinterpret_cast<const float4*>(desk)[2 * (idx - 1) + 1];
    return make_float4(a.x, a.y, a.z, bb.x);
}

// P4. gather + curvature
__device__ inline void phase_curvature(const PointXYZIRT* __restrict__ desk, PipeBufs b) {
    const uint32_t total = b.layout[3];
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < total; c += gridDim.x * blockDim.x) {
        b.cellp[c] = cell_point(desk, b.cell_idx, c);
        b.curv[c] = cell_curvature(c, total, [&](uint32_t k) { return cell_point(desk, b.cell_idx, k); });
    }
}

// ---------------------------------------------------------------------------------------------------
// Ordered (stable) compaction in two passes over contiguous per-block chunks: pass A counts the kept items of
// each block, pass B adds the counts of the preceding blocks and ranks inside the block.
__device__ inline void chunk_range(uint32_t total, uint32_t& begin, uint32_t& end) {
    uint32_t per = (total + gridDim.x - 1) / gridDim.x;
    per = (per + blockDim.x - 1) / blockDim.x * blockDim.x;  // whole tiles of blockDim.x items
    begin = min(total, blockIdx.x * per);
    end = min(total, begin + per);
}

// sum of counts[0..blockIdx.x) and of all gridDim.x counts; result valid in every thread. The (<= kMaxPipeBlocks) counts
// are read by one warp with all its loads in flight at once: one round trip to L2, not one per 32 counts (the dependent
// rounds cost ~4 k cycles per call, five calls per scan).
__device__ inline void block_prefix(const uint32_t* __restrict__ counts, uint32_t& before, uint32_t& all) {
    __shared__ uint32_t sh[2];
    if (threadIdx.x < 32) {
        uint32_t v[kMaxPipeBlocks / 32];
#pragma unroll
        for (uint32_t u = 0; u < kMaxPipeBlocks / 32; u++) {
            const uint32_t k = threadIdx.x + 32 * u;
            v[u] = k < gridDim.x ? __ldcg(counts + k) : 0u;
        }
        uint32_t b = 0, a = 0;
#pragma unroll
        for (uint32_t u = 0; u < kMaxPipeBlocks / 32; u++) {
            a += v[u];
            if (threadIdx.x + 32 * u < blockIdx.x) b += v[u];
        }
        for (int off = 16; off > 0; off >>= 1) {
            b += __shfl_xor_sync(0xFFFFFFFFu, b, off);
            a += __shfl_xor_sync(0xFFFFFFFFu, a, off);
        }
        if (threadIdx.x == 0) {
            sh[0] = b;
            sh[1] = a;
        }
    }
    __syncthreads();
    before = sh[0];
    all = sh[1];
    __syncthreads();
}

// number of threads with flag set before this thread in the block, and the block's total
__device__ inline uint32_t block_rank(bool flag, uint32_t& tile_total) {
    __shared__ uint32_t warp_cnt[32];
    const unsigned m = __ballot_sync(0xFFFFFFFFu, flag);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    if (lane == 0) warp_cnt[warp] = __popc(m);
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (int w = 0; w < nwarps; w++) {
        if (w < warp) before += warp_cnt[w];
        total += warp_cnt[w];
    }
    __syncthreads();
    tile_total = total;
    return before + __popc(m & ((1u << lane) - 1u));
}

// P5. planar test + normal + range filter -> cnormal.w flag, per-block counts
__device__ inline void phase_planar(PipeBufs b, float lo2, float hi2) {
    const uint32_t total = b.layout[3], W = b.layout[0];
    uint32_t begin, end;
    chunk_range(total, begin, end);
    uint32_t count = 0;
    for (uint32_t base = begin; base < end; base += blockDim.x) {
        const uint32_t c = base + threadIdx.x;
        bool flag = false;
        if (c < end) {
            float nv[3] = {0.f, 0.f, 0.f};
            if (cell_planar(c, W, b.cellp, b.curv, nv)) {
                const float4 p = b.cellp[c];
                flag = in_range(p.x, p.y, p.z, lo2, hi2);
            }
            b.cnormal[c] = make_float4(nv[0], nv[1], nv[2], flag ? 1.f : 0.f);
        }
        count += __syncthreads_count(flag);
    }
    if (threadIdx.x == 0) b.block_counts[blockIdx.x] = count;
}

// find-or-create the voxel of (x,y,z) in a down-sampler table and record arrival index `pos`; returns the slot.
// `created`: this call created the voxel (exactly one call per voxel does).
__device__ inline uint32_t ds_insert(DsSlot* __restrict__ slots, uint32_t mask, float voxel_size, float x, float y, float z, uint32_t pos,
                                     uint32_t* __restrict__ range_err, bool& created) {
    created = false;
    const long long ix = axis_index(x, voxel_size), iy = axis_index(y, voxel_size), iz = axis_index(z, voxel_size);
    if (!axis_in_range(ix) || !axis_in_range(iy) || !axis_in_range(iz)) {
        *range_err = 1;
        return 0xFFFFFFFFu;
    }
    const unsigned long long key = pack_key(ix, iy, iz);
    // Not probe_hash: that one gives the 8 voxels of a brick 8 consecutive slots, which pays for the key frame's 27-neighbour
    // search; a down-sampler is only inserted into, and bricks that collide make runs of occupied slots (its inserts took
    // 7-10 k cycles per 512 points; neighbouring points of a ring are neighbouring voxels). murmur3's 64-bit finaliser.
    unsigned long long mix = key;
    mix ^= mix >> 33; mix *= 0xFF51AFD7ED558CCDull;
    mix ^= mix >> 33; mix *= 0xC4CEB9FE1A85EC53ull;
    mix ^= mix >> 33;
    uint32_t h = static_cast<uint32_t>(mix) & mask;
    while (true) {
        unsigned long long k = slots[h].key;
        if (k == kKeyEmpty) {
            k = atomicCAS(&slots[h].key, kKeyEmpty, key);
            created = k == kKeyEmpty;
        }
        if (k == kKeyEmpty || k == key) break;
        h = (h + 1) & mask;
    }
    atomicMin(&slots[h].ticket, pos);   // VoxelGrid(vs,1): the first arrival owns the voxel (voxel_grid.h:82-91)
    return h;
}

// P6. filtered planar cloud in cell order (cloud_classifier.h:114-165 walks rays then columns) + both down-samplers
__device__ inline void phase_scatter(OdomDev* s, PipeBufs b, PointNormal* __restrict__ filtered, float vs_u,
                                                                     float vs_m, int with_match, uint32_t* __restrict__ range_err) {
    const uint32_t total = b.layout[3];
    uint32_t begin, end, before, all;
    chunk_range(total, begin, end);
    block_prefix(b.block_counts, before, all);
    if (blockIdx.x == 0 && threadIdx.x == 0) s->n_planar = all;
    uint32_t running = before, made_u = 0, made_m = 0;
    for (uint32_t base = begin; base < end; base += blockDim.x) {
        const uint32_t c = base + threadIdx.x;
        float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < end) nv = b.cnormal[c];
        const bool flag = nv.w != 0.f;
        uint32_t tile_total;
        const uint32_t pos = running + block_rank(flag, tile_total);
        running += tile_total;
        bool new_u = false, new_m = false;
        if (flag) {
            const float4 p = b.cellp[c];
            float4* o = reinterpret_cast<float4*>(filtered + pos);
            o[0] = make_float4(p.x, p.y, p.z, 1.f);
            o[1] = make_float4(nv.x, nv.y, nv.z, 0.f);
            o[2] = make_float4(0.f, 0.f, 0.f, 0.f);
            b.slot_u[pos] = ds_insert(b.ds_u, b.ds_cap - 1, vs_u, p.x, p.y, p.z, pos, range_err, new_u);
            if (with_match) b.slot_m[pos] = ds_insert(b.ds_m, b.ds_cap - 1, vs_m, p.x, p.y, p.z, pos, range_err, new_m);
        }
        made_u += __syncthreads_count(new_u);
        made_m += __syncthreads_count(new_m);
    }
    // the sizes of the two voxel sets = voxels created (known one phase before their winners are)
    if (threadIdx.x == 0) {
        if (made_u) atomicAdd(&s->n_update, made_u);
        if (made_m) atomicAdd(&s->n_query, made_m);
    }
}

constexpr uint32_t kNoBucket = 0xFFFFFFFFu;

__device__ inline uint32_t robin_bin_shift(uint32_t buckets) {   // log2(buckets per coarse bin)
    uint32_t shift = 0;
    while ((buckets >> shift) > kRobinBins) shift++;
    return shift;
}

__device__ inline uint32_t robin_bucket_count(uint32_t n) {  // tsl::robin_map after n insertions: 2 * pow2ceil(n)
    uint32_t c = 2;
    while (c < 2 * n) c <<= 1;
    return c;
}

// P7. winners of the two down-samplers: update winners get their ideal robin_map bucket (histogrammed), matching
// winners are counted per block for the ordered compaction of P8
// do_hist: the update winners get their ideal robin_map bucket and are histogrammed (the counting-sort order, used where the
// exact block-local replay does not apply); do_count: the winners of both down-samplers are counted per block for the
// ordered compactions of P8 (queries for align; update winners in arrival order for the replay)
__device__ inline void phase_select(const OdomDev* s, PipeBufs b, int with_match, bool do_hist, bool do_count, bool count_update = false) {
    const uint32_t total = __ldcg(&s->n_planar);   // written by block 0 in P6, at least one grid-wide step ago (L2: the line may sit stale in L1)
    const uint32_t buckets = robin_bucket_count(s->n_update), bin_shift = robin_bin_shift(buckets);
    uint32_t begin, end;
    chunk_range(total, begin, end);
    uint32_t cm = 0, cu = 0;
    for (uint32_t base = begin; base < end; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        bool wm = false, wu = false;
        if (i < end) {
            if (do_hist || count_update) {
                const uint32_t su = b.slot_u[i];
                wu = su != 0xFFFFFFFFu && b.ds_u[su].ticket == i;
            }
            if (do_hist) {
                const uint32_t su = b.slot_u[i];
                uint32_t bucket = kNoBucket;
                if (wu) {
                    long long ix, iy, iz;
                    unpack_key(b.ds_u[su].key, ix, iy, iz);
                    bucket = reference_hash22(ix, iy, iz) & (buckets - 1);
                    atomicAdd(&b.rb_hist[bucket], 1u);
                    atomicAdd(&b.rb_coarse[bucket >> bin_shift], 1u);
                }
                b.ubucket[i] = bucket;
            }
            if (with_match && do_count) {
                const uint32_t sm = b.slot_m[i];
                wm = sm != 0xFFFFFFFFu && b.ds_m[sm].ticket == i;
            }
        }
        if (do_count) cm += __syncthreads_count(wm);
        if (count_update) cu += __syncthreads_count(wu);
    }
    if (threadIdx.x == 0 && do_count) b.block_counts[2 * kMaxPipeBlocks + blockIdx.x] = cm;
    if (threadIdx.x == 0 && count_update) b.block_counts[1 * kMaxPipeBlocks + blockIdx.x] = cu;
}

// exclusive prefix of v over the block's threads, and the block total
__device__ inline uint32_t block_exclusive_scan(uint32_t v, uint32_t& total) {
    __shared__ uint32_t warp_sums[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    uint32_t incl = v;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, incl, off);
        if (lane >= off) incl += u;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t before = 0, all = 0;
    for (int w = 0; w < nwarps; w++) {
        const uint32_t ws = warp_sums[w];
        if (w < warp) before += ws;
        all += ws;
    }
    __syncthreads();
    total = all;
    return before + incl - v;
}

// ---------------------------------------------------------------------------------------------------
// P8. (a) query cloud (matching_downsampler.getCloudWithoutNormals(), cpp:50) in arrival order — its order only decides
// the order of the residual sum; (b) start of every robin_map bucket.
// keyframe_downsampler.getCloud() (cpp:38,42,69) is the update set in tsl::robin_map iteration order. For n inserted keys the
// map has C = 2*pow2ceil(n) buckets; iteration is ascending bucket position; robin-hood probing keeps elements sorted
// by ideal bucket b = hash22 & (C-1); a cluster that runs past the last bucket wraps to the front (lo_robin.cu states
// the same rule for the stage-by-stage path). Counting sort by b: P7 histogrammed the buckets, also into <= kRobinBins
// coarse bins, so every bin is scanned here by one block on its own (start of the bin = sum of the coarse counts before
// it); P9 places the winners and orders every tie group by arrival; the insert reads them through the wrap rotation.
__device__ inline void phase_emit(const OdomDev* s, PipeBufs b, const PointNormal* __restrict__ filtered,
                                  PointXYZ* __restrict__ query, int with_match, bool do_u, bool do_m, bool arrival_update = false) {
    uint32_t bm = 0, nm = 0, bu = 0, nu = 0;
    if (do_m || arrival_update) {   // ordered compactions: the query cloud (if there is a key frame to match); the update winners with
                                    // their hashes in arrival order (the reference map's insertion sequence, for the replay)
        const uint32_t total = __ldcg(&s->n_planar);
        if (do_m) block_prefix(b.block_counts + 2 * kMaxPipeBlocks, bm, nm);
        if (arrival_update) block_prefix(b.block_counts + 1 * kMaxPipeBlocks, bu, nu);
        uint32_t begin, end;
        chunk_range(total, begin, end);
        for (uint32_t base = begin; base < end; base += blockDim.x) {
            const uint32_t i = base + threadIdx.x;
            bool wm = false, wu = false;
            uint32_t su = 0xFFFFFFFFu;
            if (i < end) {
                if (arrival_update) {
                    su = b.slot_u[i];
                    wu = su != 0xFFFFFFFFu && b.ds_u[su].ticket == i;
                }
                if (with_match && do_m) {
                    const uint32_t sm = b.slot_m[i];
                    wm = sm != 0xFFFFFFFFu && b.ds_m[sm].ticket == i;
                }
            }
            uint32_t tm = 0, tu = 0, rm = 0, ru = 0;
            if (do_m) rm = block_rank(wm, tm);
            if (arrival_update) ru = block_rank(wu, tu);
            if (wm) {
                const float4 p = *reinterpret_cast<const float4*>(filtered + i);
                reinterpret_cast<float4*>(query)[bm + rm] = make_float4(p.x, p.y, p.z, 1.f);
            }
            if (wu) {
                long long ix, iy, iz;
                unpack_key(b.ds_u[su].key, ix, iy, iz);
                b.uarr[bu + ru] = i;
                b.uhash[bu + ru] = reference_hash22(ix, iy, iz);
            }
            bm += tm;
            bu += tu;
        }
    }
    const uint32_t n = s->n_update;
    if (n == 0 || !do_u) return;
    const uint32_t C = robin_bucket_count(n);
    const uint32_t shift = robin_bin_shift(C), bins = C >> shift, per = 1u << shift;   // bins <= kRobinBins <= blockDim.x
    for (uint32_t bin = blockIdx.x; bin < bins; bin += gridDim.x) {
        uint32_t run;
        block_exclusive_scan(threadIdx.x < bin ? b.rb_coarse[threadIdx.x] : 0u, run);   // winners in the bins before this one
        // the bin's buckets in chunks of blockDim.x (one chunk for a VLP16 scan; 128-beam scans have up to 2048 buckets per bin)
        for (uint32_t off = 0; off < per; off += blockDim.x) {
            const uint32_t k = off + threadIdx.x, c = (bin << shift) + k;
            const uint32_t v = k < per ? b.rb_hist[c] : 0u;
            uint32_t chunk_total;
            const uint32_t start = run + block_exclusive_scan(v, chunk_total);
            if (k < per) {
                b.rb_hist[c] = start;    // exclusive start of bucket c
                b.rb_fill[c] = start;
            }
            if (bin == bins - 1 && k == per - 1) b.rb_hist[C] = start + v;   // = n
            run += chunk_total;
        }
    }
}

// P9. every update winner goes to its bucket's segment; the member that completes a tie group orders it by arrival
// (= index in the filtered cloud). Sorted element k sits at bucket position k + max_{j<=k}(b_j - j): rb_shift collects
// the maximum, which gives the wrap of the last cluster.
__device__ inline void phase_robin_place(const OdomDev* s, PipeBufs b) {
    const uint32_t total = __ldcg(&s->n_planar);
    int shift = INT_MIN;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const uint32_t bk = b.ubucket[i];
        if (bk == kNoBucket) continue;
        const uint32_t seg_begin = b.rb_hist[bk], seg_end = b.rb_hist[bk + 1];
        if (seg_end - seg_begin == 1) {
            b.usorted[seg_begin] = i;
            shift = max(shift, static_cast<int>(bk) - static_cast<int>(seg_begin));
            continue;
        }
        b.umember[atomicAdd(&b.rb_fill[bk], 1u)] = i;
        __threadfence();
        if (atomicAdd(&b.rb_done[bk], 1u) + 1 != seg_end - seg_begin) continue;
        __threadfence();   // the last member to arrive sees all the others
        for (uint32_t a = seg_begin; a < seg_end; a++) {
            const uint32_t ma = __ldcg(b.umember + a);
            uint32_t rank = 0;
            for (uint32_t q = seg_begin; q < seg_end; q++) rank += __ldcg(b.umember + q) < ma ? 1u : 0u;
            b.usorted[seg_begin + rank] = ma;
            shift = max(shift, static_cast<int>(bk) - static_cast<int>(seg_begin + rank));
        }
    }
    for (int off = 16; off > 0; off >>= 1) shift = max(shift, __shfl_xor_sync(0xFFFFFFFFu, shift, off));
    if ((threadIdx.x & 31) == 0 && shift != INT_MIN) atomicMax(b.rb_shift, shift);
}

// the key-frame voxel an update point goes to: the arithmetic of grid_locate_item (load_point with the pose fused in, then
// the reference's float division, voxel_grid.h:60-64); kKeyEmpty for a point the insert skips
__device__ inline unsigned long long update_voxel_key(const PointNormal* __restrict__ filtered, uint32_t i, const Pose* pose, float voxel_size) {
    const PointSource one{reinterpret_cast<const float*>(filtered), 12u, 1};
    float q[3], nrm[3];
    load_point(one, i, pose, q, nrm);
    const long long ix = axis_index(q[0], voxel_size), iy = axis_index(q[1], voxel_size), iz = axis_index(q[2], voxel_size);
    const bool finite = isfinite(q[0]) && isfinite(q[1]) && isfinite(q[2]);
    if (!finite || !axis_in_range(ix) || !axis_in_range(iy) || !axis_in_range(iz)) return kKeyEmpty;
    return pack_key(ix, iy, iz);
}

// lidar_odometry.cpp:53-65: Euler-angle divergence gate; returns the pose that becomes current_transform_
__device__ inline Pose gate_pose(const Pose& nw, const Pose& cur, const Pose& guess, float threshold, uint32_t& unstable) {
    Pose a = identity_pose(), c = identity_pose();   // new.rotation * current.rotation.inverse()
    a.w = nw.w; a.x = nw.x; a.y = nw.y; a.z = nw.z;
    c.w = cur.w; c.x = cur.x; c.y = cur.y; c.z = cur.z;
    const Pose d = pose_compose(a, pose_inverse(c));
    float r[9];
    rotation_matrix_f(d.w, d.x, d.y, d.z, r);
    float ang[3];
    euler_angles_012(r, ang);
    bool ok = true;
    const float pi_f = 3.14159265358979323846f;
    for (int k = 0; k < 3; k++) {
        const float deg = (ang[k] * 180.0f) / pi_f;
        const float m = fabsf(deg);
        ok = ok && (m < threshold || m > 180 - threshold);
    }
    unstable = ok ? 0u : 1u;
    return ok ? nw : guess;
}

struct ScanArgs {
    OdomDev* s;
    const PointXYZIRT* scan;
    PointXYZIRT* desk;
    PointNormal* filtered;
    PointXYZ* query;
    uint32_t* minmax;
    PipeBufs b;
    GridView kf;                 // key-frame grid (after the host made room for this scan's insert)
    uint32_t kf_capacity;
    uint32_t* kf_counters;
    uint32_t* kf_free_list;
    uint32_t ticket_base;        // arrival ticket of the first update point
    ScanReadback* readback;      // pinned host memory
    uint32_t seq;                // this scan's number, published in readback->seq when everything else is written
    uint32_t n;
    float lo2, hi2, vs_u, vs_m, cleanup_range, gate_threshold;
    uint32_t force_replay;       // developer switch (LO_B200_REPLAY_ALWAYS): run the block replay on every scan it can take
    long long* trace;
};

// second half of the read-back block (one thread, when the key-frame insert is through); re-arms the next scan's time min/max
__device__ inline void scan_publish_tail(const ScanArgs& p, uint32_t update_order) {
    grid_fix_counters(p.kf_counters);
    ScanReadback* out = p.readback;
    for (int k = 0; k < GC_COUNT; k++) out->kf_counters[k] = p.kf_counters[k];
    out->update_order = update_order;
    p.minmax[0] = 0xFFFFFFFFu;
    p.minmax[1] = 0u;
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t*>(&out->seq) = p.seq;
}

// kTiled: align may use the tile search (scans with more than ~19 k queries); the VLP16-sized instantiation is compiled
// without it so that its code and register allocation stay what they were tuned to.
template <bool kTiled>
__device__ __forceinline__ void scan_body(const ScanArgs& p) {
    cg::grid_group grid = cg::this_grid();
    OdomDev* s = p.s;
    // the key frame is empty: bootstrap scan (cpp:40-44). Read on the device, so the host can queue a scan before the
    // previous one has reported back; nothing changes the counter before the insert at the end.
    const int first = __ldcg(p.kf_counters + GC_VOXELS) == 0 ? 1 : 0;
    const int with_match = first ? 0 : 1;
    uint32_t* range_err = p.kf_counters + GC_RANGE_ERR;
    int trace_n = 0;
#define LO_SCAN_TRACE(tag)                                                     \
    do {                                                                       \
        if (p.trace && blockIdx.x == 0 && threadIdx.x == 0 && trace_n < 500) { \
            p.trace[2 * trace_n] = (tag);                                      \
            p.trace[2 * trace_n + 1] = clock64();                              \
            trace_n++;                                                         \
        }                                                                      \
    } while (0)
    LO_SCAN_TRACE(100);
    phase_prep(s, p.scan, p.n, p.minmax, p.b);
    grid.sync();
    LO_SCAN_TRACE(101);
    phase_deskew(s, p.scan, p.n, p.minmax, p.desk, p.b.hist);
    grid.sync();
    LO_SCAN_TRACE(102);
    phase_cell_fill(p.desk, p.n, p.b);
    grid.sync();
    LO_SCAN_TRACE(103);
    phase_curvature(p.desk, p.b);
    grid.sync();
    LO_SCAN_TRACE(104);
    phase_planar(p.b, p.lo2, p.hi2);
    grid.sync();
    LO_SCAN_TRACE(105);
    phase_scatter(s, p.b, p.filtered, p.vs_u, p.vs_m, with_match, range_err);
    grid.sync();
    LO_SCAN_TRACE(106);
    // the query set align needs: winners of the matching down-sampler, in arrival order (the bootstrap scan has none)
    if (with_match) {
        phase_select(s, p.b, with_match, false, true);
        grid.sync();
        LO_SCAN_TRACE(107);
        phase_emit(s, p.b, p.filtered, p.query, with_match, false, true);
        grid.sync();
        LO_SCAN_TRACE(108);
    }
    __shared__ float next_s[7];
    __shared__ Pose current_s;
    __shared__ uint32_t unstable_s;
    const Pose* insert_pose = nullptr;   // first scan: inserted un-transformed (cpp:42)
    // first half of the read-back block: what processCloud's caller can see (getCurrentPose) is final once the gate has
    // decided; the host returns on `pose_seq` while radiusCleanup and the key-frame insert below are still running
    auto publish_pose = [&](const Pose& now, uint32_t unstable) {
        ScanReadback* out = p.readback;
        pose_to7(now, out->pose);
        out->n_planar = s->n_planar;
        out->n_update = s->n_update;
        out->n_query = s->n_query;
        out->unstable = unstable;
        out->outer_iterations = s->align.outer_iterations;
        out->bootstrap = static_cast<uint32_t>(first);
        out->overflow = p.b.layout[2];
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(&out->pose_seq) = p.seq;
    };
    if (!first) {
        AlignArgs a;                     // CloudMatcher::align(keyframe_, query, current.compose(relative))  cpp:49-51
        a.g = p.kf;
        a.queries = p.query;
        a.m = p.n;
        a.m_dev = &s->n_query;
        a.pair_o = p.b.pair_o;
        a.pair_n = p.b.pair_n;
        a.partials = p.b.partials;
        a.guess = reinterpret_cast<const float*>(&s->guess);
        a.out_pose = reinterpret_cast<float*>(&s->next);
        a.stats = &s->align;
        a.max_dist = 0.3f;               // cloud_matcher.cpp:139
        a.trace = p.trace;
        a.trace_start = trace_n;
        a.out_pose_block = next_s;
        // large query sets (a 128-beam scan has ~10^5 queries: more than 8 per warp) take the tile search; a VLP16 scan's
        // ~6 k queries stay on the per-query search. Same winners either way (lo_tile_search.cuh).
        if constexpr (kTiled) {
            extern __shared__ __align__(16) unsigned char dyn_smem[];
            align_device(a, grid, reinterpret_cast<TileScratch*>(dyn_smem));
        } else {
            align_device(a, grid);
        }
        if (p.trace && blockIdx.x == 0 && threadIdx.x == 0) trace_n = static_cast<int>(p.trace[1000]);
        // every block runs the gate on the (identical) result: no grid-wide hand-over; block 0 stores the state at the end
        if (threadIdx.x == 0) {
            Pose nw;
            nw.t[0] = next_s[0]; nw.t[1] = next_s[1]; nw.t[2] = next_s[2];
            nw.w = next_s[3]; nw.x = next_s[4]; nw.y = next_s[5]; nw.z = next_s[6];
            current_s = gate_pose(nw, s->current, s->guess, p.gate_threshold, unstable_s);   // cpp:53-65
            if (blockIdx.x == 0) publish_pose(current_s, unstable_s);
        }
        __syncthreads();
        insert_pose = &current_s;        // transformWithNormals fused into the load (cpp:69)
    } else if (blockIdx.x == 0 && threadIdx.x == 0) {
        publish_pose(s->current, s->unstable);   // bootstrap scan: the pose does not move (cpp:40-44)
    }
    // ---- behind the pose: everything only the key frame needs -------------------------------------------------
    LO_SCAN_TRACE(120);
    // keyframe_downsampler.getCloud() (cpp:38,42,69): the update set in the iteration order of the reference's tsl::robin_map.
    // (1) Counting sort by ideal bucket with the wrap rotation: exact between different ideal buckets; a tie group (equal
    // ideal bucket) comes out in arrival order, while the map's order inside it depends on its whole history. Output
    // position d of the order is sorted element (d - rot) mod n, rot = how far the last cluster runs past the last bucket.
    // radiusCleanup (cpp:67) shares the first phase.
    const uint32_t nu = min(p.n, s->n_update);
    if (!first) {
        for (uint32_t sl = blockIdx.x * blockDim.x + threadIdx.x; sl < p.kf_capacity; sl += gridDim.x * blockDim.x)
            grid_cleanup_slot(sl, p.kf, current_s.t, p.cleanup_range, p.kf_counters, p.kf_free_list);   // cpp:67
    }
    phase_select(s, p.b, with_match, true, false);
    grid.sync();
    LO_SCAN_TRACE(121);
    phase_emit(s, p.b, p.filtered, p.query, with_match, true, false);
    grid.sync();
    LO_SCAN_TRACE(125);
    phase_robin_place(s, p.b);
    grid.sync();
    LO_SCAN_TRACE(126);
    uint32_t gather_shift = 0;
    if (nu) {
        const long long over = static_cast<long long>(nu) + *p.b.rb_shift - static_cast<long long>(robin_bucket_count(nu));
        const uint32_t rot = over > 0 ? static_cast<uint32_t>(over % nu) : 0u;
        gather_shift = (nu - rot) % nu;
    }
    // (2) keyframe_.addCloud(update) (cpp:42 / cpp:70), first pass: the key-frame voxel of every update point; the update cloud
    // is read in place through the order. The order inside a tie group reaches the key frame only where two of its members
    // fall into the SAME key-frame voxel (it decides their slots there, and which is dropped once the voxel is full,
    // voxel_grid.h:82-91): the pass looks for such a pair. None (about 39 scans in 40 of the VLP16 stream): the key frame
    // comes out exactly as from the reference's order.
    uint32_t update_order = kOrderSorted;
    const bool can_replay = gridDim.x > 1;   // one block alone could not wait for itself in the replay kernel's hand-over
    {
        const PointSource src{reinterpret_cast<const float*>(p.filtered), 12u, 1, p.b.usorted, nu, gather_shift};
        bool pair_found = false;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nu; i += gridDim.x * blockDim.x) {
            grid_locate_item(i, src, insert_pose, p.kf, p.kf_counters, p.kf_free_list, p.b.kf_slot);
            uint32_t k = i + gather_shift;
            if (k >= nu) k -= nu;
            const uint32_t me = p.b.usorted[k], bucket = p.b.ubucket[me];
            if (k + 1 < nu && p.b.ubucket[p.b.usorted[k + 1]] == bucket) {   // not the last of its tie group
                const unsigned long long mine = update_voxel_key(p.filtered, me, insert_pose, p.kf.voxel_size);
                for (uint32_t k2 = k + 1; k2 < nu; k2++) {
                    const uint32_t other = p.b.usorted[k2];
                    if (p.b.ubucket[other] != bucket) break;
                    pair_found = pair_found || (mine != kKeyEmpty && update_voxel_key(p.filtered, other, insert_pose, p.kf.voxel_size) == mine);
                }
            }
        }
        if (pair_found) p.b.robin_flags[0] = 1u;
    }
    grid.sync();
    LO_SCAN_TRACE(122);
    // (3) such a pair exists: the rest of the insert is left to scan_replay_kernel, queued right behind this kernel: one block
    // that replays the map's history in shared memory (lo_robin_block.cuh: 176 KB — as dynamic shared memory of THIS kernel
    // it would leave 28 KB of L1 for its spills, +11 % on every scan) and finishes the insert in the exact order. A scan
    // too large for the replay keeps the counting-sort order and says so in its stats.
    if (__ldcg(p.b.robin_flags) || (p.force_replay && can_replay)) {
        if (can_replay) {
            if (blockIdx.x == 0 && threadIdx.x == 0) {
                if (!first) {   // the gate's outcome becomes the state of the next scan; the replay kernel inserts with it
                    s->current = current_s;
                    s->unstable = unstable_s;
                    s->center[0] = current_s.t[0]; s->center[1] = current_s.t[1]; s->center[2] = current_s.t[2];
                }
                p.b.robin_flags[2] = first ? 2u : 1u;
                if (p.trace) p.trace[1000] = trace_n;
            }
            return;
        }
        update_order = kOrderApproximate;
    }
    const PointSource src{reinterpret_cast<const float*>(p.filtered), 12u, 1, p.b.usorted, nu, gather_shift};

    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nu; i += gridDim.x * blockDim.x)
        grid_chain_item(i, p.ticket_base, p.kf, p.b.kf_slot, p.b.kf_base);
    grid.sync();
    LO_SCAN_TRACE(123);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nu; i += gridDim.x * blockDim.x)
        grid_write_item(i, src, p.ticket_base, insert_pose, p.kf, p.kf_counters, p.b.kf_slot, p.b.kf_base);
    grid.sync();
    LO_SCAN_TRACE(124);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (!first) {                            // the gate's outcome becomes the state of the next scan
            s->current = current_s;
            s->unstable = unstable_s;
            s->center[0] = current_s.t[0]; s->center[1] = current_s.t[1]; s->center[2] = current_s.t[2];
        }
        scan_publish_tail(p, update_order);
        if (p.trace) p.trace[1000] = trace_n;
    }
#undef LO_SCAN_TRACE
}

__global__ void __launch_bounds__(kPipeThreads, 1) scan_kernel(ScanArgs p) { scan_body<false>(p); }
__global__ void __launch_bounds__(kPipeThreads, 1) scan_kernel_large(ScanArgs p) { scan_body<true>(p); }
constexpr uint32_t kSmallScanPoints = 32768;   // scans up to this size (a VLP16 revolution is ~30 k) run scan_kernel

// Queued behind every scan_kernel (same cooperative grid); returns at once unless that kernel left the insert to it (a tie
// group with two members in one key-frame voxel: about one VLP16 scan in 25). Block 0 computes the exact tsl::robin_map
// iteration order of the update set in its shared memory (lo_robin_block.cuh), then all blocks run
// keyframe_.addCloud(update) (lidar_odometry.cpp:42,70) in that order and block 0 writes the second half of the read-back.
__global__ void __launch_bounds__(kPipeThreads, 1) scan_replay_kernel(ScanArgs p) {
    const uint32_t deferred = p.b.robin_flags[2];
    if (!deferred) return;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    OdomDev* s = p.s;
    const uint32_t nu = min(p.n, s->n_update);
    // the update winners with their hashes in arrival order = the reference map's insertion sequence
    phase_select(s, p.b, 0, false, false, true);
    grid.sync();
    phase_emit(s, p.b, p.filtered, p.query, 0, false, false, true);
    grid.sync();
    __shared__ Pose pose_s;
    if (threadIdx.x == 0) pose_s = s->current;
    // up to 2^14 update points: block 0 alone, in shared memory; more (a 128-beam scan): the later epochs on the whole grid
    const RobinGridBufs rg = robin_grid_carve(p.b.robin_grid, max(p.n, 1u));
    const bool exact = robin_order_any(grid, p.b.uhash, p.b.uarr, nu, dyn_smem, rg, p.b.usorted, kRobinHandover);   // refused: usorted is untouched
    __syncthreads();
    const Pose* insert_pose = deferred == 2u ? nullptr : &pose_s;
    uint32_t gather_shift = 0;
    if (!exact && nu) {   // a run of more than 2 048 occupied buckets: the counting-sort order and its rotation stand
        const long long over = static_cast<long long>(nu) + *p.b.rb_shift - static_cast<long long>(robin_bucket_count(nu));
        const uint32_t rot = over > 0 ? static_cast<uint32_t>(over % nu) : 0u;
        gather_shift = (nu - rot) % nu;
    }
    const PointSource src{reinterpret_cast<const float*>(p.filtered), 12u, 1, p.b.usorted, nu, gather_shift};
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    if (exact) {   // the first pass again, in the new order: it finds the voxels scan_kernel created
        for (uint32_t i = tid; i < nu; i += nthreads) grid_locate_item(i, src, insert_pose, p.kf, p.kf_counters, p.kf_free_list, p.b.kf_slot);
        grid.sync();
    }
    for (uint32_t i = tid; i < nu; i += nthreads) grid_chain_item(i, p.ticket_base, p.kf, p.b.kf_slot, p.b.kf_base);
    grid.sync();
    for (uint32_t i = tid; i < nu; i += nthreads)
        grid_write_item(i, src, p.ticket_base, insert_pose, p.kf, p.kf_counters, p.b.kf_slot, p.b.kf_base);
    grid.sync();
    if (tid == 0) scan_publish_tail(p, exact ? kOrderReplayed : kOrderApproximate);
}

}  // namespace lo

using namespace lo;

// Scratch of the per-scan kernel for scans of up to `n` points; called once at lo_odometry_create for a VLP16-sized scan so
// that no submit allocates (larger scans grow it on first sight).
int lo::fused_reserve(lo_odometry* h, uint32_t n) {
    PipeBufs b;
    n = std::max<uint32_t>(n, 1);
    LO_TRY(h->fused.ensure(carve(b, nullptr, n)));
    LO_TRY(h->deskewed.ensure(static_cast<size_t>(n) * sizeof(PointXYZIRT)));
    LO_TRY(h->filtered.ensure(static_cast<size_t>(n) * sizeof(PointNormal)));
    LO_TRY(h->query.ensure(static_cast<size_t>(n) * sizeof(PointXYZ)));
    return LO_OK;
}

int lo::fused_process_enqueue(lo_odometry* h, const PointXYZIRT* d_scan, uint32_t n) {
    lo_ctx* ctx = h->ctx;
    lo_grid* kf = h->keyframe;
    PipeBufs b;
    LO_TRY(fused_reserve(h, n));
    carve(b, h->fused.as<char>(), std::max<uint32_t>(n, 1));
    static int blocks_per_sm = 0;
    if (blocks_per_sm == 0) {
        int large_per_sm = 0;
        LO_CUDA(cudaFuncSetAttribute(scan_kernel_large, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kAlignTileScratchBytes)));
        LO_CUDA(cudaFuncSetAttribute(scan_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kRobinBlockSmemBytes)));
        LO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&large_per_sm, scan_kernel_large, kPipeThreads, kAlignTileScratchBytes));
        LO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, scan_kernel, kPipeThreads, 0));
        int replay_per_sm = 0;
        LO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&replay_per_sm, scan_replay_kernel, kPipeThreads, kRobinBlockSmemBytes));
        blocks_per_sm = std::min(std::min(blocks_per_sm, large_per_sm), replay_per_sm);
        if (blocks_per_sm < 1) {
            set_error("the per-scan kernel does not fit on an SM");
            return LO_E_CUDA;
        }
    }
    const int blocks = std::min<int>(ctx->sm_count, kMaxPipeBlocks);   // one block per SM
    // Room for this scan's insert and for those of the scans whose counters have not come back yet: a scan adds at most
    // one voxel per update point, and the update set is a subset of the planar points, themselves a subset of the scan
    // (the host mirror of the counters is as old as the last scan adopted). With the key frame sized at creation this
    // never grows in practice; when it does, the growth is stream-ordered (lo_grid.cu) — no synchronisation here.
    LO_TRY(grid_reserve(kf, static_cast<size_t>(n) + h->pending_points));
    ScanArgs p;
    p.s = h->d_state;
    p.scan = d_scan;
    p.desk = h->deskewed.as<PointXYZIRT>();
    p.filtered = h->filtered.as<PointNormal>();
    p.query = h->query.as<PointXYZ>();
    p.minmax = h->d_minmax;
    p.b = b;
    p.kf = kf->view();
    p.kf_capacity = kf->capacity;
    p.kf_counters = kf->counters;
    p.kf_free_list = kf->free_list;
    const uint32_t ticket_before = kf->h_counters[GC_TICKET];
    p.ticket_base = grid_take_tickets(kf, n);
    p.seq = h->scan_seq + 1;
    p.readback = h->h_readback + (p.seq % kQueueDepth);   // pinned host memory (unified addressing): no D2H copy, the host polls `seq`
    p.n = n;
    p.lo2 = h->params.lidar_min_range * h->params.lidar_min_range;
    p.hi2 = h->params.lidar_max_range * h->params.lidar_max_range;
    p.vs_u = h->params.keyframe_update_voxel_size;
    p.vs_m = h->params.keyframe_matching_voxel_size;
    p.cleanup_range = h->params.keyframe_cleanup_range;
    p.gate_threshold = h->params.angular_divergence_threshold;
    static const bool replay_always = getenv("LO_B200_REPLAY_ALWAYS") != nullptr;
    p.force_replay = replay_always ? 1u : 0u;
    static const bool want_trace = getenv("LO_SCAN_TRACE") != nullptr;   // developer aid, see fused_trace_dump
    p.trace = nullptr;
    if (want_trace) {
        LO_TRY(ctx->tmp_d.ensure(1001 * sizeof(long long)));
        p.trace = ctx->tmp_d.as<long long>();
    }
    void* kargs[] = {&p};
    const cudaError_t e =
        n <= kSmallScanPoints
            ? cudaLaunchCooperativeKernel(reinterpret_cast<void*>(scan_kernel), dim3(blocks), dim3(kPipeThreads), kargs, 0, ctx->stream)
            : cudaLaunchCooperativeKernel(reinterpret_cast<void*>(scan_kernel_large), dim3(blocks), dim3(kPipeThreads), kargs,
                                          kAlignTileScratchBytes, ctx->stream);
    if (e != cudaSuccess) {   // nothing was queued: the scan number and the tickets stay where they were
        kf->h_counters[GC_TICKET] = ticket_before;
        set_error("launching the per-scan kernel (%d cooperative blocks of %d threads) failed: %s", blocks, kPipeThreads, cudaGetErrorString(e));
        return LO_E_CUDA;
    }
    h->scan_seq = p.seq;
    ctx->launches++;
    {   // returns at once unless the scan kernel left the insert to it (see there)
        const cudaError_t e2 = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(scan_replay_kernel), dim3(blocks), dim3(kPipeThreads), kargs,
                                                           kRobinBlockSmemBytes, ctx->stream);
        if (e2 != cudaSuccess) {
            set_error("launching the replay kernel (%d cooperative blocks of %d threads) failed: %s", blocks, kPipeThreads, cudaGetErrorString(e2));
            return LO_E_CUDA;
        }
        ctx->launches++;
    }
    return LO_OK;
}

void lo::fused_release(lo_odometry* h) { h->fused.release(); }

// LO_SCAN_TRACE=1: phase timeline of block 0 of the last scan_kernel, in SM clocks (call after the stream has drained)
void lo::fused_trace_dump(lo_odometry* h) {
    if (!getenv("LO_SCAN_TRACE") || !h->ctx->tmp_d.p) return;
    static long long t[1001];
    if (cudaMemcpy(t, h->ctx->tmp_d.p, sizeof(t), cudaMemcpyDeviceToHost) != cudaSuccess) return;
    for (long long k = 1; k < t[1000] && k < 500; k++) fprintf(stderr, "trace %lld %lld\n", t[2 * k], t[2 * k + 1] - t[2 * (k - 1) + 1]);
    fprintf(stderr, "trace end %lld\n", t[1000] > 0 ? t[2 * (t[1000] - 1) + 1] - t[1] : 0);
}
