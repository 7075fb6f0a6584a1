// K1 (time normalisation + deskew), rigid transforms, range filter and K8 (CloudClassifier) kernels.
#include <cub/cub.cuh>

#include <algorithm>

#include "lo_points.cuh"
#include "lo_stages.cuh"

namespace lo {

// ---------------------------------------------------------------------------------------------------
// K1: utils::pointTimeNormalize (src/utils/point_time_normalize.h:14-39) fused with
// CloudTransformer::transformNonRigid (src/utils/cloud_transform.h:15-40).
// minmax[0] = ordered min, minmax[1] = ordered max (initialised to 0xFFFFFFFF / 0)
__global__ void time_minmax_kernel(const PointXYZIRT* __restrict__ in, uint32_t n, uint32_t* __restrict__ minmax) {
    uint32_t lo_v = 0xFFFFFFFFu, hi_v = 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float t = in[i].time;
        if (t == t) {  // std::min/std::max keep the running value when the operand is NaN
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
}

__global__ void deskew_kernel(const PointXYZIRT* __restrict__ in, uint32_t n, const uint32_t* __restrict__ minmax,
                              const Pose* __restrict__ start_p, const Pose* __restrict__ end_p, PointXYZIRT* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Pose s = *start_p, e = *end_p;
    const float t_min = time_min_of(minmax), t_max = time_max_of(minmax);
    // 32-byte record as two 16-byte loads
    const float4 a = reinterpret_cast<const float4*>(in)[2 * i];
    float4 b = reinterpret_cast<const float4*>(in)[2 * i + 1];  // intensity, ring|pad, time, pad
    const float4 o = deskew_point(s, e, t_min, t_max - t_min, a, b);
    reinterpret_cast<float4*>(out)[2 * i] = o;
    reinterpret_cast<float4*>(out)[2 * i + 1] = b;
}

// utils::pointTimeNormalize alone (point_time_normalize.h:31-36): time <- (time - min) / (max - min), everything else copied
__global__ void time_normalize_kernel(const PointXYZIRT* __restrict__ in, uint32_t n, const uint32_t* __restrict__ minmax,
                                      PointXYZIRT* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float t_min = time_min_of(minmax), range = time_max_of(minmax) - t_min;
    const float4 a = reinterpret_cast<const float4*>(in)[2 * i];
    float4 b = reinterpret_cast<const float4*>(in)[2 * i + 1];
    b.z = (b.z - t_min) / range;
    reinterpret_cast<float4*>(out)[2 * i] = a;
    reinterpret_cast<float4*>(out)[2 * i + 1] = b;
}

int deskew_dev(lo_ctx* ctx, const PointXYZIRT* d_in, uint32_t n, const Pose* d_start, const Pose* d_end, PointXYZIRT* d_out) {
    if (n == 0) return LO_OK;
    uint32_t* d_minmax = ctx->small.as<uint32_t>() + 128;
    const uint32_t init[2] = {0xFFFFFFFFu, 0u};
    LO_CUDA(cudaMemcpyAsync(d_minmax, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
    const unsigned blocks = std::min<unsigned>(blocks_for(n, 256), static_cast<unsigned>(ctx->sm_count) * 4);
    time_minmax_kernel<<<blocks, 256, 0, ctx->stream>>>(d_in, n, d_minmax);
    LO_LAUNCHED(ctx);
    deskew_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, d_minmax, d_start, d_end, d_out);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// ---------------------------------------------------------------------------------------------------
// rigid transforms (src/utils/cloud_transform.h:42-97)
__global__ void transform_xyz_kernel(const PointXYZ* __restrict__ in, uint32_t n, const Pose* __restrict__ pose, PointXYZ* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Pose p = *pose;
    float r[9], o[3];
    rotation_matrix_f(p.w, p.x, p.y, p.z, r);
    PointXYZ q = in[i];
    mat_vec_f(r, q.x, q.y, q.z, o);
    q.x = o[0] + p.t[0]; q.y = o[1] + p.t[1]; q.z = o[2] + p.t[2];
    out[i] = q;
}

__global__ void transform_pn_kernel(const PointNormal* __restrict__ in, uint32_t n, const Pose* __restrict__ pose,
                                    PointNormal* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Pose p = *pose;
    float r[9], o[3];
    rotation_matrix_f(p.w, p.x, p.y, p.z, r);
    PointNormal q = in[i];
    mat_vec_f(r, q.x, q.y, q.z, o);
    q.x = o[0] + p.t[0]; q.y = o[1] + p.t[1]; q.z = o[2] + p.t[2];
    mat_vec_f(r, q.nx, q.ny, q.nz, o);
    q.nx = o[0]; q.ny = o[1]; q.nz = o[2];
    out[i] = q;
}

int transform_xyz_dev(lo_ctx* ctx, const PointXYZ* d_in, uint32_t n, const Pose* d_pose, PointXYZ* d_out) {
    if (n == 0) return LO_OK;
    transform_xyz_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, d_pose, d_out);
    LO_LAUNCHED(ctx);
    return LO_OK;
}
int transform_pn_dev(lo_ctx* ctx, const PointNormal* d_in, uint32_t n, const Pose* d_pose, PointNormal* d_out) {
    if (n == 0) return LO_OK;
    transform_pn_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, d_pose, d_out);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// ---------------------------------------------------------------------------------------------------
// stable compaction: flags -> exclusive scan -> scatter
static int exclusive_scan(lo_ctx* ctx, const uint32_t* d_flags, uint32_t* d_offsets, uint32_t n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, d_flags, d_offsets, static_cast<int>(n), ctx->stream);
    LO_TRY(ctx->tmp_d.ensure(bytes));
    LO_CUDA(cub::DeviceScan::ExclusiveSum(ctx->tmp_d.p, bytes, d_flags, d_offsets, static_cast<int>(n), ctx->stream));
    ctx->launches += 2;
    return LO_OK;
}

__global__ void range_flags_kernel(const PointNormal* __restrict__ in, uint32_t n, float lo2, float hi2, uint32_t* __restrict__ flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = *reinterpret_cast<const float4*>(in + i);
    flags[i] = in_range(p.x, p.y, p.z, lo2, hi2) ? 1u : 0u;
}

__global__ void scatter_pn_kernel(const PointNormal* __restrict__ in, uint32_t n, const uint32_t* __restrict__ flags,
                                  const uint32_t* __restrict__ offsets, PointNormal* __restrict__ out, uint32_t* __restrict__ count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flags[i]) out[offsets[i]] = in[i];
    if (i == n - 1) *count = offsets[i] + flags[i];
}

int range_filter_dev(lo_ctx* ctx, const PointNormal* d_in, uint32_t n, float min_range, float max_range, PointNormal* d_out,
                     uint32_t* d_count) {
    LO_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), ctx->stream));
    if (n == 0) return LO_OK;
    LO_TRY(ctx->tmp_a.ensure(static_cast<size_t>(n) * sizeof(uint32_t)));
    LO_TRY(ctx->tmp_b.ensure(static_cast<size_t>(n) * sizeof(uint32_t)));
    uint32_t *flags = ctx->tmp_a.as<uint32_t>(), *offsets = ctx->tmp_b.as<uint32_t>();
    range_flags_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, min_range * min_range, max_range * max_range, flags);
    LO_LAUNCHED(ctx);
    LO_TRY(exclusive_scan(ctx, flags, offsets, n));
    scatter_pn_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, flags, offsets, d_out, d_count);
    LO_LAUNCHED(ctx);
    return LO_OK;
}

// ---------------------------------------------------------------------------------------------------
// K8: CloudClassifier::classify (src/utils/cloud_classifier.h:19-168).
// layout[0] = W (max ring population), layout[1] = H (rings present), layout[2] = overflow flag
__global__ void ring_hist_kernel(const PointXYZIRT* __restrict__ in, uint32_t n, uint32_t* __restrict__ hist) {
    __shared__ uint32_t local[256];
    local[threadIdx.x] = 0;  // blockDim.x == 256
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        atomicAdd(&local[in[i].ring & 0xFF], 1u);  // the reference keys its map by uint8_t (:21-23)
    __syncthreads();
    if (local[threadIdx.x]) atomicAdd(&hist[threadIdx.x], local[threadIdx.x]);
}

__global__ void ring_layout_kernel(const uint32_t* __restrict__ hist, uint32_t* __restrict__ row_of, uint32_t* __restrict__ layout,
                                   uint32_t cell_cap) {
    __shared__ uint32_t present[256];
    __shared__ uint32_t width;
    if (threadIdx.x == 0) width = 0;
    __syncthreads();
    const uint32_t c = hist[threadIdx.x];
    present[threadIdx.x] = c ? 1u : 0u;
    atomicMax(&width, c);
    __syncthreads();
    uint32_t rank = 0;
    for (uint32_t r = 0; r < threadIdx.x; r++) rank += present[r];
    row_of[threadIdx.x] = rank;
    if (threadIdx.x == 255) {
        const uint32_t h = rank + present[255];
        layout[0] = width;
        layout[1] = h;
        layout[2] = (static_cast<unsigned long long>(width) * h > cell_cap) ? 1u : 0u;
    }
}

__global__ void cell_fill_kernel(const PointXYZIRT* __restrict__ in, uint32_t n, const uint32_t* __restrict__ row_of,
                                 const uint32_t* __restrict__ layout, uint32_t* __restrict__ cell_idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || layout[2]) return;
    const uint32_t W = layout[0];
    const float4 a = reinterpret_cast<const float4*>(in)[2 * i];
    const uint32_t ring = in[i].ring & 0xFF;
    uint32_t col;
    if (!azimuth_column(a.x, a.y, W, col)) return;
    atomicMax(&cell_idx[row_of[ring] * W + col], i + 1);  // last writer (largest input index) wins (:52-54)
}

__global__ void cell_gather_kernel(const PointXYZIRT* __restrict__ in, const uint32_t* __restrict__ layout,
                                   const uint32_t* __restrict__ cell_idx, float4* __restrict__ cellp) {
    const uint32_t total = layout[2] ? 0u : layout[0] * layout[1];
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < total; c += gridDim.x * blockDim.x) {
        const uint32_t idx = cell_idx[c];
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);  // PointType() : value-initialised
        if (idx) {
            const float4 a = reinterpret_cast<const float4*>(in)[2 * (idx - 1)];
            const float4 b = reinterpret_cast<const float4*>(in)[2 * (idx - 1) + 1];
            v = make_float4(a.x, a.y, a.z, b.x);
        }
        cellp[c] = v;
    }
}

// :81-105 curvature over the flattened grid; first/last 4 cells keep their raw intensity
__global__ void curvature_kernel(const uint32_t* __restrict__ layout, const float4* __restrict__ cellp, float* __restrict__ curv) {
    const uint32_t total = layout[2] ? 0u : layout[0] * layout[1];
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < total; c += gridDim.x * blockDim.x) {
        curv[c] = cell_curvature(c, total, [&](uint32_t k) { return cellp[k]; });
    }
}

// :107-165 planar test + normal from two neighbours on the previous ring. Writes flags for ALL cell_cap cells.
__global__ void planar_kernel(const uint32_t* __restrict__ layout, const float4* __restrict__ cellp, const float* __restrict__ curv,
                              int filter, float lo2, float hi2, uint32_t cell_cap, uint32_t* __restrict__ flags,
                              float4* __restrict__ normals) {
    const uint32_t W = layout[0], H = layout[1];
    const uint32_t total = layout[2] ? 0u : W * H;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < cell_cap; c += gridDim.x * blockDim.x) {
        uint32_t flag = 0;
        if (c < total) {
            float nv[3];
            if (cell_planar(c, W, cellp, curv, nv)) {
                const float4 p = cellp[c];
                normals[c] = make_float4(nv[0], nv[1], nv[2], 0.f);
                flag = (!filter || in_range(p.x, p.y, p.z, lo2, hi2)) ? 1u : 0u;
            }
        }
        flags[c] = flag;
    }
}

__global__ void planar_scatter_kernel(uint32_t cell_cap, const float4* __restrict__ cellp, const float4* __restrict__ normals,
                                      const uint32_t* __restrict__ flags, const uint32_t* __restrict__ offsets,
                                      PointNormal* __restrict__ out, uint32_t* __restrict__ count) {
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < cell_cap; c += gridDim.x * blockDim.x) {
        if (flags[c]) {
            const float4 p = cellp[c], nv = normals[c];
            float4* o = reinterpret_cast<float4*>(out + offsets[c]);
            o[0] = make_float4(p.x, p.y, p.z, 1.f);
            o[1] = make_float4(nv.x, nv.y, nv.z, 0.f);
            o[2] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (c == cell_cap - 1) *count = offsets[c] + flags[c];
    }
}

// the second element of classify's pair (:156,160-162): cells of rays >= 1 inside the normals window that did not become
// planar points and whose curvature is below intensity_max
__global__ void unclassified_flag_kernel(const uint32_t* __restrict__ layout, const float4* __restrict__ cellp, const float* __restrict__ curv,
                                         uint32_t cell_cap, uint32_t* __restrict__ flags) {
    const uint32_t W = layout[0], H = layout[1];
    const uint32_t total = layout[2] ? 0u : W * H;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < cell_cap; c += gridDim.x * blockDim.x) {
        uint32_t flag = 0;
        if (c < total) {
            const uint32_t ray = c / W, col = c % W;
            if (ray >= 1 && col >= 4 && col + 4 < W) {
                float nv[3];
                flag = (!cell_planar(c, W, cellp, curv, nv) && curv[c] < 1000.0f) ? 1u : 0u;
            }
        }
        flags[c] = flag;
    }
}

__global__ void unclassified_scatter_kernel(uint32_t cell_cap, const PointXYZIRT* __restrict__ in, const uint32_t* __restrict__ cell_idx,
                                            const float* __restrict__ curv, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ offsets,
                                            PointXYZIRT* __restrict__ out, uint32_t* __restrict__ count) {
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < cell_cap; c += gridDim.x * blockDim.x) {
        if (flags[c]) {
            const uint32_t idx = cell_idx[c];
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;   // PointType(): value-initialised
            if (idx) {
                a = reinterpret_cast<const float4*>(in)[2 * (idx - 1)];
                b = reinterpret_cast<const float4*>(in)[2 * (idx - 1) + 1];
            }
            b.x = curv[c];   // the curvature pass wrote it into `intensity` (:104)
            float4* o = reinterpret_cast<float4*>(out + offsets[c]);
            o[0] = a;
            o[1] = b;
        }
        if (c == cell_cap - 1) *count = offsets[c] + flags[c];
    }
}

struct ClassifyScratch {
    uint32_t *hist, *row_of, *layout, *cell_idx, *flags, *offsets;
    float4 *cellp, *normals;
    float* curv;
};

int classify_dev(lo_ctx* ctx, const PointXYZIRT* d_in, uint32_t n, int filter, float min_range, float max_range, PointNormal* d_out,
                 uint32_t* d_count, PointXYZIRT* d_unclassified, uint32_t* d_count_unclassified, uint32_t* d_overflow) {
    LO_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), ctx->stream));
    if (d_count_unclassified) LO_CUDA(cudaMemsetAsync(d_count_unclassified, 0, sizeof(uint32_t), ctx->stream));
    if (d_overflow) LO_CUDA(cudaMemsetAsync(d_overflow, 0, sizeof(uint32_t), ctx->stream));
    if (n == 0) return LO_OK;
    const uint32_t cell_cap = std::max<uint32_t>(8u * n, 65536u);
    // scratch carved out of tmp_a: hist[256] row_of[256] layout[4] | cell_idx | flags | offsets | curv | cellp | normals
    const size_t words = 256 + 256 + 64 + 4ull * cell_cap;
    LO_TRY(ctx->tmp_a.ensure(words * sizeof(uint32_t) + 2ull * cell_cap * sizeof(float4) + 64));
    ClassifyScratch s;
    s.hist = ctx->tmp_a.as<uint32_t>();
    s.row_of = s.hist + 256;
    s.layout = s.row_of + 256;
    s.cell_idx = s.layout + 64;
    s.flags = s.cell_idx + cell_cap;
    s.offsets = s.flags + cell_cap;
    s.curv = reinterpret_cast<float*>(s.offsets + cell_cap);
    s.cellp = reinterpret_cast<float4*>(s.curv + cell_cap);
    s.normals = s.cellp + cell_cap;
    LO_CUDA(cudaMemsetAsync(s.hist, 0, (256 + 256 + 64 + static_cast<size_t>(cell_cap)) * sizeof(uint32_t), ctx->stream));
    const unsigned wide = static_cast<unsigned>(ctx->sm_count) * 8;
    ring_hist_kernel<<<std::min<unsigned>(blocks_for(n, 256), wide), 256, 0, ctx->stream>>>(d_in, n, s.hist);
    LO_LAUNCHED(ctx);
    ring_layout_kernel<<<1, 256, 0, ctx->stream>>>(s.hist, s.row_of, s.layout, cell_cap);
    LO_LAUNCHED(ctx);
    cell_fill_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_in, n, s.row_of, s.layout, s.cell_idx);
    LO_LAUNCHED(ctx);
    cell_gather_kernel<<<wide, 256, 0, ctx->stream>>>(d_in, s.layout, s.cell_idx, s.cellp);
    LO_LAUNCHED(ctx);
    curvature_kernel<<<wide, 256, 0, ctx->stream>>>(s.layout, s.cellp, s.curv);
    LO_LAUNCHED(ctx);
    planar_kernel<<<wide, 256, 0, ctx->stream>>>(s.layout, s.cellp, s.curv, filter, min_range * min_range, max_range * max_range, cell_cap,
                                                 s.flags, s.normals);
    LO_LAUNCHED(ctx);
    LO_TRY(exclusive_scan(ctx, s.flags, s.offsets, cell_cap));
    planar_scatter_kernel<<<wide, 256, 0, ctx->stream>>>(cell_cap, s.cellp, s.normals, s.flags, s.offsets, d_out, d_count);
    LO_LAUNCHED(ctx);
    if (d_unclassified && d_count_unclassified) {
        unclassified_flag_kernel<<<wide, 256, 0, ctx->stream>>>(s.layout, s.cellp, s.curv, cell_cap, s.flags);
        LO_LAUNCHED(ctx);
        LO_TRY(exclusive_scan(ctx, s.flags, s.offsets, cell_cap));
        unclassified_scatter_kernel<<<wide, 256, 0, ctx->stream>>>(cell_cap, d_in, s.cell_idx, s.curv, s.flags, s.offsets, d_unclassified,
                                                                   d_count_unclassified);
        LO_LAUNCHED(ctx);
    }
    if (d_overflow) LO_CUDA(cudaMemcpyAsync(d_overflow, s.layout + 2, sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    return LO_OK;
}

}  // namespace lo

// ---------------------------------------------------------------------------------------------------
// host-pointer C ABI
using namespace lo;

static int stage_pose(lo_ctx* ctx, const float p7[7], int slot, Pose** d_pose) {
    Pose* d = reinterpret_cast<Pose*>(ctx->small.as<char>() + 2048) + slot;
    const Pose h = pose_from7(p7);
    LO_CUDA(cudaMemcpyAsync(d, &h, sizeof(Pose), cudaMemcpyHostToDevice, ctx->stream));
    *d_pose = d;
    return LO_OK;
}

extern "C" int lo_deskew(lo_ctx* ctx, const void* in, size_t n, const float start7[7], const float end7[7], void* out) {
    if (!ctx || (!in && n) || (!out && n) || !start7 || !end7) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointXYZIRT);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    Pose *ds, *de;
    LO_TRY(stage_pose(ctx, start7, 0, &ds));
    LO_TRY(stage_pose(ctx, end7, 1, &de));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(deskew_dev(ctx, ctx->in_a.as<PointXYZIRT>(), static_cast<uint32_t>(n), ds, de, ctx->out_a.as<PointXYZIRT>()));
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" int lo_point_time_normalize(lo_ctx* ctx, const void* in, size_t n, void* out) {
    if (!ctx || (!in && n) || (!out && n)) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointXYZIRT);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    uint32_t* d_minmax = ctx->small.as<uint32_t>() + 128;
    const uint32_t init[2] = {0xFFFFFFFFu, 0u};
    LO_CUDA(cudaMemcpyAsync(d_minmax, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    const uint32_t nn = static_cast<uint32_t>(n);
    time_minmax_kernel<<<std::min<unsigned>(blocks_for(n, 256), static_cast<unsigned>(ctx->sm_count) * 4), 256, 0, ctx->stream>>>(
        ctx->in_a.as<PointXYZIRT>(), nn, d_minmax);
    LO_LAUNCHED(ctx);
    time_normalize_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(ctx->in_a.as<PointXYZIRT>(), nn, d_minmax, ctx->out_a.as<PointXYZIRT>());
    LO_LAUNCHED(ctx);
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

// CloudTransformer::transformNonRigid on its own: the `time` field is used as it is (cloud_transform.h:26-30). The
// deskew kernel is run with min = 0 and max = 1, for which its normalisation (t - 0) / 1 is exact.
extern "C" int lo_transform_non_rigid(lo_ctx* ctx, const void* in, size_t n, const float start7[7], const float end7[7], void* out) {
    if (!ctx || (!in && n) || (!out && n) || !start7 || !end7) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointXYZIRT);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    Pose *ds, *de;
    LO_TRY(stage_pose(ctx, start7, 0, &ds));
    LO_TRY(stage_pose(ctx, end7, 1, &de));
    uint32_t* d_minmax = ctx->small.as<uint32_t>() + 128;
    const uint32_t unit[2] = {0x80000000u, 0xBF800000u};   // float_to_ordered(0.0f), float_to_ordered(1.0f)
    LO_CUDA(cudaMemcpyAsync(d_minmax, unit, sizeof(unit), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    deskew_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(ctx->in_a.as<PointXYZIRT>(), static_cast<uint32_t>(n), d_minmax, ds, de,
                                                               ctx->out_a.as<PointXYZIRT>());
    LO_LAUNCHED(ctx);
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" int lo_transform_xyz(lo_ctx* ctx, const float* in, size_t n, const float pose7[7], float* out) {
    if (!ctx || (!in && n) || (!out && n) || !pose7) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointXYZ);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    Pose* dp;
    LO_TRY(stage_pose(ctx, pose7, 0, &dp));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(transform_xyz_dev(ctx, ctx->in_a.as<PointXYZ>(), static_cast<uint32_t>(n), dp, ctx->out_a.as<PointXYZ>()));
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" int lo_transform_with_normals(lo_ctx* ctx, const float* in, size_t n, const float pose7[7], float* out) {
    if (!ctx || (!in && n) || (!out && n) || !pose7) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointNormal);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    Pose* dp;
    LO_TRY(stage_pose(ctx, pose7, 0, &dp));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(transform_pn_dev(ctx, ctx->in_a.as<PointNormal>(), static_cast<uint32_t>(n), dp, ctx->out_a.as<PointNormal>()));
    LO_CUDA(cudaMemcpyAsync(out, ctx->out_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" int lo_range_filter(lo_ctx* ctx, const float* in, size_t n, float min_range, float max_range, float* out, size_t* n_out) {
    if (!ctx || (!in && n) || (!out && n) || !n_out) return LO_E_ARG;
    *n_out = 0;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = n * sizeof(PointNormal);
    LO_TRY(ctx->in_a.ensure(bytes));
    LO_TRY(ctx->out_a.ensure(bytes));
    uint32_t* d_count = ctx->small.as<uint32_t>() + 192;
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(range_filter_dev(ctx, ctx->in_a.as<PointNormal>(), static_cast<uint32_t>(n), min_range, max_range, ctx->out_a.as<PointNormal>(),
                            d_count));
    uint32_t h_count = 0;
    LO_CUDA(cudaMemcpyAsync(&h_count, d_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_count) LO_CUDA(cudaMemcpy(out, ctx->out_a.p, static_cast<size_t>(h_count) * sizeof(PointNormal), cudaMemcpyDeviceToHost));
    *n_out = h_count;
    return LO_OK;
}

extern "C" int lo_classify_pair(lo_ctx* ctx, const void* in, size_t n, float* planar_out, size_t cap, size_t* n_planar, void* unclassified_out,
                                size_t cap_unclassified, size_t* n_unclassified) {
    if (!ctx || (!in && n) || !n_planar) return LO_E_ARG;
    *n_planar = 0;
    if (n_unclassified) *n_unclassified = 0;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    const bool want_un = n_unclassified != nullptr;
    LO_TRY(ctx->in_a.ensure(n * sizeof(PointXYZIRT)));
    LO_TRY(ctx->out_a.ensure(n * sizeof(PointNormal)));
    if (want_un) LO_TRY(ctx->out_b.ensure(n * sizeof(PointXYZIRT)));
    uint32_t* d_count = ctx->small.as<uint32_t>() + 192;   // [0] planar, [1] unclassified, [2] overflow
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, in, n * sizeof(PointXYZIRT), cudaMemcpyHostToDevice, ctx->stream));
    LO_TRY(classify_dev(ctx, ctx->in_a.as<PointXYZIRT>(), static_cast<uint32_t>(n), 0, 0.f, 0.f, ctx->out_a.as<PointNormal>(), d_count,
                        want_un ? ctx->out_b.as<PointXYZIRT>() : nullptr, want_un ? d_count + 1 : nullptr, d_count + 2));
    uint32_t h_count[3] = {0, 0, 0};
    LO_CUDA(cudaMemcpyAsync(h_count, d_count, sizeof(h_count), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_count[2]) {   // never silently: the reference handles such a scan
        set_error("the classifier's ring x azimuth grid (widest ring x rings present) exceeds the scratch sized for %zu points", n);
        return LO_E_CAPACITY;
    }
    *n_planar = h_count[0];
    if (want_un) *n_unclassified = h_count[1];
    if (h_count[0] > cap || (want_un && unclassified_out && h_count[1] > cap_unclassified)) return LO_E_CAPACITY;
    if (h_count[0] && planar_out)
        LO_CUDA(cudaMemcpy(planar_out, ctx->out_a.p, static_cast<size_t>(h_count[0]) * sizeof(PointNormal), cudaMemcpyDeviceToHost));
    if (want_un && h_count[1] && unclassified_out)
        LO_CUDA(cudaMemcpy(unclassified_out, ctx->out_b.p, static_cast<size_t>(h_count[1]) * sizeof(PointXYZIRT), cudaMemcpyDeviceToHost));
    return LO_OK;
}

extern "C" int lo_classify(lo_ctx* ctx, const void* in, size_t n, float* planar_out, size_t cap, size_t* n_planar) {
    return lo_classify_pair(ctx, in, n, planar_out, cap, n_planar, nullptr, 0, nullptr);
}
