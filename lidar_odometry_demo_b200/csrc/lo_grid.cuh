// Device side of the voxel grid's mutations (insert passes, radius clean-up), shared by the stage kernels of
// lo_grid.cu and the per-scan kernel of lo_pipeline.cu. See lo_grid.cu for the data-structure description.
#pragma once

#include "lo_host.h"

namespace lo {

constexpr uint32_t kInvalid = 0xFFFFFFFFu;

struct PointSource {       // AoS input: pcl::PointNormal (48 B) or pcl::PointXYZ (16 B)
    const float* base;
    uint32_t stride_f;     // floats per record: 12 or 4
    int has_normals;
    const uint32_t* gather = nullptr;   // optional: item i is record gather[(i + gather_shift) % gather_n]
    uint32_t gather_n = 0, gather_shift = 0;
};

__device__ inline void load_point(const PointSource& src, uint32_t i, const Pose* pose, float p[3], float nrm[3]) {
    if (src.gather) {
        uint32_t at = i + src.gather_shift;
        if (at >= src.gather_n) at -= src.gather_n;
        i = src.gather[at];
    }
    const float4 a = *reinterpret_cast<const float4*>(src.base + static_cast<size_t>(i) * src.stride_f);
    p[0] = a.x; p[1] = a.y; p[2] = a.z;
    if (src.has_normals) {
        const float4 b = *reinterpret_cast<const float4*>(src.base + static_cast<size_t>(i) * src.stride_f + 4);
        nrm[0] = b.x; nrm[1] = b.y; nrm[2] = b.z;
    } else {
        nrm[0] = nrm[1] = nrm[2] = 0.f;
    }
    if (pose) {  // transformWithNormals fused into the load (src/utils/cloud_transform.h:68-97)
        float r[9], o[3];
        rotation_matrix_f(pose->w, pose->x, pose->y, pose->z, r);
        mat_vec_f(r, p[0], p[1], p[2], o);
        p[0] = o[0] + pose->t[0]; p[1] = o[1] + pose->t[1]; p[2] = o[2] + pose->t[2];
        mat_vec_f(r, nrm[0], nrm[1], nrm[2], o);
        nrm[0] = o[0]; nrm[1] = o[1]; nrm[2] = o[2];
    }
}

__device__ inline uint32_t alloc_vid(uint32_t* counters, const uint32_t* free_list) {
    const int t = atomicSub(reinterpret_cast<int*>(&counters[GC_FREE_TOP]), 1);
    if (t > 0) return free_list[t - 1];
    return atomicAdd(&counters[GC_BUMP], 1u);
}

// pass 1: find or create the voxel of point i
__device__ inline void grid_locate_item(uint32_t i, const PointSource& src, const Pose* __restrict__ pose, const GridView& g,
                                        uint32_t* __restrict__ counters, const uint32_t* __restrict__ free_list,
                                        uint32_t* __restrict__ slot_of) {
    float p[3], nrm[3];
    load_point(src, i, pose, p, nrm);
    const long long ix = axis_index(p[0], g.voxel_size), iy = axis_index(p[1], g.voxel_size), iz = axis_index(p[2], g.voxel_size);
    const bool finite = isfinite(p[0]) && isfinite(p[1]) && isfinite(p[2]);
    if (!finite || !axis_in_range(ix) || !axis_in_range(iy) || !axis_in_range(iz)) {
        counters[GC_RANGE_ERR] = 1;
        slot_of[i] = kInvalid;
        return;
    }
    if (g.shard_world > 1 && shard_owner(ix, iy, iz, g.shard_world) != g.shard_rank) {  // another rank's voxel
        slot_of[i] = kInvalid;
        return;
    }
    const unsigned long long key = pack_key(ix, iy, iz);
    uint32_t h = probe_hash(key) & g.mask;
    while (true) {
        unsigned long long k = g.slots[h].key;
        if (k == kKeyEmpty) {
            k = atomicCAS(&g.slots[h].key, kKeyEmpty, key);
            if (k == kKeyEmpty) {  // this thread created the voxel
                g.slots[h].vid = alloc_vid(counters, free_list);
                g.slots[h].count = 0;
                atomicAdd(&counters[GC_VOXELS], 1u);
                break;
            }
        }
        if (k == key) break;
        h = (h + 1) & g.mask;
    }
    slot_of[i] = h;
}

// pass 2: ticket chain. Entry e >= base of the voxel's slab ends up with the (e-base)-th smallest ticket.
__device__ inline void grid_chain_item(uint32_t i, uint32_t ticket_base, const GridView& g, const uint32_t* __restrict__ slot_of,
                                       uint32_t* __restrict__ base_of) {
    const uint32_t s = slot_of[i];
    if (s == kInvalid) return;
    const uint32_t vid = g.slots[s].vid;
    const uint32_t base = g.slots[s].count;
    base_of[i] = base;
    uint32_t carry = ticket_base + i;
    uint32_t* lane_w = reinterpret_cast<uint32_t*>(g.points + static_cast<size_t>(vid) * g.width) + 3;
    for (uint32_t e = base; e < g.cap; e++) {
        const uint32_t old = atomicMin(lane_w + 4 * e, carry);
        if (old == kInvalid) break;     // entry was free: carry placed, nothing displaced
        if (old > carry) carry = old;   // carry placed, the displaced ticket moves on
    }
}

// pass 3: a point whose ticket survived writes its coordinates and normal
__device__ inline void grid_write_item(uint32_t i, const PointSource& src, uint32_t ticket_base, const Pose* __restrict__ pose, const GridView& g,
                                       uint32_t* __restrict__ counters, const uint32_t* __restrict__ slot_of,
                                       const uint32_t* __restrict__ base_of) {
    const uint32_t s = slot_of[i];
    if (s == kInvalid) return;
    const uint32_t vid = g.slots[s].vid;
    const uint32_t ticket = ticket_base + i;
    float4* slab = g.points + static_cast<size_t>(vid) * g.width;
    for (uint32_t e = base_of[i]; e < g.cap; e++) {
        const uint32_t w = __float_as_uint(slab[e].w);
        if (w == ticket) {
            float p[3], nrm[3];
            load_point(src, i, pose, p, nrm);
            slab[e] = make_float4(p[0], p[1], p[2], __uint_as_float(ticket));
            g.normals[static_cast<size_t>(vid) * g.width + e] = make_float4(nrm[0], nrm[1], nrm[2], 0.f);
            atomicMax(&g.slots[s].count, e + 1);
            atomicAdd(&counters[GC_POINTS], 1u);
            break;
        }
        if (w > ticket) break;  // entries are sorted: the ticket was dropped (voxel full)
    }
}

__device__ inline void grid_fix_counters(uint32_t* counters) {
    if (reinterpret_cast<int*>(counters)[GC_FREE_TOP] < 0) counters[GC_FREE_TOP] = 0;
}

// radiusCleanup (voxel_grid.h:236-246): the voxel's position is its FIRST stored point
__device__ inline void grid_cleanup_slot(uint32_t s, const GridView& g, const float* __restrict__ center, float radius,
                                         uint32_t* __restrict__ counters, uint32_t* __restrict__ free_list) {
    const Slot slot = g.slots[s];
    if (slot.key == kKeyEmpty || slot.key == kKeyTomb) return;
    const float4 o = g.points[static_cast<size_t>(slot.vid) * g.width];
    const float r2 = radius * radius;
    const float dx = o.x - center[0], dy = o.y - center[1], dz = o.z - center[2];
    const float d2 = dx * dx + (dy * dy + dz * dz);
    if (d2 > r2) {
        g.slots[s].key = kKeyTomb;
        g.slots[s].count = 0;
        uint32_t* lane_w = reinterpret_cast<uint32_t*>(g.points + static_cast<size_t>(slot.vid) * g.width) + 3;
        for (uint32_t e = 0; e < slot.count; e++) lane_w[4 * e] = kInvalid;
        const uint32_t top = atomicAdd(&counters[GC_FREE_TOP], 1u);
        free_list[top] = slot.vid;
        atomicSub(&counters[GC_VOXELS], 1u);
        atomicSub(&counters[GC_POINTS], slot.count);
        atomicAdd(&counters[GC_TOMBS], 1u);
    }
}

}  // namespace lo
