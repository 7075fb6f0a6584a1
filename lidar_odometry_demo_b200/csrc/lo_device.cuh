// Device-side building blocks shared by every kernel of liblo_b200: reference-exact float math, pose
// algebra, voxel keys and the open-addressing voxel table.
//
// Numerics contract (DESIGN.md "numerics"): the whole library is compiled with -fmad=false, so every
// `a*b+c` below is an unfused IEEE multiply followed by an IEEE add, exactly like the reference, which is
// built for baseline x86-64 without FMA (/root/reference/CMakeLists.txt:4-8). Association orders follow
// Eigen 3.4 for the fixed-size types the reference uses (3-float reductions a0+(a1+a2); Matrix3*Vector3
// (m0*x+m1*y)+m2*z; quaternion coefficient reductions (x+z)+(y+w)).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lo {

// ---------------------------------------------------------------------------------------------------
// wire structs (identical to the reference's PCL point types)
struct alignas(16) PointXYZ {
    float x, y, z, pad;
};
struct alignas(16) PointNormal {
    float x, y, z, pad0;
    float nx, ny, nz, pad1;
    float curvature, pad2[3];
};
struct alignas(16) PointXYZIRT {
    float x, y, z, pad0;
    float intensity;
    uint16_t ring;
    uint16_t pad1;
    float time;
    float pad2;
};
static_assert(sizeof(PointXYZ) == 16 && sizeof(PointNormal) == 48 && sizeof(PointXYZIRT) == 32, "wire layout");

struct Pose {  // Pose3D (src/pose_3d.h:10-21): float translation + float quaternion (w,x,y,z)
    float t[3];
    float w, x, y, z;
};

// ---------------------------------------------------------------------------------------------------
// float pose algebra, restating Eigen 3.4's formulas (see oracle/shim for the same list)
__host__ __device__ inline void cross3f(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

// Quaternionf * Vector3f: uv = q.vec x v; uv += uv; (v + w*uv) + q.vec x uv
__host__ __device__ inline void quat_rotate_f(float w, float x, float y, float z, const float v[3], float o[3]) {
    const float qv[3] = {x, y, z};
    float uv[3], c[3];
    cross3f(qv, v, uv);
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    cross3f(qv, uv, c);
    o[0] = (v[0] + uv[0] * w) + c[0];
    o[1] = (v[1] + uv[1] * w) + c[1];
    o[2] = (v[2] + uv[2] * w) + c[2];
}

__host__ __device__ inline float quat_dot_f(const Pose& a, const Pose& b) {
    return (a.x * b.x + a.z * b.z) + (a.y * b.y + a.w * b.w);
}

// Quaternionf::toRotationMatrix, row-major (src/pose_3d.h:41-44)
__host__ __device__ inline void rotation_matrix_f(float w, float x, float y, float z, float r[9]) {
    const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x;
    const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
    r[0] = 1.f - (tyy + tzz); r[1] = txy - twz;         r[2] = txz + twy;
    r[3] = txy + twz;         r[4] = 1.f - (txx + tzz); r[5] = tyz - twx;
    r[6] = txz - twy;         r[7] = tyz + twx;         r[8] = 1.f - (txx + tyy);
}

__host__ __device__ inline Pose pose_compose(const Pose& a, const Pose& b) {  // src/pose_3d.h:29-32
    Pose o;
    float rt[3];
    quat_rotate_f(a.w, a.x, a.y, a.z, b.t, rt);
    o.t[0] = a.t[0] + rt[0]; o.t[1] = a.t[1] + rt[1]; o.t[2] = a.t[2] + rt[2];
    o.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    o.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    o.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    o.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    return o;
}

__host__ __device__ inline Pose pose_inverse(const Pose& a) {  // src/pose_3d.h:34-39
    Pose o;
    const float n2 = quat_dot_f(a, a);
    if (n2 > 0.f) {
        o.w = a.w / n2; o.x = -a.x / n2; o.y = -a.y / n2; o.z = -a.z / n2;
    } else {
        o.w = o.x = o.y = o.z = 0.f;
    }
    const float neg[3] = {-a.t[0], -a.t[1], -a.t[2]};
    quat_rotate_f(o.w, o.x, o.y, o.z, neg, o.t);
    return o;
}

__host__ __device__ inline Pose pose_relative_to(const Pose& a, const Pose& target) {  // src/pose_3d.h:23-27
    return pose_compose(pose_inverse(a), target);
}

__host__ __device__ inline Pose pose_from7(const float p[7]) {
    Pose o;
    o.t[0] = p[0]; o.t[1] = p[1]; o.t[2] = p[2];
    o.w = p[3]; o.x = p[4]; o.y = p[5]; o.z = p[6];
    return o;
}
__host__ __device__ inline void pose_to7(const Pose& p, float o[7]) {
    o[0] = p.t[0]; o[1] = p.t[1]; o[2] = p.t[2];
    o[3] = p.w; o[4] = p.x; o[5] = p.y; o[6] = p.z;
}

// Matrix3f * Vector3f
__host__ __device__ inline void mat_vec_f(const float r[9], float x, float y, float z, float o[3]) {
    o[0] = (r[0] * x + r[1] * y) + r[2] * z;
    o[1] = (r[3] * x + r[4] * y) + r[5] * z;
    o[2] = (r[6] * x + r[7] * y) + r[8] * z;
}

// The query transform of findMatchingPairs (src/voxel_grid.h:212-223): R (float) widened to double,
// R*p + t in double, result narrowed to float.
struct QueryTransform {
    double r[9];
    double t[3];
};
__host__ __device__ inline QueryTransform make_query_transform(const Pose& p) {
    float rf[9];
    rotation_matrix_f(p.w, p.x, p.y, p.z, rf);
    QueryTransform q;
    for (int i = 0; i < 9; i++) q.r[i] = static_cast<double>(rf[i]);
    for (int i = 0; i < 3; i++) q.t[i] = static_cast<double>(p.t[i]);
    return q;
}
__device__ inline void transform_query(const QueryTransform& q, float x, float y, float z, float w[3]) {
    const double px = x, py = y, pz = z;
    w[0] = static_cast<float>(((q.r[0] * px + q.r[1] * py) + q.r[2] * pz) + q.t[0]);
    w[1] = static_cast<float>(((q.r[3] * px + q.r[4] * py) + q.r[5] * pz) + q.t[1]);
    w[2] = static_cast<float>(((q.r[6] * px + q.r[7] * py) + q.r[8] * pz) + q.t[2]);
}

// ---------------------------------------------------------------------------------------------------
// voxel keys (src/voxel_grid.h:68-75): trunc(x / voxel_size) with an IEEE float divide.
__host__ __device__ inline long long axis_index(float v, float voxel_size) {
    return static_cast<long long>(v / voxel_size);
}

// the reference's 22-bit hash (src/voxel_grid.h:31-38); only needed for robin-order emulation / parity dumps
__host__ __device__ inline uint32_t reference_hash22(long long ix, long long iy, long long iz) {
    const unsigned long long a = static_cast<unsigned long long>(ix) * 73856093ull;
    const unsigned long long b = static_cast<unsigned long long>(iy) * 19349669ull;
    const unsigned long long c = static_cast<unsigned long long>(iz) * 83492791ull;
    return static_cast<uint32_t>((a ^ b ^ c) & ((1ull << 22) - 1ull));
}

// Packed 64-bit table key: 21 bits per axis, biased by 2^20. Indices outside [-2^20, 2^20) are not
// representable (LO_E_RANGE): at 0.05 m voxels that is +-52 km.
constexpr long long kAxisBias = 1ll << 20;
constexpr unsigned long long kKeyEmpty = ~0ull;
constexpr unsigned long long kKeyTomb = ~0ull - 1ull;

__host__ __device__ inline bool axis_in_range(long long i) { return i >= -kAxisBias && i < kAxisBias; }
__host__ __device__ inline unsigned long long pack_key(long long ix, long long iy, long long iz) {
    return (static_cast<unsigned long long>(ix + kAxisBias) << 42) | (static_cast<unsigned long long>(iy + kAxisBias) << 21) |
           static_cast<unsigned long long>(iz + kAxisBias);
}
__host__ __device__ inline void unpack_key(unsigned long long k, long long& ix, long long& iy, long long& iz) {
    ix = static_cast<long long>((k >> 42) & 0x1FFFFFull) - kAxisBias;
    iy = static_cast<long long>((k >> 21) & 0x1FFFFFull) - kAxisBias;
    iz = static_cast<long long>(k & 0x1FFFFFull) - kAxisBias;
}

// probe hash of the device table (free choice: lookups are order independent). Brick-coherent: the 2x2x2 voxels that
// share (ix>>1, iy>>1, iz>>1) get 8 consecutive home slots — one 128-byte line — chosen by a 32-bit mix of the
// brick coordinates; the slot inside the line is the three parity bits. A query's 27 neighbours
// therefore touch at most 8 lines instead of 27 scattered sectors, and the 27 lanes' probes coalesce.
__host__ __device__ inline uint32_t probe_hash(unsigned long long k) {
    const uint32_t local = static_cast<uint32_t>(((k >> 42) & 1ull) << 2 | ((k >> 21) & 1ull) << 1 | (k & 1ull));
    // brick coordinates (20 bits each) mixed by three odd multipliers and murmur3's 32-bit finaliser
    const uint32_t bx = static_cast<uint32_t>(k >> 43) & 0xFFFFFu, by = static_cast<uint32_t>(k >> 22) & 0xFFFFFu,
                   bz = static_cast<uint32_t>(k >> 1) & 0xFFFFFu;
    uint32_t h = (bx * 0x9E3779B1u) ^ (by * 0x85EBCA77u) ^ (bz * 0xC2B2AE3Du);
    h ^= h >> 16; h *= 0x85EBCA6Bu;
    h ^= h >> 13; h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return (h << 3) | local;
}

// ---------------------------------------------------------------------------------------------------
// Voxel table: open addressing, linear probing, 16-byte slots. A voxel's payload (point and normal
// slabs) is addressed by `vid`, so rebuilding the table moves 16 B per voxel, not the payload.
struct alignas(16) Slot {
    unsigned long long key;  // packed key, kKeyEmpty or kKeyTomb
    uint32_t vid;            // payload id
    uint32_t count;          // points stored (<= width)
};

struct GridView {            // passed by value to kernels
    Slot* slots;
    float4* points;          // [vid * width + s]  xyz + w (w = arrival ticket while inserting)
    float4* normals;         // [vid * width + s]
    uint32_t mask;           // capacity - 1
    uint32_t width;          // entries per voxel slab (stride); >= cap
    uint32_t cap;            // max(max_points, 1): a voxel accepts points while it holds fewer (voxel_grid.h:89-91)
    float voxel_size;
    uint32_t shard_rank;     // this grid keeps only the voxels whose hash range it owns (lo_grid_set_shard);
    uint32_t shard_world;    // world <= 1: unsharded
};

// Owner of a voxel in a key-frame map partitioned by voxel-hash range over `world` ranks (BASELINE.json config 4):
// the reference's 22-bit voxel hash (src/voxel_grid.h:31-38) cut into `world` equal contiguous ranges.
__host__ __device__ inline uint32_t shard_owner(long long ix, long long iy, long long iz, uint32_t world) {
    return static_cast<uint32_t>((static_cast<unsigned long long>(reference_hash22(ix, iy, iz)) * world) >> 22);
}

// returns slot index or 0xFFFFFFFF
__device__ inline uint32_t table_find(const GridView& g, unsigned long long key) {
    uint32_t h = probe_hash(key) & g.mask;
    while (true) {
        const unsigned long long k = g.slots[h].key;
        if (k == key) return h;
        if (k == kKeyEmpty) return 0xFFFFFFFFu;
        h = (h + 1) & g.mask;
    }
}

}  // namespace lo
