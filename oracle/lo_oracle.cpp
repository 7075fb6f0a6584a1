// TEST INFRASTRUCTURE (oracle/) — the "port": a self-contained CPU restatement of the reference's
// scan-to-keyframe registration path. It is the checker the GPU parity tests compare against; it is
// never linked, imported or executed by the product (lidar_odometry_demo_b200/). Allowed users: tests/,
// __graft_entry__.smoke(), bench.py's cpu_baseline and --impl reference legs.
//
// What is restated, and from where (paths relative to /root/reference):
//   voxel keys / hash / VoxelGrid            src/voxel_grid.h:20-38,56-110,112-162,164-251
//   VoxelWithPlanes payload                  src/voxel_with_planes.h:10-36
//   tsl::robin_map bucket behaviour          third-party/robin-map/include/tsl/robin_hash.h:1163-1410,
//                                            robin_growth_policy.h:90-128 (v1.2.1 @851a59e, vendored)
//   Pose3D algebra                           src/pose_3d.h:10-59
//   deskew / rigid transforms                src/utils/cloud_transform.h:15-97
//   time normalisation, range filter         src/utils/point_time_normalize.h:14-39, range_filter.h:12-28
//   CloudClassifier                          src/utils/cloud_classifier.h:19-168
//   CloudMatcher::align, LidarOdometry       restated_pipeline.inc (src/cloud_matcher.cpp, lidar_odometry.cpp)
//   ceres::Solve                             mini_ceres.h (third-party, absent: PARITY UNPINNED for that part)
//   PCL NormalEstimation (test only)         orc_pca_normals below (third-party, absent: PARITY UNPINNED)
//
// Pinning: tests/test_oracle_cpu.py compares every function below with oracle/_ref/liblo_ref.so (the
// reference's own headers compiled from /root/reference) bit for bit on randomised inputs, and
// tests/golden/ holds vectors generated from that library. Float evaluation orders follow Eigen 3.4 as
// documented in oracle/shim/eigen3/Eigen/Eigen (unverifiable here: Eigen is not installed).
// Build: make -C oracle port   (g++ -O3 -ffp-contract=off, no -march: as CMakeLists.txt:4-8)
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <numbers>
#include <vector>

#include "mini_ceres.h"
#include "oracle_api.h"
#include "orc_types.h"

namespace port {

// ------------------------------------------------------------------------------------------------
// Voxel key and hash (voxel_grid.h:20-38,68-75)
struct Key {
    int64_t ix, iy, iz;
    bool operator==(const Key& o) const { return ix == o.ix && iy == o.iy && iz == o.iz; }
};

inline int64_t axis_index(float v, float voxel_size) {
    return static_cast<int64_t>(v / voxel_size);  // IEEE float divide, truncation toward zero
}
inline Key key_of(float x, float y, float z, float voxel_size) {
    return {axis_index(x, voxel_size), axis_index(y, voxel_size), axis_index(z, voxel_size)};
}
inline uint64_t hash_of(const Key& k) {
    const uint64_t a = static_cast<uint64_t>(k.ix) * 73856093ull;
    const uint64_t b = static_cast<uint64_t>(k.iy) * 19349669ull;
    const uint64_t c = static_cast<uint64_t>(k.iz) * 83492791ull;
    return (a ^ b ^ c) & ((1ull << 22) - 1ull);
}

// ------------------------------------------------------------------------------------------------
// Robin-hood bucket array with the observable behaviour of tsl::robin_map<Key, V, Hash> as configured
// by the reference (default ctor: 0 buckets, power-of-two growth x2, max load 0.5, no shrinking):
// lookup, insertion with displacement, growth/rehash order, backward-shift erase and bucket-order
// iteration. Values are 32-bit payload ids.
class RobinIndex {
public:
    static constexpr int32_t kEmpty = -1;
    static constexpr int32_t kDistLimit = 8192;  // robin_hash.h:327
    struct Bucket {
        int32_t dist = kEmpty;
        uint32_t hash = 0;
        Key key{0, 0, 0};
        uint32_t payload = 0;
    };

    size_t size() const { return count_; }
    size_t bucket_count() const { return buckets_.size(); }
    const std::vector<Bucket>& buckets() const { return buckets_; }

    void clear() {  // robin_hash.h:725-736 with min_load_factor == 0: buckets kept, emptied
        for (auto& b : buckets_) b.dist = kEmpty;
        count_ = 0;
        grow_on_next_insert_ = false;
    }

    // returns bucket position or SIZE_MAX
    size_t find(const Key& key, uint64_t hash) const {
        if (buckets_.empty()) return SIZE_MAX;
        size_t ib = hash & mask_;
        int32_t dist = 0;
        while (dist <= buckets_[ib].dist) {
            if (buckets_[ib].key == key) return ib;
            ib = (ib + 1) & mask_;
            dist++;
        }
        return SIZE_MAX;
    }

    // Insert a key known to be absent (the reference always calls find() first, voxel_grid.h:82-87).
    void insert_new(const Key& key, uint64_t hash, uint32_t payload) {
        size_t ib = 0;
        int32_t dist = 0;
        if (!buckets_.empty()) probe_to_slot(hash, ib, dist);
        while (grow_if_needed(dist)) probe_to_slot(hash, ib, dist);
        Bucket carry;
        carry.dist = dist;
        carry.hash = static_cast<uint32_t>(hash);
        carry.key = key;
        carry.payload = payload;
        if (buckets_[ib].dist == kEmpty) {
            buckets_[ib] = carry;
        } else {
            displace_from(ib, carry, /*may_flag_growth=*/true);
        }
        count_++;
    }

    // Erase the element in bucket `pos` (backward shift, robin_hash.h:1184-1213). Returns true when a
    // following element moved into `pos` (the reference's iterator then stays on the same bucket).
    bool erase_at(size_t pos) {
        buckets_[pos].dist = kEmpty;
        count_--;
        size_t prev = pos;
        size_t ib = (pos + 1) & mask_;
        while (buckets_[ib].dist > 0) {
            buckets_[prev] = buckets_[ib];
            buckets_[prev].dist -= 1;
            buckets_[ib].dist = kEmpty;
            prev = ib;
            ib = (ib + 1) & mask_;
        }
        return buckets_[pos].dist != kEmpty;
    }

private:
    void probe_to_slot(uint64_t hash, size_t& ib, int32_t& dist) const {
        ib = hash & mask_;
        dist = 0;
        while (dist <= buckets_[ib].dist) {
            ib = (ib + 1) & mask_;
            dist++;
        }
    }

    // robin_hash.h:1283-1317 / :1352-1370: the carried element takes the bucket of any element that is
    // strictly closer to its ideal bucket; that element is carried on.
    void displace_from(size_t ib, Bucket carry, bool may_flag_growth) {
        while (true) {
            Bucket& b = buckets_[ib];
            if (b.dist == kEmpty) {
                b = carry;
                return;
            }
            if (carry.dist > b.dist) {
                if (may_flag_growth && carry.dist > kDistLimit) grow_on_next_insert_ = true;
                std::swap(carry, b);
            }
            ib = (ib + 1) & mask_;
            carry.dist++;
        }
    }

    bool grow_if_needed(int32_t dist) {  // robin_hash.h:1379-1403 (shrink branch inert: min load 0)
        if (grow_on_next_insert_ || dist > kDistLimit || count_ >= load_threshold_) {
            rehash(buckets_.empty() ? 2 : buckets_.size() * 2);  // power_of_two_growth_policy<2>
            grow_on_next_insert_ = false;
            return true;
        }
        return false;
    }

    void rehash(size_t new_count) {  // robin_hash.h:1319-1341: old buckets in order, stored hash reused
        std::vector<Bucket> old;
        old.swap(buckets_);
        buckets_.assign(new_count, Bucket());
        mask_ = new_count - 1;
        load_threshold_ = static_cast<size_t>(static_cast<float>(new_count) * 0.5f);
        for (const Bucket& b : old) {
            if (b.dist == kEmpty) continue;
            Bucket carry = b;
            carry.dist = 0;
            displace_from(b.hash & mask_, carry, /*may_flag_growth=*/false);
        }
    }

    std::vector<Bucket> buckets_;
    size_t mask_ = 0;
    size_t count_ = 0;
    size_t load_threshold_ = 0;
    bool grow_on_next_insert_ = false;
};

// ------------------------------------------------------------------------------------------------
// VoxelGrid restated over RobinIndex. A voxel is the list of (point, normal) it accepted, in arrival
// order (voxel_with_planes.h:22,26-30); its "origin" is the first of them (:32-35).
struct StoredPoint {
    float p[3];
    float n[3];
};

class Grid {
public:
    Grid(float voxel_size, size_t max_points) : voxel_size_(voxel_size), max_points_(max_points) {}

    float voxel_size() const { return voxel_size_; }
    size_t size() const { return index_.size(); }
    void set_max_points(size_t m) { max_points_ = m; }   // voxel_grid.h:56-59

    void add(float x, float y, float z, float nx, float ny, float nz) {  // voxel_grid.h:77-110
        const Key k = key_of(x, y, z, voxel_size_);
        const uint64_t h = hash_of(k);
        const size_t pos = index_.find(k, h);
        if (pos == SIZE_MAX) {
            uint32_t id;
            if (!free_.empty()) {
                id = free_.back();
                free_.pop_back();
                voxels_[id].clear();
            } else {
                id = static_cast<uint32_t>(voxels_.size());
                voxels_.emplace_back();
            }
            voxels_[id].push_back({{x, y, z}, {nx, ny, nz}});
            index_.insert_new(k, h, id);
        } else {
            auto& v = voxels_[index_.buckets()[pos].payload];
            if (v.size() < max_points_) v.push_back({{x, y, z}, {nx, ny, nz}});
        }
    }

    template <typename F>
    void for_each_voxel(F&& f) const {  // robin_map iteration = ascending bucket position
        for (const auto& b : index_.buckets())
            if (b.dist != RobinIndex::kEmpty) f(voxels_[b.payload]);
    }

    const std::vector<StoredPoint>* lookup(const Key& k) const {
        const size_t pos = index_.find(k, hash_of(k));
        return pos == SIZE_MAX ? nullptr : &voxels_[index_.buckets()[pos].payload];
    }

    void radius_cleanup(const float c[3], float radius) {  // voxel_grid.h:236-246
        const float r2 = radius * radius;
        size_t pos = 0;
        while (pos < index_.bucket_count()) {
            const auto& b = index_.buckets()[pos];
            if (b.dist == RobinIndex::kEmpty) { pos++; continue; }
            const StoredPoint& o = voxels_[b.payload].front();
            const float dx = o.p[0] - c[0], dy = o.p[1] - c[1], dz = o.p[2] - c[2];
            const float d2 = dx * dx + (dy * dy + dz * dz);  // Vector3f::squaredNorm order
            if (d2 > r2) {
                free_.push_back(b.payload);
                if (!index_.erase_at(pos)) pos++;
            } else {
                pos++;
            }
        }
    }

private:
    float voxel_size_;
    size_t max_points_;
    RobinIndex index_;
    std::vector<std::vector<StoredPoint>> voxels_;
    std::vector<uint32_t> free_;
};

}  // namespace port

namespace port {

// ------------------------------------------------------------------------------------------------
// Float pose algebra (pose_3d.h:10-59) with Eigen 3.4's evaluation orders written out.
// pose7 = {tx,ty,tz,qw,qx,qy,qz}.
struct Quat {
    float w, x, y, z;
};

inline void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

inline void rotate(const Quat& q, const float v[3], float o[3]) {  // Quaternionf * Vector3f
    const float qv[3] = {q.x, q.y, q.z};
    float uv[3], c[3];
    cross3(qv, v, uv);
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    cross3(qv, uv, c);
    for (int i = 0; i < 3; i++) o[i] = (v[i] + uv[i] * q.w) + c[i];
}

inline Quat qmul(const Quat& a, const Quat& b) {  // Quaternionf * Quaternionf (scalar formula)
    return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
}

inline float qdot(const Quat& a, const Quat& b) { return (a.x * b.x + a.z * b.z) + (a.y * b.y + a.w * b.w); }

inline Quat qinverse(const Quat& q) {
    const float n2 = qdot(q, q);
    if (n2 > 0.f) return {q.w / n2, -q.x / n2, -q.y / n2, -q.z / n2};
    return {0.f, 0.f, 0.f, 0.f};
}

inline void rotation_matrix(const Quat& q, float r[9]) {  // Quaternionf::toRotationMatrix, row-major
    const float tx = 2.f * q.x, ty = 2.f * q.y, tz = 2.f * q.z;
    const float twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    const float txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const float tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    r[0] = 1.f - (tyy + tzz); r[1] = txy - twz;         r[2] = txz + twy;
    r[3] = txy + twz;         r[4] = 1.f - (txx + tzz); r[5] = tyz - twx;
    r[6] = txz - twy;         r[7] = tyz + twx;         r[8] = 1.f - (txx + tyy);
}

inline Quat quat_of(const float p[7]) { return {p[3], p[4], p[5], p[6]}; }
inline void set_pose(float out[7], const float t[3], const Quat& q) {
    out[0] = t[0]; out[1] = t[1]; out[2] = t[2];
    out[3] = q.w; out[4] = q.x; out[5] = q.y; out[6] = q.z;
}

inline void pose_compose(const float a[7], const float b[7], float out[7]) {  // pose_3d.h:29-32
    float rt[3];
    rotate(quat_of(a), b, rt);
    const float t[3] = {a[0] + rt[0], a[1] + rt[1], a[2] + rt[2]};
    set_pose(out, t, qmul(quat_of(a), quat_of(b)));
}
inline void pose_inverse(const float a[7], float out[7]) {  // pose_3d.h:34-39
    const Quat qi = qinverse(quat_of(a));
    const float neg[3] = {-a[0], -a[1], -a[2]};
    float t[3];
    rotate(qi, neg, t);
    set_pose(out, t, qi);
}
inline void pose_relative_to(const float a[7], const float target[7], float out[7]) {  // pose_3d.h:23-27
    float inv[7];
    pose_inverse(a, inv);
    pose_compose(inv, target, out);
}

// Quaternionf::slerp (Eigen 3.4 Quaternion.h), used by the deskew.
inline Quat slerp(const Quat& a, float t, const Quat& b) {
    const float one = 1.f - std::numeric_limits<float>::epsilon();
    const float d = qdot(a, b);
    const float ad = std::abs(d);
    float s0, s1;
    if (ad >= one) {
        s0 = 1.f - t;
        s1 = t;
    } else {
        const float theta = std::acos(ad);
        const float sin_theta = std::sin(theta);
        s0 = std::sin((1.f - t) * theta) / sin_theta;
        s1 = std::sin(t * theta) / sin_theta;
    }
    if (d < 0.f) s1 = -s1;
    return {s0 * a.w + s1 * b.w, s0 * a.x + s1 * b.x, s0 * a.y + s1 * b.y, s0 * a.z + s1 * b.z};
}

// ------------------------------------------------------------------------------------------------
// Point utilities
inline void time_normalize(const OrcXYZIRT* in, size_t n, OrcXYZIRT* out) {  // point_time_normalize.h:14-39
    float lo = std::numeric_limits<float>::max();
    float hi = std::numeric_limits<float>::lowest();
    for (size_t i = 0; i < n; i++) {
        lo = std::min(lo, in[i].time);
        hi = std::max(hi, in[i].time);
    }
    const float range = hi - lo;
    for (size_t i = 0; i < n; i++) {
        out[i] = in[i];
        out[i].time = (in[i].time - lo) / range;
    }
}

inline void deskew(const OrcXYZIRT* in, size_t n, const float start[7], const float end[7], OrcXYZIRT* out) {
    // cloud_transform.h:15-40. Note the translation weights: start.t * time + end.t * (1.0 - time),
    // where (1.0 - time) is evaluated in double and then narrowed to float.
    const Quat qs = quat_of(start), qe = quat_of(end);
    for (size_t i = 0; i < n; i++) {
        const float t = in[i].time;
        const float v[3] = {in[i].x, in[i].y, in[i].z};
        float r[3];
        rotate(slerp(qs, t, qe), v, r);
        const float w_end = static_cast<float>(1.0 - static_cast<double>(t));
        out[i] = in[i];
        out[i].x = (r[0] + start[0] * t) + end[0] * w_end;
        out[i].y = (r[1] + start[1] * t) + end[1] * w_end;
        out[i].z = (r[2] + start[2] * t) + end[2] * w_end;
    }
}

inline void mat_vec(const float r[9], const float v[3], float o[3]) {  // Matrix3f * Vector3f
    for (int i = 0; i < 3; i++) o[i] = (r[i * 3] * v[0] + r[i * 3 + 1] * v[1]) + r[i * 3 + 2] * v[2];
}

inline void rigid_xyz(const OrcXYZ* in, size_t n, const float pose[7], OrcXYZ* out) {  // cloud_transform.h:42-66
    float r[9];
    rotation_matrix(quat_of(pose), r);
    for (size_t i = 0; i < n; i++) {
        const float v[3] = {in[i].x, in[i].y, in[i].z};
        float o[3];
        mat_vec(r, v, o);
        out[i] = in[i];
        out[i].x = o[0] + pose[0]; out[i].y = o[1] + pose[1]; out[i].z = o[2] + pose[2];
    }
}

inline void rigid_pn(const OrcPN* in, size_t n, const float pose[7], OrcPN* out) {  // cloud_transform.h:68-97
    float r[9];
    rotation_matrix(quat_of(pose), r);
    for (size_t i = 0; i < n; i++) {
        const float v[3] = {in[i].x, in[i].y, in[i].z};
        const float nv[3] = {in[i].nx, in[i].ny, in[i].nz};
        float o[3], on[3];
        mat_vec(r, v, o);
        mat_vec(r, nv, on);
        out[i] = in[i];
        out[i].x = o[0] + pose[0]; out[i].y = o[1] + pose[1]; out[i].z = o[2] + pose[2];
        out[i].nx = on[0]; out[i].ny = on[1]; out[i].nz = on[2];
    }
}

inline size_t range_filter(const OrcPN* in, size_t n, float mn, float mx, OrcPN* out) {  // range_filter.h:12-28
    const float lo = mn * mn, hi = mx * mx;
    size_t k = 0;
    for (size_t i = 0; i < n; i++) {
        const float r2 = in[i].x * in[i].x + in[i].y * in[i].y + in[i].z * in[i].z;  // plain C++: left to right
        if (r2 >= lo && r2 <= hi) out[k++] = in[i];
    }
    return k;
}

// findMatchingPairs for one query (voxel_grid.h:164-204,212-226)
inline bool correspondence(const Grid& g, const float rot9[9], const float pose[7], const OrcXYZ& q, float max_d2,
                           double out9[9]) {
    const double pl[3] = {q.x, q.y, q.z};
    float w[3];
    for (int i = 0; i < 3; i++) {
        const double v = (static_cast<double>(rot9[i * 3]) * pl[0] + static_cast<double>(rot9[i * 3 + 1]) * pl[1]) +
                         static_cast<double>(rot9[i * 3 + 2]) * pl[2];
        w[i] = static_cast<float>(v + static_cast<double>(pose[i]));
    }
    const Key o = key_of(w[0], w[1], w[2], g.voxel_size());
    const double gate = static_cast<double>(max_d2);
    double best = std::numeric_limits<double>::max();
    const StoredPoint* hit = nullptr;
    for (int64_t ix = o.ix - 1; ix <= o.ix + 1; ix++)
        for (int64_t iy = o.iy - 1; iy <= o.iy + 1; iy++)
            for (int64_t iz = o.iz - 1; iz <= o.iz + 1; iz++) {
                const auto* v = g.lookup({ix, iy, iz});
                if (!v) continue;
                for (const StoredPoint& s : *v) {
                    const float dx = w[0] - s.p[0], dy = w[1] - s.p[1], dz = w[2] - s.p[2];
                    const double d2 = static_cast<double>(dx * dx + (dy * dy + dz * dz));
                    if (d2 < gate && d2 < best) {
                        best = d2;
                        hit = &s;
                    }
                }
            }
    if (!(best < gate)) return false;
    for (int i = 0; i < 3; i++) {
        out9[i] = pl[i];
        out9[3 + i] = static_cast<double>(hit->p[i]);
        out9[6 + i] = static_cast<double>(hit->n[i]);
    }
    return true;
}

}  // namespace port

namespace port {

// ------------------------------------------------------------------------------------------------
// CloudClassifier::classify (cloud_classifier.h:19-168): ring x azimuth organisation, LOAM-style
// curvature over the FLATTENED grid, normals from two neighbours on the previous ring.
// The reference calls atan2/sqrt/powf unqualified on floats; this restatement (and the oracle/_ref
// build, which includes <math.h> first) resolves them to the float overloads, as a translation unit
// that has seen libstdc++'s <math.h> does. See DESIGN.md "numerics".
inline size_t classify(const OrcXYZIRT* in, size_t n, OrcPN* planar_out, size_t* n_unclassified, OrcXYZIRT* unclassified_out = nullptr) {
    // :21-31 group by ring (narrowed to uint8_t), input order kept inside a ring
    std::map<uint8_t, std::vector<OrcXYZIRT>> rings;
    for (size_t i = 0; i < n; i++) rings[static_cast<uint8_t>(in[i].ring)].push_back(in[i]);
    size_t width = 0;  // :34-40
    for (const auto& r : rings) width = std::max(width, r.second.size());
    const size_t height = rings.size();

    // :42-64 azimuth binning, last writer wins, empty cells are zero-initialised points
    std::vector<OrcXYZIRT> grid(height * width);
    std::memset(static_cast<void*>(grid.data()), 0, grid.size() * sizeof(OrcXYZIRT));
    size_t row = 0;
    for (const auto& r : rings) {
        for (const OrcXYZIRT& p : r.second) {
            const float azimuth = static_cast<float>(static_cast<double>(std::atan2(-p.y, p.x)) + std::numbers::pi);
            const double scaled = static_cast<double>(azimuth * static_cast<float>(width)) / (2.0 * std::numbers::pi);
            const size_t col = static_cast<size_t>(std::abs(scaled));
            if (col < width) grid[row * width + col] = p;
        }
        row++;
    }

    // :81-105 curvature written into the intensity field; the first and last 4 flattened cells keep theirs
    const int win = 4;
    const float intensity_max = 1000.0f;
    const long total = static_cast<long>(grid.size());
    for (long i = win; i < total - win; i++) {
        OrcXYZIRT& c = grid[i];
        const float range = c.x * c.x + c.y * c.y + c.z * c.z;
        if (static_cast<double>(range) < 0.1) {
            c.intensity = intensity_max;
            continue;
        }
        float dx = static_cast<float>(static_cast<double>(-c.x) * 9.0);
        float dy = static_cast<float>(static_cast<double>(-c.y) * 9.0);
        float dz = static_cast<float>(static_cast<double>(-c.z) * 9.0);
        for (int w = -win; w <= win; w++) {
            dx += grid[i + w].x;
            dy += grid[i + w].y;
            dz += grid[i + w].z;
        }
        c.intensity = std::sqrt(dx * dx + dy * dy + dz * dz) / range;
    }

    // :107-165 normals
    const float flat = 0.05f;
    const double neighbour_gate = static_cast<double>(flat) * 10.0;
    size_t n_planar = 0, n_other = 0;
    const long W = static_cast<long>(width), H = static_cast<long>(height);
    for (long ray = 1; ray < H; ray++) {
        for (long col = win; col < W - win; col++) {
            const OrcXYZIRT& p = grid[ray * W + col];
            if (p.intensity < flat) {
                const OrcXYZIRT* left = nullptr;
                const OrcXYZIRT* right = nullptr;
                for (long c = col - win; c < col; c++) {
                    const OrcXYZIRT& q = grid[(ray - 1) * W + c];
                    if (static_cast<double>(q.intensity) < neighbour_gate) { left = &q; break; }
                }
                for (long c = col + win; c > col; c--) {
                    const OrcXYZIRT& q = grid[(ray - 1) * W + c];
                    if (static_cast<double>(q.intensity) < neighbour_gate) { right = &q; break; }
                }
                if (left && right) {
                    const float a[3] = {left->x - p.x, left->y - p.y, left->z - p.z};
                    const float b[3] = {right->x - p.x, right->y - p.y, right->z - p.z};
                    float nrm[3];
                    cross3(a, b, nrm);
                    const float n2 = nrm[0] * nrm[0] + (nrm[1] * nrm[1] + nrm[2] * nrm[2]);
                    if (n2 > 0.f) {
                        const float len = std::sqrt(n2);
                        nrm[0] /= len; nrm[1] /= len; nrm[2] /= len;
                    }
                    OrcPN o;
                    std::memset(static_cast<void*>(&o), 0, sizeof(o));
                    o.x = p.x; o.y = p.y; o.z = p.z; o.pad0 = 1.0f;
                    o.nx = nrm[0]; o.ny = nrm[1]; o.nz = nrm[2];
                    planar_out[n_planar++] = o;
                } else {
                    if (unclassified_out) unclassified_out[n_other] = p;   // :156 the organised point, intensity = curvature
                    n_other++;
                }
            } else if (p.intensity < intensity_max) {
                if (unclassified_out) unclassified_out[n_other] = p;       // :160-162
                n_other++;
            }
        }
    }
    if (n_unclassified) *n_unclassified = n_other;
    return n_planar;
}

// ------------------------------------------------------------------------------------------------
// PCA normals: what test/test.cpp:196-205 asks of pcl::NormalEstimation (radius search 0.25 m,
// viewpoint at the origin). PCL is absent, so this is a plain restatement of the textbook procedure
// PCL documents: neighbours within `radius` (the point itself included), fewer than 3 -> NaN; centroid
// and 3x3 covariance; eigenvector of the smallest eigenvalue; flipped so that it faces the viewpoint.
// Computed in double with cyclic Jacobi sweeps. PARITY UNPINNED.
inline void smallest_eigenvector(const double cov[6] /*xx,xy,xz,yy,yz,zz*/, double out[3]) {
    double a[3][3] = {{cov[0], cov[1], cov[2]}, {cov[1], cov[3], cov[4]}, {cov[2], cov[4], cov[5]}};
    double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 32; sweep++) {
        const double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
        if (off < 1e-300) break;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                if (a[p][q] == 0.0) continue;
                const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 3; k++) {
                    const double akp = a[k][p], akq = a[k][q];
                    a[k][p] = c * akp - s * akq;
                    a[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 3; k++) {
                    const double apk = a[p][k], aqk = a[q][k];
                    a[p][k] = c * apk - s * aqk;
                    a[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 3; k++) {
                    const double vkp = v[k][p], vkq = v[k][q];
                    v[k][p] = c * vkp - s * vkq;
                    v[k][q] = s * vkp + c * vkq;
                }
            }
    }
    int m = 0;
    if (a[1][1] < a[m][m]) m = 1;
    if (a[2][2] < a[m][m]) m = 2;
    const double len = std::sqrt(v[0][m] * v[0][m] + v[1][m] * v[1][m] + v[2][m] * v[2][m]);
    for (int k = 0; k < 3; k++) out[k] = v[k][m] / len;
}

inline void pca_normals(const OrcXYZ* pts, size_t n, float radius, float* normals3) {
    // uniform cells of edge `radius`; a neighbour lies in the 27 cells around the query's cell
    struct CellKey {
        int64_t x, y, z;
        bool operator<(const CellKey& o) const { return x != o.x ? x < o.x : (y != o.y ? y < o.y : z < o.z); }
    };
    auto cell = [&](const OrcXYZ& p) {
        return CellKey{static_cast<int64_t>(std::floor(p.x / radius)), static_cast<int64_t>(std::floor(p.y / radius)),
                       static_cast<int64_t>(std::floor(p.z / radius))};
    };
    std::map<CellKey, std::vector<uint32_t>> cells;
    for (size_t i = 0; i < n; i++) cells[cell(pts[i])].push_back(static_cast<uint32_t>(i));
    const double r2 = static_cast<double>(radius) * static_cast<double>(radius);
    const float nanv = std::numeric_limits<float>::quiet_NaN();
    for (size_t i = 0; i < n; i++) {
        const CellKey c = cell(pts[i]);
        double sx = 0, sy = 0, sz = 0;
        size_t cnt = 0;
        std::vector<uint32_t> nb;
        for (int64_t ax = c.x - 1; ax <= c.x + 1; ax++)
            for (int64_t ay = c.y - 1; ay <= c.y + 1; ay++)
                for (int64_t az = c.z - 1; az <= c.z + 1; az++) {
                    auto it = cells.find({ax, ay, az});
                    if (it == cells.end()) continue;
                    for (uint32_t j : it->second) {
                        const double dx = double(pts[j].x) - double(pts[i].x), dy = double(pts[j].y) - double(pts[i].y),
                                     dz = double(pts[j].z) - double(pts[i].z);
                        if (dx * dx + dy * dy + dz * dz <= r2) {
                            nb.push_back(j);
                            sx += dx; sy += dy; sz += dz;
                            cnt++;
                        }
                    }
                }
        if (cnt < 3) {
            normals3[i * 3] = normals3[i * 3 + 1] = normals3[i * 3 + 2] = nanv;
            continue;
        }
        const double mx = sx / cnt, my = sy / cnt, mz = sz / cnt;  // centroid relative to the query point
        double cov[6] = {0, 0, 0, 0, 0, 0};
        for (uint32_t j : nb) {
            const double dx = double(pts[j].x) - double(pts[i].x) - mx, dy = double(pts[j].y) - double(pts[i].y) - my,
                         dz = double(pts[j].z) - double(pts[i].z) - mz;
            cov[0] += dx * dx; cov[1] += dx * dy; cov[2] += dx * dz;
            cov[3] += dy * dy; cov[4] += dy * dz; cov[5] += dz * dz;
        }
        for (double& v : cov) v /= static_cast<double>(cnt);
        double nv[3];
        smallest_eigenvector(cov, nv);
        // face the viewpoint (0,0,0): (vp - p) . n >= 0
        const double d = -(double(pts[i].x) * nv[0] + double(pts[i].y) * nv[1] + double(pts[i].z) * nv[2]);
        if (d < 0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
        normals3[i * 3] = static_cast<float>(nv[0]);
        normals3[i * 3 + 1] = static_cast<float>(nv[1]);
        normals3[i * 3 + 2] = static_cast<float>(nv[2]);
    }
}

}  // namespace port

// ------------------------------------------------------------------------------------------------
// C API (oracle_api.h)
using port::Grid;

extern "C" const char* orc_kind(void) { return "port"; }

extern "C" void* orc_grid_create(float voxel_size, size_t max_points) { return new Grid(voxel_size, max_points); }
extern "C" void* orc_grid_clone(const void* g) { return new Grid(*static_cast<const Grid*>(g)); }
extern "C" void orc_grid_destroy(void* g) { delete static_cast<Grid*>(g); }
extern "C" void orc_grid_add_cloud(void* g, const float* pn48, size_t n) {
    const auto* p = reinterpret_cast<const OrcPN*>(pn48);
    auto* grid = static_cast<Grid*>(g);
    for (size_t i = 0; i < n; i++) grid->add(p[i].x, p[i].y, p[i].z, p[i].nx, p[i].ny, p[i].nz);
}
extern "C" void orc_grid_add_cloud_xyz(void* g, const float* xyz16, size_t n) {
    const auto* p = reinterpret_cast<const OrcXYZ*>(xyz16);
    auto* grid = static_cast<Grid*>(g);
    for (size_t i = 0; i < n; i++) grid->add(p[i].x, p[i].y, p[i].z, 0.f, 0.f, 0.f);
}
extern "C" size_t orc_grid_size(const void* g) { return static_cast<const Grid*>(g)->size(); }
extern "C" size_t orc_grid_num_points(const void* g) {
    size_t k = 0;
    static_cast<const Grid*>(g)->for_each_voxel([&](const std::vector<port::StoredPoint>& v) { k += v.size(); });
    return k;
}
extern "C" size_t orc_grid_get_cloud(const void* g, float* out, size_t cap) {  // voxel_grid.h:112-129
    auto* o = reinterpret_cast<OrcPN*>(out);
    size_t k = 0;
    static_cast<const Grid*>(g)->for_each_voxel([&](const std::vector<port::StoredPoint>& v) {
        for (const auto& s : v) {
            if (o && k < cap) {
                OrcPN p;
                std::memset(static_cast<void*>(&p), 0, sizeof(p));
                p.x = s.p[0]; p.y = s.p[1]; p.z = s.p[2]; p.pad0 = 1.0f;
                p.nx = s.n[0]; p.ny = s.n[1]; p.nz = s.n[2];
                o[k] = p;
            }
            k++;
        }
    });
    return k;
}
extern "C" size_t orc_grid_get_cloud_xyz(const void* g, float* out, size_t cap) {  // voxel_grid.h:132-146
    auto* o = reinterpret_cast<OrcXYZ*>(out);
    size_t k = 0;
    static_cast<const Grid*>(g)->for_each_voxel([&](const std::vector<port::StoredPoint>& v) {
        for (const auto& s : v) {
            if (o && k < cap) o[k] = OrcXYZ{s.p[0], s.p[1], s.p[2], 1.0f};
            k++;
        }
    });
    return k;
}
extern "C" size_t orc_grid_get_sparse_xyz(const void* g, float* out, size_t cap) {  // voxel_grid.h:149-162
    auto* o = reinterpret_cast<OrcXYZ*>(out);
    size_t k = 0;
    static_cast<const Grid*>(g)->for_each_voxel([&](const std::vector<port::StoredPoint>& v) {
        const auto& s = v.front();
        if (o && k < cap) o[k] = OrcXYZ{s.p[0], s.p[1], s.p[2], 1.0f};
        k++;
    });
    return k;
}
extern "C" void orc_grid_radius_cleanup(void* g, const float c[3], float radius) {
    static_cast<Grid*>(g)->radius_cleanup(c, radius);
}
extern "C" void orc_get_indices(float voxel_size, const float* xyz, size_t stride, size_t n, int64_t* out3) {
    for (size_t i = 0; i < n; i++) {
        const port::Key k = port::key_of(xyz[i * stride], xyz[i * stride + 1], xyz[i * stride + 2], voxel_size);
        out3[i * 3] = k.ix; out3[i * 3 + 1] = k.iy; out3[i * 3 + 2] = k.iz;
    }
}
extern "C" uint64_t orc_hash(int64_t ix, int64_t iy, int64_t iz) { return port::hash_of({ix, iy, iz}); }

extern "C" void orc_grid_set_max_points(void* g, size_t max_points) { static_cast<Grid*>(g)->set_max_points(max_points); }
// getCorrespondence (voxel_grid.h:164-204) for one query in the grid's frame: the per-query code of findMatchingPairs under
// the identity transform (exact), with the double bound T turned into the float bound that decides the same: d2 < T for a
// float d2 <=> d2 < T rounded up to float.
extern "C" int orc_grid_get_correspondence(const void* g, const float q3[3], double max_sq, double origin3[3], double normal3[3]) {
    const auto* grid = static_cast<const Grid*>(g);
    float bound = static_cast<float>(max_sq);
    if (static_cast<double>(bound) < max_sq) bound = std::nextafter(bound, std::numeric_limits<float>::infinity());
    if (!(max_sq > 0.0)) bound = 0.f;
    const float ident[7] = {0, 0, 0, 1, 0, 0, 0};
    float r9[9];
    port::rotation_matrix(port::quat_of(ident), r9);
    const OrcXYZ q = {q3[0], q3[1], q3[2], 1.f};
    double out9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool ok = port::correspondence(*grid, r9, ident, q, bound, out9);
    for (int k = 0; k < 3; k++) {
        origin3[k] = ok ? out9[3 + k] : 0.0;
        normal3[k] = ok ? out9[6 + k] : 0.0;
    }
    return ok ? 1 : 0;
}

extern "C" size_t orc_grid_find_matching_pairs(const void* g, const float* xyz16, size_t n, const float pose7[7],
                                               float max_dist, double* out9, uint8_t* valid) {
    const auto* grid = static_cast<const Grid*>(g);
    const auto* q = reinterpret_cast<const OrcXYZ*>(xyz16);
    float r9[9];
    port::rotation_matrix(port::quat_of(pose7), r9);  // voxel_grid.h:212
    const float max_d2 = max_dist * max_dist;         // :215
    size_t nv = 0;
    for (size_t i = 0; i < n; i++) {
        for (int k = 0; k < 9; k++) out9[i * 9 + k] = 0.0;
        valid[i] = port::correspondence(*grid, r9, pose7, q[i], max_d2, out9 + i * 9) ? 1 : 0;
        nv += valid[i];
    }
    return nv;
}

extern "C" void orc_time_normalize(const void* in, size_t n, void* out) {
    port::time_normalize(static_cast<const OrcXYZIRT*>(in), n, static_cast<OrcXYZIRT*>(out));
}
extern "C" void orc_transform_nonrigid(const void* in, size_t n, const float s7[7], const float e7[7], void* out) {
    port::deskew(static_cast<const OrcXYZIRT*>(in), n, s7, e7, static_cast<OrcXYZIRT*>(out));
}
extern "C" void orc_transform_xyz(const float* in, size_t n, const float p7[7], float* out) {
    port::rigid_xyz(reinterpret_cast<const OrcXYZ*>(in), n, p7, reinterpret_cast<OrcXYZ*>(out));
}
extern "C" void orc_transform_with_normals(const float* in, size_t n, const float p7[7], float* out) {
    port::rigid_pn(reinterpret_cast<const OrcPN*>(in), n, p7, reinterpret_cast<OrcPN*>(out));
}
extern "C" size_t orc_range_filter(const float* in, size_t n, float mn, float mx, float* out) {
    return port::range_filter(reinterpret_cast<const OrcPN*>(in), n, mn, mx, reinterpret_cast<OrcPN*>(out));
}
extern "C" size_t orc_classify_full(const void* in, size_t n, float* planar_out, void* unclassified_out, size_t* n_unclassified) {
    return port::classify(static_cast<const OrcXYZIRT*>(in), n, reinterpret_cast<OrcPN*>(planar_out), n_unclassified,
                          static_cast<OrcXYZIRT*>(unclassified_out));
}
extern "C" size_t orc_classify(const void* in, size_t n, float* planar_out, size_t* n_unclassified) {
    return port::classify(static_cast<const OrcXYZIRT*>(in), n, reinterpret_cast<OrcPN*>(planar_out), n_unclassified);
}
extern "C" void orc_pose_compose(const float a[7], const float b[7], float out[7]) { port::pose_compose(a, b, out); }
extern "C" void orc_pose_inverse(const float a[7], float out[7]) { port::pose_inverse(a, out); }
extern "C" void orc_pose_relative_to(const float a[7], const float b[7], float out[7]) { port::pose_relative_to(a, b, out); }
extern "C" void orc_pose_rotation_matrix(const float a[7], float r9[9]) { port::rotation_matrix(port::quat_of(a), r9); }
extern "C" void orc_pca_normals(const float* xyz16, size_t n, float radius, float* normals3_out) {
    port::pca_normals(reinterpret_cast<const OrcXYZ*>(xyz16), n, radius, normals3_out);
}

#include "restated_pipeline.inc"
