// lo_b200.hpp — C++ host facade over the C ABI (include/lo_b200.h): the reference's class names, method
// names, argument meaning and return conventions, so that code written against
//   src/pose_3d.h, src/voxel_grid.h, src/cloud_matcher.h, src/lidar_odometry.h, src/utils/*.h
// of vovo-4K/lidar_odometry_demo compiles against this header and runs on the GPU.
//
// Differences a caller can observe (DESIGN.md "boundary"):
//   * Eigen and PCL are optional. Without them the header supplies layout-compatible minimal types
//     (lo::Vector3f, lo::Quaternionf, pcl::PointXYZ / PointNormal / PointCloud<T>, lidar_point::PointXYZIRT).
//     With -DLO_B200_WITH_EIGEN / -DLO_B200_WITH_PCL the real ones are used (same memory layout).
//   * VoxelGrid owns device memory: it is movable, and copying it deep-copies through the host
//     (the reference's VoxelGrid is implicitly copyable, src/lidar_odometry.h:82).
//   * errors of the C ABI become lo::Error exceptions on this side of the boundary (the reference signals none).
//   * there is no CPU fallback: constructing any of these objects without a CUDA device throws.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/lo_b200.h"

#ifdef LO_B200_WITH_EIGEN
#include <eigen3/Eigen/Eigen>
#endif
#ifdef LO_B200_WITH_PCL
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#endif

namespace lo {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline void check(int code) {
    if (code != LO_OK) throw Error(code, std::string("liblo_b200: ") + lo_last_error());
}

#ifdef LO_B200_WITH_EIGEN
using Vector3f = Eigen::Vector3f;
using Quaternionf = Eigen::Quaternionf;
using Matrix3f = Eigen::Matrix3f;
#else
// just enough of Eigen's interface for Pose3D users: x()/y()/z()/w(), construction in Eigen's (w,x,y,z) order
struct Vector3f {
    float v[3] = {0.f, 0.f, 0.f};
    Vector3f() = default;
    Vector3f(float x, float y, float z) : v{x, y, z} {}
    float& x() { return v[0]; }
    float& y() { return v[1]; }
    float& z() { return v[2]; }
    float x() const { return v[0]; }
    float y() const { return v[1]; }
    float z() const { return v[2]; }
    float& operator[](int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
    void setZero() { v[0] = v[1] = v[2] = 0.f; }
    float norm() const { return std::sqrt(v[0] * v[0] + (v[1] * v[1] + v[2] * v[2])); }
};
struct Matrix3f {                       // row-major 3x3, Eigen's element access
    float m[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
    float& operator()(int r, int c) { return m[3 * r + c]; }
    float operator()(int r, int c) const { return m[3 * r + c]; }
    Vector3f operator*(const Vector3f& v) const {
        return Vector3f((m[0] * v.x() + m[1] * v.y()) + m[2] * v.z(), (m[3] * v.x() + m[4] * v.y()) + m[5] * v.z(),
                        (m[6] * v.x() + m[7] * v.y()) + m[8] * v.z());
    }
};
struct Quaternionf {
    float qw = 1.f, qx = 0.f, qy = 0.f, qz = 0.f;
    Quaternionf() = default;
    Quaternionf(float w, float x, float y, float z) : qw(w), qx(x), qy(y), qz(z) {}
    static Quaternionf Identity() { return Quaternionf(); }
    float w() const { return qw; }
    float x() const { return qx; }
    float y() const { return qy; }
    float z() const { return qz; }
    void setIdentity() { *this = Quaternionf(); }
    float dot(const Quaternionf& o) const { return (qx * o.qx + qz * o.qz) + (qy * o.qy + qw * o.qw); }
};
#endif

#ifdef LO_B200_WITH_ROS2
}  // namespace lo
#include <rclcpp/rclcpp.hpp>
namespace lo {
using ParameterValue = rclcpp::ParameterValue;
#else
// stand-in for rclcpp::ParameterValue in builds without ROS 2: a number with the same get<T>() accessor
struct ParameterValue {
    double v = 0.0;
    ParameterValue() = default;
    explicit ParameterValue(double d) : v(d) {}
    explicit ParameterValue(int i) : v(i) {}
    template <typename T>
    T get() const { return static_cast<T>(v); }
};
#endif

// process-wide default context (device 0, or LO_B200_DEVICE); objects may also be bound to an explicit one
class Context {
public:
    explicit Context(int device = 0) { check(lo_ctx_create(device, &h_)); }
    ~Context() { lo_ctx_destroy(h_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    lo_ctx* handle() const { return h_; }
    static Context& instance() {
        static Context ctx(std::getenv("LO_B200_DEVICE") ? std::atoi(std::getenv("LO_B200_DEVICE")) : 0);
        return ctx;
    }

private:
    lo_ctx* h_ = nullptr;
};

}  // namespace lo

// ---- point types (layout of pcl/point_types.h and src/lidar_point_type.h:13-31) --------------------------
#ifndef LO_B200_WITH_PCL
namespace pcl {
struct alignas(16) PointXYZ {
    float x = 0.f, y = 0.f, z = 0.f, pad = 1.f;
    PointXYZ() = default;
    PointXYZ(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct alignas(16) PointNormal {
    float x = 0.f, y = 0.f, z = 0.f, pad0 = 1.f;
    float normal_x = 0.f, normal_y = 0.f, normal_z = 0.f, pad1 = 0.f;
    float curvature = 0.f, pad2[3] = {0.f, 0.f, 0.f};
    PointNormal() = default;
    PointNormal(float x_, float y_, float z_, float nx = 0.f, float ny = 0.f, float nz = 0.f)
        : x(x_), y(y_), z(z_), normal_x(nx), normal_y(ny), normal_z(nz) {}
};
template <typename PointT>
struct PointCloud {
    using Ptr = std::shared_ptr<PointCloud<PointT>>;
    using ConstPtr = std::shared_ptr<const PointCloud<PointT>>;
    std::vector<PointT> points;
    uint32_t width = 0, height = 1;
    size_t size() const { return points.size(); }
    bool empty() const { return points.empty(); }
    PointT& at(size_t i) { return points.at(i); }
    const PointT& at(size_t i) const { return points.at(i); }
    PointT& operator[](size_t i) { return points[i]; }
    const PointT& operator[](size_t i) const { return points[i]; }
    auto begin() { return points.begin(); }
    auto end() { return points.end(); }
    auto begin() const { return points.begin(); }
    auto end() const { return points.end(); }
    void push_back(const PointT& p) { points.push_back(p); width = static_cast<uint32_t>(points.size()); }
    void resize(size_t n) { points.resize(n); width = static_cast<uint32_t>(n); }
};
}  // namespace pcl
#endif
namespace lidar_point {
struct alignas(16) PointXYZIRT {
    float x = 0.f, y = 0.f, z = 0.f, pad0 = 1.f;
    float intensity = 0.f;
    uint16_t ring = 0, pad1 = 0;
    float time = 0.f, pad2 = 0.f;
};
}  // namespace lidar_point
static_assert(sizeof(pcl::PointXYZ) == 16 && sizeof(pcl::PointNormal) == 48 && sizeof(lidar_point::PointXYZIRT) == 32,
              "wire layout of include/lo_b200.h");

// ---- Pose3D (src/pose_3d.h:10-59) --------------------------------------------------------------------------
// Host-side float algebra in Eigen 3.4's operation order (same formulas as csrc/lo_device.cuh, which the
// device uses; the CPU suite checks both against the reference's own test vectors, test/test.cpp:77-149).
class Pose3D {
public:
    lo::Vector3f translation;
    lo::Quaternionf rotation;

    Pose3D() : translation(0.f, 0.f, 0.f), rotation(1.f, 0.f, 0.f, 0.f) {}
    Pose3D(const lo::Vector3f& t, const lo::Quaternionf& q) : translation(t), rotation(q) {}

    Pose3D relativeTo(const Pose3D& target) const { return inverse().compose(target); }

    Pose3D compose(const Pose3D& b) const {
        float rt[3];
        rotate(rotation, b.translation, rt);
        const float aw = rotation.w(), ax = rotation.x(), ay = rotation.y(), az = rotation.z();
        const float bw = b.rotation.w(), bx = b.rotation.x(), by = b.rotation.y(), bz = b.rotation.z();
        return Pose3D(lo::Vector3f(translation.x() + rt[0], translation.y() + rt[1], translation.z() + rt[2]),
                      lo::Quaternionf(aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                                      aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx));
    }

    Pose3D inverse() const {
        const float w = rotation.w(), x = rotation.x(), y = rotation.y(), z = rotation.z();
        const float n2 = (x * x + z * z) + (y * y + w * w);
        lo::Quaternionf inv(0.f, 0.f, 0.f, 0.f);
        if (n2 > 0.f) inv = lo::Quaternionf(w / n2, -x / n2, -y / n2, -z / n2);
        float t[3];
        rotate(inv, lo::Vector3f(-translation.x(), -translation.y(), -translation.z()), t);
        return Pose3D(lo::Vector3f(t[0], t[1], t[2]), inv);
    }

    // Quaternionf::toRotationMatrix() by value, as src/pose_3d.h:41-44 returns it
    lo::Matrix3f rotationMatrix() const {
        float r[9];
        rotationMatrix(r);
        lo::Matrix3f out;
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) out(i, j) = r[3 * i + j];
        return out;
    }
    // the same, row-major into a plain array
    void rotationMatrix(float r[9]) const {
        const float w = rotation.w(), x = rotation.x(), y = rotation.y(), z = rotation.z();
        const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
        const float twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y,
                    tzz = tz * z;
        r[0] = 1.f - (tyy + tzz); r[1] = txy - twz; r[2] = txz + twy;
        r[3] = txy + twz; r[4] = 1.f - (txx + tzz); r[5] = tyz - twx;
        r[6] = txz - twy; r[7] = tyz + twx; r[8] = 1.f - (txx + tyy);
    }

    void to7(float p[7]) const {
        p[0] = translation.x(); p[1] = translation.y(); p[2] = translation.z();
        p[3] = rotation.w(); p[4] = rotation.x(); p[5] = rotation.y(); p[6] = rotation.z();
    }
    static Pose3D from7(const float p[7]) { return Pose3D(lo::Vector3f(p[0], p[1], p[2]), lo::Quaternionf(p[3], p[4], p[5], p[6])); }

private:
    static void rotate(const lo::Quaternionf& q, const lo::Vector3f& v, float o[3]) {
        const float x = q.x(), y = q.y(), z = q.z(), w = q.w();
        float uv[3] = {y * v.z() - z * v.y(), z * v.x() - x * v.z(), x * v.y() - y * v.x()};
        uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
        const float c[3] = {y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]};
        o[0] = (v.x() + uv[0] * w) + c[0];
        o[1] = (v.y() + uv[1] * w) + c[1];
        o[2] = (v.z() + uv[2] * w) + c[2];
    }
};

// ---- VoxelGrid (src/voxel_grid.h:17-257) -------------------------------------------------------------------
class VoxelGrid {
public:
    struct Indices { int64_t x, y, z; };
    struct Correspondence {            // src/voxel_grid.h:40-46
        double source_point_local[3];
        double plane_origin[3];
        double plane_normal[3];
        bool valid = false;
    };

    VoxelGrid() : VoxelGrid(0.5f, 10) {}
    VoxelGrid(float voxel_size, size_t max_points, lo::Context& ctx = lo::Context::instance())
        : ctx_(&ctx), voxel_size_(voxel_size), max_points_(max_points) {
        lo::check(lo_grid_create(ctx.handle(), voxel_size, max_points, 0, &h_));
    }
    ~VoxelGrid() { lo_grid_destroy(h_); }
    VoxelGrid(VoxelGrid&& o) noexcept : ctx_(o.ctx_), h_(o.h_), voxel_size_(o.voxel_size_), max_points_(o.max_points_) { o.h_ = nullptr; }
    VoxelGrid& operator=(VoxelGrid&& o) noexcept {
        if (this != &o) { lo_grid_destroy(h_); ctx_ = o.ctx_; h_ = o.h_; voxel_size_ = o.voxel_size_; max_points_ = o.max_points_; o.h_ = nullptr; }
        return *this;
    }
    // deep copy through a snapshot of the table (lo_grid_snapshot): the copy is in the same state as the original, clean-up
    // history included, which re-inserting getCloud() would not give
    VoxelGrid(const VoxelGrid& o) : VoxelGrid(o.voxel_size_, o.max_points_, *o.ctx_) { restore(o.snapshot()); }
    VoxelGrid& operator=(const VoxelGrid& o) { if (this != &o) { VoxelGrid t(o); *this = std::move(t); } return *this; }

    std::vector<uint8_t> snapshot() const {
        size_t n = 0;
        lo::check(lo_grid_snapshot(h_, nullptr, 0, &n));
        std::vector<uint8_t> blob(n);
        lo::check(lo_grid_snapshot(h_, blob.data(), blob.size(), &n));
        return blob;
    }
    void restore(const std::vector<uint8_t>& blob) { lo::check(lo_grid_restore(h_, blob.data(), blob.size())); }

    void setMaxPoints(size_t max_points) { max_points_ = max_points; lo::check(lo_grid_set_max_points(h_, max_points)); }   // :56-59
    void setVoxelSize(float voxel_size) { voxel_size_ = voxel_size; lo::check(lo_grid_set_voxel_size(h_, voxel_size)); }
    Indices getIndices(float x, float y, float z) const {
        return {static_cast<int64_t>(x / voxel_size_), static_cast<int64_t>(y / voxel_size_), static_cast<int64_t>(z / voxel_size_)};
    }
    void addCloud(const pcl::PointCloud<pcl::PointNormal>& cloud) { lo::check(lo_grid_add_cloud(h_, cloud.points.data(), cloud.points.size(), 1)); }
    void addCloudWithoutNormals(const pcl::PointCloud<pcl::PointXYZ>& cloud) { lo::check(lo_grid_add_cloud(h_, cloud.points.data(), cloud.points.size(), 0)); }

    // reference_order: emit in the reference's tsl::robin_map iteration order (what its getCloud* callers see)
    pcl::PointCloud<pcl::PointNormal>::Ptr getCloud(bool reference_order = true) const { return get<pcl::PointNormal>(0, reference_order); }
    pcl::PointCloud<pcl::PointXYZ>::Ptr getCloudWithoutNormals(bool reference_order = true) const { return get<pcl::PointXYZ>(1, reference_order); }
    pcl::PointCloud<pcl::PointXYZ>::Ptr getSparseCloudWithoutNormals(bool reference_order = true) const { return get<pcl::PointXYZ>(2, reference_order); }

    // :164-204. `query` is a point in the grid's frame; source_point_local stays zero as in the reference.
    Correspondence getCorrespondence(const lo::Vector3f& query, double max_correspondence_distance_sq) const {
        const float q[3] = {query.x(), query.y(), query.z()};
        lo_pair pair;
        int valid = 0;
        lo::check(lo_grid_get_correspondence(h_, q, max_correspondence_distance_sq, &pair, &valid));
        Correspondence c;
        c.valid = valid != 0;
        if (c.valid) std::memcpy(c.source_point_local, pair.source_point_local, sizeof(double) * 9);
        else std::memset(c.source_point_local, 0, sizeof(double) * 9);
        return c;
    }

    std::vector<Correspondence> findMatchingPairs(const pcl::PointCloud<pcl::PointXYZ>& cloud, const Pose3D& pose, float max_distance) const {
        const size_t m = cloud.points.size();
        std::vector<lo_pair> pairs(m);
        std::vector<uint8_t> valid(m);
        size_t n_valid = 0;
        float p7[7];
        pose.to7(p7);
        lo::check(lo_grid_find_matching_pairs(h_, reinterpret_cast<const float*>(cloud.points.data()), m, p7, max_distance, pairs.data(),
                                              valid.data(), &n_valid));
        std::vector<Correspondence> out;   // serial (input) order, valid pairs only, as the reference without TBB
        out.reserve(n_valid);
        for (size_t i = 0; i < m; i++) {
            if (!valid[i]) continue;
            Correspondence c;
            std::memcpy(c.source_point_local, pairs[i].source_point_local, sizeof(double) * 9);
            c.valid = true;
            out.push_back(c);
        }
        return out;
    }
    void radiusCleanup(const lo::Vector3f& pose, float radius) {
        const float c[3] = {pose.x(), pose.y(), pose.z()};
        lo::check(lo_grid_radius_cleanup(h_, c, radius));
    }
    size_t size() const { size_t v = 0, p = 0; lo::check(lo_grid_size(h_, &v, &p)); return v; }
    lo_grid* handle() const { return h_; }

private:
    template <typename PointT>
    typename pcl::PointCloud<PointT>::Ptr get(int mode, bool reference_order) const {
        auto out = std::make_shared<pcl::PointCloud<PointT>>();
        size_t n = 0;
        lo::check(lo_grid_get_cloud(h_, mode, 0, nullptr, 0, &n));
        out->points.resize(n);
        out->width = static_cast<uint32_t>(n);
        if (n) {
            int st = lo_grid_get_cloud(h_, mode, reference_order ? 1 : 0, out->points.data(), n, &n);
            if (st == LO_E_ARG && reference_order) st = lo_grid_get_cloud(h_, mode, 0, out->points.data(), n, &n);  // after radiusCleanup
            lo::check(st);
        }
        return out;
    }
    lo::Context* ctx_;
    lo_grid* h_ = nullptr;
    float voxel_size_;
    size_t max_points_;
};

// ---- CloudMatcher (src/cloud_matcher.h:13-17) ----------------------------------------------------------------
class CloudMatcher {
public:
    Pose3D align(const VoxelGrid& keyframe, const pcl::PointCloud<pcl::PointXYZ>& planar_cloud, const Pose3D& position_guess,
                 lo_align_stats* stats = nullptr) const {
        float guess[7], out[7];
        position_guess.to7(guess);
        lo::check(lo_align(keyframe.handle(), reinterpret_cast<const float*>(planar_cloud.points.data()), planar_cloud.points.size(), guess, out,
                           stats));
        return Pose3D::from7(out);
    }
};

// ---- CloudMatcher over a key-frame map sharded by voxel-hash range, one process per GPU (BASELINE.json config 4) ------
// Extension (the reference is single-process): every rank constructs one with the same NCCL id (ShardedCloudMatcher::
// UniqueId() on rank 0, shipped by the job's own channel), adds the replicated key-frame cloud to keyframe() (only the
// voxels the rank owns are kept) and calls align() with the same queries and guess; every rank receives the same pose.
class ShardedCloudMatcher {
public:
    using UniqueIdBytes = std::vector<uint8_t>;
    static UniqueIdBytes UniqueId() {
        UniqueIdBytes id(LO_SHARDED_ID_BYTES);
        lo::check(lo_sharded_unique_id(id.data()));
        return id;
    }
    ShardedCloudMatcher(const UniqueIdBytes& id, int rank, int world, float voxel_size, size_t max_points, lo::Context& ctx = lo::Context::instance()) {
        lo::check(lo_sharded_create(ctx.handle(), id.empty() ? nullptr : id.data(), rank, world, voxel_size, max_points, 0, &h_));
    }
    ~ShardedCloudMatcher() { lo_sharded_destroy(h_); }
    ShardedCloudMatcher(const ShardedCloudMatcher&) = delete;
    ShardedCloudMatcher& operator=(const ShardedCloudMatcher&) = delete;

    // VoxelGrid::addCloud on this rank's shard
    void addCloud(const pcl::PointCloud<pcl::PointNormal>& cloud) { lo::check(lo_grid_add_cloud(lo_sharded_grid(h_), cloud.points.data(), cloud.points.size(), 1)); }
    size_t size() const { size_t v = 0, p = 0; lo::check(lo_grid_size(lo_sharded_grid(h_), &v, &p)); return v; }
    // CloudMatcher::align (src/cloud_matcher.h:13-17) against the union of all shards; findMatchingPairs' 0.3 m gate
    Pose3D align(const pcl::PointCloud<pcl::PointXYZ>& planar_cloud, const Pose3D& position_guess, float max_distance = 0.3f) const {
        float guess[7], out[7];
        position_guess.to7(guess);
        lo::check(lo_sharded_align(h_, reinterpret_cast<const float*>(planar_cloud.points.data()), planar_cloud.points.size(), guess, max_distance, out, nullptr));
        return Pose3D::from7(out);
    }
    lo_sharded_info info() const {
        lo_sharded_info i;
        lo::check(lo_sharded_info_get(h_, &i));
        return i;
    }

private:
    lo_sharded* h_ = nullptr;
};

// ---- CloudTransformer (src/utils/cloud_transform.h:13-98) and utils:: (point_time_normalize.h, range_filter.h) ---
class CloudTransformer {
public:
    // transformNonRigid (:15-40): `time` is used as the caller left it. (lo_deskew is the fused form of
    // utils::pointTimeNormalize + transformNonRigid that src/lidar_odometry.cpp:25-30 calls back to back.)
    static pcl::PointCloud<lidar_point::PointXYZIRT>::Ptr transformNonRigid(const pcl::PointCloud<lidar_point::PointXYZIRT>& cloud, const Pose3D& start,
                                                                  const Pose3D& end, lo::Context& ctx = lo::Context::instance()) {
        auto out = std::make_shared<pcl::PointCloud<lidar_point::PointXYZIRT>>();
        out->points.resize(cloud.points.size());
        out->width = static_cast<uint32_t>(cloud.points.size());
        float s7[7], e7[7];
        start.to7(s7);
        end.to7(e7);
        lo::check(lo_transform_non_rigid(ctx.handle(), cloud.points.data(), cloud.points.size(), s7, e7, out->points.data()));
        return out;
    }
    static pcl::PointCloud<pcl::PointXYZ>::Ptr transform(const pcl::PointCloud<pcl::PointXYZ>& cloud, const Pose3D& pose,
                                                          lo::Context& ctx = lo::Context::instance()) {
        auto out = std::make_shared<pcl::PointCloud<pcl::PointXYZ>>();
        out->points.resize(cloud.points.size());
        out->width = static_cast<uint32_t>(cloud.points.size());
        float p7[7];
        pose.to7(p7);
        lo::check(lo_transform_xyz(ctx.handle(), reinterpret_cast<const float*>(cloud.points.data()), cloud.points.size(), p7,
                                   reinterpret_cast<float*>(out->points.data())));
        return out;
    }
    static pcl::PointCloud<pcl::PointNormal>::Ptr transformWithNormals(const pcl::PointCloud<pcl::PointNormal>& cloud, const Pose3D& pose,
                                                                        lo::Context& ctx = lo::Context::instance()) {
        auto out = std::make_shared<pcl::PointCloud<pcl::PointNormal>>();
        out->points.resize(cloud.points.size());
        out->width = static_cast<uint32_t>(cloud.points.size());
        float p7[7];
        pose.to7(p7);
        lo::check(lo_transform_with_normals(ctx.handle(), reinterpret_cast<const float*>(cloud.points.data()), cloud.points.size(), p7,
                                            reinterpret_cast<float*>(out->points.data())));
        return out;
    }
};

namespace utils {
// pointTimeNormalize (src/utils/point_time_normalize.h:14-39); the reference's template is only ever instantiated for
// lidar_point::PointXYZIRT (src/lidar_odometry.cpp:25)
template <typename PointType>
inline typename pcl::PointCloud<PointType>::Ptr pointTimeNormalize(const pcl::PointCloud<PointType>& input, lo::Context& ctx = lo::Context::instance()) {
    static_assert(sizeof(PointType) == 32, "pointTimeNormalize takes lidar_point::PointXYZIRT records");
    auto out = std::make_shared<pcl::PointCloud<PointType>>();
    out->points.resize(input.points.size());
    out->width = static_cast<uint32_t>(input.points.size());
    lo::check(lo_point_time_normalize(ctx.handle(), input.points.data(), input.points.size(), out->points.data()));
    return out;
}
inline pcl::PointCloud<pcl::PointNormal>::Ptr rangeFilter(const pcl::PointCloud<pcl::PointNormal>& cloud, float min_range, float max_range,
                                                           lo::Context& ctx = lo::Context::instance()) {
    auto out = std::make_shared<pcl::PointCloud<pcl::PointNormal>>();
    out->points.resize(cloud.points.size());
    size_t n = 0;
    lo::check(lo_range_filter(ctx.handle(), reinterpret_cast<const float*>(cloud.points.data()), cloud.points.size(), min_range, max_range,
                              reinterpret_cast<float*>(out->points.data()), &n));
    out->points.resize(n);
    out->width = static_cast<uint32_t>(n);
    return out;
}
}  // namespace utils

// CloudClassifier::classify (src/utils/cloud_classifier.h:19-168): {planar points with normals, unclassified points}
class CloudClassifier {
public:
    template <typename PointType>
    static std::pair<pcl::PointCloud<pcl::PointNormal>::Ptr, typename pcl::PointCloud<PointType>::Ptr> classify(
        const pcl::PointCloud<PointType>& cloud, lo::Context& ctx = lo::Context::instance()) {
        static_assert(sizeof(PointType) == 32, "classify takes lidar_point::PointXYZIRT records");
        auto planar = std::make_shared<pcl::PointCloud<pcl::PointNormal>>();
        auto unclassified = std::make_shared<pcl::PointCloud<PointType>>();
        const size_t cap = cloud.points.size() ? cloud.points.size() : 1;
        planar->points.resize(cap);
        unclassified->points.resize(cap);
        size_t n = 0, nu = 0;
        lo::check(lo_classify_pair(ctx.handle(), cloud.points.data(), cloud.points.size(), reinterpret_cast<float*>(planar->points.data()), cap, &n,
                                   unclassified->points.data(), cap, &nu));
        planar->points.resize(n);
        planar->width = static_cast<uint32_t>(n);
        unclassified->points.resize(nu);
        unclassified->width = static_cast<uint32_t>(nu);
        return {planar, unclassified};
    }
};

// ---- LidarOdometry (src/lidar_odometry.h:18-86) ----------------------------------------------------------------
class LidarOdometry {
public:
    using CloudType = pcl::PointCloud<lidar_point::PointXYZIRT>;

    class Params {   // same 8 public fields and defaults (src/lidar_odometry.h:25-48); yaml instead of the rclcpp map
    public:
        float lidar_min_range = 4.0f, lidar_max_range = 80.0f, keyframe_voxel_size = 0.2f;
        size_t keyframe_max_points_cnt = 20;
        float keyframe_matching_voxel_size = 0.3f, keyframe_update_voxel_size = 0.1f, keyframe_cleanup_range = 80.0f,
              angular_divergence_threshold = 5.0f;
        // the reference's parameter plumbing, same names (src/lidar_odometry.h:34-62); rclcpp::ParameterValue with -DLO_B200_WITH_ROS2
        using ROS2ParametersMap = std::map<std::string, lo::ParameterValue>;
        static ROS2ParametersMap GetROSDeclaration() {
            return {{"lidar_min_range", lo::ParameterValue(4.0)}, {"lidar_max_range", lo::ParameterValue(80.0)},
                    {"keyframe_voxel_size", lo::ParameterValue(0.2)}, {"keyframe_max_points_cnt", lo::ParameterValue(20)},
                    {"keyframe_matching_voxel_size", lo::ParameterValue(0.3)}, {"keyframe_update_voxel_size", lo::ParameterValue(0.1)},
                    {"keyframe_cleanup_range", lo::ParameterValue(80.0)}, {"angular_divergence_threshold", lo::ParameterValue(5.0)}};
        }
        static Params FromROS2Params(const ROS2ParametersMap& params) {
            Params o;
            o.lidar_min_range = params.at("lidar_min_range").get<float>();
            o.lidar_max_range = params.at("lidar_max_range").get<float>();
            o.keyframe_voxel_size = params.at("keyframe_voxel_size").get<float>();
            o.keyframe_max_points_cnt = params.at("keyframe_max_points_cnt").get<size_t>();
            o.keyframe_matching_voxel_size = params.at("keyframe_matching_voxel_size").get<float>();
            o.keyframe_update_voxel_size = params.at("keyframe_update_voxel_size").get<float>();
            o.keyframe_cleanup_range = params.at("keyframe_cleanup_range").get<float>();
            o.angular_divergence_threshold = params.at("angular_divergence_threshold").get<float>();
            return o;
        }
        static Params FromYaml(const std::string& path) {   // config/params.yaml layout
            lo_params p;
            lo::check(lo_params_from_yaml(path.c_str(), &p));
            Params o;
            o.lidar_min_range = p.lidar_min_range; o.lidar_max_range = p.lidar_max_range; o.keyframe_voxel_size = p.keyframe_voxel_size;
            o.keyframe_max_points_cnt = static_cast<size_t>(p.keyframe_max_points_cnt);
            o.keyframe_matching_voxel_size = p.keyframe_matching_voxel_size; o.keyframe_update_voxel_size = p.keyframe_update_voxel_size;
            o.keyframe_cleanup_range = p.keyframe_cleanup_range; o.angular_divergence_threshold = p.angular_divergence_threshold;
            return o;
        }
    };

    explicit LidarOdometry(const Params& config, lo::Context& ctx = lo::Context::instance()) : config_(config) {
        lo_params p;
        p.lidar_min_range = config.lidar_min_range; p.lidar_max_range = config.lidar_max_range; p.keyframe_voxel_size = config.keyframe_voxel_size;
        p.keyframe_max_points_cnt = config.keyframe_max_points_cnt; p.keyframe_matching_voxel_size = config.keyframe_matching_voxel_size;
        p.keyframe_update_voxel_size = config.keyframe_update_voxel_size; p.keyframe_cleanup_range = config.keyframe_cleanup_range;
        p.angular_divergence_threshold = config.angular_divergence_threshold;
        lo::check(lo_odometry_create(ctx.handle(), &p, &h_));
    }
    ~LidarOdometry() { lo_odometry_destroy(h_); }
    LidarOdometry(const LidarOdometry&) = delete;
    LidarOdometry& operator=(const LidarOdometry&) = delete;

    void processCloud(const CloudType& input_cloud) { lo::check(lo_odometry_process(h_, input_cloud.points.data(), input_cloud.points.size())); }
    // Extension: process `input_cloud` while `next_cloud` (the scan the caller will pass next, unchanged until then; pinned
    // memory for a real overlap) is uploaded behind it.
    void processCloud(const CloudType& input_cloud, const CloudType& next_cloud) {
        lo::check(lo_odometry_process_next(h_, input_cloud.points.data(), input_cloud.points.size(), next_cloud.points.data(),
                                           next_cloud.points.size()));
    }

    // Extension for recorded streams: queue a scan (at most LO_ODOMETRY_QUEUE_DEPTH in flight; its buffer must stay untouched until the matching
    // wait) and collect the results in order. submit + wait == processCloud.
    void submit(const CloudType& input_cloud) { lo::check(lo_odometry_submit(h_, input_cloud.points.data(), input_cloud.points.size())); }
    void wait() { lo::check(lo_odometry_wait(h_)); }

    pcl::PointCloud<pcl::PointXYZ>::Ptr getKeyFrameCloud() const { return keyframe(0); }
    pcl::PointCloud<pcl::PointXYZ>::Ptr getFullKeyFrameCloud() const { return keyframe(1); }
    Pose3D getCurrentPose() const {
        float p[7];
        lo::check(lo_odometry_pose(h_, p));
        return Pose3D::from7(p);
    }
    CloudType::Ptr getTempCloud() const {
        auto out = std::make_shared<CloudType>();
        size_t n = 0;
        lo::check(lo_odometry_deskewed(h_, nullptr, 0, &n));
        out->points.resize(n);
        out->width = static_cast<uint32_t>(n);
        if (n) lo::check(lo_odometry_deskewed(h_, out->points.data(), n, &n));
        return out;
    }
    lo_scan_stats stats() const {
        lo_scan_stats s;
        lo::check(lo_odometry_stats(h_, &s));
        return s;
    }

    // Extension: checkpoint of previous_transform_ / current_transform_ / keyframe_ (src/lidar_odometry.h:82-85); an object
    // built with the same Params and restored from it continues with bit-identical poses.
    std::vector<uint8_t> snapshot() const {
        size_t n = 0;
        lo::check(lo_odometry_snapshot(h_, nullptr, 0, &n));
        std::vector<uint8_t> blob(n);
        lo::check(lo_odometry_snapshot(h_, blob.data(), blob.size(), &n));
        return blob;
    }
    void restore(const std::vector<uint8_t>& blob) { lo::check(lo_odometry_restore(h_, blob.data(), blob.size())); }
    void saveCheckpoint(const std::string& path) const {
        const std::vector<uint8_t> blob = snapshot();
        std::FILE* f = std::fopen(path.c_str(), "wb");
        if (!f || std::fwrite(blob.data(), 1, blob.size(), f) != blob.size()) {
            if (f) std::fclose(f);
            throw lo::Error(LO_E_IO, "cannot write checkpoint " + path);
        }
        std::fclose(f);
    }
    void loadCheckpoint(const std::string& path) {
        std::FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) throw lo::Error(LO_E_IO, "cannot open checkpoint " + path);
        std::vector<uint8_t> blob;
        uint8_t chunk[1 << 16];
        for (size_t got; (got = std::fread(chunk, 1, sizeof(chunk), f)) > 0;) blob.insert(blob.end(), chunk, chunk + got);
        std::fclose(f);
        restore(blob);
    }

private:
    pcl::PointCloud<pcl::PointXYZ>::Ptr keyframe(int full) const {
        auto out = std::make_shared<pcl::PointCloud<pcl::PointXYZ>>();
        size_t n = 0;
        lo::check(lo_odometry_keyframe(h_, full, nullptr, 0, &n));
        out->points.resize(n);
        out->width = static_cast<uint32_t>(n);
        if (n) lo::check(lo_odometry_keyframe(h_, full, reinterpret_cast<float*>(out->points.data()), n, &n));
        return out;
    }
    const Params config_;
    lo_odometry* h_ = nullptr;
};
