// TEST INFRASTRUCTURE (oracle/_ref). Compiles the reference's OWN, UNMODIFIED headers where they lie under
// /root/reference (src/pose_3d.h, src/voxel_grid.h, src/voxel_with_planes.h, src/lidar_point_type.h,
// src/utils/{cloud_transform,point_time_normalize,range_filter,cloud_classifier}.h and the vendored
// third-party/robin-map) against the shim headers in oracle/shim (Eigen and PCL are not installed) and
// exposes them through oracle_api.h. Nothing from /root/reference is copied into this repository: the
// sources are reached with -I at build time (oracle/Makefile), output goes to oracle/_ref/liblo_ref.so.
// src/cloud_matcher.cpp and src/lidar_odometry.cpp need Ceres / rclcpp and are restated in
// restated_pipeline.inc + mini_ceres.h.
// <math.h> first: brings the float overloads of atan2/sqrt into the global namespace, which the
// reference's cloud_classifier.h calls unqualified (see lo_oracle.cpp classify()).
#include <math.h>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <execution>
#include <map>
#include <numbers>
#include <vector>

// include order mirrors src/lidar_odometry.cpp:6-11 (voxel_grid.h relies on Pose3D being declared first)
#include "pose_3d.h"
#include "lidar_point_type.h"
#include "utils/cloud_classifier.h"
#include "utils/cloud_transform.h"
#include "utils/point_time_normalize.h"
#include "utils/range_filter.h"
#include "voxel_grid.h"

#include "../mini_ceres.h"
#include "../oracle_api.h"
#include "../orc_types.h"

static_assert(sizeof(lidar_point::PointXYZIRT) == 32, "PointXYZIRT ABI");
static_assert(offsetof(lidar_point::PointXYZIRT, ring) == 20 && offsetof(lidar_point::PointXYZIRT, time) == 24, "PointXYZIRT ABI");

namespace {
Pose3D to_pose(const float p[7]) {
    return Pose3D(Eigen::Vector3f(p[0], p[1], p[2]), Eigen::Quaternionf(p[3], p[4], p[5], p[6]));
}
void from_pose(const Pose3D& p, float out[7]) {
    out[0] = p.translation.x(); out[1] = p.translation.y(); out[2] = p.translation.z();
    out[3] = p.rotation.w(); out[4] = p.rotation.x(); out[5] = p.rotation.y(); out[6] = p.rotation.z();
}
template <typename PointT>
pcl::PointCloud<PointT> to_cloud(const void* data, size_t n) {
    static_assert(std::is_trivially_copyable_v<PointT>);
    pcl::PointCloud<PointT> c;
    c.points.resize(n);
    if (n) std::memcpy(static_cast<void*>(c.points.data()), data, n * sizeof(PointT));
    c.width = static_cast<std::uint32_t>(n);
    c.height = 1;
    return c;
}
template <typename PointT>
size_t from_cloud(const pcl::PointCloud<PointT>& c, void* out, size_t cap) {
    const size_t n = c.points.size();
    if (out) std::memcpy(out, static_cast<const void*>(c.points.data()), std::min(n, cap) * sizeof(PointT));
    return n;
}
}  // namespace

extern "C" const char* orc_kind(void) { return "reference"; }

extern "C" void* orc_grid_create(float voxel_size, size_t max_points) { return new VoxelGrid(voxel_size, max_points); }
extern "C" void* orc_grid_clone(const void* g) { return new VoxelGrid(*static_cast<const VoxelGrid*>(g)); }
extern "C" void orc_grid_destroy(void* g) { delete static_cast<VoxelGrid*>(g); }
extern "C" void orc_grid_add_cloud(void* g, const float* pn48, size_t n) {
    static_cast<VoxelGrid*>(g)->addCloud(to_cloud<pcl::PointNormal>(pn48, n));
}
extern "C" void orc_grid_add_cloud_xyz(void* g, const float* xyz16, size_t n) {
    static_cast<VoxelGrid*>(g)->addCloudWithoutNormals(to_cloud<pcl::PointXYZ>(xyz16, n));
}
extern "C" size_t orc_grid_size(const void* g) { return static_cast<const VoxelGrid*>(g)->size(); }
extern "C" size_t orc_grid_num_points(const void* g) { return static_cast<const VoxelGrid*>(g)->getCloudWithoutNormals()->size(); }
extern "C" size_t orc_grid_get_cloud(const void* g, float* out, size_t cap) {
    return from_cloud(*static_cast<const VoxelGrid*>(g)->getCloud(), out, cap);
}
extern "C" size_t orc_grid_get_cloud_xyz(const void* g, float* out, size_t cap) {
    return from_cloud(*static_cast<const VoxelGrid*>(g)->getCloudWithoutNormals(), out, cap);
}
extern "C" size_t orc_grid_get_sparse_xyz(const void* g, float* out, size_t cap) {
    return from_cloud(*static_cast<const VoxelGrid*>(g)->getSparseCloudWithoutNormals(), out, cap);
}
extern "C" void orc_grid_radius_cleanup(void* g, const float c[3], float radius) {
    static_cast<VoxelGrid*>(g)->radiusCleanup(Eigen::Vector3f(c[0], c[1], c[2]), radius);
}
extern "C" void orc_get_indices(float voxel_size, const float* xyz, size_t stride, size_t n, int64_t* out3) {
    VoxelGrid g(voxel_size, 1);
    for (size_t i = 0; i < n; i++) {
        const auto idx = g.getIndices(xyz[i * stride], xyz[i * stride + 1], xyz[i * stride + 2]);
        out3[i * 3] = idx.ix; out3[i * 3 + 1] = idx.iy; out3[i * 3 + 2] = idx.iz;
    }
}
extern "C" uint64_t orc_hash(int64_t ix, int64_t iy, int64_t iz) {
    return VoxelGrid::IndicesHash()(VoxelGrid::Indices{ix, iy, iz});
}
extern "C" void orc_grid_set_max_points(void* g, size_t max_points) { static_cast<VoxelGrid*>(g)->setMaxPoints(max_points); }
extern "C" int orc_grid_get_correspondence(const void* g, const float q3[3], double max_sq, double origin3[3], double normal3[3]) {
    const auto c = static_cast<const VoxelGrid*>(g)->getCorrespondence(Eigen::Vector3f(q3[0], q3[1], q3[2]), max_sq);   // voxel_grid.h:164
    for (int k = 0; k < 3; k++) {
        origin3[k] = c.valid ? c.plane_origin(k) : 0.0;
        normal3[k] = c.valid ? c.plane_normal(k) : 0.0;
    }
    return c.valid ? 1 : 0;
}
extern "C" size_t orc_grid_find_matching_pairs(const void* g, const float* xyz16, size_t n, const float pose7[7],
                                               float max_dist, double* out9, uint8_t* valid) {
    // The reference's findMatchingPairs drops invalid queries and (with TBB) scrambles the order
    // (voxel_grid.h:217-231). To keep the query index, the same per-query statements of :212-226 are
    // driven here one query at a time through the reference's own getCorrespondence (:164-204).
    const auto* grid = static_cast<const VoxelGrid*>(g);
    const Pose3D transform = to_pose(pose7);
    Eigen::Matrix3d R = transform.rotationMatrix().cast<double>();
    Eigen::Vector3d t = transform.translation.cast<double>();
    float max_sq = max_dist * max_dist;
    const auto* pts = reinterpret_cast<const OrcXYZ*>(xyz16);
    size_t nv = 0;
    for (size_t i = 0; i < n; i++) {
        Eigen::Vector3d local(pts[i].x, pts[i].y, pts[i].z);
        Eigen::Vector3d world = R * local + t;
        auto c = grid->getCorrespondence(world.cast<float>(), max_sq);
        valid[i] = c.valid ? 1 : 0;
        for (int k = 0; k < 9; k++) out9[i * 9 + k] = 0.0;
        if (c.valid) {
            nv++;
            for (int k = 0; k < 3; k++) {
                out9[i * 9 + k] = local(k);
                out9[i * 9 + 3 + k] = c.plane_origin(k);
                out9[i * 9 + 6 + k] = c.plane_normal(k);
            }
        }
    }
    return nv;
}

extern "C" void orc_time_normalize(const void* in, size_t n, void* out) {
    auto r = utils::pointTimeNormalize(to_cloud<lidar_point::PointXYZIRT>(in, n));
    from_cloud(*r, out, n);
}
extern "C" void orc_transform_nonrigid(const void* in, size_t n, const float s7[7], const float e7[7], void* out) {
    auto r = CloudTransformer::transformNonRigid(to_cloud<lidar_point::PointXYZIRT>(in, n), to_pose(s7), to_pose(e7));
    from_cloud(*r, out, n);
}
extern "C" void orc_transform_xyz(const float* in, size_t n, const float p7[7], float* out) {
    auto r = CloudTransformer::transform(to_cloud<pcl::PointXYZ>(in, n), to_pose(p7));
    from_cloud(*r, out, n);
}
extern "C" void orc_transform_with_normals(const float* in, size_t n, const float p7[7], float* out) {
    auto r = CloudTransformer::transformWithNormals(to_cloud<pcl::PointNormal>(in, n), to_pose(p7));
    from_cloud(*r, out, n);
}
extern "C" size_t orc_range_filter(const float* in, size_t n, float mn, float mx, float* out) {
    auto r = utils::rangeFilter(to_cloud<pcl::PointNormal>(in, n), mn, mx);
    return from_cloud(*r, out, n);
}
extern "C" size_t orc_classify(const void* in, size_t n, float* planar_out, size_t* n_unclassified) {
    auto [planar, unclassified] = CloudClassifier::classify(to_cloud<lidar_point::PointXYZIRT>(in, n));
    if (n_unclassified) *n_unclassified = unclassified->size();
    return from_cloud(*planar, planar_out, n);
}

extern "C" size_t orc_classify_full(const void* in, size_t n, float* planar_out, void* unclassified_out, size_t* n_unclassified) {
    auto [planar, unclassified] = CloudClassifier::classify(to_cloud<lidar_point::PointXYZIRT>(in, n));   // cloud_classifier.h:167
    if (n_unclassified) *n_unclassified = unclassified->size();
    if (unclassified_out && !unclassified->points.empty())
        std::memcpy(unclassified_out, unclassified->points.data(), unclassified->points.size() * sizeof(lidar_point::PointXYZIRT));
    return from_cloud(*planar, planar_out, n);
}

extern "C" void orc_pose_compose(const float a[7], const float b[7], float out[7]) { from_pose(to_pose(a).compose(to_pose(b)), out); }
extern "C" void orc_pose_inverse(const float a[7], float out[7]) { from_pose(to_pose(a).inverse(), out); }
extern "C" void orc_pose_relative_to(const float a[7], const float b[7], float out[7]) { from_pose(to_pose(a).relativeTo(to_pose(b)), out); }
extern "C" void orc_pose_rotation_matrix(const float a[7], float r9[9]) {
    const Eigen::Matrix3f R = to_pose(a).rotationMatrix();
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) r9[r * 3 + c] = R(r, c);
}

// PCL NormalEstimation is test-only third-party code (test/test.cpp:196-205) with no source here; the
// restatement lives in the port (lo_oracle.cpp). The reference build does not provide it.
extern "C" void orc_pca_normals(const float*, size_t n, float, float* normals3_out) {
    for (size_t i = 0; i < n * 3; i++) normals3_out[i] = std::numeric_limits<float>::quiet_NaN();
}

#include "../restated_pipeline.inc"
