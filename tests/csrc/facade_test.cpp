// The reference's own gtest cases (test/test.cpp) re-expressed against the C++ facade
// (lidar_odometry_demo_b200/host/lo_b200.hpp), without gtest: each CHECK failure prints and counts.
// Run on a GPU box by tests/test_gpu_facade.py:  facade_test <xyz_f32_file> <n_points> <normals_f32_file>
#include <algorithm>
#include <cstdio>
#include <fstream>
#include <numbers>

#include "../../lidar_odometry_demo_b200/host/lo_b200.hpp"

static int failures = 0;
#define CHECK(cond)                                                          \
    do {                                                                     \
        if (!(cond)) {                                                       \
            failures++;                                                      \
            std::printf("CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
        }                                                                    \
    } while (0)

static lo::Quaternionf angle_axis(float angle, float ax, float ay, float az) {
    const float n = std::sqrt(ax * ax + ay * ay + az * az);
    const float s = std::sin(0.5f * angle) / n;
    return lo::Quaternionf(std::cos(0.5f * angle), s * ax, s * ay, s * az);
}

static void voxel_grid_unique_points() {   // test/test.cpp:26-55
    pcl::PointCloud<pcl::PointNormal> in;
    const float pts[7][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {-1, 0, 0}, {0, -1, 0}, {0, 0, -1}};
    for (auto& p : pts) in.points.emplace_back(p[0], p[1], p[2]);
    VoxelGrid grid(0.5f, 1);
    grid.addCloud(in);
    CHECK(grid.size() == in.size());
    auto out = grid.getCloud();
    CHECK(out->size() == in.size());
    auto remaining = in.points;
    for (const auto& o : out->points) {
        auto it = std::find_if(remaining.begin(), remaining.end(), [&](const auto& i) { return i.x == o.x && i.y == o.y && i.z == o.z; });
        CHECK(it != remaining.end());
        if (it != remaining.end()) remaining.erase(it);
    }
}

static void voxel_grid_duplicate_points() {   // test/test.cpp:57-75
    pcl::PointCloud<pcl::PointNormal> in;
    in.points.emplace_back(0, 0, 0); in.points.emplace_back(1, 0, 0); in.points.emplace_back(0, 0, 0); in.points.emplace_back(1, 0, 0);
    VoxelGrid grid(0.5f, 1);
    grid.addCloud(in);
    CHECK(grid.size() == 2);
    auto out = grid.getCloud();
    CHECK(out->size() == 2);
    CHECK(!(out->at(0).x == out->at(1).x && out->at(0).y == out->at(1).y && out->at(0).z == out->at(1).z));
}

static void rigid_transform() {   // test/test.cpp:151-189, checked against a double-precision transform
    pcl::PointCloud<pcl::PointXYZ> cloud;
    const float pts[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    for (auto& p : pts) cloud.points.emplace_back(p[0], p[1], p[2]);
    const float d = static_cast<float>(std::numbers::pi / 180.0);
    const Pose3D cases[] = {Pose3D({0, 0, 0}, angle_axis(0, 0, 0, 1)), Pose3D({0, 0, 0}, angle_axis(80 * d, 0, 0, 1)),
                            Pose3D({0, 0, 0}, angle_axis(80 * d, 0, 1, 0)), Pose3D({0, 0, 0}, angle_axis(90 * d, 1, 0, 1)),
                            Pose3D({1, 1, 1}, angle_axis(45 * d, 0, 0.5f, 0.5f)), Pose3D({-2, 2, 0}, angle_axis(45 * d, 0.5f, 0.5f, 0))};
    for (const auto& pose : cases) {
        auto out = CloudTransformer::transform(cloud, pose);
        const double w = pose.rotation.w(), x = pose.rotation.x(), y = pose.rotation.y(), z = pose.rotation.z();
        const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z), 1 - 2 * (x * x + z * z),
                             2 * (y * z - w * x), 2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
        for (size_t i = 0; i < cloud.size(); i++) {
            const auto& p = cloud.at(i);
            const double e[3] = {R[0] * p.x + R[1] * p.y + R[2] * p.z + pose.translation.x(), R[3] * p.x + R[4] * p.y + R[5] * p.z + pose.translation.y(),
                                 R[6] * p.x + R[7] * p.y + R[8] * p.z + pose.translation.z()};
            CHECK(std::abs(out->at(i).x - e[0]) < 3e-7 && std::abs(out->at(i).y - e[1]) < 3e-7 && std::abs(out->at(i).z - e[2]) < 3e-7);
        }
    }
}

static std::vector<float> read_f32(const char* path, size_t count) {
    std::vector<float> v(count);
    std::ifstream f(path, std::ios::binary);
    f.read(reinterpret_cast<char*>(v.data()), static_cast<std::streamsize>(count * sizeof(float)));
    if (!f) v.clear();
    return v;
}

// test/test.cpp:191-264 on intersection00056.pcd (the original crop is missing from the reference mount);
// normals come from the file written by the Python side (lo_pca_normals, NaN rows dropped here as :219-221 does)
static void matching_test(const char* xyz_path, size_t n, const char* normals_path) {
    const auto xyz = read_f32(xyz_path, n * 3), nrm = read_f32(normals_path, n * 3);
    CHECK(!xyz.empty() && !nrm.empty());
    if (xyz.empty() || nrm.empty()) return;
    pcl::PointCloud<pcl::PointNormal> with_normals;
    pcl::PointCloud<pcl::PointXYZ> full;
    for (size_t i = 0; i < n; i++) {
        full.points.emplace_back(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
        if (std::isnan(nrm[3 * i]) || std::isnan(nrm[3 * i + 1]) || std::isnan(nrm[3 * i + 2])) continue;
        with_normals.points.emplace_back(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
    }
    VoxelGrid keyframe(0.25f, 20);
    keyframe.addCloud(with_normals);
    VoxelGrid voxel_filter(0.5f, 1);
    voxel_filter.addCloudWithoutNormals(full);
    auto sub = voxel_filter.getCloudWithoutNormals();
    CloudMatcher matcher;
    const float d = static_cast<float>(std::numbers::pi / 180.0);
    const Pose3D guesses[] = {Pose3D({0, 0, 0}, lo::Quaternionf::Identity()), Pose3D({0, 0, 0.1f}, lo::Quaternionf::Identity()),
                              Pose3D({0.1f, 0.1f, 0.1f}, lo::Quaternionf::Identity()), Pose3D({-0.1f, -0.1f, -0.1f}, lo::Quaternionf::Identity()),
                              Pose3D({0.1f, -0.1f, 0}, lo::Quaternionf::Identity()), Pose3D({0, 0, 0}, angle_axis(-1.0f * d, 0, 0, 1)),
                              Pose3D({-0.2f, 0, 0}, angle_axis(2.0f * d, 0, 0, 1))};
    for (const auto& guess : guesses) {
        auto guess_cloud = CloudTransformer::transform(*sub, guess.inverse());
        const Pose3D fin = matcher.align(keyframe, *guess_cloud, Pose3D());
        const Pose3D diff = fin.relativeTo(guess);
        std::printf("guess t=(%g %g %g) -> final t=(%g %g %g) |diff t|=%g\n", guess.translation.x(), guess.translation.y(), guess.translation.z(),
                    fin.translation.x(), fin.translation.y(), fin.translation.z(), diff.translation.norm());
        CHECK(diff.translation.norm() < 0.05f);                          // test.cpp:261
        CHECK(1.0f - std::abs(fin.rotation.dot(guess.rotation)) < 0.01f);  // test.cpp:262
    }
}

// A small VLP16-like sweep of a street canyon (ground z = -1.8, walls y = +-8) seen from x = shift: enough structure for
// LidarOdometry::processCloud; used to check that the facade's one-call and queued entry points agree.
static LidarOdometry::CloudType street_scan(float shift) {
    LidarOdometry::CloudType cloud;
    for (int ring = 0; ring < 16; ring++) {
        const float el = (-15.0f + 2.0f * ring) * static_cast<float>(std::numbers::pi / 180.0);
        for (int k = 0; k < 1800; k++) {
            const float az = k * static_cast<float>(2.0 * std::numbers::pi / 1800.0);
            const float dx = std::cos(el) * std::cos(az), dy = std::cos(el) * std::sin(az), dz = std::sin(el);
            float range = 1e9f;
            if (dz < -1e-3f) range = std::min(range, -1.8f / dz);
            if (std::abs(dy) > 1e-3f) range = std::min(range, 8.0f / std::abs(dy));
            if (range > 100.f) continue;
            lidar_point::PointXYZIRT p;
            p.x = range * dx - shift; p.y = range * dy; p.z = range * dz;
            p.intensity = 10.f; p.ring = static_cast<uint16_t>(ring); p.time = 0.1f * k / 1800.0f;
            cloud.points.push_back(p);
        }
    }
    cloud.width = static_cast<uint32_t>(cloud.points.size());
    cloud.height = 1;
    return cloud;
}

static void odometry_one_call_and_queued() {
    std::vector<LidarOdometry::CloudType> scans;
    for (int k = 0; k < 5; k++) scans.push_back(street_scan(0.1f * k));
    LidarOdometry one(LidarOdometry::Params{}), queued(LidarOdometry::Params{});
    std::vector<Pose3D> ref;
    for (const auto& sc : scans) {
        one.processCloud(sc);
        ref.push_back(one.getCurrentPose());
    }
    CHECK(one.getKeyFrameCloud()->size() > 0);
    CHECK(std::abs(ref.back().translation.y()) < 0.05f);    // the sensor moved along x only
    size_t done = 0;
    queued.submit(scans[0]);
    for (size_t k = 1; k < scans.size(); k++) {
        queued.submit(scans[k]);                              // scan k queued behind scan k-1
        queued.wait();
        const Pose3D p = queued.getCurrentPose();
        CHECK(p.translation.x() == ref[done].translation.x() && p.rotation.w() == ref[done].rotation.w());
        done++;
    }
    queued.wait();
    CHECK(queued.getCurrentPose().translation.x() == ref[done].translation.x());
    std::printf("odometry: %zu scans, final x = %g (one call) / %g (queued)\n", scans.size(), ref.back().translation.x(),
                queued.getCurrentPose().translation.x());
}

int main(int argc, char** argv) {
    try {
        voxel_grid_unique_points();
        voxel_grid_duplicate_points();
        rigid_transform();
        if (argc >= 4) matching_test(argv[1], std::strtoul(argv[2], nullptr, 10), argv[3]);
        LidarOdometry odom(LidarOdometry::Params{});
        CHECK(odom.getKeyFrameCloud()->size() == 0);
        odometry_one_call_and_queued();
    } catch (const lo::Error& e) {
        std::printf("lo::Error %d: %s\n", e.code, e.what());
        return 2;
    }
    std::printf("%s (%d failures)\n", failures ? "FAILED" : "PASSED", failures);
    return failures ? 1 : 0;
}
