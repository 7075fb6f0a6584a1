// TEST INFRASTRUCTURE. Runs the reference's own LidarOdometry::processCloud (src/lidar_odometry.cpp, compiled unmodified
// against the GPU facade, see dropin.hpp) and the library's fused lo_odometry_process side by side on the same scans.
//   dropin_test <scans.bin>      scans.bin = u32 count, then per scan u32 n + n x 32-byte PointXYZIRT records
// Prints one line per scan: both poses and their difference; exit code 0 iff every scan agrees within the tolerances
// north_star states (1e-4 m, 1e-4 rad).
#include <cstdio>
#include <fstream>
#include <vector>

int main(int argc, char** argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s scans.bin\n", argv[0]);
        return 2;
    }
    std::ifstream f(argv[1], std::ios::binary);
    uint32_t count = 0;
    f.read(reinterpret_cast<char*>(&count), 4);
    std::vector<LidarOdometry::CloudType> scans(count);
    for (auto& s : scans) {
        uint32_t n = 0;
        f.read(reinterpret_cast<char*>(&n), 4);
        s.points.resize(n);
        s.width = n;
        f.read(reinterpret_cast<char*>(s.points.data()), static_cast<std::streamsize>(n) * 32);
    }
    if (!f) {
        std::fprintf(stderr, "short read\n");
        return 2;
    }
    const LidarOdometry::Params params = LidarOdometry::Params::FromROS2Params(LidarOdometry::Params::GetROSDeclaration());
    LidarOdometry reference_code(params);        // the reference's processCloud, every stage through the facade
    LoB200LidarOdometry fused(params);           // one scan_kernel per scan
    double worst_t = 0.0, worst_r = 0.0;
    for (uint32_t k = 0; k < count; k++) {
        reference_code.processCloud(scans[k]);
        fused.processCloud(scans[k]);
        const Pose3D a = reference_code.getCurrentPose(), b = fused.getCurrentPose();
        const double dt = std::sqrt(std::pow(double(a.translation.x()) - b.translation.x(), 2) + std::pow(double(a.translation.y()) - b.translation.y(), 2) +
                                    std::pow(double(a.translation.z()) - b.translation.z(), 2));
        // rotation angle of a^-1 * b from the vector part of the relative quaternion (acos of a dot product near 1 would
        // turn the float quaternions' distance from unit norm into 1e-4 rad of noise)
        double qa[4] = {a.rotation.w(), a.rotation.x(), a.rotation.y(), a.rotation.z()}, qb[4] = {b.rotation.w(), b.rotation.x(), b.rotation.y(), b.rotation.z()};
        const double na = std::sqrt(qa[0] * qa[0] + qa[1] * qa[1] + qa[2] * qa[2] + qa[3] * qa[3]);
        const double nb = std::sqrt(qb[0] * qb[0] + qb[1] * qb[1] + qb[2] * qb[2] + qb[3] * qb[3]);
        for (int i = 0; i < 4; i++) { qa[i] /= na; qb[i] /= nb; }
        const double vx = qa[0] * qb[1] - qa[1] * qb[0] - qa[2] * qb[3] + qa[3] * qb[2];
        const double vy = qa[0] * qb[2] + qa[1] * qb[3] - qa[2] * qb[0] - qa[3] * qb[1];
        const double vz = qa[0] * qb[3] - qa[1] * qb[2] + qa[2] * qb[1] - qa[3] * qb[0];
        const double dr = 2.0 * std::asin(std::min(1.0, std::sqrt(vx * vx + vy * vy + vz * vz)));
        worst_t = std::max(worst_t, dt);
        worst_r = std::max(worst_r, dr);
        std::printf("scan %u: reference-code pose t=(%.6f %.6f %.6f)  fused t=(%.6f %.6f %.6f)  |dt|=%.3e |dr|=%.3e  keyframe %zu voxels\n", k,
                    a.translation.x(), a.translation.y(), a.translation.z(), b.translation.x(), b.translation.y(), b.translation.z(), dt, dr,
                    reference_code.getKeyFrameCloud()->size());
    }
    const size_t kf_a = reference_code.getFullKeyFrameCloud()->size(), kf_b = fused.getFullKeyFrameCloud()->size();
    std::printf("worst |dt|=%.3e m |dr|=%.3e rad; key-frame points %zu vs %zu\n", worst_t, worst_r, kf_a, kf_b);
    const bool ok = worst_t < 1e-4 && worst_r < 1e-4 && kf_a == kf_b;
    std::printf(ok ? "PASSED\n" : "FAILED\n");
    return ok ? 0 : 1;
}
