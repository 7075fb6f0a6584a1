// Host-only checks of lidar_odometry_demo_b200/host/lo_io.hpp, driven by tests/test_io_cpu.py:
//   io_test roundtrip <dir>            write + read every wire struct as ascii and binary PCD and as raw records
//   io_test xyz <in.pcd> <out.f32>     dump x,y,z of a PCD as float32 triples
//   io_test xyzirt <in> <out.bin>      read a scan (.pcd or raw) and dump the 32-byte records
//   io_test traj <tum.txt> <kitti.txt> write two known poses in both trajectory formats
//   io_test bad <file>                 must fail with lo::io::Error (exit code 3)
#include <cstdio>
#include <cstring>
#include <string>

#include "../../lidar_odometry_demo_b200/host/lo_io.hpp"

static int failures = 0;
#define CHECK(cond)                                                          \
    do {                                                                     \
        if (!(cond)) {                                                       \
            failures++;                                                      \
            std::printf("CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
        }                                                                    \
    } while (0)

static float noise(uint32_t i) {   // deterministic, awkward floats
    uint32_t x = i * 2654435761u + 12345u;
    x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
    return (static_cast<float>(x >> 8) / 16777216.0f - 0.5f) * 237.123f;
}

static int roundtrip(const std::string& dir) {
    using namespace lo::io;
    const size_t n = 1237;
    pcl::PointCloud<pcl::PointXYZ> a;
    pcl::PointCloud<pcl::PointNormal> b;
    pcl::PointCloud<lidar_point::PointXYZIRT> c;
    for (uint32_t i = 0; i < n; i++) {
        a.push_back(pcl::PointXYZ(noise(3 * i), noise(3 * i + 1), noise(3 * i + 2)));
        pcl::PointNormal pn(noise(7 * i), noise(7 * i + 1), noise(7 * i + 2), noise(7 * i + 3) / 200.f, noise(7 * i + 4) / 200.f, noise(7 * i + 5) / 200.f);
        pn.curvature = noise(7 * i + 6);
        b.push_back(pn);
        lidar_point::PointXYZIRT p;
        p.x = noise(5 * i); p.y = noise(5 * i + 1); p.z = noise(5 * i + 2); p.intensity = noise(5 * i + 3);
        p.ring = static_cast<uint16_t>(i % 16); p.time = static_cast<float>(i) * 0.1f / n;
        c.push_back(p);
    }
    for (int binary = 0; binary < 2; binary++) {
        const std::string tag = binary ? "bin" : "ascii";
        write_pcd(dir + "/a_" + tag + ".pcd", a, binary);
        write_pcd(dir + "/b_" + tag + ".pcd", b, binary);
        write_pcd(dir + "/c_" + tag + ".pcd", c, binary);
        pcl::PointCloud<pcl::PointXYZ> a2;
        pcl::PointCloud<pcl::PointNormal> b2;
        pcl::PointCloud<lidar_point::PointXYZIRT> c2;
        read_pcd(dir + "/a_" + tag + ".pcd", a2);
        read_pcd(dir + "/b_" + tag + ".pcd", b2);
        read_pcd(dir + "/c_" + tag + ".pcd", c2);
        CHECK(a2.size() == n && b2.size() == n && c2.size() == n && a2.width == n && a2.height == 1);
        // 9 significant digits round-trip a float exactly, so ascii is bit-exact too
        CHECK(std::memcmp(a.points.data(), a2.points.data(), n * sizeof(pcl::PointXYZ)) == 0);
        CHECK(std::memcmp(b.points.data(), b2.points.data(), n * sizeof(pcl::PointNormal)) == 0);
        CHECK(std::memcmp(c.points.data(), c2.points.data(), n * sizeof(lidar_point::PointXYZIRT)) == 0);
        // a PointNormal file read as PointXYZ keeps the coordinates
        pcl::PointCloud<pcl::PointXYZ> bx;
        read_pcd(dir + "/b_" + tag + ".pcd", bx);
        CHECK(bx.size() == n && bx[17].x == b[17].x && bx[17].z == b[17].z);
    }
    write_scan_bin(dir + "/c.bin", c);
    pcl::PointCloud<lidar_point::PointXYZIRT> c3, c4, empty, e2;
    read_scan(dir + "/c.bin", c3);
    read_scan(dir + "/c_bin.pcd", c4);
    CHECK(c3.size() == n && std::memcmp(c.points.data(), c3.points.data(), n * 32) == 0);
    CHECK(c4.size() == n && std::memcmp(c.points.data(), c4.points.data(), n * 32) == 0);
    write_pcd(dir + "/empty.pcd", empty, true);
    read_pcd(dir + "/empty.pcd", e2);
    CHECK(e2.size() == 0);
    write_scan_bin(dir + "/empty.bin", empty);
    read_scan(dir + "/empty.bin", e2);
    CHECK(e2.size() == 0);
    // a scan file without ring / time cannot feed the classifier and the deskew
    bool threw = false;
    try {
        read_pcd(dir + "/a_bin.pcd", c3);
    } catch (const Error&) {
        threw = true;
    }
    CHECK(threw);
    std::printf(failures ? "FAILED %d\n" : "PASSED\n", failures);
    return failures ? 1 : 0;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const std::string cmd = argv[1];
    try {
        if (cmd == "roundtrip") return roundtrip(argv[2]);
        if (cmd == "xyz" && argc == 4) {
            pcl::PointCloud<pcl::PointXYZ> cloud;
            lo::io::read_pcd(argv[2], cloud);
            std::FILE* f = std::fopen(argv[3], "wb");
            for (const auto& p : cloud.points) std::fwrite(&p.x, sizeof(float), 3, f);
            std::fclose(f);
            std::printf("%zu\n", cloud.size());
            return 0;
        }
        if (cmd == "xyzirt" && argc == 4) {
            pcl::PointCloud<lidar_point::PointXYZIRT> cloud;
            lo::io::read_scan(argv[2], cloud);
            lo::io::write_scan_bin(argv[3], cloud);
            std::printf("%zu\n", cloud.size());
            return 0;
        }
        if (cmd == "traj" && argc == 4) {
            lo::io::TrajectoryWriter tum(argv[2], lo::io::TrajectoryWriter::TUM), kitti(argv[3], lo::io::TrajectoryWriter::KITTI);
            const Pose3D poses[2] = {Pose3D(), Pose3D(lo::Vector3f(1.5f, -2.25f, 0.125f), lo::Quaternionf(0.70710678f, 0.f, 0.f, 0.70710678f))};
            for (int k = 0; k < 2; k++) {
                tum.add(0.1 * k, poses[k]);
                kitti.add(0.1 * k, poses[k]);
            }
            return 0;
        }
        if (cmd == "bad") {
            pcl::PointCloud<pcl::PointXYZ> cloud;
            lo::io::read_pcd(argv[2], cloud);
            return 0;
        }
    } catch (const lo::io::Error& e) {
        std::printf("io error: %s\n", e.what());
        return 3;
    }
    return 2;
}
