// lo_replay — headless driver: recorded scans in, trajectory out.
//
// Stands where the reference's ROS 2 node stands (src/lidar_odometry_node.cpp:45-107: one processCloud per incoming
// PointCloud2, then TF + nav_msgs/Odometry from getCurrentPose and, on demand, the key-frame and deskewed clouds), with
// files instead of topics: scans are .pcd (x y z intensity ring time) or raw 32-byte PointXYZIRT records
// (host/lo_io.hpp), the pose of every scan goes to a TUM or KITTI trajectory file, the final key frame and the last
// deskewed scan can be written as PCD, and the odometry state can be check-pointed / resumed.
//
//   lo_replay [--params params.yaml] [--out traj.txt] [--format tum|kitti] [--period 0.1] [--queue]
//             [--keyframe-pcd kf.pcd] [--full-keyframe-pcd kf_full.pcd] [--deskewed-pcd last.pcd]
//             [--checkpoint-in state.bin] [--checkpoint-out state.bin] [--verbose]
//             (--list scans.txt | scan files ...)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

#include "../host/lo_io.hpp"

namespace {

struct Options {
    std::string params, out, keyframe_pcd, full_keyframe_pcd, deskewed_pcd, checkpoint_in, checkpoint_out;
    lo::io::TrajectoryWriter::Format format = lo::io::TrajectoryWriter::TUM;
    double period = 0.1;   // the 10 Hz sweep of the sensor the node is written for
    bool queue = false, verbose = false;
    std::vector<std::string> scans;
};

[[noreturn]] void usage(const char* why) {
    std::fprintf(stderr,
                 "lo_replay: %s\n"
                 "usage: lo_replay [--params params.yaml] [--out traj.txt] [--format tum|kitti] [--period s] [--queue]\n"
                 "                 [--keyframe-pcd f] [--full-keyframe-pcd f] [--deskewed-pcd f]\n"
                 "                 [--checkpoint-in f] [--checkpoint-out f] [--verbose] (--list scans.txt | scan files ...)\n",
                 why);
    std::exit(2);
}

Options parse(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto value = [&]() -> std::string {
            if (i + 1 >= argc) usage(("missing value after " + a).c_str());
            return argv[++i];
        };
        if (a == "--params") o.params = value();
        else if (a == "--out") o.out = value();
        else if (a == "--format") {
            const std::string f = value();
            if (f == "tum") o.format = lo::io::TrajectoryWriter::TUM;
            else if (f == "kitti") o.format = lo::io::TrajectoryWriter::KITTI;
            else usage("--format takes tum or kitti");
        } else if (a == "--period") o.period = std::atof(value().c_str());
        else if (a == "--queue") o.queue = true;
        else if (a == "--verbose") o.verbose = true;
        else if (a == "--keyframe-pcd") o.keyframe_pcd = value();
        else if (a == "--full-keyframe-pcd") o.full_keyframe_pcd = value();
        else if (a == "--deskewed-pcd") o.deskewed_pcd = value();
        else if (a == "--checkpoint-in") o.checkpoint_in = value();
        else if (a == "--checkpoint-out") o.checkpoint_out = value();
        else if (a == "--list") {
            std::ifstream in(value());
            if (!in) usage("cannot open the --list file");
            for (std::string line; std::getline(in, line);)
                if (!line.empty() && line[0] != '#') o.scans.push_back(line);
        } else if (a.rfind("--", 0) == 0) usage(("unknown option " + a).c_str());
        else o.scans.push_back(a);
    }
    if (o.scans.empty()) usage("no scans given");
    return o;
}

void report(const Options& o, size_t k, const LidarOdometry& odom) {
    if (!o.verbose) return;
    const lo_scan_stats s = odom.stats();
    const Pose3D p = odom.getCurrentPose();
    std::fprintf(stderr, "scan %zu: %u points, %u planar, %u update, %u queries, %d outer iterations%s, key frame %u voxels, t = %.4f %.4f %.4f\n", k,
                 s.n_points, s.n_planar, s.n_update, s.n_query, s.outer_iterations, s.unstable ? " (unstable rotation: prediction kept)" : "",
                 s.keyframe_voxels, p.translation.x(), p.translation.y(), p.translation.z());
}

}  // namespace

int main(int argc, char** argv) {
    const Options o = parse(argc, argv);
    try {
        const LidarOdometry::Params params = o.params.empty() ? LidarOdometry::Params() : LidarOdometry::Params::FromYaml(o.params);
        LidarOdometry odom(params);
        if (!o.checkpoint_in.empty()) odom.loadCheckpoint(o.checkpoint_in);
        std::unique_ptr<lo::io::TrajectoryWriter> traj;
        if (!o.out.empty()) traj = std::make_unique<lo::io::TrajectoryWriter>(o.out, o.format);

        // scans are read one ahead of the one being registered; the buffers alternate so that a queued scan stays untouched
        const size_t n = o.scans.size();
        LidarOdometry::CloudType buf[3];
        double seconds = 0.0;
        size_t points = 0;
        auto clock = [] { return std::chrono::steady_clock::now(); };
        lo::io::read_scan(o.scans[0], buf[0]);
        if (o.queue) {
            // two scans in flight: submit k+1 before waiting for k (lo_odometry_submit / lo_odometry_wait)
            const auto t0 = clock();
            odom.submit(buf[0]);
            seconds += std::chrono::duration<double>(clock() - t0).count();
            for (size_t k = 0; k < n; k++) {
                if (k + 1 < n) lo::io::read_scan(o.scans[k + 1], buf[(k + 1) % 3]);
                const auto t1 = clock();
                if (k + 1 < n) odom.submit(buf[(k + 1) % 3]);
                odom.wait();
                seconds += std::chrono::duration<double>(clock() - t1).count();
                points += buf[k % 3].size();
                if (traj) traj->add(static_cast<double>(k) * o.period, odom.getCurrentPose());
                report(o, k, odom);
            }
        } else {
            for (size_t k = 0; k < n; k++) {
                if (k + 1 < n) lo::io::read_scan(o.scans[k + 1], buf[(k + 1) % 2]);
                const auto t1 = clock();
                if (k + 1 < n) odom.processCloud(buf[k % 2], buf[(k + 1) % 2]);   // the next scan's upload runs behind this one
                else odom.processCloud(buf[k % 2]);
                seconds += std::chrono::duration<double>(clock() - t1).count();
                points += buf[k % 2].size();
                if (traj) traj->add(static_cast<double>(k) * o.period, odom.getCurrentPose());
                report(o, k, odom);
            }
        }
        if (traj) traj->flush();
        if (!o.keyframe_pcd.empty()) lo::io::write_pcd(o.keyframe_pcd, *odom.getKeyFrameCloud());
        if (!o.full_keyframe_pcd.empty()) lo::io::write_pcd(o.full_keyframe_pcd, *odom.getFullKeyFrameCloud());
        if (!o.deskewed_pcd.empty()) lo::io::write_pcd(o.deskewed_pcd, *odom.getTempCloud());
        if (!o.checkpoint_out.empty()) odom.saveCheckpoint(o.checkpoint_out);
        const Pose3D p = odom.getCurrentPose();
        std::printf("{\"scans\": %zu, \"points\": %zu, \"seconds_in_process_cloud\": %.6f, \"scans_per_s\": %.1f, \"final_pose\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g, %.9g]}\n",
                    n, points, seconds, seconds > 0 ? n / seconds : 0.0, p.translation.x(), p.translation.y(), p.translation.z(), p.rotation.w(),
                    p.rotation.x(), p.rotation.y(), p.rotation.z());
    } catch (const std::exception& e) {
        std::fprintf(stderr, "lo_replay: %s\n", e.what());
        return 1;
    }
    return 0;
}
