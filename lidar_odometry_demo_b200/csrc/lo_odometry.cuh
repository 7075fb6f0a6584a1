// Shared between lo_odometry.cu (entry points, generic stage-by-stage pipeline) and lo_pipeline.cu (the fused
// per-scan pipeline): device-resident odometry state and the lo_odometry object.
#pragma once

#include "lo_points.cuh"

namespace lo {

struct OdomDev {          // device-resident state of one LidarOdometry
    Pose previous;        // previous_transform_  (lidar_odometry.h:84)
    Pose current;         // current_transform_   (lidar_odometry.h:85)
    Pose relative, relative_inv, guess, next, identity;
    float center[4];      // current translation for radiusCleanup
    uint32_t n_planar, n_update, n_query, unstable;
    uint32_t pad[4];
    lo_align_stats align;
};

__device__ inline Pose identity_pose() {
    Pose p;
    p.t[0] = p.t[1] = p.t[2] = 0.f;
    p.w = 1.f; p.x = p.y = p.z = 0.f;
    return p;
}

// Matrix3f::eulerAngles(0,1,2) (Eigen 3.4) of a row-major rotation matrix, negated as the gate uses them
// (lidar_odometry.cpp:55); libm float functions evaluated in double and rounded once (only a 5 degree threshold
// depends on them).
__device__ inline void euler_angles_012(const float r[9], float ang[3]) {
    const float pi_f = 3.14159265358979323846f;
    float e0 = static_cast<float>(atan2(static_cast<double>(r[5]), static_cast<double>(r[8])));
    const float c2 = sqrtf(r[0] * r[0] + r[1] * r[1]);
    float e1;
    if (e0 > 0.f) {
        e0 -= pi_f;
        e1 = static_cast<float>(atan2(static_cast<double>(-r[2]), static_cast<double>(-c2)));
    } else {
        e1 = static_cast<float>(atan2(static_cast<double>(-r[2]), static_cast<double>(c2)));
    }
    const float s1 = static_cast<float>(sin(static_cast<double>(e0))), c1 = static_cast<float>(cos(static_cast<double>(e0)));
    const float e2 = static_cast<float>(atan2(static_cast<double>(s1 * r[6] - c1 * r[3]), static_cast<double>(c1 * r[4] - s1 * r[7])));
    ang[0] = -e0; ang[1] = -e1; ang[2] = -e2;
}

// Per scan, device -> host. The per-scan kernel writes it straight into pinned host memory in two steps: the pose and
// the scan's counts as soon as the divergence gate has decided (`pose_seq` last), and the key-frame counters when
// radiusCleanup and the key-frame insert are through (`seq` last). lo_odometry_wait / the blocking process calls return
// on `pose_seq` — the reference's processCloud has nothing else to hand back — while the kernel's tail (cpp:67-70) runs on
// behind them; whatever needs the key frame waits for `seq`. The stage-by-stage pipeline fills a device copy that is
// fetched with one D2H.
// lo_scan_stats.update_order (include/lo_b200.h)
constexpr uint32_t kOrderSorted = 0, kOrderReplayed = 1, kOrderApproximate = 2;

struct ScanReadback {
    float pose[7];
    uint32_t n_planar, n_update, n_query, unstable;
    int32_t outer_iterations;
    uint32_t bootstrap;   // the key frame was empty: the scan was inserted un-aligned (cpp:40-44)
    uint32_t overflow;    // the classifier's ring x azimuth grid did not fit the scratch: no planar points were produced
    uint32_t pose_seq;    // scan number, written after a system-scope fence once everything above is final
    uint32_t kf_counters[GC_COUNT];
    uint32_t seq;         // scan number, written last (after a system-scope fence)
    uint32_t update_order;   // kOrder*: how the update cloud's order was obtained (lo_scan_stats.update_order)
    uint32_t pad[2];
};

}  // namespace lo

namespace lo {
constexpr uint32_t kQueueDepth = LO_ODOMETRY_QUEUE_DEPTH;   // scans in flight through lo_odometry_submit*: rings of result blocks and staging buffers
}

namespace lo { class CopyPool; }

struct lo_odometry {
    lo_ctx* ctx = nullptr;
    lo_params params;
    lo_grid* keyframe = nullptr;
    lo_grid* update_ds = nullptr;    // VoxelGrid(keyframe_update_voxel_size, 1)   lidar_odometry.cpp:37
    lo_grid* match_ds = nullptr;     // VoxelGrid(keyframe_matching_voxel_size, 1) lidar_odometry.cpp:46
    lo::OdomDev* d_state = nullptr;
    lo::ScanReadback* d_readback = nullptr;
    lo::ScanReadback* h_readback = nullptr;  // [kQueueDepth + 1] pinned, mapped: the per-scan kernel of scan s writes slot s % kQueueDepth directly
    lo::ScanReadback* h_stage_readback = nullptr;   // the extra block: D2H target of the stage-by-stage pipeline
    uint32_t scan_seq = 0;                   // scans submitted to the per-scan kernel
    uint32_t wait_seq = 0;                   // ... waited for, i.e. their pose has been handed out (at most kQueueDepth in flight)
    uint32_t adopt_seq = 0;                  // ... and whose key-frame counters (end of the kernel) have been taken over
    uint32_t queued_n[lo::kQueueDepth] = {};      // per slot: points of the scan, key-frame ticket after it
    uint32_t queued_ticket[lo::kQueueDepth] = {};
    uint64_t queued_rebuilds[lo::kQueueDepth] = {};   // per slot: key-frame table rebuilds before the scan was queued
    size_t pending_points = 0;               // points of the scans whose key-frame counters are not on the host yet
    lo::PinnedBuf stage;                     // lo_odometry_process: pinned staging of a pageable host scan
    lo::CopyPool* copy_pool = nullptr;       // ... and the helper threads of that copy (created with the first such scan)
    lo::DevBuf scan_q[lo::kQueueDepth];      // lo_odometry_submit: device copies of the queued host scans
    cudaEvent_t scan_q_done[lo::kQueueDepth] = {};
    lo::DevBuf scan, deskewed, filtered, update, query;
    lo::DevBuf scan_next;                    // lo_odometry_process_next: the scan being uploaded behind the current one
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_done = nullptr;
    const void* next_ptr = nullptr;          // host pointer / count whose copy sits (or is arriving) in scan_next
    size_t next_n = 0;
    lo::DevBuf fused;                // scratch of the fused pipeline (lo_pipeline.cu)
    uint32_t* d_minmax = nullptr;    // [2] ordered min / max of `time`, re-armed by the read-back kernel
    uint32_t n_last = 0;             // points in the deskewed cloud
    lo_scan_stats stats;
    float pose[7] = {0, 0, 0, 1, 0, 0, 0};
};


namespace lo {
// lo_pipeline.cu: LidarOdometry::processCloud as one persistent cooperative kernel. Requires n <= kFusedMaxPoints;
// returns LO_OK after enqueueing the kernel, which also fills the device-side ScanReadback block.
constexpr uint32_t kFusedMaxPoints = 1u << 19;   // 524 288: a 128-beam revolution (262 144 returns) twice over
uint32_t fused_max_points();   // kFusedMaxPoints, or the developer override LO_B200_FUSED_MAX
int fused_process_enqueue(lo_odometry* h, const PointXYZIRT* d_scan, uint32_t n);
int fused_reserve(lo_odometry* h, uint32_t n);   // scratch for scans of up to n points (no-op when it is there already)
constexpr uint32_t kPresizedScanPoints = 32768;   // what lo_odometry_create reserves for: a VLP16 revolution is ~30 k returns
void fused_release(lo_odometry* h);
void fused_trace_dump(lo_odometry* h);
}  // namespace lo
