// LidarOdometry::processCloud (src/lidar_odometry.cpp:22-76) as a device-resident pipeline: every stage is
// enqueued on the context's stream, poses and counts stay in device memory, and the host synchronises once
// per scan to read back the pose and the per-scan counters (28 + 64 bytes).
#include <algorithm>
#include <cstdlib>
#include <atomic>
#include <cstring>

#include <chrono>

#include "lo_copy_pool.h"
#include "lo_odometry.cuh"

namespace lo {

__global__ void odom_init_kernel(OdomDev* s) {
    s->previous = identity_pose();
    s->current = identity_pose();
    s->identity = identity_pose();
    s->n_planar = s->n_update = s->n_query = s->unstable = 0;
}

// lidar_odometry.cpp:27-30,51: relative = previous.relativeTo(current); previous = current;
// deskew start pose = relative.inverse(); guess = current.compose(relative)
__global__ void odom_predict_kernel(OdomDev* s) {
    const Pose rel = pose_relative_to(s->previous, s->current);
    s->previous = s->current;
    s->relative = rel;
    s->relative_inv = pose_inverse(rel);
    s->guess = pose_compose(s->current, rel);
    s->unstable = 0;
    s->n_query = 0;
    s->align.outer_iterations = 0;
}

__device__ inline float as_float(double v) { return static_cast<float>(v); }

// lidar_odometry.cpp:53-65: Euler-angle divergence gate, then current = new
__global__ void odom_gate_kernel(OdomDev* s, float threshold) {
    const Pose nw = s->next, cur = s->current;
    // new.rotation * current.rotation.inverse()
    Pose a = identity_pose(), b = identity_pose();
    a.w = nw.w; a.x = nw.x; a.y = nw.y; a.z = nw.z;
    b.w = cur.w; b.x = cur.x; b.y = cur.y; b.z = cur.z;
    const Pose d = pose_compose(a, pose_inverse(b));
    float r[9];
    rotation_matrix_f(d.w, d.x, d.y, d.z, r);
    const float pi_f = 3.14159265358979323846f;
    float ang[3];
    euler_angles_012(r, ang);
    bool ok = true;
    for (int k = 0; k < 3; k++) {
        const float deg = (ang[k] * 180.0f) / pi_f;
        const float m = fabsf(deg);
        ok = ok && (m < threshold || m > 180 - threshold);
    }
    Pose out = nw;
    if (!ok) {
        out = s->guess;
        s->unstable = 1;
    }
    s->current = out;
    s->center[0] = out.t[0]; s->center[1] = out.t[1]; s->center[2] = out.t[2];
}

__global__ void odom_readback_kernel(const OdomDev* s, const uint32_t* kf_counters, ScanReadback* out, uint32_t* minmax) {
    minmax[0] = 0xFFFFFFFFu;  // re-arm the time min/max of the fused pipeline's next scan
    minmax[1] = 0u;
    pose_to7(s->current, out->pose);
    out->n_planar = s->n_planar;
    out->n_update = s->n_update;
    out->n_query = s->n_query;
    out->unstable = s->unstable;
    out->outer_iterations = s->align.outer_iterations;
    out->bootstrap = 0;
    out->overflow = 0;   // the stage-by-stage classifier reports an overflow through classify_dev's own flag
    out->pose_seq = out->seq = 0;
    for (int k = 0; k < GC_COUNT; k++) out->kf_counters[k] = kf_counters[k];
}

}  // namespace lo

uint32_t lo::fused_max_points() {
    static const uint32_t v = [] {
        const char* e = getenv("LO_B200_FUSED_MAX");
        return e && atoll(e) > 0 ? static_cast<uint32_t>(atoll(e)) : kFusedMaxPoints;
    }();
    return v;
}

using namespace lo;

static int run_from_filtered(lo_odometry* h, uint32_t bound, const uint32_t* d_P) {
    lo_ctx* ctx = h->ctx;
    OdomDev* s = h->d_state;
    LO_TRY(h->update.ensure(std::max<size_t>(bound, 1) * sizeof(PointNormal)));
    LO_TRY(h->query.ensure(std::max<size_t>(bound, 1) * sizeof(PointXYZ)));
    // keyframe_downsampler (cpp:37-38) and its getCloud() in the reference's iteration order
    LO_TRY(grid_clear(h->update_ds));
    LO_TRY(grid_add_dev(h->update_ds, h->filtered.p, bound, 1, nullptr, d_P));
    LO_TRY(robin_export_bounded(h->update_ds, 0, h->update.p, &s->n_update, bound));
    const bool first = h->keyframe->h_counters[GC_VOXELS] == 0;  // cpp:40 (host mirror from the last scan's readback)
    if (first) {
        LO_TRY(grid_add_dev(h->keyframe, h->update.p, bound, 1, nullptr, &s->n_update));  // cpp:42
        return LO_OK;
    }
    // matching_downsampler (cpp:46-47) and getCloudWithoutNormals()
    LO_TRY(grid_clear(h->match_ds));
    LO_TRY(grid_add_dev(h->match_ds, h->filtered.p, bound, 1, nullptr, d_P));
    LO_TRY(robin_export_bounded(h->match_ds, 1, h->query.p, &s->n_query, bound));
    // CloudMatcher::align(keyframe_, query, current.compose(relative))  cpp:49-51
    LO_TRY(align_enqueue(h->keyframe, h->query.as<PointXYZ>(), bound, &s->n_query, reinterpret_cast<const float*>(&s->guess),
                         reinterpret_cast<float*>(&s->next), &s->align));
    odom_gate_kernel<<<1, 1, 0, ctx->stream>>>(s, h->params.angular_divergence_threshold);  // cpp:53-65
    LO_LAUNCHED(ctx);
    LO_TRY(grid_cleanup_dev(h->keyframe, s->center, h->params.keyframe_cleanup_range));  // cpp:67
    // transformWithNormals fused into the insert (cpp:69-70)
    LO_TRY(grid_add_dev(h->keyframe, h->update.p, bound, 1, &s->current, &s->n_update));
    return LO_OK;
}

// second half of a scan's result block: host mirror of the key-frame counters
static int adopt_counters(lo_odometry* h, const uint32_t kf_counters[GC_COUNT], bool predates_rebuild) {
    lo_grid* kf = h->keyframe;
    const uint32_t ticket = kf->h_counters[GC_TICKET], tombs = kf->h_counters[GC_TOMBS];   // tickets are tracked on the host
    std::memcpy(kf->h_counters, kf_counters, sizeof(kf->h_counters));
    kf->h_counters[GC_TICKET] = ticket;
    // a block written before the table was last rebuilt still counts the tombstones that rebuild purged
    if (predates_rebuild) kf->h_counters[GC_TOMBS] = tombs;
    h->stats.keyframe_voxels = kf_counters[GC_VOXELS];
    if (kf_counters[GC_RANGE_ERR]) {
        set_error("a key-frame point was not finite or outside the +-2^20 voxel index range");
        return LO_E_RANGE;
    }
    return LO_OK;
}

// first half: pose and per-scan counts
static int adopt_pose(lo_odometry* h, const ScanReadback& r, uint32_t n_points) {
    std::memcpy(h->pose, r.pose, sizeof(h->pose));
    h->stats.n_points = n_points;
    h->stats.n_planar = r.n_planar;
    h->stats.n_update = r.n_update;
    h->stats.n_query = r.n_query;
    h->stats.outer_iterations = r.outer_iterations;
    h->stats.unstable = static_cast<int32_t>(r.unstable);
    if (r.overflow) {
        set_error("the classifier's ring x azimuth grid (widest ring x rings present) exceeds the scratch sized for %u points: "
                  "the scan produced no planar points", n_points);
        return LO_E_CAPACITY;
    }
    return LO_OK;
}

// stage-by-stage pipeline: read-back kernel + one D2H
static int finish_scan(lo_odometry* h, uint32_t n_points) {
    lo_ctx* ctx = h->ctx;
    odom_readback_kernel<<<1, 1, 0, ctx->stream>>>(h->d_state, h->keyframe->counters, h->d_readback, h->d_minmax);
    LO_LAUNCHED(ctx);
    LO_CUDA(cudaMemcpyAsync(h->h_stage_readback, h->d_readback, sizeof(ScanReadback), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    h->keyframe->cleaned = true;
    const int st = adopt_pose(h, *h->h_stage_readback, n_points);
    const int st2 = adopt_counters(h, h->h_stage_readback->kf_counters, false);
    h->stats.update_order = kOrderReplayed;   // the staged pipeline replays the whole insertion history (lo_robin.cu)
    return st != LO_OK ? st : st2;
}

// spin until *word == want; the stream is checked now and then so that a failed launch cannot hang the caller.
// `pause` between polls: the word lives in pinned host memory, the store that ends the wait comes over PCIe.
static int poll_word(lo_ctx* ctx, const volatile uint32_t* word, uint32_t want, const char* what) {
    for (uint32_t spins = 0; *word != want; spins++) {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
        if ((spins & 0x3FFFFu) == 0x3FFFFu && cudaStreamQuery(ctx->stream) != cudaErrorNotReady) break;
    }
    if (*word != want) LO_CUDA(cudaStreamSynchronize(ctx->stream));
    if (*word != want) {
        set_error("the per-scan kernel finished without publishing its %s", what);
        return LO_E_CUDA;
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return LO_OK;
}

// Take over the key-frame counters of the scans that have been waited for, oldest first. block = false: only those whose
// kernel has ended already. Whatever looks at the key frame from the host (stats, export, snapshot, the next reserve
// when the ring slot is needed) calls this with block = true; the blocking process calls do not, so they return while
// the kernel's tail is still running.
static int adopt_finished(lo_odometry* h, uint32_t upto, bool block) {
    int st = LO_OK;
    while (h->adopt_seq < upto) {
        const uint32_t want = h->adopt_seq + 1, slot = want % kQueueDepth;
        const volatile uint32_t* seq = &h->h_readback[slot].seq;
        if (*seq != want) {
            if (!block) break;
            LO_TRY(poll_word(h->ctx, seq, want, "key-frame counters"));
        }
        std::atomic_thread_fence(std::memory_order_acquire);
        h->adopt_seq = want;
        h->pending_points -= h->queued_n[slot];
        if (h->adopt_seq == h->scan_seq) fused_trace_dump(h);
        const int s2 = adopt_counters(h, h->h_readback[slot].kf_counters, h->queued_rebuilds[slot] != h->keyframe->rebuild_count);
        h->stats.update_order = h->h_readback[slot].update_order;   // known once the kernel's tail has run
        if (st == LO_OK) st = s2;
    }
    return st;
}

// per-scan kernel: wait for the pose of the oldest scan in flight. The kernel wrote the block into pinned memory; poll
// its sequence word (the result is there a few microseconds before the stream would report anything).
static int wait_oldest(lo_odometry* h) {
    lo_ctx* ctx = h->ctx;
    const uint32_t want = h->wait_seq + 1, slot = want % kQueueDepth;
    static const char* wait_mode = getenv("LO_B200_WAIT");   // developer switch: "sync" waits on the stream first
    if (wait_mode && wait_mode[0] == 's') LO_CUDA(cudaStreamSynchronize(ctx->stream));
    LO_TRY(poll_word(ctx, &h->h_readback[slot].pose_seq, want, "pose"));
    h->wait_seq = want;
    const ScanReadback& r = h->h_readback[slot];
    if (!r.bootstrap) h->keyframe->cleaned = true;   // radiusCleanup ran: reference iteration order can no longer be emulated
    const int st = adopt_pose(h, r, h->queued_n[slot]);
    const int st2 = adopt_finished(h, h->wait_seq, false);   // whatever has ended by now, without waiting
    return st != LO_OK ? st : st2;
}

static int submit_fused(lo_odometry* h, const PointXYZIRT* d_scan, uint32_t n) {
    if (h->scan_seq - h->wait_seq >= kQueueDepth) {
        set_error("%u scans are already in flight: call lo_odometry_wait first", kQueueDepth);
        return LO_E_BUSY;
    }
    // the ring slot this scan will write must have been read out completely (it is kQueueDepth scans old: long done)
    if (h->scan_seq + 1 > kQueueDepth) LO_TRY(adopt_finished(h, h->scan_seq + 1 - kQueueDepth, true));
    LO_TRY(adopt_finished(h, h->wait_seq, false));
    const uint64_t rebuilds = h->keyframe->rebuild_count;
    LO_TRY(fused_process_enqueue(h, d_scan, n));
    const uint32_t slot = h->scan_seq % kQueueDepth;
    h->queued_n[slot] = n;
    h->queued_ticket[slot] = h->keyframe->h_counters[GC_TICKET];
    h->queued_rebuilds[slot] = rebuilds;   // a rebuild queued by this very submit runs before the scan's kernel
    if (h->keyframe->rebuild_count != rebuilds) h->queued_rebuilds[slot] = h->keyframe->rebuild_count;
    h->pending_points += n;
    h->n_last = n;
    return LO_OK;
}

// every scan submitted has handed out its pose: wait for the kernels' tails too and mirror the key-frame counters
static int settle(lo_odometry* h) { return adopt_finished(h, h->wait_seq, true); }

// keep the host-tracked ticket counter across the device counter readback
static void preserve_ticket(lo_grid* g, uint32_t ticket) { g->h_counters[GC_TICKET] = ticket; }

// Key-frame capacity (voxels) a LidarOdometry is created with. The table and both slabs are sized ONCE, here (SURVEY 7.2:
// "power-of-two capacity sized once, no rehash on device"): 2^19 voxel slabs of max_points points (2 x 168 MB at the
// defaults) behind 2^20 slots (16 MB) hold three times the ~1.6e5 voxels the VLP16 street stream settles at inside its
// 80 m clean-up range; nothing on a 180 GB part. A map that outgrows it is regrown stream-ordered (lo_grid.cu), never
// by a synchronising call. LO_B200_KEYFRAME_VOXELS / lo_odometry_create_sized override the default.
static size_t default_keyframe_voxels() {
    const char* e = getenv("LO_B200_KEYFRAME_VOXELS");
    return e && atoll(e) > 0 ? static_cast<size_t>(atoll(e)) : (static_cast<size_t>(1) << 19);
}

extern "C" int lo_odometry_create_sized(lo_ctx* ctx, const lo_params* params, size_t keyframe_voxels, size_t scan_points, lo_odometry** out) {
    if (!ctx || !params || !out) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (keyframe_voxels == 0) keyframe_voxels = default_keyframe_voxels();
    if (scan_points == 0) scan_points = kPresizedScanPoints;
    if (scan_points > fused_max_points()) scan_points = fused_max_points();
    lo_odometry* h = new lo_odometry();
    h->ctx = ctx;
    h->params = *params;
    std::memset(&h->stats, 0, sizeof(h->stats));
    int st = lo_grid_create(ctx, params->keyframe_voxel_size, static_cast<size_t>(params->keyframe_max_points_cnt), keyframe_voxels, &h->keyframe);
    if (st == LO_OK) st = lo_grid_create(ctx, params->keyframe_update_voxel_size, 1, 1 << 15, &h->update_ds);
    if (st == LO_OK) st = lo_grid_create(ctx, params->keyframe_matching_voxel_size, 1, 1 << 15, &h->match_ds);
    // everything a scan of up to scan_points points needs, so that no submit / process call allocates
    if (st == LO_OK) st = fused_reserve(h, static_cast<uint32_t>(scan_points));
    if (st == LO_OK) st = h->scan.ensure(scan_points * sizeof(PointXYZIRT));
    if (st == LO_OK) st = h->stage.ensure(scan_points * sizeof(PointXYZIRT));
    if (st != LO_OK) {
        lo_odometry_destroy(h);
        return st;
    }
    LO_CUDA(cudaMalloc(&h->d_state, sizeof(OdomDev)));
    LO_CUDA(cudaMemsetAsync(h->d_state, 0, sizeof(OdomDev), ctx->stream));
    LO_CUDA(cudaMalloc(&h->d_readback, sizeof(ScanReadback)));
    LO_CUDA(cudaMallocHost(&h->h_readback, (kQueueDepth + 1) * sizeof(ScanReadback)));
    std::memset(h->h_readback, 0, (kQueueDepth + 1) * sizeof(ScanReadback));
    h->h_stage_readback = h->h_readback + kQueueDepth;   // D2H target of the stage-by-stage pipeline
    LO_CUDA(cudaMalloc(&h->d_minmax, 2 * sizeof(uint32_t)));
    const uint32_t minmax_init[2] = {0xFFFFFFFFu, 0u};
    LO_CUDA(cudaMemcpyAsync(h->d_minmax, minmax_init, sizeof(minmax_init), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    LO_CUDA(cudaEventCreateWithFlags(&h->copy_done, cudaEventDisableTiming));
    for (uint32_t k = 0; k < kQueueDepth; k++) LO_CUDA(cudaEventCreateWithFlags(&h->scan_q_done[k], cudaEventDisableTiming));
    odom_init_kernel<<<1, 1, 0, ctx->stream>>>(h->d_state);
    LO_LAUNCHED(ctx);
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = h;
    return LO_OK;
}

extern "C" int lo_odometry_create(lo_ctx* ctx, const lo_params* params, lo_odometry** out) {
    return lo_odometry_create_sized(ctx, params, 0, 0, out);
}

extern "C" void lo_odometry_destroy(lo_odometry* h) {
    if (!h) return;
    cudaSetDevice(h->ctx->device);
    cudaStreamSynchronize(h->ctx->stream);
    lo_grid_destroy(h->keyframe);
    lo_grid_destroy(h->update_ds);
    lo_grid_destroy(h->match_ds);
    if (h->d_state) cudaFree(h->d_state);
    if (h->d_readback) cudaFree(h->d_readback);
    if (h->h_readback) cudaFreeHost(h->h_readback);
    if (h->d_minmax) cudaFree(h->d_minmax);
    fused_release(h);
    if (h->copy_stream) {
        cudaStreamSynchronize(h->copy_stream);
        cudaStreamDestroy(h->copy_stream);
    }
    if (h->copy_done) cudaEventDestroy(h->copy_done);
    delete h->copy_pool;
    for (DevBuf* b : {&h->scan, &h->scan_next, &h->deskewed, &h->filtered, &h->update, &h->query}) b->release();
    h->stage.release();
    for (DevBuf& b : h->scan_q) b.release();
    for (cudaEvent_t e : h->scan_q_done)
        if (e) cudaEventDestroy(e);
    delete h;
}

extern "C" int lo_odometry_process_dev(lo_odometry* h, const void* d_xyzirt32, size_t n) {
    if (!h || (!d_xyzirt32 && n)) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    OdomDev* s = h->d_state;
    const uint32_t nn = static_cast<uint32_t>(n);
    static const bool staged = getenv("LO_B200_STAGED") != nullptr;  // developer switch: stage-by-stage pipeline
    if (h->scan_seq != h->wait_seq) {
        set_error("scans submitted with lo_odometry_submit are still in flight: call lo_odometry_wait first");
        return LO_E_BUSY;
    }
    if (!staged && nn <= fused_max_points()) {
        LO_TRY(submit_fused(h, static_cast<const PointXYZIRT*>(d_xyzirt32), nn));
        return wait_oldest(h);
    }
    LO_TRY(settle(h));
    const uint32_t kf_ticket = h->keyframe->h_counters[GC_TICKET];
    LO_TRY(h->deskewed.ensure(std::max<size_t>(n, 1) * sizeof(PointXYZIRT)));
    LO_TRY(h->filtered.ensure(std::max<size_t>(n, 1) * sizeof(PointNormal)));
    odom_predict_kernel<<<1, 1, 0, ctx->stream>>>(s);  // cpp:27-28
    LO_LAUNCHED(ctx);
    LO_TRY(deskew_dev(ctx, static_cast<const PointXYZIRT*>(d_xyzirt32), nn, &s->relative_inv, &s->identity,
                      h->deskewed.as<PointXYZIRT>()));  // cpp:25,30
    h->n_last = nn;
    // classify (cpp:33) fused with rangeFilter (cpp:35)
    LO_TRY(classify_dev(ctx, h->deskewed.as<PointXYZIRT>(), nn, 1, h->params.lidar_min_range, h->params.lidar_max_range,
                        h->filtered.as<PointNormal>(), &s->n_planar));
    LO_TRY(run_from_filtered(h, nn, &s->n_planar));
    const uint32_t ticket_after = h->keyframe->h_counters[GC_TICKET];
    (void)kf_ticket;
    const int st = finish_scan(h, nn);
    preserve_ticket(h->keyframe, ticket_after);
    return st;
}

// Upload of a scan that sits in ordinary (pageable) host memory, as processCloud's caller has it: the scan is copied
// through the handle's pinned staging buffer in pieces, each piece's DMA running under the memcpy of the next one.
// (cudaMemcpyAsync on pageable memory stages too, but serially and with a synchronising tail.) The DMA runs on the copy
// stream: the previous process call returned when its pose was final, so its kernel may still be in its tail
// (radiusCleanup + key-frame insert, a fifth of the kernel) — which reads the filtered cloud, never the raw scan (last read
// by the deskew phase) — and the upload goes on underneath it instead of queueing behind it.
// helper threads of the staging copy: LO_B200_COPY_THREADS, else up to 3, leaving every GPU of the box (one process per GPU
// is the normal deployment) half of its share of the cores
static int copy_helpers() {
    if (const char* e = getenv("LO_B200_COPY_THREADS")) return std::max(0, std::min(15, atoi(e)));
    int gpus = 1;
    if (cudaGetDeviceCount(&gpus) != cudaSuccess || gpus < 1) gpus = 1;
    const int cores = static_cast<int>(std::thread::hardware_concurrency());
    return std::max(0, std::min(3, cores / (2 * gpus) - 1));
}

static int upload_scan(lo_odometry* h, const void* xyzirt32, size_t n) {
    lo_ctx* ctx = h->ctx;
    const size_t bytes = n * sizeof(PointXYZIRT);
    LO_TRY(h->scan.ensure(std::max<size_t>(bytes, sizeof(PointXYZIRT))));
    if (!n) return LO_OK;
    static const int mode = [] { const char* e = getenv("LO_B200_UPLOAD"); return e ? atoi(e) : 0; }();   // developer switch: 1 = plain
    if (mode == 1) {                                                                                    // cudaMemcpyAsync, 2 = main stream
        LO_CUDA(cudaMemcpyAsync(h->scan.p, xyzirt32, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return LO_OK;
    }
    cudaStream_t copy = mode == 2 ? ctx->stream : h->copy_stream;
    // the previous scan's DMA out of the staging buffer ended before that call returned (its kernel was waited for)
    LO_TRY(h->stage.ensure(bytes));
    if (!h->copy_pool) h->copy_pool = new CopyPool(copy_helpers());
    // the helpers copy chunk after chunk; this thread hands every completed piece of the prefix to the DMA engine at once and
    // copies chunks itself while there is nothing to hand over
    CopyPool& pool = *h->copy_pool;
    pool.begin(h->stage.p, xyzirt32, bytes);
    constexpr size_t kPiece = 256 * 1024, kFirstPiece = 64 * 1024;   // the DMA (~40 us for a VLP16 scan) is the longer leg: start it early
    size_t sent = 0;
    while (sent < bytes) {
        const size_t ready = pool.ready_bytes();
        if (ready - sent >= (sent ? kPiece : kFirstPiece) || (ready == bytes && ready > sent)) {
            const cudaError_t e = cudaMemcpyAsync(h->scan.as<char>() + sent, h->stage.as<char>() + sent, ready - sent, cudaMemcpyHostToDevice, copy);
            if (e != cudaSuccess) {
                pool.finish();   // the job must be complete before the pool sees the next one
                set_error("uploading the scan failed: %s", cudaGetErrorString(e));
                return LO_E_CUDA;
            }
            sent = ready;
        } else if (!pool.help_one()) {
            CopyPool::cpu_relax();
        }
    }
    if (copy != ctx->stream) {
        LO_CUDA(cudaEventRecord(h->copy_done, copy));
        LO_CUDA(cudaStreamWaitEvent(ctx->stream, h->copy_done, 0));
    }
    return LO_OK;
}

extern "C" int lo_odometry_host_threads(const lo_odometry* h) {
    if (!h) return 0;
    return 1 + (h->copy_pool ? h->copy_pool->helpers() : copy_helpers());
}

extern "C" int lo_odometry_process(lo_odometry* h, const void* xyzirt32, size_t n) {
    if (!h || (!xyzirt32 && n)) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (h->scan_seq != h->wait_seq) {
        set_error("scans submitted with lo_odometry_submit are still in flight: call lo_odometry_wait first");
        return LO_E_BUSY;
    }
    static const bool timing = getenv("LO_B200_TIMING") != nullptr;   // developer switch: host microseconds per part of the call, on stderr
    if (!timing) {
        LO_TRY(upload_scan(h, xyzirt32, n));
        return lo_odometry_process_dev(h, h->scan.p, n);
    }
    static double t_upload = 0, t_submit = 0, t_wait = 0;
    static long calls = 0;
    auto now = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    LO_TRY(upload_scan(h, xyzirt32, n));
    const double t1 = now();
    int st = LO_OK;
    if (n <= fused_max_points()) {
        st = submit_fused(h, h->scan.as<PointXYZIRT>(), static_cast<uint32_t>(n));
        const double t2 = now();
        if (st == LO_OK) st = wait_oldest(h);
        const double t3 = now();
        t_upload += t1 - t0; t_submit += t2 - t1; t_wait += t3 - t2;
        if (++calls % 100 == 0) {
            fprintf(stderr, "lo_odometry_process: per call over the last 100: upload (staging memcpy + DMA enqueue) %.1f us, submit %.1f us, wait for the pose %.1f us\n",
                    t_upload / 100, t_submit / 100, t_wait / 100);
            t_upload = t_submit = t_wait = 0;
        }
    } else {
        st = lo_odometry_process_dev(h, h->scan.p, n);
    }
    return st;
}

extern "C" int lo_odometry_process_next(lo_odometry* h, const void* xyzirt32, size_t n, const void* next_xyzirt32, size_t next_n) {
    if (!h || (!xyzirt32 && n) || (!next_xyzirt32 && next_n)) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (h->next_ptr && h->next_ptr == xyzirt32 && h->next_n == n) {
        std::swap(h->scan, h->scan_next);                                   // uploaded (or arriving) behind the previous scan
        LO_CUDA(cudaStreamWaitEvent(ctx->stream, h->copy_done, 0));
    } else {
        LO_TRY(h->scan.ensure(std::max<size_t>(n, 1) * sizeof(PointXYZIRT)));
        if (n) LO_CUDA(cudaMemcpyAsync(h->scan.p, xyzirt32, n * sizeof(PointXYZIRT), cudaMemcpyHostToDevice, ctx->stream));
    }
    h->next_ptr = nullptr;
    if (next_xyzirt32 && next_n) {
        // scan_next was last read by a scan that has completed (every process call returns after its scan)
        LO_TRY(h->scan_next.ensure(next_n * sizeof(PointXYZIRT)));
        LO_CUDA(cudaMemcpyAsync(h->scan_next.p, next_xyzirt32, next_n * sizeof(PointXYZIRT), cudaMemcpyHostToDevice, h->copy_stream));
        LO_CUDA(cudaEventRecord(h->copy_done, h->copy_stream));
        h->next_ptr = next_xyzirt32;
        h->next_n = next_n;
    }
    return lo_odometry_process_dev(h, h->scan.p, n);
}

extern "C" int lo_odometry_submit_dev(lo_odometry* h, const void* d_xyzirt32, size_t n) {
    if (!h || (!d_xyzirt32 && n)) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    if (n > fused_max_points()) {
        set_error("lo_odometry_submit takes scans of at most %u points; larger ones go through lo_odometry_process", fused_max_points());
        return LO_E_ARG;
    }
    return submit_fused(h, static_cast<const PointXYZIRT*>(d_xyzirt32), static_cast<uint32_t>(n));
}

extern "C" int lo_odometry_submit(lo_odometry* h, const void* xyzirt32, size_t n) {
    if (!h || (!xyzirt32 && n)) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (n > fused_max_points()) {
        set_error("lo_odometry_submit takes scans of at most %u points; larger ones go through lo_odometry_process", fused_max_points());
        return LO_E_ARG;
    }
    if (h->scan_seq - h->wait_seq >= kQueueDepth) {
        set_error("%u scans are already in flight: call lo_odometry_wait first", kQueueDepth);
        return LO_E_BUSY;
    }
    // the slot's device copy was last read by the scan kQueueDepth submits ago, whose pose has been waited for — but its
    // kernel may still be reading the scan in its tail only if it were the scan in flight, which the ring excludes;
    // the upload runs on the copy stream, i.e. under the kernel of the scan in flight
    const uint32_t slot = (h->scan_seq + 1) % kQueueDepth;
    if (h->scan_seq + 1 > kQueueDepth) LO_TRY(adopt_finished(h, h->scan_seq + 1 - kQueueDepth, true));
    LO_TRY(h->scan_q[slot].ensure(std::max<size_t>(std::max<size_t>(n, kPresizedScanPoints), 1) * sizeof(PointXYZIRT)));
    if (n) LO_CUDA(cudaMemcpyAsync(h->scan_q[slot].p, xyzirt32, n * sizeof(PointXYZIRT), cudaMemcpyHostToDevice, h->copy_stream));
    LO_CUDA(cudaEventRecord(h->scan_q_done[slot], h->copy_stream));
    LO_CUDA(cudaStreamWaitEvent(ctx->stream, h->scan_q_done[slot], 0));
    return submit_fused(h, h->scan_q[slot].as<PointXYZIRT>(), static_cast<uint32_t>(n));
}

extern "C" int lo_odometry_wait(lo_odometry* h) {
    if (!h) return LO_E_ARG;
    if (h->scan_seq == h->wait_seq) {
        set_error("lo_odometry_wait: no scan in flight");
        return LO_E_ARG;
    }
    LO_CUDA(cudaSetDevice(h->ctx->device));
    return wait_oldest(h);
}

extern "C" int lo_odometry_process_planar(lo_odometry* h, const float* pn48_planar, size_t n) {
    if (!h || (!pn48_planar && n)) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    OdomDev* s = h->d_state;
    const uint32_t nn = static_cast<uint32_t>(n);
    if (h->scan_seq != h->wait_seq) {
        set_error("scans submitted with lo_odometry_submit are still in flight: call lo_odometry_wait first");
        return LO_E_BUSY;
    }
    LO_TRY(settle(h));
    LO_TRY(h->scan.ensure(std::max<size_t>(n, 1) * sizeof(PointNormal)));
    LO_TRY(h->filtered.ensure(std::max<size_t>(n, 1) * sizeof(PointNormal)));
    if (n) LO_CUDA(cudaMemcpyAsync(h->scan.p, pn48_planar, n * sizeof(PointNormal), cudaMemcpyHostToDevice, ctx->stream));
    odom_predict_kernel<<<1, 1, 0, ctx->stream>>>(s);
    LO_LAUNCHED(ctx);
    LO_TRY(range_filter_dev(ctx, h->scan.as<PointNormal>(), nn, h->params.lidar_min_range, h->params.lidar_max_range,
                            h->filtered.as<PointNormal>(), &s->n_planar));
    h->n_last = 0;
    LO_TRY(run_from_filtered(h, nn, &s->n_planar));
    const uint32_t ticket_after = h->keyframe->h_counters[GC_TICKET];
    const int st = finish_scan(h, nn);
    preserve_ticket(h->keyframe, ticket_after);
    return st;
}

extern "C" int lo_odometry_pose(lo_odometry* h, float out7[7]) {
    if (!h || !out7) return LO_E_ARG;
    std::memcpy(out7, h->pose, sizeof(h->pose));
    return LO_OK;
}

extern "C" int lo_odometry_stats(lo_odometry* h, lo_scan_stats* out) {
    if (!h || !out) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    const int st = settle(h);   // keyframe_voxels is known when the scan's key-frame insert is through
    *out = h->stats;
    return st;
}

extern "C" int lo_odometry_keyframe_info(lo_odometry* h, lo_keyframe_info* out) {
    if (!h || !out) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    const int st = settle(h);
    const lo_grid* kf = h->keyframe;
    out->voxels = kf->h_counters[GC_VOXELS];
    out->points = kf->h_counters[GC_POINTS];
    out->tombstones = kf->h_counters[GC_TOMBS];
    out->capacity_slots = kf->capacity;
    out->capacity_voxels = kf->payload_cap;
    out->table_rebuilds = kf->rebuild_count;
    out->payload_growths = kf->grow_count;
    return st;
}

extern "C" int lo_odometry_keyframe(lo_odometry* h, int full, float* xyz16_out, size_t cap, size_t* n) {
    if (!h || !n) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    LO_TRY(settle(h));
    return lo_grid_get_cloud(h->keyframe, full ? 1 : 2, 0, xyz16_out, cap, n);
}

extern "C" int lo_odometry_keyframe_pn(lo_odometry* h, float* pn48_out, size_t cap, size_t* n) {
    if (!h || !n) return LO_E_ARG;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    LO_TRY(settle(h));
    return lo_grid_get_cloud(h->keyframe, 0, 0, pn48_out, cap, n);
}

extern "C" int lo_odometry_deskewed(lo_odometry* h, void* xyzirt32_out, size_t cap, size_t* n) {
    if (!h || !n) return LO_E_ARG;
    *n = h->n_last;
    if (!xyzirt32_out) return LO_OK;
    if (cap < h->n_last) return LO_E_CAPACITY;
    if (h->n_last == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(h->ctx->device));
    LO_CUDA(cudaMemcpyAsync(xyzirt32_out, h->deskewed.p, static_cast<size_t>(h->n_last) * sizeof(PointXYZIRT), cudaMemcpyDeviceToHost,
                            h->ctx->stream));
    LO_CUDA(cudaStreamSynchronize(h->ctx->stream));
    return LO_OK;
}

// ---- snapshot / restore of a running odometry. The reference keeps this state in plain value
// members (keyframe_, previous_transform_, current_transform_: src/lidar_odometry.h:82-85) and has no checkpoint;
// here it is the device-resident pose block plus the key-frame table as it lies in HBM, so a restored object
// continues with bit-identical poses. The two per-scan down-samplers are rebuilt by every scan and are not saved;
// getTempCloud is empty until the next scan.
namespace {
struct OdomSnapHeader {
    uint64_t magic;
    uint32_t header_bytes, state_bytes;
    lo_params params;
    lo_scan_stats stats;
    float pose[7];
    uint32_t reserved;
};
constexpr uint64_t kOdomMagic = 0x00014D4F444F4F4CULL;   // "LOODOM\1\0"
bool same_params(const lo_params& a, const lo_params& b) {
    return a.lidar_min_range == b.lidar_min_range && a.lidar_max_range == b.lidar_max_range && a.keyframe_voxel_size == b.keyframe_voxel_size &&
           a.keyframe_max_points_cnt == b.keyframe_max_points_cnt && a.keyframe_matching_voxel_size == b.keyframe_matching_voxel_size &&
           a.keyframe_update_voxel_size == b.keyframe_update_voxel_size && a.keyframe_cleanup_range == b.keyframe_cleanup_range &&
           a.angular_divergence_threshold == b.angular_divergence_threshold;
}
}  // namespace

extern "C" int lo_odometry_snapshot(lo_odometry* h, void* out, size_t cap, size_t* bytes) {
    if (!h || !bytes) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (h->scan_seq != h->wait_seq) {
        set_error("scans submitted with lo_odometry_submit are still in flight: call lo_odometry_wait first");
        return LO_E_BUSY;
    }
    LO_TRY(settle(h));
    size_t grid_bytes = 0;
    LO_TRY(grid_snapshot(h->keyframe, nullptr, 0, &grid_bytes));
    const size_t fixed = sizeof(OdomSnapHeader) + sizeof(OdomDev);
    *bytes = fixed + grid_bytes;
    if (!out) return LO_OK;
    if (cap < *bytes) return LO_E_CAPACITY;
    OdomSnapHeader hd{};
    hd.magic = kOdomMagic;
    hd.header_bytes = sizeof(OdomSnapHeader);
    hd.state_bytes = sizeof(OdomDev);
    hd.params = h->params;
    hd.stats = h->stats;
    std::memcpy(hd.pose, h->pose, sizeof(hd.pose));
    char* o = static_cast<char*>(out);
    std::memcpy(o, &hd, sizeof(hd));
    LO_CUDA(cudaMemcpyAsync(o + sizeof(hd), h->d_state, sizeof(OdomDev), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return grid_snapshot(h->keyframe, o + fixed, cap - fixed, &grid_bytes);
}

extern "C" int lo_odometry_restore(lo_odometry* h, const void* blob, size_t bytes) {
    if (!h || !blob) return LO_E_ARG;
    lo_ctx* ctx = h->ctx;
    LO_CUDA(cudaSetDevice(ctx->device));
    if (h->scan_seq != h->wait_seq) {
        set_error("scans submitted with lo_odometry_submit are still in flight: call lo_odometry_wait first");
        return LO_E_BUSY;
    }
    LO_TRY(settle(h));
    OdomSnapHeader hd;
    const size_t fixed = sizeof(OdomSnapHeader) + sizeof(OdomDev);
    if (bytes < fixed) {
        set_error("odometry snapshot: %zu bytes is shorter than the header", bytes);
        return LO_E_IO;
    }
    std::memcpy(&hd, blob, sizeof(hd));
    if (hd.magic != kOdomMagic || hd.header_bytes != sizeof(OdomSnapHeader) || hd.state_bytes != sizeof(OdomDev)) {
        set_error("odometry snapshot: bad magic or version");
        return LO_E_IO;
    }
    if (!same_params(hd.params, h->params)) {
        set_error("odometry snapshot was taken with different LidarOdometry::Params");
        return LO_E_ARG;
    }
    const char* in = static_cast<const char*>(blob);
    LO_TRY(grid_restore(h->keyframe, in + fixed, bytes - fixed));
    LO_CUDA(cudaMemcpyAsync(h->d_state, in + sizeof(hd), sizeof(OdomDev), cudaMemcpyHostToDevice, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    h->stats = hd.stats;
    std::memcpy(h->pose, hd.pose, sizeof(h->pose));
    h->n_last = 0;
    h->next_ptr = nullptr;
    return LO_OK;
}
