/*
 * lo_b200.h — C ABI of the B200-native scan-to-keyframe registration library (liblo_b200.so).
 *
 * This is the drop-in boundary for the reference's registration hot path. The reference
 * (vovo-4K/lidar_odometry_demo) has no FFI layer: its seam is three plain C++ classes. Each entry point
 * below names the reference interface it replaces (paths relative to the reference tree); the C++ facade
 * in lidar_odometry_demo_b200/host/ re-creates those classes on top of this ABI, and INTEGRATION.md shows
 * the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns LO_OK (0) or a negative LO_E_* code; lo_last_error() gives the text.
 *   - no C++ exceptions cross the ABI; all device memory is owned by the opaque handles.
 *   - point layouts are the reference's wire structs, so `cloud.points.data()` can be passed as is:
 *       xyz16    pcl::PointXYZ      16 B  {x,y,z,pad}
 *       pn48     pcl::PointNormal   48 B  {x,y,z,pad, nx,ny,nz,pad, curvature,pad[3]}
 *       xyzirt32 lidar_point::PointXYZIRT 32 B {x,y,z,pad, intensity, ring(u16)@20, time@24}
 *                (src/lidar_point_type.h:13-31)
 *   - pose7 = {tx,ty,tz, qw,qx,qy,qz} float  (Pose3D, src/pose_3d.h:10-21)
 *   - handles are not thread-safe; distinct lo_ctx are independent (one per GPU / per sequence).
 *   - calls taking host pointers are synchronous: results are on the host when they return.
 *     `_dev` variants take device pointers (cudaMalloc'ed or a torch tensor's data_ptr) on the context's
 *     device and only enqueue work on the context's stream.
 *   - there is no CPU fallback: if no CUDA device is present lo_ctx_create fails with LO_E_CUDA.
 */
#ifndef LO_B200_H
#define LO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    LO_OK = 0,
    LO_E_CUDA = -1,      /* CUDA runtime error (text in lo_last_error) */
    LO_E_ARG = -2,       /* bad argument */
    LO_E_CAPACITY = -3,  /* caller buffer too small; the needed count is returned through the size out-param */
    LO_E_RANGE = -4,     /* a voxel index fell outside the supported +-2^20 range */
    LO_E_IO = -5,        /* file could not be read / parsed */
    LO_E_BUSY = -6       /* lo_odometry_submit: the queue is full / scans in flight during a synchronous call */
};

typedef struct lo_ctx lo_ctx;
typedef struct lo_grid lo_grid;
typedef struct lo_odometry lo_odometry;

/* ---- context: one CUDA device + one stream + scratch ---------------------------------------- */
int lo_ctx_create(int device, lo_ctx** out);
void lo_ctx_destroy(lo_ctx* ctx);
const char* lo_last_error(void);
/* cudaStream_t of the context (for CUDA-event timing by the caller) */
void* lo_ctx_stream(lo_ctx* ctx);
int lo_ctx_synchronize(lo_ctx* ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t lo_ctx_kernel_launches(const lo_ctx* ctx);

/* ---- voxel keys: VoxelGrid::getIndices + IndicesHash (src/voxel_grid.h:31-38,68-75) ---------- */
/* keys3: n x 3 int64 (ix,iy,iz); hash22: n uint32, the reference's 22-bit hash; either may be NULL */
int lo_voxel_keys(lo_ctx* ctx, float voxel_size, const float* xyz16, size_t n, int64_t* keys3, uint32_t* hash22);

/* ---- VoxelGrid (src/voxel_grid.h:17-257) ------------------------------------------------------ */
/* VoxelGrid(voxel_size, max_points). capacity_hint = expected number of voxels (0: default); the table
 * grows on demand. The grid keeps, per voxel, the first max_points points in arrival order (:77-93). */
int lo_grid_create(lo_ctx* ctx, float voxel_size, size_t max_points, size_t capacity_hint, lo_grid** out);
void lo_grid_destroy(lo_grid* g);
/* setMaxPoints (:56-59): only the cap changes; voxels keep what they hold */
int lo_grid_set_max_points(lo_grid* g, size_t max_points);
/* setVoxelSize (:61-66): clears the grid */
int lo_grid_set_voxel_size(lo_grid* g, float voxel_size);
/* size() (:248-251): number of voxels */
int lo_grid_size(lo_grid* g, size_t* n_voxels, size_t* n_points);
/* addCloud (:77-93) for pn48 (has_normals=1) / addCloudWithoutNormals (:95-110) for xyz16 (has_normals=0) */
int lo_grid_add_cloud(lo_grid* g, const void* points, size_t n, int has_normals);
int lo_grid_add_cloud_dev(lo_grid* g, const void* d_points, size_t n, int has_normals);
/* getCloud (:112-129) mode 0 -> pn48; getCloudWithoutNormals (:132-146) mode 1 -> xyz16;
 * getSparseCloudWithoutNormals (:149-162) mode 2 -> xyz16 (first point of every voxel).
 * order 0: unspecified (fast); order 1: the reference's tsl::robin_map iteration order, valid for grids
 * that were filled by add_cloud calls only (no radius_cleanup) — the down-sampler use of lidar_odometry.cpp:37-50.
 * *n receives the number of points; LO_E_CAPACITY if cap is too small. out may be NULL to query *n. */
int lo_grid_get_cloud(lo_grid* g, int mode, int order, void* out, size_t cap, size_t* n);
/* The iteration order of the reference's container on its own (third-party/robin-map robin_hash.h, as src/voxel_grid.h:82-87
 * fills it): order_out[d] = index of the key at iteration position d after n DISTINCT keys with the 22-bit hashes hash22[0..n)
 * (IndicesHash, src/voxel_grid.h:31-38) were inserted in index order into an empty map. This is the block-local replay
 * (csrc/lo_robin_block.cuh) the per-scan pipeline runs for keyframe_downsampler.getCloud() (src/lidar_odometry.cpp:38,69) when
 * the order inside a group of equal ideal bucket matters; n <= 524 288 (LO_E_CAPACITY; one block replays the first
 * 4 096 keys in shared memory, the later epochs run on the whole grid, csrc/lo_robin_grid.cuh); LO_E_RANGE for a key set with a run of
 * more than 2 048 occupied buckets (degenerate hashes: lo_grid_get_cloud with order 1 replays any key set, sequentially). */
int lo_robin_iteration_order(lo_ctx* ctx, const uint32_t* hash22, size_t n, uint32_t* order_out);
/* radiusCleanup (:236-246): erase voxels whose FIRST stored point is farther than radius from center */
int lo_grid_radius_cleanup(lo_grid* g, const float center[3], float radius);

/* Snapshot / restore (the reference's VoxelGrid is an implicitly copyable value,
 * src/lidar_odometry.h:82, and has no serialisation). The blob is the table as it lies in HBM (slots, slabs, free
 * list, counters) behind a small header; a restored grid answers every call exactly like the original. `out` may be
 * NULL to query *bytes; LO_E_CAPACITY if cap is too small; restore wants a grid created with the same max_points and
 * returns LO_E_IO for a blob it does not recognise. */
int lo_grid_snapshot(lo_grid* g, void* out, size_t cap, size_t* bytes);
int lo_grid_restore(lo_grid* g, const void* blob, size_t bytes);

typedef struct lo_pair {   /* VoxelGrid::Correspondence (:40-46) */
    double source_point_local[3];
    double plane_origin[3];
    double plane_normal[3];
} lo_pair;
/* findMatchingPairs (:206-234) with the query index kept: pairs[i] / valid[i] for every query i
 * (entries of invalid queries are zero); *n_valid = number of valid correspondences. */
int lo_grid_find_matching_pairs(lo_grid* g, const float* xyz16, size_t m, const float pose7[7], float max_dist,
                                lo_pair* pairs, uint8_t* valid, size_t* n_valid);

/* getCorrespondence (:164-204): nearest stored point among the 27 voxels around `query3` (a point in the grid's frame, no
 * transform) with squared distance below max_correspondence_distance_sq; plane_origin / plane_normal of *out and *valid as
 * the reference fills them, source_point_local zero (findMatchingPairs sets it, :225). */
int lo_grid_get_correspondence(lo_grid* g, const float query3[3], double max_correspondence_distance_sq, lo_pair* out, int* valid);

/* ---- residual + normal equations (src/cloud_matcher.cpp:38-103,134-154) ----------------------- */
typedef struct lo_normal_eq {
    double cost;        /* sum of 0.5*rho(r^2) + prior */
    double g[6];        /* J'r, tangent order: rotation(3), translation(3) */
    double jtj[36];     /* J'J row-major */
    uint64_t n_valid;   /* correspondences found */
    uint64_t candidates; /* stored points visited in the 27 neighbours, summed over queries (roofline bytes) */
    uint64_t occupied;   /* occupied neighbour voxels, summed over queries */
} lo_normal_eq;
/* One fused pass: correspondence search at pose7 (float, as findMatchingPairs sees it) then
 * PointToPlaneErrorAnalytic + HuberLoss(0.15) corrector evaluated at x7d = {qw,qx,qy,qz,tx,ty,tz} (double),
 * reduced to the 6x6 normal equations; prior_b = NormalPrior mean (the guess translation). */
int lo_nn_residual(lo_grid* g, const float* xyz16, size_t m, const float pose7[7], const double x7d[7],
                   const double prior_b[3], float max_dist, lo_normal_eq* out);
int lo_nn_residual_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], const double x7d[7],
                       const double prior_b[3], float max_dist, lo_normal_eq* out_host_or_null);

/* The same fused pass for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). It has no counterpart in the reference.
 * It calls no cudaMalloc / cudaFree and no allocating library routine, copies nothing and does not synchronise: it only
 * enqueues its two kernels on the context's stream. The per-block partial sums go to a fixed 256 KiB device array inside
 * the library (one per GPU that loads it, held until the process exits) instead of the context's growable scratch, so:
 *   - at most LO_NN_STATIC_MAX_QUERIES queries per call, LO_E_CAPACITY beyond that;
 *   - every call on a device shares that array: calls that overlap on the GPU (two streams / contexts) race; calls on
 *     one stream, one after the other, are fine (nothing is carried over from the previous call);
 *   - d_xyz16 and d_out are DEVICE pointers the caller owns; *d_out is valid once the stream has reached that point. */
#define LO_NN_STATIC_MAX_QUERIES (1u << 22)
int lo_nn_residual_static_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], const double x7d[7],
                              const double prior_b[3], float max_dist, lo_normal_eq* d_out);

/* ---- CloudMatcher::align (src/cloud_matcher.h:13-17, src/cloud_matcher.cpp:105-178) ----------- */
typedef struct lo_align_stats {
    int32_t outer_iterations;        /* 1..35 */
    int32_t lm_iterations;           /* recorded LM iterations summed over all outer iterations */
    int32_t valid_pairs;             /* correspondences in the last outer iteration */
    int32_t termination;             /* of the last solve: 0 no_convergence, 1 function tol, 2 parameter tol, 3 gradient tol, 4 failure */
    double final_cost;
    double last_step_norm;
    float pose_log[35][7];           /* pose after each outer iteration's float round-trip */
    int32_t pairs_log[35];
} lo_align_stats;
int lo_align(lo_grid* keyframe, const float* xyz16, size_t m, const float guess7[7], float out7[7],
             lo_align_stats* stats /* nullable */);
int lo_align_dev(lo_grid* keyframe, const void* d_xyz16, size_t m, const float guess7[7], float out7[7],
                 lo_align_stats* stats /* nullable */);

/* ---- key-frame map sharded by voxel-hash range over several GPUs (BASELINE.json config 4) ----
 * The reference is single-process; these phases split CloudMatcher::align (src/cloud_matcher.cpp:105-178) and
 * VoxelGrid::findMatchingPairs (src/voxel_grid.h:206-234) at the two points where ranks must exchange data. The
 * library issues no collective itself: the host places an all-reduce(MIN, u64[m]) after lo_shard_search_dev and an
 * all-reduce(SUM, f64[32]) after lo_shard_evaluate_dev, on the context's stream
 * (lidar_odometry_demo_b200/sharded.py: ShardedCloudMatcher over torch.distributed / NCCL; a C++ caller does the
 * same with two ncclAllReduce calls). */
/* Call on an empty grid: add_cloud then keeps only the voxels whose 22-bit hash falls in this rank's range. */
int lo_grid_set_shard(lo_grid* g, uint32_t rank, uint32_t world);
/* d_packed[i] = (float bits of d2 << 32 | probe << 24 | slot) of query i's nearest stored point among THIS rank's
 * voxels, ~0 if none; d_hit16[i] = that point (x,y,z) + its slab index (device scratch the caller owns, 16 B each). */
int lo_shard_search_dev(lo_grid* g, const void* d_xyz16, size_t m, const float pose7[7], float max_dist, uint64_t* d_packed, void* d_hit16);
/* Sums (21 upper JtJ, 6 Jtr, cost, count; 32 doubles, no prior) over the queries whose reduced packed value equals
 * this rank's local one, evaluated at x7d = {qw,qx,qy,qz,tx,ty,tz}. */
int lo_shard_evaluate_dev(lo_grid* g, const void* d_xyz16, size_t m, const uint64_t* d_packed_local, const uint64_t* d_packed_global,
                          const void* d_hit16, const double x7d[7], double* d_acc32);
/* ceres::Solve's trust-region logic on the host for the sharded path (every rank runs it on identical sums). */
typedef struct lo_lm lo_lm;
lo_lm* lo_lm_create(void);
void lo_lm_destroy(lo_lm* h);
int lo_lm_begin(lo_lm* h, const double x0[7], const double acc32[32], const double prior_b[3]);   /* 1: continue */
int lo_lm_propose(lo_lm* h, double cand7[7]);                                                      /* 1: evaluate cand7 */
int lo_lm_update(lo_lm* h, const double acc32[32]);                                                /* 1: continue */
int lo_lm_result(const lo_lm* h, double x7[7], double* last_step_norm, double* cost, int32_t* termination, int32_t* recorded);

/* ---- the same, device-resident and with the collectives inside the library (C++ / C callers; one process per GPU) ----
 * lo_sharded holds this rank's shard (a lo_grid with lo_grid_set_shard applied), an NCCL communicator and the control
 * block of the solve. lo_sharded_align is CloudMatcher::align (src/cloud_matcher.cpp:105-178) over the union of all
 * ranks' shards; every rank calls it with the same queries and guess and receives the same pose. Per outer iteration:
 * search -> ncclAllReduce(MIN, u64[m]) -> evaluate -> one kernel that sums, exchanges the 32 doubles with the peers
 * through NVLink peer memory and runs the LM step (LO_SHARDED_SUM=nccl: ncclAllReduce(SUM, f64[32]) instead); the host
 * synchronises once per align in the usual case. NCCL is bound at run time (libnccl.so.2, LO_B200_NCCL_LIB overrides),
 * and not at all for world = 1. */
#define LO_SHARDED_ID_BYTES 128
typedef struct lo_sharded lo_sharded;
typedef struct lo_sharded_info {
    int32_t rank, world;
    int32_t fused_sum;        /* 1: exchange 2 runs inside the step kernel over peer memory; 0: ncclAllReduce */
    int32_t nccl_version;     /* ncclGetVersion, 0 for world = 1 */
    uint64_t min_allreduces, sum_allreduces, host_syncs;   /* since creation */
    uint64_t evaluations;     /* cost / Jacobian evaluations of the last align */
    int32_t fused_min;        /* 1: exchange 1 runs inside the search kernel over peer memory; 0: ncclAllReduce(MIN) */
    int32_t reserved;
    uint64_t min_exchanges_fused;   /* exchanges 1 done through peer memory since creation */
} lo_sharded_info;
/* rank 0 makes the id (ncclGetUniqueId) and ships it to the other ranks by whatever channel the job has */
int lo_sharded_unique_id(uint8_t id[LO_SHARDED_ID_BYTES]);
/* collective over all ranks; id may be NULL for world = 1. VoxelGrid(voxel_size, max_points) semantics for the shard. */
int lo_sharded_create(lo_ctx* ctx, const uint8_t id[LO_SHARDED_ID_BYTES], int rank, int world, float voxel_size, size_t max_points,
                      size_t capacity_hint, lo_sharded** out);
void lo_sharded_destroy(lo_sharded* s);
/* this rank's shard: lo_grid_add_cloud(_dev) with the replicated update cloud keeps only the voxels this rank owns */
lo_grid* lo_sharded_grid(lo_sharded* s);
int lo_sharded_info_get(const lo_sharded* s, lo_sharded_info* out);
/* choose exchange 2 at run time: 1 = inside the step kernel over peer memory, 0 = ncclAllReduce. Every rank must make
 * the same choice before the same align; LO_E_ARG if the peers' memory could not be mapped at creation. */
int lo_sharded_set_fused_sum(lo_sharded* s, int enable);
/* the same choice for exchange 1: 1 = every rank folds its packed keys into the query's home rank with atomicMin through
 * peer memory inside the search kernel, one cross-rank barrier, one pull of the winners; 0 = ncclAllReduce(MIN, u64[m]).
 * The peer-memory form takes at most LO_SHARDED_MAX_QUERIES (env, default 2^20) queries per align. */
int lo_sharded_set_fused_min(lo_sharded* s, int enable);
/* The shards of ONE process, without NCCL: `world` contexts, each with a stream of its own, on different GPUs of the box or
 * several on one GPU (how the one-GPU test rig exercises the exchanges). out[world] receives the shards in rank order.
 * They are driven together: lo_sharded_align_group enqueues every shard's schedule before it waits for any (a shard's
 * kernels wait on the device for its peers). d_xyz16[k] = the replicated queries on shard k's device; out7s = world x 7. */
int lo_sharded_create_group(lo_ctx* const* ctxs, int world, float voxel_size, size_t max_points, size_t capacity_hint, lo_sharded** out);
int lo_sharded_align_group(lo_sharded* const* shards, int world, const void* const* d_xyz16, size_t m, const float guess7[7],
                           float max_dist, float* out7s, lo_align_stats* stats /* nullable */);
int lo_sharded_align_group_host(lo_sharded* const* shards, int world, const float* xyz16, size_t m, const float guess7[7],
                                float max_dist, float* out7s, lo_align_stats* stats /* nullable */);
int lo_sharded_align(lo_sharded* s, const float* xyz16, size_t m, const float guess7[7], float max_dist, float out7[7],
                     lo_align_stats* stats /* nullable */);
int lo_sharded_align_dev(lo_sharded* s, const void* d_xyz16, size_t m, const float guess7[7], float max_dist, float out7[7],
                         lo_align_stats* stats /* nullable */);

/* ---- CloudTransformer / utils (src/utils) ------------------------------------------------------ */
/* pointTimeNormalize (point_time_normalize.h:14-39) followed by transformNonRigid
 * (cloud_transform.h:15-40) with (start7, end7); out may alias in. */
int lo_deskew(lo_ctx* ctx, const void* xyzirt32_in, size_t n, const float start7[7], const float end7[7],
              void* xyzirt32_out);
/* utils::pointTimeNormalize alone (point_time_normalize.h:14-39); out may alias in */
int lo_point_time_normalize(lo_ctx* ctx, const void* xyzirt32_in, size_t n, void* xyzirt32_out);
/* CloudTransformer::transformNonRigid alone (cloud_transform.h:15-40): `time` is used as it is */
int lo_transform_non_rigid(lo_ctx* ctx, const void* xyzirt32_in, size_t n, const float start7[7], const float end7[7],
                           void* xyzirt32_out);
/* transform<PointXYZ> (cloud_transform.h:42-66) */
int lo_transform_xyz(lo_ctx* ctx, const float* xyz16_in, size_t n, const float pose7[7], float* xyz16_out);
/* transformWithNormals (cloud_transform.h:68-97) */
int lo_transform_with_normals(lo_ctx* ctx, const float* pn48_in, size_t n, const float pose7[7], float* pn48_out);
/* rangeFilter (range_filter.h:12-28), order preserving; *n_out = points kept */
int lo_range_filter(lo_ctx* ctx, const float* pn48_in, size_t n, float min_range, float max_range, float* pn48_out,
                    size_t* n_out);
/* CloudClassifier::classify (cloud_classifier.h:19-168): planar points with normals, reference order */
int lo_classify(lo_ctx* ctx, const void* xyzirt32_in, size_t n, float* pn48_planar_out, size_t cap, size_t* n_planar);
/* the pair classify returns (cloud_classifier.h:19,167): {planar points with normals, unclassified points}. The second
 * cloud holds the organised points that were looked at and did not become planar (intensity = curvature, :104);
 * at most n of each. LO_E_CAPACITY if the ring x azimuth grid does not fit the scratch or an out buffer is short. */
int lo_classify_pair(lo_ctx* ctx, const void* xyzirt32_in, size_t n, float* pn48_planar_out, size_t cap, size_t* n_planar,
                     void* xyzirt32_unclassified_out, size_t cap_unclassified, size_t* n_unclassified);

/* ---- per-point PCA normals: the test's pcl::NormalEstimation (test/test.cpp:196-205) ---------- */
int lo_pca_normals(lo_ctx* ctx, const float* xyz16, size_t n, float radius, float* normals3_out);

/* ---- LidarOdometry (src/lidar_odometry.h:18-86, src/lidar_odometry.cpp:14-89) ----------------- */
typedef struct lo_params {               /* LidarOdometry::Params (lidar_odometry.h:25-32), config/params.yaml:4-11 */
    float lidar_min_range;
    float lidar_max_range;
    float keyframe_voxel_size;
    uint64_t keyframe_max_points_cnt;
    float keyframe_matching_voxel_size;
    float keyframe_update_voxel_size;
    float keyframe_cleanup_range;
    float angular_divergence_threshold;
} lo_params;
void lo_params_default(lo_params* p);                      /* GetROSDeclaration defaults (lidar_odometry.h:36-48) */
int lo_params_from_yaml(const char* path, lo_params* p);  /* reads config/params.yaml's lidar_odometry block */

#define LO_ORDER_SORTED 0
#define LO_ORDER_REPLAYED 1
#define LO_ORDER_APPROXIMATE 2
typedef struct lo_scan_stats {
    uint32_t n_points, n_planar, n_update, n_query;  /* N, P, U, M of SURVEY.md section 3.1 */
    int32_t outer_iterations;
    int32_t unstable;            /* 1 if the divergence gate fell back to the prediction */
    uint32_t keyframe_voxels;
    uint32_t update_order;       /* how the order in which the update cloud entered the key frame (tsl::robin_map iteration order,
                                    lidar_odometry.cpp:38,42,69) was obtained. LO_ORDER_SORTED: counting sort by ideal bucket; no
                                    group of equal ideal bucket had two points in one key-frame voxel, so the key frame is
                                    exactly what the reference's order gives. LO_ORDER_REPLAYED: such a pair existed and the
                                    map's insertion history was replayed: the exact order. LO_ORDER_APPROXIMATE: such a pair
                                    existed but the replay refused the key set (a run of more than 2 048 consecutive occupied
                                    buckets: degenerate hashes): groups of equal ideal bucket went in in arrival order */
} lo_scan_stats;

/* LidarOdometry(const Params&) (lidar_odometry.cpp:14-20). All device memory a scan needs is allocated here, once: the
 * key-frame table and slabs for 2^19 voxels (env LO_B200_KEYFRAME_VOXELS overrides) and the per-scan scratch for scans of
 * up to 32 768 points, so that no process / submit call allocates or synchronises the device. A key frame that outgrows
 * its capacity, or a larger scan, is accommodated by a stream-ordered re-allocation (no host synchronisation). */
int lo_odometry_create(lo_ctx* ctx, const lo_params* params, lo_odometry** out);
/* the same with explicit sizes: keyframe_voxels = expected voxels of the key frame inside keyframe_cleanup_range,
 * scan_points = largest scan expected (each 0: the default). Config 3 of BASELINE.json (128 beams, 0.1 m voxels) wants
 * e.g. (1 << 21, 1 << 18). */
int lo_odometry_create_sized(lo_ctx* ctx, const lo_params* params, size_t keyframe_voxels, size_t scan_points, lo_odometry** out);
void lo_odometry_destroy(lo_odometry* h);
/* processCloud (lidar_odometry.cpp:22-76). The host-pointer form takes ordinary (pageable) memory, as the reference's
 * caller has it, and stages it through a pinned buffer of the handle. Both return as soon as the scan's pose is final
 * (after the divergence gate, cpp:53-65); radiusCleanup and the key-frame insert (cpp:67-70) of the scan finish on the
 * device behind the call, in front of whatever is queued next. Calls that look at the key frame (lo_odometry_stats,
 * _keyframe*, _snapshot) wait for that tail; an LO_E_RANGE raised by the insert is reported by the call that does. */
int lo_odometry_process(lo_odometry* h, const void* xyzirt32, size_t n);
int lo_odometry_process_dev(lo_odometry* h, const void* d_xyzirt32, size_t n);
/* processCloud on `xyzirt32` while the NEXT scan (`next_xyzirt32`, nullable; pinned host memory for a real overlap) is
 * uploaded on a copy stream behind it; a later call that passes that same pointer and count as its current scan
 * finds it in HBM. The caller must not modify the next scan's buffer until that call. Same results as
 * lo_odometry_process. */
int lo_odometry_process_next(lo_odometry* h, const void* xyzirt32, size_t n, const void* next_xyzirt32, size_t next_n);
/* processCloud split in two for recorded streams: lo_odometry_submit queues a scan (its upload and its kernel run behind
 * the scan in flight, so the GPU goes from one scan to the next without waiting for the host) and returns at once;
 * lo_odometry_wait blocks until the OLDEST queued scan is done, after which lo_odometry_pose / _stats describe that
 * scan. At most LO_ODOMETRY_QUEUE_DEPTH scans in flight (LO_E_BUSY otherwise); a host buffer must stay untouched until the wait that
 * matches its submit; scans of more than 524 288 points are not accepted here. submit + wait == lo_odometry_process:
 * same poses, same key frame. */
#define LO_ODOMETRY_QUEUE_DEPTH 8
int lo_odometry_submit(lo_odometry* h, const void* xyzirt32, size_t n);
int lo_odometry_submit_dev(lo_odometry* h, const void* d_xyzirt32, size_t n);
int lo_odometry_wait(lo_odometry* h);
/* same pipeline entered after CloudClassifier::classify (lidar_odometry.cpp:35 onwards) with a planar cloud */
int lo_odometry_process_planar(lo_odometry* h, const float* pn48_planar, size_t n);
int lo_odometry_pose(lo_odometry* h, float out7[7]);                       /* getCurrentPose */
/* size and allocation history of keyframe_ (src/lidar_odometry.h:82): what bench.py prints next to its timings so that a
 * slow step can be attributed. Waits for the tail of the last scan. */
typedef struct lo_keyframe_info {
    uint64_t voxels, points, tombstones;        /* after the last scan's radiusCleanup + insert */
    uint64_t capacity_slots, capacity_voxels;   /* table slots / voxel slabs allocated */
    uint64_t table_rebuilds, payload_growths;   /* stream-ordered re-allocations since creation (0 in a sized run) */
} lo_keyframe_info;
int lo_odometry_keyframe_info(lo_odometry* h, lo_keyframe_info* out);
/* Host threads lo_odometry_process uses for a scan in pageable memory: the caller plus the helpers of the staging copy into
 * pinned memory (csrc/lo_copy_pool.h; up to 3, env LO_B200_COPY_THREADS overrides, 0 = none). The helpers spin for 1 ms after
 * a copy and sleep otherwise. No counterpart in the reference (its processCloud takes the cloud where it lies). */
int lo_odometry_host_threads(const lo_odometry* h);
int lo_odometry_stats(lo_odometry* h, lo_scan_stats* out);
/* getKeyFrameCloud (full=0, sparse) / getFullKeyFrameCloud (full=1): xyz16 */
int lo_odometry_keyframe(lo_odometry* h, int full, float* xyz16_out, size_t cap, size_t* n);
int lo_odometry_keyframe_pn(lo_odometry* h, float* pn48_out, size_t cap, size_t* n);
int lo_odometry_deskewed(lo_odometry* h, void* xyzirt32_out, size_t cap, size_t* n);  /* getTempCloud */

/* Checkpoint of a running LidarOdometry: previous_transform_ / current_transform_ / keyframe_
 * (src/lidar_odometry.h:82-85) + the last scan's stats. A fresh lo_odometry created with the same Params and restored
 * from the blob continues with bit-identical poses. No scan may be in flight (LO_E_BUSY). */
int lo_odometry_snapshot(lo_odometry* h, void* out, size_t cap, size_t* bytes);
int lo_odometry_restore(lo_odometry* h, const void* blob, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* LO_B200_H */
