/* TEST INFRASTRUCTURE (oracle/). C API shared by the two CPU checkers:
 *   oracle/liblo_oracle.so      "port": a from-scratch CPU restatement (oracle/lo_oracle.cpp), builds anywhere
 *   oracle/_ref/liblo_ref.so    "reference": the reference's own unmodified headers compiled from
 *                               /root/reference against oracle/shim + the vendored tsl::robin_map
 *                               (oracle/ref_build/ref_api.cpp); only Ceres is restated (mini_ceres.h)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs load these.
 *
 * Layouts: xyz16 = pcl::PointXYZ (16 B: x,y,z,pad); pn48 = pcl::PointNormal (48 B: x,y,z,pad,nx,ny,nz,pad,
 * curvature,pad3); xyzirt32 = lidar_point::PointXYZIRT (32 B: x,y,z,pad,intensity,ring(u16)@20,time@24).
 * pose7 = {tx,ty,tz,qw,qx,qy,qz} float.
 */
#ifndef LO_ORACLE_API_H
#define LO_ORACLE_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* orc_kind(void); /* "reference" or "port" */

/* ---- VoxelGrid (src/voxel_grid.h) ---- */
void* orc_grid_create(float voxel_size, size_t max_points);
void* orc_grid_clone(const void* g);
void orc_grid_destroy(void* g);
void orc_grid_add_cloud(void* g, const float* pn48, size_t n);
void orc_grid_add_cloud_xyz(void* g, const float* xyz16, size_t n);
size_t orc_grid_size(const void* g);
size_t orc_grid_num_points(const void* g);
size_t orc_grid_get_cloud(const void* g, float* pn48_out, size_t cap);          /* robin iteration order */
size_t orc_grid_get_cloud_xyz(const void* g, float* xyz16_out, size_t cap);
size_t orc_grid_get_sparse_xyz(const void* g, float* xyz16_out, size_t cap);
void orc_grid_radius_cleanup(void* g, const float center[3], float radius);
void orc_get_indices(float voxel_size, const float* xyz, size_t stride_floats, size_t n, int64_t* out3);
uint64_t orc_hash(int64_t ix, int64_t iy, int64_t iz);
/* findMatchingPairs in serial (input) order. out9[i*9..] = src(3), origin(3), normal(3) for every query,
 * valid[i] = 1 if a correspondence was found (entries of invalid queries are zero). Returns #valid. */
size_t orc_grid_find_matching_pairs(const void* g, const float* xyz16, size_t n, const float pose7[7],
                                    float max_dist, double* out9, uint8_t* valid);

void orc_grid_set_max_points(void* g, size_t max_points);   /* setMaxPoints (voxel_grid.h:56-59) */
/* getCorrespondence (voxel_grid.h:164-204): 1 if valid; origin / normal of the nearest stored point */
int orc_grid_get_correspondence(const void* g, const float query3[3], double max_correspondence_distance_sq, double origin3[3],
                                double normal3[3]);

/* ---- point utilities (src/utils headers) ---- */
void orc_time_normalize(const void* xyzirt32_in, size_t n, void* xyzirt32_out);
void orc_transform_nonrigid(const void* xyzirt32_in, size_t n, const float start7[7], const float end7[7],
                            void* xyzirt32_out);
void orc_transform_xyz(const float* xyz16_in, size_t n, const float pose7[7], float* xyz16_out);
void orc_transform_with_normals(const float* pn48_in, size_t n, const float pose7[7], float* pn48_out);
size_t orc_range_filter(const float* pn48_in, size_t n, float min_range, float max_range, float* pn48_out);
/* CloudClassifier::classify: planar points with normals; returns #planar, *n_unclassified optional */
size_t orc_classify(const void* xyzirt32_in, size_t n, float* pn48_planar_out, size_t* n_unclassified);
/* the same with the second element of the returned pair (cloud_classifier.h:167): xyzirt32 records, capacity n */
size_t orc_classify_full(const void* xyzirt32_in, size_t n, float* pn48_planar_out, void* xyzirt32_unclassified_out, size_t* n_unclassified);

/* ---- Pose3D (src/pose_3d.h) ---- */
void orc_pose_compose(const float a7[7], const float b7[7], float out7[7]);
void orc_pose_inverse(const float a7[7], float out7[7]);
void orc_pose_relative_to(const float a7[7], const float target7[7], float out7[7]);
void orc_pose_rotation_matrix(const float a7[7], float r9_rowmajor[9]);

/* ---- residual / solver (src/cloud_matcher.cpp + restated Ceres) ---- */
/* Evaluate the Ceres program at x7d = {qw,qx,qy,qz,tx,ty,tz} (double) on n pairs (pairs9 as above):
 * cost, gradient g[6] = J'r and JtJ[36] (row-major) of the corrected, unscaled tangent-space Jacobian,
 * prior included. residuals_out (n, corrected) and jac_out (n*6) are optional. */
void orc_evaluate(const double* pairs9, size_t n, const double x7d[7], const double prior_b[3], double* cost,
                  double g6[6], double jtj36[36], double* residuals_out, double* jac_out);
/* One ceres::Solve on fixed pairs. x7d in/out. log: per recorded iteration 8 doubles
 * {iteration, cost, step_norm, relative_decrease, radius, valid, successful, gradient_max_norm};
 * returns number of recorded iterations; *termination as mini_ceres::Termination. */
int orc_solve(const double* pairs9, size_t n, const double prior_b[3], double x7d[7], double* log8,
              int log_cap, int* termination);

#define ORC_ALIGN_LOG_STRIDE 16
/* CloudMatcher::align. log (optional, 35*ORC_ALIGN_LOG_STRIDE doubles), per outer iteration:
 * {0:num_pairs, 1:lm_iterations_recorded, 2:last_step_norm, 3:final_cost, 4:termination,
 *  5..11: pose7 after the float round-trip, 12: initial_cost, 13: recorded LM iterations with a rejected step,
 *  14: ... with an invalid step, 15: trust-region radius after the last recorded iteration}. Returns outer iterations run. */
int orc_align(const void* grid, const float* xyz16, size_t n, const float guess7[7], float out7[7], double* log);

/* ---- PCA normals: the test's pcl::NormalEstimation radius search (test/test.cpp:196-205), restated ----
 * normals_out: n x 3 floats (NaN where fewer than 3 neighbours). PARITY UNPINNED (PCL absent). */
void orc_pca_normals(const float* xyz16, size_t n, float radius, float* normals3_out);

/* ---- LidarOdometry (src/lidar_odometry.{h,cpp}) ---- */
typedef struct orc_params {
    float lidar_min_range;
    float lidar_max_range;
    float keyframe_voxel_size;
    uint64_t keyframe_max_points_cnt;
    float keyframe_matching_voxel_size;
    float keyframe_update_voxel_size;
    float keyframe_cleanup_range;
    float angular_divergence_threshold;
} orc_params;
void orc_params_default(orc_params* p);
void* orc_odom_create(const orc_params* p);
void orc_odom_destroy(void* h);
/* counts4 (optional): {N, P(planar in range), U(update set), M(query set)}; stats (optional, 4 doubles):
 * {outer_iterations, unstable(0/1), ms_total, ms_align}. use_classifier=1 runs CloudClassifier; 0 expects the
 * caller to use orc_odom_process_planar instead. */
void orc_odom_process(void* h, const void* xyzirt32, size_t n, uint64_t counts4[4], double stats4[4]);
void orc_odom_process_planar(void* h, const float* pn48_planar, size_t n, uint64_t counts4[4], double stats4[4]);
void orc_odom_pose(const void* h, float out7[7]);
size_t orc_odom_keyframe(const void* h, int full, float* xyz16_out, size_t cap);
size_t orc_odom_keyframe_pn(const void* h, float* pn48_out, size_t cap);
size_t orc_odom_deskewed(const void* h, void* xyzirt32_out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
