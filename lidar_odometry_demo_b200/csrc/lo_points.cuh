// Device-side entry points of the per-point stages (K1, rigid transforms, range filter, K8 classifier),
// callable with device pointers from the odometry pipeline (lo_odometry.cu) and from the host-pointer
// C ABI wrappers (lo_points.cu).
#pragma once

#include "lo_host.h"

namespace lo {

// pointTimeNormalize + transformNonRigid(start, end) (K1). start/end are DEVICE pointers to Pose.
int deskew_dev(lo_ctx* ctx, const PointXYZIRT* d_in, uint32_t n, const Pose* d_start, const Pose* d_end, PointXYZIRT* d_out);
int transform_xyz_dev(lo_ctx* ctx, const PointXYZ* d_in, uint32_t n, const Pose* d_pose, PointXYZ* d_out);
int transform_pn_dev(lo_ctx* ctx, const PointNormal* d_in, uint32_t n, const Pose* d_pose, PointNormal* d_out);
// order-preserving rangeFilter; *d_count receives the number kept
int range_filter_dev(lo_ctx* ctx, const PointNormal* d_in, uint32_t n, float min_range, float max_range, PointNormal* d_out,
                     uint32_t* d_count);
// CloudClassifier::classify; planar points (reference order) that also pass [min_range,max_range] when
// filter != 0 (the pipeline fuses classify + rangeFilter). d_out must hold n records.
// d_unclassified / d_count_unclassified (nullable): the second element of classify's pair; d_overflow (nullable): set to 1 when
// the ring x azimuth grid does not fit the scratch (no output is produced then)
int classify_dev(lo_ctx* ctx, const PointXYZIRT* d_in, uint32_t n, int filter, float min_range, float max_range, PointNormal* d_out,
                 uint32_t* d_count, PointXYZIRT* d_unclassified = nullptr, uint32_t* d_count_unclassified = nullptr, uint32_t* d_overflow = nullptr);

}  // namespace lo
