// TEST INFRASTRUCTURE (oracle/): the reference's voxel_grid.h includes <pcl/io/pcd_io.h> only to
// pull in the point/cloud types; no PCD function is used on the hot path.
#ifndef LO_ORACLE_SHIM_PCL_PCD_IO_H
#define LO_ORACLE_SHIM_PCL_PCD_IO_H
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#endif
