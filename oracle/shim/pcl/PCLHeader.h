// TEST INFRASTRUCTURE (oracle/): placeholder for <pcl/PCLHeader.h> (lidar_point_type.h:8).
#ifndef LO_ORACLE_SHIM_PCL_HEADER_H
#define LO_ORACLE_SHIM_PCL_HEADER_H
#endif
