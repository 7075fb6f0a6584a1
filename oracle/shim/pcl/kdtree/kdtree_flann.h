// TEST INFRASTRUCTURE (oracle/): included but unused by the reference's cloud_classifier.h:9.
#ifndef LO_ORACLE_SHIM_PCL_KDTREE_H
#define LO_ORACLE_SHIM_PCL_KDTREE_H
#endif
