// TEST INFRASTRUCTURE (oracle/): minimal stand-in for the PCL point types the reference's headers use
// (PCL is not installed in this image). Field layout follows PCL 1.12's documented
// structs: PointXYZ 16 B {x,y,z,pad}; PointNormal 48 B {x,y,z,pad, nx,ny,nz,pad, curvature,pad[3]}.
#ifndef LO_ORACLE_SHIM_PCL_POINT_TYPES_H
#define LO_ORACLE_SHIM_PCL_POINT_TYPES_H

#include <cstdint>
#include <eigen3/Eigen/Eigen>

#define PCL_ADD_POINT4D \
    union {             \
        float data[4];  \
        struct {        \
            float x;    \
            float y;    \
            float z;    \
        };              \
    };

#define POINT_CLOUD_REGISTER_POINT_STRUCT(name, fseq)

namespace pcl {

struct alignas(16) PointXYZ {
    PCL_ADD_POINT4D
    PointXYZ() : data{0.f, 0.f, 0.f, 1.f} {}
    PointXYZ(float x_, float y_, float z_) : data{x_, y_, z_, 1.f} {}
};

struct alignas(16) Normal {
    union {
        float data_n[4];
        float normal[3];
        struct {
            float normal_x;
            float normal_y;
            float normal_z;
        };
    };
    union {
        struct {
            float curvature;
        };
        float data_c[4];
    };
    Normal() : data_n{0.f, 0.f, 0.f, 0.f}, data_c{0.f, 0.f, 0.f, 0.f} {}
};

struct alignas(16) PointNormal {
    PCL_ADD_POINT4D
    union {
        float data_n[4];
        float normal[3];
        struct {
            float normal_x;
            float normal_y;
            float normal_z;
        };
    };
    union {
        struct {
            float curvature;
        };
        float data_c[4];
    };
    PointNormal() : data{0.f, 0.f, 0.f, 1.f}, data_n{0.f, 0.f, 0.f, 0.f}, data_c{0.f, 0.f, 0.f, 0.f} {}
    PointNormal(float x_, float y_, float z_) : data{x_, y_, z_, 1.f}, data_n{0.f, 0.f, 0.f, 0.f}, data_c{0.f, 0.f, 0.f, 0.f} {}
};

static_assert(sizeof(PointXYZ) == 16, "PointXYZ ABI");
static_assert(sizeof(PointNormal) == 48, "PointNormal ABI");

}  // namespace pcl

#endif
