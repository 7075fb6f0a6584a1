// TEST INFRASTRUCTURE (oracle/): plain wire structs shared by both oracle builds.
// Layouts follow PCL's documented point types and /root/reference/src/lidar_point_type.h:13-20.
#ifndef LO_ORACLE_TYPES_H
#define LO_ORACLE_TYPES_H
#include <cstdint>

struct alignas(16) OrcXYZ {
    float x, y, z, pad;
};
struct alignas(16) OrcPN {
    float x, y, z, pad0;
    float nx, ny, nz, pad1;
    float curvature, pad2[3];
};
struct alignas(16) OrcXYZIRT {
    float x, y, z, pad0;
    float intensity;
    std::uint16_t ring;
    std::uint16_t pad1;
    float time;
    float pad2;
};
static_assert(sizeof(OrcXYZ) == 16 && sizeof(OrcPN) == 48 && sizeof(OrcXYZIRT) == 32, "wire layout");

#endif
