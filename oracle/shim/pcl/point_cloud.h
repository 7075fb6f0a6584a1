// TEST INFRASTRUCTURE (oracle/): minimal pcl::PointCloud<T> stand-in (PCL is absent).
#ifndef LO_ORACLE_SHIM_PCL_POINT_CLOUD_H
#define LO_ORACLE_SHIM_PCL_POINT_CLOUD_H

#include <cstddef>
#include <cstdint>
#include <map>
#include <memory>
#include <numbers>
#include <vector>
#include <cmath>

#include <pcl/point_types.h>

namespace pcl {

template <typename PointT>
class PointCloud {
public:
    using PointType = PointT;
    using VectorType = std::vector<PointT>;
    using Ptr = std::shared_ptr<PointCloud<PointT>>;
    using ConstPtr = std::shared_ptr<const PointCloud<PointT>>;
    using iterator = typename VectorType::iterator;
    using const_iterator = typename VectorType::const_iterator;

    VectorType points;
    std::uint32_t width = 0;
    std::uint32_t height = 0;
    bool is_dense = true;

    std::size_t size() const { return points.size(); }
    bool empty() const { return points.empty(); }
    PointT& at(std::size_t i) { return points.at(i); }
    const PointT& at(std::size_t i) const { return points.at(i); }
    PointT& operator[](std::size_t i) { return points[i]; }
    const PointT& operator[](std::size_t i) const { return points[i]; }
    iterator begin() { return points.begin(); }
    iterator end() { return points.end(); }
    const_iterator begin() const { return points.begin(); }
    const_iterator end() const { return points.end(); }
    const_iterator cbegin() const { return points.cbegin(); }
    const_iterator cend() const { return points.cend(); }
    void push_back(const PointT& p) { points.push_back(p); }
    void clear() { points.clear(); width = height = 0; }
};

}  // namespace pcl

#endif
