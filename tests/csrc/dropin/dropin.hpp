// TEST INFRASTRUCTURE. Force-included (-include) in front of the reference's OWN, UNMODIFIED src/lidar_odometry.cpp, which is
// compiled from where it lies under /root/reference (never copied). It pre-defines the include guards of the reference's
// headers, so that their #include lines become empty, and supplies the same names from the GPU facade
// (lidar_odometry_demo_b200/host/lo_b200.hpp): Pose3D, VoxelGrid, CloudMatcher, CloudTransformer, CloudClassifier,
// utils::pointTimeNormalize / rangeFilter. The only declaration restated here is the LidarOdometry class itself
// (/root/reference/src/lidar_odometry.h:18-86: its public interface and five data members), because the translation
// unit under test defines that class's member functions.
//
// Built with the oracle's Eigen / PCL shims standing in for the real libraries (-DLO_B200_WITH_EIGEN -DLO_B200_WITH_PCL
// -Ioracle/shim): the reference source uses Eigen expressions (eulerAngles, setZero, ...) on the pose members.
#pragma once

// guards of /root/reference/src/*.h
#define BUILD_LIDARODOMETRY_H
#define BUILD_CLOUDMATCHER_H
#define BUILD_LIDAR_POINT_TYPE_H
#define BUILD_POSE3D_H
#define BUILD_VOXELGRID_H
#define BUILD_VOXEL_WITH_PLANES_H
#define BUILD_CLOUD_CLASSIFIER_H
#define BUILD_CLOUD_TRANSFORM_H
#define BUILD_POINT_TIME_NORMALIZE_H
#define BUILD_RANGEFILTER_H

#include <chrono>
#include <cmath>
#include <iostream>
#include <numbers>

// the facade's own LidarOdometry (a thin wrapper of lo_odometry_process) keeps out of the way of the reference's class
#define LidarOdometry LoB200LidarOdometry
#include "../../../lidar_odometry_demo_b200/host/lo_b200.hpp"
#undef LidarOdometry

class LidarOdometry {   // src/lidar_odometry.h:18-86
public:
    using CloudType = pcl::PointCloud<lidar_point::PointXYZIRT>;
    using Params = LoB200LidarOdometry::Params;

    explicit LidarOdometry(const Params& config);
    void processCloud(const CloudType& input_cloud);
    pcl::PointCloud<pcl::PointXYZ>::Ptr getKeyFrameCloud() const;
    pcl::PointCloud<pcl::PointXYZ>::Ptr getFullKeyFrameCloud() const;
    Pose3D getCurrentPose() const;
    auto getTempCloud() const { return temp_cloud_; }

private:
    pcl::PointCloud<lidar_point::PointXYZIRT>::Ptr temp_cloud_;
    const Params config_;
    VoxelGrid keyframe_;
    Pose3D previous_transform_;
    Pose3D current_transform_;
};
