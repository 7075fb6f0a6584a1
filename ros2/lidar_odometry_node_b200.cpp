// ROS 2 adapter over the C++ facade: the topics, frames and parameters of the reference's node
// (src/lidar_odometry_node.cpp:9-130, config/params.yaml) with liblo_b200 underneath.
//
// NOT BUILT OR TESTED ON THIS BOX: there is no ROS 2 installation (rclcpp, sensor_msgs, nav_msgs, tf2_ros) in the image
// or on the GPU machines. Everything below the message conversion is the code path of apps/lo_replay.cpp, which is
// tested (tests/test_gpu_replay.py). Build recipe: ros2/CMakeLists.txt.
//
//   in   /lidar_points     sensor_msgs/PointCloud2 with x y z intensity ring time (src/lidar_point_type.h:24-31)
//   out  /odometry         nav_msgs/Odometry        pose of every scan
//        TF                output_frame_parent -> output_frame_child
//        /keyframe_cloud   sparse or full key frame (publish_full_keyframe), only while somebody listens
//        /deskewed_cloud   the deskewed scan, only while somebody listens
// PCL is not needed: PointCloud2 fields are located by name and copied into the 32-byte wire records.
#include <nav_msgs/msg/odometry.hpp>
#include <rclcpp/rclcpp.hpp>
#include <sensor_msgs/msg/point_cloud2.hpp>
#include <tf2_ros/transform_broadcaster.h>

#include <cstring>
#include <memory>
#include <string>

#include <lo_b200.hpp>

namespace {

using sensor_msgs::msg::PointCloud2;
using sensor_msgs::msg::PointField;

const PointField* find_field(const PointCloud2& msg, const char* name) {
    for (const auto& f : msg.fields)
        if (f.name == name) return &f;
    return nullptr;
}

double read_scalar(const uint8_t* p, uint8_t datatype) {
    switch (datatype) {
        case PointField::INT8: { int8_t v; std::memcpy(&v, p, 1); return v; }
        case PointField::UINT8: { uint8_t v; std::memcpy(&v, p, 1); return v; }
        case PointField::INT16: { int16_t v; std::memcpy(&v, p, 2); return v; }
        case PointField::UINT16: { uint16_t v; std::memcpy(&v, p, 2); return v; }
        case PointField::INT32: { int32_t v; std::memcpy(&v, p, 4); return v; }
        case PointField::UINT32: { uint32_t v; std::memcpy(&v, p, 4); return v; }
        case PointField::FLOAT32: { float v; std::memcpy(&v, p, 4); return v; }
        case PointField::FLOAT64: { double v; std::memcpy(&v, p, 8); return v; }
    }
    return 0.0;
}

// PointCloud2 -> PointXYZIRT records; false if a field the pipeline needs is missing
bool to_scan(const PointCloud2& msg, LidarOdometry::CloudType& scan) {
    const PointField *fx = find_field(msg, "x"), *fy = find_field(msg, "y"), *fz = find_field(msg, "z");
    const PointField *fi = find_field(msg, "intensity"), *fr = find_field(msg, "ring"), *ft = find_field(msg, "time");
    if (!ft) ft = find_field(msg, "t");
    if (!fx || !fy || !fz || !fr || !ft) return false;
    const size_t n = static_cast<size_t>(msg.width) * msg.height;
    scan.points.resize(n);
    scan.width = static_cast<uint32_t>(n);
    scan.height = 1;
    for (size_t i = 0; i < n; i++) {
        const uint8_t* rec = msg.data.data() + i * msg.point_step;
        lidar_point::PointXYZIRT p;
        p.x = static_cast<float>(read_scalar(rec + fx->offset, fx->datatype));
        p.y = static_cast<float>(read_scalar(rec + fy->offset, fy->datatype));
        p.z = static_cast<float>(read_scalar(rec + fz->offset, fz->datatype));
        if (fi) p.intensity = static_cast<float>(read_scalar(rec + fi->offset, fi->datatype));
        p.ring = static_cast<uint16_t>(read_scalar(rec + fr->offset, fr->datatype));
        p.time = static_cast<float>(read_scalar(rec + ft->offset, ft->datatype));
        scan.points[i] = p;
    }
    return true;
}

PointField make_field(const std::string& name, uint32_t offset, uint8_t datatype) {
    PointField f;
    f.name = name;
    f.offset = offset;
    f.datatype = datatype;
    f.count = 1;
    return f;
}

template <typename PointT>
PointCloud2 to_msg(const pcl::PointCloud<PointT>& cloud, const std_msgs::msg::Header& header) {
    PointCloud2 msg;
    msg.header = header;
    msg.height = 1;
    msg.width = static_cast<uint32_t>(cloud.size());
    msg.is_bigendian = false;
    msg.is_dense = false;
    msg.point_step = sizeof(PointT);
    msg.row_step = msg.point_step * msg.width;
    msg.fields = {make_field("x", 0, PointField::FLOAT32), make_field("y", 4, PointField::FLOAT32), make_field("z", 8, PointField::FLOAT32)};
    if constexpr (sizeof(PointT) == sizeof(lidar_point::PointXYZIRT)) {
        msg.fields.push_back(make_field("intensity", 16, PointField::FLOAT32));
        msg.fields.push_back(make_field("ring", 20, PointField::UINT16));
        msg.fields.push_back(make_field("time", 24, PointField::FLOAT32));
    }
    msg.data.resize(static_cast<size_t>(msg.row_step));
    if (!cloud.points.empty()) std::memcpy(msg.data.data(), cloud.points.data(), msg.data.size());   // records go out unchanged
    return msg;
}

}  // namespace

class LidarOdometryNodeB200 : public rclcpp::Node {
public:
    LidarOdometryNodeB200() : Node("lidar_odometry_node") {
        // same names and defaults as LidarOdometry::Params::GetROSDeclaration (src/lidar_odometry.h:36-48)
        LidarOdometry::Params p;
        p.lidar_min_range = static_cast<float>(declare_parameter("lidar_odometry.lidar_min_range", 4.0));
        p.lidar_max_range = static_cast<float>(declare_parameter("lidar_odometry.lidar_max_range", 80.0));
        p.keyframe_voxel_size = static_cast<float>(declare_parameter("lidar_odometry.keyframe_voxel_size", 0.2));
        p.keyframe_max_points_cnt = static_cast<size_t>(declare_parameter("lidar_odometry.keyframe_max_points_cnt", 20));
        p.keyframe_matching_voxel_size = static_cast<float>(declare_parameter("lidar_odometry.keyframe_matching_voxel_size", 0.3));
        p.keyframe_update_voxel_size = static_cast<float>(declare_parameter("lidar_odometry.keyframe_update_voxel_size", 0.1));
        p.keyframe_cleanup_range = static_cast<float>(declare_parameter("lidar_odometry.keyframe_cleanup_range", 80.0));
        p.angular_divergence_threshold = static_cast<float>(declare_parameter("lidar_odometry.angular_divergence_threshold", 5.0));
        parent_frame_ = declare_parameter("output_frame_parent", std::string("odom"));
        child_frame_ = declare_parameter("output_frame_child", std::string("base_scan"));
        full_keyframe_ = declare_parameter("publish_full_keyframe", false);
        odometry_ = std::make_unique<LidarOdometry>(p);

        keyframe_pub_ = create_publisher<PointCloud2>("/keyframe_cloud", rclcpp::QoS(1).keep_last(1));
        deskewed_pub_ = create_publisher<PointCloud2>("/deskewed_cloud", rclcpp::QoS(10));
        odom_pub_ = create_publisher<nav_msgs::msg::Odometry>("/odometry", rclcpp::QoS(10));
        tf_ = std::make_unique<tf2_ros::TransformBroadcaster>(*this);
        scan_sub_ = create_subscription<PointCloud2>("/lidar_points", rclcpp::QoS(50), [this](PointCloud2::ConstSharedPtr msg) { on_scan(*msg); });
    }

private:
    void on_scan(const PointCloud2& msg) {
        if (!to_scan(msg, scan_)) {
            RCLCPP_ERROR_THROTTLE(get_logger(), *get_clock(), 5000, "scan without x/y/z/ring/time fields dropped");
            return;
        }
        try {
            odometry_->processCloud(scan_);
        } catch (const lo::Error& e) {
            RCLCPP_ERROR(get_logger(), "%s", e.what());
            return;
        }
        if (odometry_->stats().unstable) RCLCPP_WARN(get_logger(), "unstable rotation: constant-velocity prediction kept");
        std_msgs::msg::Header world = msg.header, body = msg.header;
        world.frame_id = parent_frame_;
        body.frame_id = child_frame_;
        if (keyframe_pub_->get_subscription_count() > 0)
            keyframe_pub_->publish(to_msg(full_keyframe_ ? *odometry_->getFullKeyFrameCloud() : *odometry_->getKeyFrameCloud(), world));
        if (deskewed_pub_->get_subscription_count() > 0) deskewed_pub_->publish(to_msg(*odometry_->getTempCloud(), body));

        const Pose3D pose = odometry_->getCurrentPose();
        geometry_msgs::msg::TransformStamped tf;
        tf.header = world;
        tf.child_frame_id = child_frame_;
        tf.transform.translation.x = pose.translation.x();
        tf.transform.translation.y = pose.translation.y();
        tf.transform.translation.z = pose.translation.z();
        tf.transform.rotation.w = pose.rotation.w();
        tf.transform.rotation.x = pose.rotation.x();
        tf.transform.rotation.y = pose.rotation.y();
        tf.transform.rotation.z = pose.rotation.z();
        tf_->sendTransform(tf);

        nav_msgs::msg::Odometry odom;
        odom.header = world;
        odom.child_frame_id = child_frame_;
        odom.pose.pose.position.x = tf.transform.translation.x;
        odom.pose.pose.position.y = tf.transform.translation.y;
        odom.pose.pose.position.z = tf.transform.translation.z;
        odom.pose.pose.orientation = tf.transform.rotation;
        odom_pub_->publish(odom);
    }

    std::unique_ptr<LidarOdometry> odometry_;
    LidarOdometry::CloudType scan_;
    std::string parent_frame_, child_frame_;
    bool full_keyframe_ = false;
    rclcpp::Subscription<PointCloud2>::SharedPtr scan_sub_;
    rclcpp::Publisher<PointCloud2>::SharedPtr keyframe_pub_, deskewed_pub_;
    rclcpp::Publisher<nav_msgs::msg::Odometry>::SharedPtr odom_pub_;
    std::unique_ptr<tf2_ros::TransformBroadcaster> tf_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<LidarOdometryNodeB200>());
    rclcpp::shutdown();
    return 0;
}
