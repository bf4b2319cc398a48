// pointcloud_filter_node.cpp — executable `pointcloud_filter`: the rclcpp surface of
// SimplifiedScanMatcher (/root/reference/src/pointcloud_filter.cpp:8-62) around plc_filter_scan
// (crop box + voxel-grid downsample on the GPU).  Same node name, topics (depth 1), parameters;
// the published cloud keeps the input header (:79) and the PointXYZ wire layout of pcl::toROSMsg.
#include <memory>
#include <string>

#include <rclcpp/rclcpp.hpp>

#include "cloud_msg.hpp"
#include "plc_node.h"

using sensor_msgs::msg::PointCloud2;

class PointcloudFilterGpuNode : public rclcpp::Node {
public:
    PointcloudFilterGpuNode() : Node("simplified_scan_matcher") {
        for (const char* who : {"robot", "backhoe"})
            for (const char* what : {"front_range", "side_range", "height_range"})
                declare_parameter(std::string(who) + "_" + what, what[0] == 'f' ? 15.0 : 10.0);
        declare_parameter("voxel_leaf_size", 0.2);
        declare_parameter("cuda_device", 0);
        if (plc_ctx_create(static_cast<int>(get_parameter("cuda_device").as_int()), nullptr, &ctx_) != PLC_OK)
            throw std::runtime_error("no usable CUDA device: pointcloud_filter has no CPU fallback");
        robot_pub_ = create_publisher<PointCloud2>("/four_wheel_robot/filtered_points", 1);
        backhoe_pub_ = create_publisher<PointCloud2>("/zx120/filtered_points", 1);
        robot_sub_ = create_subscription<PointCloud2>("/four_wheel_robot/velodyne_points", 1,
            [this](PointCloud2::SharedPtr m) { robot_pub_->publish(process(*m, "robot")); });
        backhoe_sub_ = create_subscription<PointCloud2>("/zx120/velodyne_points", 1,
            [this](PointCloud2::SharedPtr m) { backhoe_pub_->publish(process(*m, "backhoe")); });
        RCLCPP_INFO(get_logger(), "Simplified Scan Matcher initialized");
        RCLCPP_INFO(get_logger(), "Function: Front area cropping and downsampling only");
    }
    ~PointcloudFilterGpuNode() override { plc_ctx_destroy(ctx_); }

private:
    PointCloud2 process(const PointCloud2& in, const std::string& vehicle) {
        PointCloud2 out;
        out.header = in.header;
        plc_cloud_view v{};
        uint64_t n = 0;
        if (plc_ros::view_of(in, &v)) {
            plc_ros::shape_output(out, v.n_points, false);        // capacity: never more points than came in
            const plc_status st = plc_filter_scan(
                ctx_, &v, get_parameter(vehicle + "_front_range").as_double(),
                get_parameter(vehicle + "_side_range").as_double(), get_parameter(vehicle + "_height_range").as_double(),
                get_parameter("voxel_leaf_size").as_double(), out.data.data(), v.n_points, &n);
            if (st != PLC_OK) {
                RCLCPP_ERROR(get_logger(), "plc_filter_scan: %s (%s)", plc_strerror(st), plc_last_error(ctx_));
                n = 0;
            }
            RCLCPP_DEBUG(get_logger(), "%s cloud: %zu -> %zu points", vehicle.c_str(), static_cast<size_t>(v.n_points),
                         static_cast<size_t>(n));
        }
        plc_ros::shape_output(out, n, false);                     // shrink to what was produced
        return out;
    }

    plc_ctx* ctx_ = nullptr;
    rclcpp::Subscription<PointCloud2>::SharedPtr robot_sub_, backhoe_sub_;
    rclcpp::Publisher<PointCloud2>::SharedPtr robot_pub_, backhoe_pub_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<PointcloudFilterGpuNode>());
    rclcpp::shutdown();
    return 0;
}
