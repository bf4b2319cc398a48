// calc_drivable_area_node.cpp — executable `calc_drivable_area`: the rclcpp surface of SimplifiedScanMatcher
// (/root/reference/src/calc_drivable_area.cpp:14-60, :69-232) around plc_drivable_area (transform into the map
// frame, per-cell height range and the occupancy decision on the GPU).  Same node name, topic, parameters, TF
// lookups and OccupancyGrid layout.
#include <memory>
#include <string>

#include <nav_msgs/msg/occupancy_grid.hpp>
#include <rclcpp/rclcpp.hpp>
#include <tf2_ros/buffer.h>
#include <tf2_ros/transform_listener.h>

#include "cloud_msg.hpp"
#include "plc.h"

using sensor_msgs::msg::PointCloud2;

class DrivableAreaGpuNode : public rclcpp::Node {
public:
    DrivableAreaGpuNode() : Node("simplified_scan_matcher") {
        plc_drivable_params_default(&prm_);
        prm_.grid_resolution = declare_parameter("grid_resolution", prm_.grid_resolution);
        prm_.map_width = declare_parameter("map_width", prm_.map_width);
        prm_.map_height = declare_parameter("map_height", prm_.map_height);
        prm_.max_gradient = declare_parameter("max_gradient", prm_.max_gradient);
        prm_.min_points_per_cell = static_cast<int32_t>(declare_parameter("min_points_per_cell", 10));
        prm_.start_clear_radius = declare_parameter("start_clear_radius", prm_.start_clear_radius);
        declare_parameter("cuda_device", 0);
        if (plc_ctx_create(static_cast<int>(get_parameter("cuda_device").as_int()), nullptr, &ctx_) != PLC_OK)
            throw std::runtime_error("no usable CUDA device: calc_drivable_area has no CPU fallback");
        tf_buffer_ = std::make_shared<tf2_ros::Buffer>(get_clock());
        tf_listener_ = std::make_shared<tf2_ros::TransformListener>(*tf_buffer_);
        sub_ = create_subscription<PointCloud2>("/four_wheel_robot/filtered_points", 1,
                                                [this](PointCloud2::SharedPtr m) { on_cloud(*m); });
        pub_ = create_publisher<nav_msgs::msg::OccupancyGrid>("/occupancy_grid", 10);
        RCLCPP_INFO(get_logger(), "Occupancy Grid Map Generator initialized");
    }
    ~DrivableAreaGpuNode() override { plc_ctx_destroy(ctx_); }

private:
    void on_cloud(const PointCloud2& msg) {                      // robotCloudCallback, :69-232
        geometry_msgs::msg::TransformStamped to_map, robot;
        try {
            if (!tf_buffer_->canTransform("map", msg.header.frame_id, tf2::TimePointZero, tf2::durationFromSec(0.5))) {
                RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Transform from %s to map not available yet",
                                     msg.header.frame_id.c_str());
                return;
            }
            to_map = tf_buffer_->lookupTransform("map", msg.header.frame_id, tf2::TimePointZero, tf2::durationFromSec(0.5));
            robot = tf_buffer_->lookupTransform("map", "four_wheel_robot/base_link", tf2::TimePointZero, tf2::durationFromSec(0.1));
        } catch (tf2::TransformException& ex) {
            RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Could not transform point cloud: %s", ex.what());
            return;
        }
        plc_cloud_view v{};
        if (!plc_ros::view_of(msg, &v)) return;
        if (v.n_points == 0) {
            RCLCPP_WARN(get_logger(), "Received empty point cloud");      // :108-112
            return;
        }
        const double rx = robot.transform.translation.x, ry = robot.transform.translation.y;
        if (!start_set_) {                                       // :127-134
            start_x_ = rx;
            start_y_ = ry;
            start_set_ = true;
            RCLCPP_INFO(get_logger(), "Start position set at (%.2f, %.2f)", start_x_, start_y_);
        }
        const auto& tr = to_map.transform.translation;
        const auto& ro = to_map.transform.rotation;
        const double t[3] = {tr.x, tr.y, tr.z}, q[4] = {ro.x, ro.y, ro.z, ro.w};
        nav_msgs::msg::OccupancyGrid grid;
        const size_t cells = static_cast<size_t>(prm_.map_width / prm_.grid_resolution) *
                             static_cast<size_t>(prm_.map_height / prm_.grid_resolution);
        grid.data.resize(cells);
        plc_drivable_info info{};
        const plc_status st = plc_drivable_area(ctx_, &v, t, q, rx, ry, start_x_, start_y_, &prm_, grid.data.data(), cells, &info);
        if (st != PLC_OK || !info.published) {
            if (st != PLC_OK) RCLCPP_ERROR(get_logger(), "plc_drivable_area: %s (%s)", plc_strerror(st), plc_last_error(ctx_));
            return;
        }
        grid.header.stamp = now();                               // :164-176
        grid.header.frame_id = "map";
        grid.info.resolution = static_cast<float>(info.resolution);
        grid.info.width = static_cast<uint32_t>(info.width);
        grid.info.height = static_cast<uint32_t>(info.height);
        grid.info.origin.position.x = info.origin_x;
        grid.info.origin.position.y = info.origin_y;
        grid.info.origin.position.z = 0.0;
        grid.info.origin.orientation.w = 1.0;
        pub_->publish(grid);
    }

    plc_ctx* ctx_ = nullptr;
    plc_drivable_params prm_{};
    bool start_set_ = false;
    double start_x_ = 0.0, start_y_ = 0.0;
    std::shared_ptr<tf2_ros::Buffer> tf_buffer_;
    std::shared_ptr<tf2_ros::TransformListener> tf_listener_;
    rclcpp::Subscription<PointCloud2>::SharedPtr sub_;
    rclcpp::Publisher<nav_msgs::msg::OccupancyGrid>::SharedPtr pub_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<DrivableAreaGpuNode>());
    rclcpp::shutdown();
    return 0;
}
