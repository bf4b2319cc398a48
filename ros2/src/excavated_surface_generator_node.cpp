// excavated_surface_generator_node.cpp — executable `excavated_surface_generator`: the rclcpp surface of
// ExcavationTerrainGenerator (/root/reference/src/excavated_surface_generator.cpp:24-76, :263-325) around
// plc_excavate (terrain heights, carve, synthetic pit surface and /excavation_area on the GPU).  Same node name,
// topics (depth 10), parameters (re-read every second, :69-71) and TF lookup; outputs keep the input stamp with
// frame "map" (:312-323).  The marker array (:586-633) is parameter-only drawing and is left to the reference node.
#include <chrono>
#include <memory>
#include <string>

#include <rclcpp/rclcpp.hpp>
#include <tf2_ros/buffer.h>
#include <tf2_ros/transform_listener.h>

#include "cloud_msg.hpp"
#include "plc.h"

using sensor_msgs::msg::PointCloud2;

class ExcavationTerrainGeneratorGpuNode : public rclcpp::Node {
public:
    ExcavationTerrainGeneratorGpuNode() : Node("excavation_terrain_generator") {
        plc_excavation_params_default(&prm_);
        declare_parameter("excavation.depth", prm_.depth);
        declare_parameter("excavation.slope_angle", prm_.slope_angle_deg);
        declare_parameter("excavation.offset_x", prm_.offset_x);
        declare_parameter("excavation.offset_y", prm_.offset_y);
        declare_parameter("excavation.point_density", prm_.point_density);
        declare_parameter("excavation.enabled", true);
        declare_parameter("excavation.terrain_search_radius", prm_.terrain_search_radius);
        declare_parameter("excavation.l_shape_enabled", true);
        declare_parameter("excavation.arm1_length", prm_.arm1_length);
        declare_parameter("excavation.arm1_width", prm_.arm1_width);
        declare_parameter("excavation.arm2_length", prm_.arm2_length);
        declare_parameter("excavation.arm2_width", prm_.arm2_width);
        declare_parameter("excavation.width", prm_.width);
        declare_parameter("excavation.length", prm_.length);
        declare_parameter("cuda_device", 0);
        update_parameters();
        if (plc_ctx_create(static_cast<int>(get_parameter("cuda_device").as_int()), nullptr, &ctx_) != PLC_OK)
            throw std::runtime_error("no usable CUDA device: excavated_surface_generator has no CPU fallback");
        tf_buffer_ = std::make_shared<tf2_ros::Buffer>(get_clock());
        tf_listener_ = std::make_shared<tf2_ros::TransformListener>(*tf_buffer_);
        sub_ = create_subscription<PointCloud2>("/matched_point_cloud", 10,
                                                [this](PointCloud2::SharedPtr m) { on_cloud(*m); });
        terrain_pub_ = create_publisher<PointCloud2>("/excavated_terrain", 10);
        area_pub_ = create_publisher<PointCloud2>("/excavation_area", 10);
        timer_ = create_wall_timer(std::chrono::seconds(1), [this] { update_parameters(); });
        RCLCPP_INFO(get_logger(), "Excavation Terrain Generator initialized");
    }
    ~ExcavationTerrainGeneratorGpuNode() override { plc_ctx_destroy(ctx_); }

private:
    void update_parameters() {                                   // :115-135
        prm_.depth = get_parameter("excavation.depth").as_double();
        prm_.slope_angle_deg = get_parameter("excavation.slope_angle").as_double();
        prm_.offset_x = get_parameter("excavation.offset_x").as_double();
        prm_.offset_y = get_parameter("excavation.offset_y").as_double();
        prm_.point_density = get_parameter("excavation.point_density").as_double();
        enabled_ = get_parameter("excavation.enabled").as_bool();
        prm_.terrain_search_radius = get_parameter("excavation.terrain_search_radius").as_double();
        prm_.l_shape_enabled = get_parameter("excavation.l_shape_enabled").as_bool() ? 1 : 0;
        prm_.arm1_length = get_parameter("excavation.arm1_length").as_double();
        prm_.arm1_width = get_parameter("excavation.arm1_width").as_double();
        prm_.arm2_length = get_parameter("excavation.arm2_length").as_double();
        prm_.arm2_width = get_parameter("excavation.arm2_width").as_double();
        prm_.width = get_parameter("excavation.width").as_double();
        prm_.length = get_parameter("excavation.length").as_double();
    }

    void on_cloud(const PointCloud2& msg) {                      // matchedCloudCallback, :263-325
        if (!enabled_) {
            terrain_pub_->publish(msg);                          // :264-267
            return;
        }
        geometry_msgs::msg::TransformStamped tf;
        try {
            tf = tf_buffer_->lookupTransform("map", "zx120/base_link", tf2::TimePointZero, tf2::durationFromSec(0.1));
        } catch (tf2::TransformException& ex) {
            RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Could not get zx120 transform: %s", ex.what());
            terrain_pub_->publish(msg);                          // :274-279
            return;
        }
        plc_cloud_view v{};
        if (!plc_ros::view_of(msg, &v)) return;
        uint32_t off_rgb = UINT32_MAX;
        for (const auto& f : msg.fields)
            if (f.name == "rgb" || f.name == "rgba") off_rgb = f.offset;
        uint64_t n_surface = 0, n_area = 0;
        if (plc_excavation_counts(&prm_, &n_surface, &n_area) != PLC_OK) return;
        PointCloud2 terrain, area;
        plc_ros::shape_output(terrain, v.n_points + n_surface, true);
        plc_ros::shape_output(area, n_area, true);
        const auto& tr = tf.transform.translation;
        const auto& ro = tf.transform.rotation;
        const double t[3] = {tr.x, tr.y, tr.z}, q[4] = {ro.x, ro.y, ro.z, ro.w};
        uint64_t nt = 0, na = 0;
        plc_excavation_info info{};
        const plc_status st = plc_excavate(ctx_, &v, off_rgb, t, q, &prm_, terrain.data.data(), v.n_points + n_surface, &nt,
                                           area.data.data(), n_area, &na, &info);
        if (st != PLC_OK) {
            RCLCPP_ERROR(get_logger(), "plc_excavate: %s (%s)", plc_strerror(st), plc_last_error(ctx_));
            return;
        }
        plc_ros::shape_output(terrain, nt, true);
        plc_ros::shape_output(area, na, true);
        terrain.header = msg.header;                             // :312-316
        terrain.header.frame_id = "map";
        area.header = msg.header;                                // :318-322
        area.header.frame_id = "map";
        terrain_pub_->publish(terrain);
        area_pub_->publish(area);
        RCLCPP_INFO_THROTTLE(get_logger(), *get_clock(), 5000, "Excavation processed - Input: %zu, Output: %zu points",
                             static_cast<size_t>(v.n_points), static_cast<size_t>(nt));      // :476-478
    }

    plc_ctx* ctx_ = nullptr;
    plc_excavation_params prm_{};
    bool enabled_ = true;
    std::shared_ptr<tf2_ros::Buffer> tf_buffer_;
    std::shared_ptr<tf2_ros::TransformListener> tf_listener_;
    rclcpp::Subscription<PointCloud2>::SharedPtr sub_;
    rclcpp::Publisher<PointCloud2>::SharedPtr terrain_pub_, area_pub_;
    rclcpp::TimerBase::SharedPtr timer_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<ExcavationTerrainGeneratorGpuNode>());
    rclcpp::shutdown();
    return 0;
}
