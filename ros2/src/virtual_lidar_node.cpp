// virtual_lidar_node.cpp — executable `virtual_lidar`: the rclcpp surface of
// SimplifiedDualLidarOptimizer (/root/reference/src/virtual_lidar.cpp:60-97, 802-970) around the
// GPU-backed node core plc_vlnode (include/plc_node.h).  Same node name, topics, depths, parameters,
// TF lookup, 3 s wall timer, log text and marker arrays; none of the computation happens here.
#include <chrono>
#include <memory>
#include <sstream>
#include <vector>

#include <geometry_msgs/msg/point_stamped.hpp>
#include <rclcpp/rclcpp.hpp>
#include <tf2_ros/buffer.h>
#include <tf2_ros/transform_listener.h>
#include <visualization_msgs/msg/marker_array.hpp>

#include "cloud_msg.hpp"
#include "plc_node.h"

using sensor_msgs::msg::PointCloud2;
using visualization_msgs::msg::Marker;
using visualization_msgs::msg::MarkerArray;

namespace {

struct Rgba { float r, g, b, a; };

Marker shaped_marker(const rclcpp::Time& stamp, const char* ns, int id, int type, double x, double y, double z,
                     double sx, double sy, double sz, Rgba c, double lifetime_s) {
    Marker m;
    m.header.stamp = stamp;
    m.header.frame_id = "map";
    m.ns = ns;
    m.id = id;
    m.type = type;
    m.action = Marker::ADD;
    m.pose.position.x = x;
    m.pose.position.y = y;
    m.pose.position.z = z;
    m.pose.orientation.w = 1.0;
    m.scale.x = sx;
    m.scale.y = sy;
    m.scale.z = sz;
    m.color.r = c.r;
    m.color.g = c.g;
    m.color.b = c.b;
    m.color.a = c.a;
    m.lifetime = rclcpp::Duration::from_seconds(lifetime_s);
    return m;
}

Marker delete_all() {
    Marker m;
    m.action = Marker::DELETEALL;
    return m;
}

}  // namespace

class VirtualLidarGpuNode : public rclcpp::Node {
public:
    VirtualLidarGpuNode() : Node("simplified_dual_lidar_optimizer") {
        tf_buffer_ = std::make_shared<tf2_ros::Buffer>(get_clock());
        tf_listener_ = std::make_shared<tf2_ros::TransformListener>(*tf_buffer_);
        for (const char* name : kDoubleParams) declare_parameter(name, default_of(name));
        declare_parameter("num_candidates", 100);
        declare_parameter("vertical_layers", 10);
        // the two additions: which GPU this node runs on, or how many GPUs (0..n-1) its sweeps are sharded over
        declare_parameter("cuda_device", 0);
        declare_parameter("cuda_devices", 1);

        const int device = static_cast<int>(get_parameter("cuda_device").as_int());
        const int n_devices = static_cast<int>(get_parameter("cuda_devices").as_int());
        plc_status st;
        if (n_devices > 1) {
            std::vector<int> devs(n_devices);
            for (int i = 0; i < n_devices; ++i) devs[i] = i;
            st = plc_group_create(devs.data(), n_devices, &group_);
            if (st == PLC_OK) st = plc_vlnode_create_on_group(group_, &core_);
            if (st == PLC_OK) ctx_ = plc_group_ctx(group_, 0);      // for error text and the per-context queries
        } else {
            st = plc_ctx_create(device, nullptr, &ctx_);
            if (st == PLC_OK) st = plc_vlnode_create(ctx_, &core_);
        }
        if (st != PLC_OK) {
            RCLCPP_FATAL(get_logger(), "no usable CUDA device(s) (%s): this node has no CPU fallback", plc_strerror(st));
            throw std::runtime_error("plc context creation failed");
        }
        push_parameters();
        plc_vlnode_update_parameters(core_);        // updateParameters() at the end of the constructor (:73)

        auto sub = [this](const char* topic, plc_status (*cb)(plc_vlnode*, const plc_cloud_view*), const char* what) {
            return create_subscription<PointCloud2>(topic, 10, [this, cb, what](PointCloud2::SharedPtr msg) {
                plc_cloud_view v{};
                if (!plc_ros::view_of(*msg, &v)) {
                    RCLCPP_ERROR(get_logger(), "%s: cloud has no FLOAT32 x/y/z fields", what);
                    return;
                }
                const plc_status st = cb(core_, &v);
                if (st != PLC_OK) RCLCPP_ERROR(get_logger(), "Failed to process %s: %s (%s)", what, plc_strerror(st),
                                               plc_last_error(ctx_));
                flush_log();
                if (cb == plc_vlnode_excavation_area_callback && st == PLC_OK) publish_grid();   // :286
            });
        };
        area_sub_ = sub("/excavation_area", plc_vlnode_excavation_area_callback, "excavation area");
        terrain_sub_ = sub("/excavated_terrain", plc_vlnode_terrain_callback, "terrain cloud");
        zx120_sub_ = sub("/zx120/filtered_points", plc_vlnode_zx120_points_callback, "ZX120 cloud");

        best_pub_ = create_publisher<geometry_msgs::msg::PointStamped>("/optimal_mobile_lidar_position", 10);
        candidates_pub_ = create_publisher<MarkerArray>("/mobile_lidar_candidate_positions", 10);
        grid_pub_ = create_publisher<MarkerArray>("/excavation_grid_visualization", 10);
        timer_ = create_wall_timer(std::chrono::seconds(3), [this] { run(); });
    }

    ~VirtualLidarGpuNode() override {
        plc_vlnode_destroy(core_);
        if (group_) plc_group_destroy(group_);      // owns its contexts
        else plc_ctx_destroy(ctx_);
    }

private:
    static constexpr const char* kDoubleParams[] = {"grid_resolution", "sensor_height", "search_radius", "max_distance"};
    static double default_of(const std::string& name) {
        if (name == "grid_resolution") return 0.1;
        if (name == "sensor_height") return 1.1;
        if (name == "search_radius") return 3.0;
        return 15.0;   // max_distance
    }

    void push_parameters() {
        for (const char* name : kDoubleParams) plc_vlnode_set_parameter_double(core_, name, get_parameter(name).as_double());
        plc_vlnode_set_parameter_int(core_, "num_candidates", get_parameter("num_candidates").as_int());
        plc_vlnode_set_parameter_int(core_, "vertical_layers", get_parameter("vertical_layers").as_int());
    }

    void flush_log() {
        std::istringstream lines(plc_vlnode_take_log(core_));
        for (std::string line; std::getline(lines, line);) RCLCPP_INFO(get_logger(), "%s", line.c_str());
    }

    void run() {
        try {
            auto tf = tf_buffer_->lookupTransform("map", "zx120/base_link", tf2::TimePointZero, tf2::durationFromSec(0.1));
            const auto& t = tf.transform.translation;
            plc_vlnode_set_base_transform(core_, 1, t.x, t.y, t.z);
        } catch (const tf2::TransformException&) {
            plc_vlnode_set_base_transform(core_, 0, 0, 0, 0);
        }
        push_parameters();
        plc_vlnode_result r{};
        const plc_status st = plc_vlnode_run_optimization(core_, &r);
        if (st != PLC_OK) {
            RCLCPP_ERROR(get_logger(), "sweep failed: %s (%s)", plc_strerror(st), plc_last_error(ctx_));
            return;
        }
        if (!r.ran) return;
        flush_log();
        geometry_msgs::msg::PointStamped best;
        best.header.stamp = now();
        best.header.frame_id = "map";
        best.point.x = r.best.x;
        best.point.y = r.best.y;
        best.point.z = r.best.z;
        best_pub_->publish(best);
        publish_candidates(r);
        publish_grid();
    }

    void publish_candidates(const plc_vlnode_result& r) {
        uint64_t n = 0;
        plc_vlnode_get_candidates(core_, nullptr, 0, &n);
        std::vector<plc_pose> cands(n);
        if (n) plc_vlnode_get_candidates(core_, cands.data(), n, &n);
        MarkerArray out;
        out.markers.push_back(delete_all());
        const auto stamp = now();
        out.markers.push_back(shaped_marker(stamp, "zx120_lidar", 0, Marker::CUBE, r.fixed.x, r.fixed.y, r.fixed.z,
                                            0.5, 0.5, 0.5, {0.f, 1.f, 1.f, 1.f}, 10.0));
        for (size_t i = 0; i < cands.size(); ++i)
            out.markers.push_back(shaped_marker(stamp, "mobile_lidar_candidates", static_cast<int>(i), Marker::SPHERE,
                                                cands[i].x, cands[i].y, cands[i].z, 0.3, 0.3, 0.3,
                                                {1.f, 1.f, 0.f, 0.7f}, 8.0));
        out.markers.push_back(shaped_marker(stamp, "optimal_mobile_lidar", 0, Marker::CYLINDER, r.best.x, r.best.y,
                                            r.best.z, 1.0, 1.0, 2.0, {0.f, 0.f, 1.f, 0.9f}, 10.0));
        candidates_pub_->publish(out);
    }

    void publish_grid() {
        uint64_t n = 0;
        plc_vlnode_get_cells(core_, nullptr, nullptr, 0, &n);
        std::vector<plc_cell> cells(n);
        std::vector<uint8_t> flags(n);
        if (n) plc_vlnode_get_cells(core_, cells.data(), flags.data(), n, &n);
        const double edge = get_parameter("grid_resolution").as_double() * 0.6;
        MarkerArray out;
        out.markers.push_back(delete_all());
        const auto stamp = now();
        for (size_t i = 0; i < cells.size(); ++i) {
            const uint8_t f = flags[i];
            Rgba c{0.f, 1.f, 0.f, 0.5f};                                                    // observable
            if (!(f & (PLC_FLAG_RANGE_FIXED | PLC_FLAG_RANGE_MOBILE))) c = {0.f, 0.f, 1.f, 0.5f};       // out of range
            else if (!(f & (PLC_FLAG_FOV_FIXED | PLC_FLAG_FOV_MOBILE))) c = {1.f, 1.f, 0.f, 0.5f};      // out of FOV
            else if (!(f & (PLC_FLAG_VIS_FIXED | PLC_FLAG_VIS_MOBILE))) c = {1.f, 0.f, 0.f, 0.5f};      // occluded
            out.markers.push_back(shaped_marker(stamp, "excavation_grid_3d", static_cast<int>(i), Marker::CUBE,
                                                cells[i].x, cells[i].y, cells[i].z, edge, edge, edge, c, 15.0));
        }
        grid_pub_->publish(out);
    }

    plc_ctx* ctx_ = nullptr;
    plc_group* group_ = nullptr;
    plc_vlnode* core_ = nullptr;
    std::shared_ptr<tf2_ros::Buffer> tf_buffer_;
    std::shared_ptr<tf2_ros::TransformListener> tf_listener_;
    rclcpp::Subscription<PointCloud2>::SharedPtr area_sub_, terrain_sub_, zx120_sub_;
    rclcpp::Publisher<geometry_msgs::msg::PointStamped>::SharedPtr best_pub_;
    rclcpp::Publisher<MarkerArray>::SharedPtr candidates_pub_, grid_pub_;
    rclcpp::TimerBase::SharedPtr timer_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<VirtualLidarGpuNode>());
    rclcpp::shutdown();
    return 0;
}
