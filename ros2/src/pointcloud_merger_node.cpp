// pointcloud_merger_node.cpp — executable `pointcloud_merger`: the rclcpp surface of GnssGicpMatcher
// (/root/reference/src/pointcloud_merger.cpp:20-395) around plc_merge_clouds.  Same node name, topics (depth 10),
// 100 ms wall timer, TF frames and log lines.  The data path of each tick — rigid transform of both filtered clouds
// into `map`, colouring, concatenation (:308-394) — runs on the GPU; the GNSS -> TF bookkeeping (:108-305, a few flops
// per message) is plain host code here as it is upstream, with tf2's quaternion formulas written out in doubles.
#include <array>
#include <chrono>
#include <memory>
#include <string>

#include <GeographicLib/LocalCartesian.hpp>
#include <geometry_msgs/msg/quaternion_stamped.hpp>
#include <geometry_msgs/msg/transform_stamped.hpp>
#include <rclcpp/rclcpp.hpp>
#include <sensor_msgs/msg/nav_sat_fix.hpp>
#include <tf2_ros/buffer.h>
#include <tf2_ros/transform_broadcaster.h>
#include <tf2_ros/transform_listener.h>

#include "cloud_msg.hpp"
#include "plc_node.h"

using geometry_msgs::msg::QuaternionStamped;
using geometry_msgs::msg::TransformStamped;
using sensor_msgs::msg::NavSatFix;
using sensor_msgs::msg::PointCloud2;

namespace {

using Quat = std::array<double, 4>;   // x, y, z, w
using Vec = std::array<double, 3>;

// tf2::Quaternion operator*(q1, q2)
Quat mul(const Quat& a, const Quat& b) {
    return {a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1], a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2],
            a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0], a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]};
}
Quat inverse(const Quat& q) { return {-q[0], -q[1], -q[2], q[3]}; }
// tf2::quatRotate: (q * v) * q^-1 with v a pure quaternion
Vec rotate(const Quat& q, const Vec& v) {
    const Quat qv{q[3] * v[0] + q[1] * v[2] - q[2] * v[1], q[3] * v[1] + q[2] * v[0] - q[0] * v[2],
                  q[3] * v[2] + q[0] * v[1] - q[1] * v[0], -q[0] * v[0] - q[1] * v[1] - q[2] * v[2]};
    const Quat r = mul(qv, inverse(q));
    return {r[0], r[1], r[2]};
}
Quat quat_of(const geometry_msgs::msg::Quaternion& m) { return {m.x, m.y, m.z, m.w}; }

}  // namespace

class PointcloudMergerGpuNode : public rclcpp::Node {
    struct Vehicle {
        std::string name;                       // TF prefix and topic namespace
        NavSatFix::SharedPtr fix;
        QuaternionStamped::SharedPtr heading;
        PointCloud2::SharedPtr cloud;
        rclcpp::Subscription<NavSatFix>::SharedPtr fix_sub;
        rclcpp::Subscription<QuaternionStamped>::SharedPtr heading_sub;
        rclcpp::Subscription<PointCloud2>::SharedPtr cloud_sub;
        rclcpp::Publisher<PointCloud2>::SharedPtr coloured_pub;
    };

public:
    PointcloudMergerGpuNode() : Node("gnss_gicp_matcher") {
        declare_parameter("cuda_device", 0);
        if (plc_ctx_create(static_cast<int>(get_parameter("cuda_device").as_int()), nullptr, &ctx_) != PLC_OK)
            throw std::runtime_error("no usable CUDA device: pointcloud_merger has no CPU fallback");
        tf_buffer_ = std::make_shared<tf2_ros::Buffer>(get_clock());
        tf_listener_ = std::make_shared<tf2_ros::TransformListener>(*tf_buffer_);
        tf_broadcaster_ = std::make_unique<tf2_ros::TransformBroadcaster>(*this);
        robot_.name = "four_wheel_robot";
        backhoe_.name = "zx120";
        for (Vehicle* v : {&robot_, &backhoe_}) {
            v->fix_sub = create_subscription<NavSatFix>("/" + v->name + "/gnss_compass_front/fix", 10,
                                                        [this, v](NavSatFix::SharedPtr m) { on_fix(*v, m); });
            v->heading_sub = create_subscription<QuaternionStamped>("/" + v->name + "/gnss_compass/calc_heading", 10,
                                                                    [this, v](QuaternionStamped::SharedPtr m) { on_heading(*v, m); });
            v->cloud_sub = create_subscription<PointCloud2>("/" + v->name + "/filtered_points", 10,
                                                            [v](PointCloud2::SharedPtr m) { v->cloud = m; });
        }
        merged_pub_ = create_publisher<PointCloud2>("/matched_point_cloud", 10);
        robot_.coloured_pub = create_publisher<PointCloud2>("/robot_colored_point_cloud", 10);
        backhoe_.coloured_pub = create_publisher<PointCloud2>("/backhoe_colored_point_cloud", 10);
        timer_ = create_wall_timer(std::chrono::milliseconds(100), [this]() { on_tick(); });
        RCLCPP_INFO(get_logger(), "GNSS GICP Matcher node initialized");
    }
    ~PointcloudMergerGpuNode() override { plc_ctx_destroy(ctx_); }

private:
    static bool usable(const NavSatFix::SharedPtr& f) { return f && f->status.status >= 0; }

    void set_origin(const NavSatFix& f, const char* whose) {
        enu_.Reset(f.latitude, f.longitude, f.altitude);
        origin_set_ = true;
        RCLCPP_INFO(get_logger(), "Origin set at %s position: lat=%f, lon=%f, alt=%f", whose, f.latitude, f.longitude, f.altitude);
    }

    // :108-135, :146-164 — the first usable fix fixes the ENU origin; the backhoe's position wins when it is known
    void on_fix(Vehicle& v, const NavSatFix::SharedPtr& m) {
        v.fix = m;
        if (!origin_set_ && m->status.status >= 0) {
            if (&v == &robot_ && usable(backhoe_.fix)) set_origin(*backhoe_.fix, "zx120");
            else set_origin(*m, v.name.c_str());
        }
        if (origin_set_ && v.heading) broadcast(v, *m, *v.heading);
    }
    // :137-144, :166-173
    void on_heading(Vehicle& v, const QuaternionStamped::SharedPtr& m) {
        v.heading = m;
        if (origin_set_ && v.fix) broadcast(v, *v.fix, *m);
    }

    // publishGnssTransform (:185-305): map -> <name>/base_link
    void broadcast(const Vehicle& v, const NavSatFix& fix, const QuaternionStamped& heading) {
        if (!origin_set_) return;
        TransformStamped out;
        out.header.stamp = now();
        out.header.frame_id = "map";
        out.child_frame_id = v.name + "/base_link";
        Vec backhoe_antenna{0.0, 0.0, 0.0};                                  // zx120/base_link -> zx120/gnss_link
        try {
            const auto& t = tf_buffer_->lookupTransform("zx120/base_link", "zx120/gnss_link", tf2::TimePointZero, tf2::durationFromSec(0.1))
                                .transform.translation;
            backhoe_antenna = {t.x, t.y, t.z};
            RCLCPP_INFO_ONCE(get_logger(), "ZX120 GNSS offset from base_link: [%.3f, %.3f, %.3f]", t.x, t.y, t.z);
        } catch (tf2::TransformException& ex) {
            RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Could not get zx120 base_link to gnss_link transform: %s", ex.what());
        }
        if (&v == &backhoe_) {                                               // pinned to the map origin (:218-231)
            out.transform.translation.x = out.transform.translation.y = out.transform.translation.z = 0.0;
            out.transform.rotation.x = out.transform.rotation.y = out.transform.rotation.z = 0.0;
            out.transform.rotation.w = 1.0;
            RCLCPP_INFO_ONCE(get_logger(), "zx120/base_link fixed at map origin (0,0,0) with no rotation");
        } else {
            Vec me{}, other{0.0, 0.0, 0.0};
            enu_.Forward(fix.latitude, fix.longitude, fix.altitude, me[0], me[1], me[2]);
            if (usable(backhoe_.fix))
                enu_.Forward(backhoe_.fix->latitude, backhoe_.fix->longitude, backhoe_.fix->altitude, other[0], other[1], other[2]);
            TransformStamped antenna;                                        // gnss_link <- base_link of this vehicle (:253-263)
            try {
                antenna = tf_buffer_->lookupTransform(v.name + "/gnss_link", v.name + "/base_link", tf2::TimePointZero);
            } catch (tf2::TransformException& ex) {
                RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Could not get transform: %s", ex.what());
                return;
            }
            const Quat q = quat_of(heading.quaternion);
            const auto& a = antenna.transform.translation;
            const Vec lever = rotate(q, Vec{-a.x, -a.y, -a.z});
            out.transform.translation.x = (me[0] - other[0]) + lever[0] + backhoe_antenna[0];
            out.transform.translation.y = (me[1] - other[1]) + lever[1] + backhoe_antenna[1];
            out.transform.translation.z = (me[2] - other[2]) + lever[2] + backhoe_antenna[2];
            if (backhoe_.heading) {                                          // heading relative to the backhoe's (:279-291)
                const Quat rel = mul(inverse(quat_of(backhoe_.heading->quaternion)), q);
                out.transform.rotation.x = rel[0];
                out.transform.rotation.y = rel[1];
                out.transform.rotation.z = rel[2];
                out.transform.rotation.w = rel[3];
            } else {
                out.transform.rotation = heading.quaternion;
            }
            RCLCPP_INFO_THROTTLE(get_logger(), *get_clock(), 5000,
                                 "four_wheel_robot relative to zx120 (with ZX120 GNSS offset compensation) - Pos: [%.2f, %.2f, %.2f]",
                                 out.transform.translation.x, out.transform.translation.y, out.transform.translation.z);
        }
        tf_broadcaster_->sendTransform(out);
    }

    // One source of the tick: the latest cloud and map <- <name>/velodyne_link (:362-366; failure: warn and skip, :389-393)
    plc_merger_input input_of(const Vehicle& v) {
        plc_merger_input in{};
        if (!v.cloud) return in;
        if (!plc_ros::view_of(*v.cloud, &in.cloud)) {                       // upstream: fromROSMsg warns and reads zeros
            RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "cloud of %s has no FLOAT32 x, y, z fields: skipped", v.name.c_str());
            return in;
        }
        in.has_cloud = 1;
        try {
            const auto tf = tf_buffer_->lookupTransform("map", v.name + "/velodyne_link", tf2::TimePointZero, tf2::durationFromSec(0.1)).transform;
            in.has_transform = 1;
            in.t[0] = tf.translation.x; in.t[1] = tf.translation.y; in.t[2] = tf.translation.z;
            in.q[0] = tf.rotation.x; in.q[1] = tf.rotation.y; in.q[2] = tf.rotation.z; in.q[3] = tf.rotation.w;
        } catch (tf2::TransformException& ex) {
            RCLCPP_WARN_THROTTLE(get_logger(), *get_clock(), 1000, "Could not transform point cloud for %s: %s", v.name.c_str(), ex.what());
        }
        return in;
    }

    // processPointClouds (:308-352)
    void on_tick() {
        if (!origin_set_) return;
        const plc_merger_input r = input_of(robot_), b = input_of(backhoe_);
        const uint64_t cap_r = r.has_cloud && r.has_transform ? r.cloud.n_points : 0, cap_b = b.has_cloud && b.has_transform ? b.cloud.n_points : 0;
        PointCloud2 merged, robot, backhoe;
        plc_ros::shape_output(merged, cap_r + cap_b, true);
        plc_ros::shape_output(robot, cap_r, true);
        plc_ros::shape_output(backhoe, cap_b, true);
        uint64_t n_m = 0, n_r = 0, n_b = 0;
        const plc_status st = plc_merge_clouds(ctx_, &r, &b, merged.data.data(), cap_r + cap_b, &n_m, robot.data.data(), cap_r, &n_r,
                                               backhoe.data.data(), cap_b, &n_b);
        if (st != PLC_OK) {
            RCLCPP_ERROR(get_logger(), "plc_merge_clouds: %s (%s)", plc_strerror(st), plc_last_error(ctx_));
            return;
        }
        send(*merged_pub_, merged, n_m);
        send(*robot_.coloured_pub, robot, n_r);
        send(*backhoe_.coloured_pub, backhoe, n_b);
    }
    void send(rclcpp::Publisher<PointCloud2>& pub, PointCloud2& msg, uint64_t n) {
        if (n == 0) return;                                                  // an empty cloud is not published (:328, :337, :345)
        plc_ros::shape_output(msg, n, true);
        msg.header.frame_id = "map";
        msg.header.stamp = now();
        pub.publish(msg);
    }

    plc_ctx* ctx_ = nullptr;
    std::shared_ptr<tf2_ros::Buffer> tf_buffer_;
    std::shared_ptr<tf2_ros::TransformListener> tf_listener_;
    std::unique_ptr<tf2_ros::TransformBroadcaster> tf_broadcaster_;
    GeographicLib::LocalCartesian enu_;
    bool origin_set_ = false;
    Vehicle robot_, backhoe_;
    rclcpp::Publisher<PointCloud2>::SharedPtr merged_pub_;
    rclcpp::TimerBase::SharedPtr timer_;
};

int main(int argc, char** argv) {
    rclcpp::init(argc, argv);
    rclcpp::spin(std::make_shared<PointcloudMergerGpuNode>());
    rclcpp::shutdown();
    return 0;
}
