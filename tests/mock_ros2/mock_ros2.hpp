// mock_ros2.hpp — just enough of rclcpp / tf2_ros / the message packages to COMPILE and DRIVE the wrappers in ros2/src
// without a ROS2 installation (there is none in this image).  Test infrastructure only: nothing in the product
// includes it.  Messages travel through an in-process bus; rclcpp::spin() plays a script (MOCK_ROS_SCRIPT) of
// transforms, messages, parameter changes and timer ticks, and every published message is dumped to MOCK_ROS_OUT.
#pragma once
#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeindex>
#include <vector>

// ---------------------------------------------------------------- messages
namespace builtin_interfaces { namespace msg {
struct Time { int32_t sec = 0; uint32_t nanosec = 0; };
struct Duration { int32_t sec = 0; uint32_t nanosec = 0; };
}}
namespace std_msgs { namespace msg {
struct Header { builtin_interfaces::msg::Time stamp; std::string frame_id; };
struct ColorRGBA { float r = 0, g = 0, b = 0, a = 0; };
}}
namespace geometry_msgs { namespace msg {
struct Point { double x = 0, y = 0, z = 0; };
struct Vector3 { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
struct Pose { Point position; Quaternion orientation; };
struct Transform { Vector3 translation; Quaternion rotation; };
struct PointStamped { std_msgs::msg::Header header; Point point; using SharedPtr = std::shared_ptr<PointStamped>; };
struct TransformStamped { std_msgs::msg::Header header; std::string child_frame_id; Transform transform; };
struct QuaternionStamped { std_msgs::msg::Header header; Quaternion quaternion; using SharedPtr = std::shared_ptr<QuaternionStamped>; };
}}
namespace sensor_msgs { namespace msg {
struct PointField {
    enum : uint8_t { INT8 = 1, UINT8 = 2, INT16 = 3, UINT16 = 4, INT32 = 5, UINT32 = 6, FLOAT32 = 7, FLOAT64 = 8 };
    std::string name; uint32_t offset = 0; uint8_t datatype = 0; uint32_t count = 0;
};
struct NavSatStatus { int8_t status = 0; uint16_t service = 0; };
struct NavSatFix {
    std_msgs::msg::Header header; NavSatStatus status; double latitude = 0, longitude = 0, altitude = 0;
    using SharedPtr = std::shared_ptr<NavSatFix>;
};
struct PointCloud2 {
    std_msgs::msg::Header header;
    uint32_t height = 0, width = 0;
    std::vector<PointField> fields;
    bool is_bigendian = false;
    uint32_t point_step = 0, row_step = 0;
    std::vector<uint8_t> data;
    bool is_dense = false;
    using SharedPtr = std::shared_ptr<PointCloud2>;
};
}}
namespace visualization_msgs { namespace msg {
struct Marker {
    enum : int32_t { ARROW = 0, CUBE = 1, SPHERE = 2, CYLINDER = 3 };
    enum : int32_t { ADD = 0, MODIFY = 0, DELETE = 2, DELETEALL = 3 };
    std_msgs::msg::Header header; std::string ns; int32_t id = 0, type = 0, action = 0;
    geometry_msgs::msg::Pose pose; geometry_msgs::msg::Vector3 scale; std_msgs::msg::ColorRGBA color;
    builtin_interfaces::msg::Duration lifetime;
};
struct MarkerArray { std::vector<Marker> markers; using SharedPtr = std::shared_ptr<MarkerArray>; };
}}
namespace nav_msgs { namespace msg {
struct MapMetaData { builtin_interfaces::msg::Time map_load_time; float resolution = 0; uint32_t width = 0, height = 0; geometry_msgs::msg::Pose origin; };
struct OccupancyGrid { std_msgs::msg::Header header; MapMetaData info; std::vector<int8_t> data; using SharedPtr = std::shared_ptr<OccupancyGrid>; };
}}

// ---------------------------------------------------------------- the bus and the script (mock_ros2.cpp)
namespace mock_ros {
void dump(const std::string& topic, const sensor_msgs::msg::PointCloud2& m);
void dump(const std::string& topic, const geometry_msgs::msg::PointStamped& m);
void dump(const std::string& topic, const visualization_msgs::msg::MarkerArray& m);
void dump(const std::string& topic, const nav_msgs::msg::OccupancyGrid& m);
void dump(const std::string& topic, const geometry_msgs::msg::TransformStamped& m);      // TransformBroadcaster: topic "/tf"
void log_line(const char* level, const std::string& text);
struct ParamValue { enum Kind { INT, DOUBLE, BOOL } kind = INT; long long i = 0; double d = 0; bool b = false; };
bool override_of(const std::string& name, ParamValue* out);          // `param` lines of the script
using Erased = std::function<void(const std::shared_ptr<void>&)>;
void subscribe(const std::string& topic, std::type_index type, Erased cb);
void add_timer(std::function<void()> cb);
bool lookup_tf(const std::string& target, const std::string& source, geometry_msgs::msg::TransformStamped* out);
void run_script(const std::function<void(const std::string&, const ParamValue&)>& set_param);
int64_t now_ns();
}

// ---------------------------------------------------------------- rclcpp
namespace rclcpp {
class Time {
public:
    Time() = default;
    explicit Time(int64_t ns) : ns_(ns) {}
    Time(const builtin_interfaces::msg::Time& t) : ns_(static_cast<int64_t>(t.sec) * 1000000000LL + t.nanosec) {}
    operator builtin_interfaces::msg::Time() const {
        builtin_interfaces::msg::Time t;
        t.sec = static_cast<int32_t>(ns_ / 1000000000LL);
        t.nanosec = static_cast<uint32_t>(ns_ % 1000000000LL);
        return t;
    }
    int64_t nanoseconds() const { return ns_; }
private:
    int64_t ns_ = 0;
};
class Duration {
public:
    static Duration from_seconds(double s) { Duration d; d.ns_ = static_cast<int64_t>(s * 1e9); return d; }
    operator builtin_interfaces::msg::Duration() const {
        builtin_interfaces::msg::Duration d;
        d.sec = static_cast<int32_t>(ns_ / 1000000000LL);
        d.nanosec = static_cast<uint32_t>(ns_ % 1000000000LL);
        return d;
    }
private:
    int64_t ns_ = 0;
};
class Clock { public: using SharedPtr = std::shared_ptr<Clock>; Time now() const { return Time(mock_ros::now_ns()); } };
class Logger { public: std::string name; };
class Parameter {
public:
    explicit Parameter(mock_ros::ParamValue v) : v_(v) {}
    int64_t as_int() const { if (v_.kind != mock_ros::ParamValue::INT) throw std::runtime_error("parameter is not an integer"); return v_.i; }
    double as_double() const { if (v_.kind != mock_ros::ParamValue::DOUBLE) throw std::runtime_error("parameter is not a double"); return v_.d; }
    bool as_bool() const { if (v_.kind != mock_ros::ParamValue::BOOL) throw std::runtime_error("parameter is not a bool"); return v_.b; }
private:
    mock_ros::ParamValue v_;
};
class TimerBase { public: using SharedPtr = std::shared_ptr<TimerBase>; };
template <class T> class Subscription { public: using SharedPtr = std::shared_ptr<Subscription<T>>; };
template <class T> class Publisher {
public:
    using SharedPtr = std::shared_ptr<Publisher<T>>;
    explicit Publisher(std::string topic) : topic_(std::move(topic)) {}
    void publish(const T& m) { mock_ros::dump(topic_, m); }
private:
    std::string topic_;
};

class Node {
public:
    explicit Node(const std::string& name) { logger_.name = name; clock_ = std::make_shared<Clock>(); }
    virtual ~Node() = default;
    Logger get_logger() const { return logger_; }
    Clock::SharedPtr get_clock() const { return clock_; }
    Time now() const { return clock_->now(); }

    int64_t declare_parameter(const std::string& name, int v) { return declare(name, from_int(v)).i; }
    int64_t declare_parameter(const std::string& name, int64_t v) { return declare(name, from_int(v)).i; }
    double declare_parameter(const std::string& name, double v) { mock_ros::ParamValue p; p.kind = p.DOUBLE; p.d = v; return declare(name, p).d; }
    bool declare_parameter(const std::string& name, bool v) { mock_ros::ParamValue p; p.kind = p.BOOL; p.b = v; return declare(name, p).b; }
    Parameter get_parameter(const std::string& name) const {
        auto it = params_.find(name);
        if (it == params_.end()) throw std::runtime_error("parameter '" + name + "' has not been declared");
        return Parameter(it->second);
    }
    void mock_set_parameter(const std::string& name, const mock_ros::ParamValue& v) {       // `setparam` lines of the script
        auto it = params_.find(name);
        if (it == params_.end()) throw std::runtime_error("setparam: '" + name + "' has not been declared");
        if (it->second.kind != v.kind) throw std::runtime_error("setparam: '" + name + "' has another type");
        it->second = v;
    }

    template <class T, class Cb>
    typename Subscription<T>::SharedPtr create_subscription(const std::string& topic, size_t /*depth*/, Cb cb) {
        std::function<void(std::shared_ptr<T>)> f = cb;
        mock_ros::subscribe(topic, std::type_index(typeid(T)),
                            [f](const std::shared_ptr<void>& m) { f(std::static_pointer_cast<T>(m)); });
        return std::make_shared<Subscription<T>>();
    }
    template <class T>
    typename Publisher<T>::SharedPtr create_publisher(const std::string& topic, size_t /*depth*/) {
        return std::make_shared<Publisher<T>>(topic);
    }
    template <class Rep, class Period, class Cb>
    TimerBase::SharedPtr create_wall_timer(std::chrono::duration<Rep, Period>, Cb cb) {
        mock_ros::add_timer(std::function<void()>(cb));
        return std::make_shared<TimerBase>();
    }

private:
    static mock_ros::ParamValue from_int(int64_t v) { mock_ros::ParamValue p; p.kind = p.INT; p.i = v; return p; }
    const mock_ros::ParamValue& declare(const std::string& name, mock_ros::ParamValue def) {
        mock_ros::ParamValue ov;
        if (mock_ros::override_of(name, &ov)) {
            if (ov.kind == mock_ros::ParamValue::INT && def.kind == mock_ros::ParamValue::DOUBLE) { ov.kind = ov.DOUBLE; ov.d = static_cast<double>(ov.i); }
            if (ov.kind != def.kind) throw std::runtime_error("parameter override of '" + name + "' has another type");
            def = ov;
        }
        return params_[name] = def;
    }
    Logger logger_;
    Clock::SharedPtr clock_;
    std::map<std::string, mock_ros::ParamValue> params_;
};

inline void init(int, char**) {}
inline void shutdown() {}
inline void spin(const std::shared_ptr<Node>& node) {
    mock_ros::run_script([node](const std::string& name, const mock_ros::ParamValue& v) { node->mock_set_parameter(name, v); });
}
inline std::string mock_format(const char* fmt, ...) {
    char buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    return buf;
}
}  // namespace rclcpp

#define RCLCPP_DEBUG(logger, ...) mock_ros::log_line("DEBUG", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_INFO(logger, ...) mock_ros::log_line("INFO", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_WARN(logger, ...) mock_ros::log_line("WARN", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_ERROR(logger, ...) mock_ros::log_line("ERROR", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_FATAL(logger, ...) mock_ros::log_line("FATAL", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_INFO_ONCE(logger, ...) do { static bool mock_once = false; if (!mock_once) { mock_once = true; mock_ros::log_line("INFO", rclcpp::mock_format(__VA_ARGS__)); } } while (0)
#define RCLCPP_INFO_THROTTLE(logger, clock, ms, ...) mock_ros::log_line("INFO", rclcpp::mock_format(__VA_ARGS__))
#define RCLCPP_WARN_THROTTLE(logger, clock, ms, ...) mock_ros::log_line("WARN", rclcpp::mock_format(__VA_ARGS__))

// ---------------------------------------------------------------- tf2 / tf2_ros
namespace tf2 {
struct TimePoint {};
static const TimePoint TimePointZero{};
using Duration = std::chrono::nanoseconds;
inline Duration durationFromSec(double s) { return Duration(static_cast<int64_t>(s * 1e9)); }
class TransformException : public std::runtime_error { public: using std::runtime_error::runtime_error; };
class LookupException : public TransformException { public: using TransformException::TransformException; };
}
namespace tf2_ros {
class Buffer {
public:
    explicit Buffer(rclcpp::Clock::SharedPtr) {}
    geometry_msgs::msg::TransformStamped lookupTransform(const std::string& target, const std::string& source, tf2::TimePoint,
                                                         tf2::Duration) const {
        geometry_msgs::msg::TransformStamped t;
        if (!mock_ros::lookup_tf(target, source, &t))
            throw tf2::LookupException("\"" + target + "\" passed to lookupTransform argument target_frame does not exist (mock)");
        return t;
    }
    geometry_msgs::msg::TransformStamped lookupTransform(const std::string& target, const std::string& source, tf2::TimePoint t) const {
        return lookupTransform(target, source, t, tf2::Duration(0));
    }
    bool canTransform(const std::string& target, const std::string& source, tf2::TimePoint, tf2::Duration) const {
        geometry_msgs::msg::TransformStamped t;
        return mock_ros::lookup_tf(target, source, &t);
    }
};
class TransformListener { public: explicit TransformListener(Buffer&) {} };
class TransformBroadcaster {       // the broadcast is recorded, not fed back into lookupTransform: the script owns the TF tree
public:
    template <class NodeT> explicit TransformBroadcaster(NodeT&&) {}
    void sendTransform(const geometry_msgs::msg::TransformStamped& t) { mock_ros::dump("/tf", t); }
};
}
