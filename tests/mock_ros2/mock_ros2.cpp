// mock_ros2.cpp — the in-process bus, the script player and the message dumps behind mock_ros2.hpp (test
// infrastructure; see the header).  Script (MOCK_ROS_SCRIPT), one command per line:
//   param NAME int|double|bool VALUE        parameter override, seen by declare_parameter (like a launch file's)
//   setparam NAME int|double|bool VALUE     parameter change while the node runs
//   tf TARGET SOURCE tx ty tz qx qy qz qw   make a transform available;   notf TARGET SOURCE   withdraw it
//   cloud TOPIC FILE N STEP FRAME SEC NSEC name:offset ...   deliver a PointCloud2 (FLOAT32 fields) read from FILE
//   fix TOPIC LAT LON ALT STATUS            deliver a NavSatFix;   heading TOPIC qx qy qz qw   deliver a QuaternionStamped
//   timers                                  fire every wall timer once
// Every published message becomes NNN.json (+ NNN.bin for clouds and grids) under MOCK_ROS_OUT, log lines go to log.txt.
#include "mock_ros2.hpp"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace mock_ros {
namespace {

struct Sub { std::type_index type; Erased cb; };
std::multimap<std::string, Sub>& subs() { static std::multimap<std::string, Sub> s; return s; }
std::vector<std::function<void()>>& timers() { static std::vector<std::function<void()>> t; return t; }
std::map<std::string, geometry_msgs::msg::TransformStamped>& tfs() { static std::map<std::string, geometry_msgs::msg::TransformStamped> t; return t; }
std::map<std::string, ParamValue>& overrides() { static std::map<std::string, ParamValue> o; return o; }
std::vector<std::string>& script() { static std::vector<std::string> s; return s; }
int64_t g_now = 1700000000LL * 1000000000LL;
int g_seq = 0;
bool g_loaded = false;

std::string out_dir() { const char* d = getenv("MOCK_ROS_OUT"); return d ? d : "."; }

bool parse_value(std::istringstream& in, ParamValue* v) {
    std::string kind, val;
    if (!(in >> kind >> val)) return false;
    if (kind == "int") { v->kind = ParamValue::INT; v->i = atoll(val.c_str()); }
    else if (kind == "double") { v->kind = ParamValue::DOUBLE; v->d = atof(val.c_str()); }
    else if (kind == "bool") { v->kind = ParamValue::BOOL; v->b = val == "true" || val == "1"; }
    else return false;
    return true;
}

void load() {
    if (g_loaded) return;
    g_loaded = true;
    const char* path = getenv("MOCK_ROS_SCRIPT");
    if (!path) return;
    std::ifstream f(path);
    for (std::string line; std::getline(f, line);) {
        std::istringstream in(line);
        std::string cmd;
        if (!(in >> cmd) || cmd[0] == '#') continue;
        if (cmd == "param") {
            std::string name;
            ParamValue v;
            in >> name;
            if (!parse_value(in, &v)) throw std::runtime_error("bad param line: " + line);
            overrides()[name] = v;
        } else {
            script().push_back(line);
        }
    }
}

std::string esc(const std::string& s) {
    std::string o;
    for (char c : s) {
        if (c == '"' || c == '\\') o += '\\';
        if (c == '\n') { o += "\\n"; continue; }
        o += c;
    }
    return o;
}

std::string header_json(const std_msgs::msg::Header& h) {
    std::ostringstream o;
    o << "\"frame_id\": \"" << esc(h.frame_id) << "\", \"stamp\": [" << h.stamp.sec << ", " << h.stamp.nanosec << "]";
    return o.str();
}

void write_files(const std::string& json, const void* bin, size_t bytes) {
    char name[64];
    snprintf(name, sizeof name, "/%03d", g_seq++);
    std::ofstream(out_dir() + name + ".json") << json << "\n";
    if (bin) std::ofstream(out_dir() + name + ".bin", std::ios::binary).write(static_cast<const char*>(bin), (std::streamsize)bytes);
}

}  // namespace

int64_t now_ns() { return g_now += 1000000; }       // a millisecond per reading: stamps differ and are reproducible

void log_line(const char* level, const std::string& text) {
    std::ofstream(out_dir() + "/log.txt", std::ios::app) << level << " " << text << "\n";
}

bool override_of(const std::string& name, ParamValue* out) {
    load();
    auto it = overrides().find(name);
    if (it == overrides().end()) return false;
    *out = it->second;
    return true;
}

void subscribe(const std::string& topic, std::type_index type, Erased cb) { subs().insert({topic, Sub{type, std::move(cb)}}); }
void add_timer(std::function<void()> cb) { timers().push_back(std::move(cb)); }

bool lookup_tf(const std::string& target, const std::string& source, geometry_msgs::msg::TransformStamped* out) {
    auto it = tfs().find(target + "<-" + source);
    if (it == tfs().end()) return false;
    *out = it->second;
    return true;
}

void dump(const std::string& topic, const sensor_msgs::msg::PointCloud2& m) {
    std::ostringstream o;
    o << "{\"topic\": \"" << topic << "\", \"type\": \"PointCloud2\", " << header_json(m.header) << ", \"width\": " << m.width
      << ", \"height\": " << m.height << ", \"point_step\": " << m.point_step << ", \"row_step\": " << m.row_step
      << ", \"is_dense\": " << (m.is_dense ? "true" : "false") << ", \"is_bigendian\": " << (m.is_bigendian ? "true" : "false")
      << ", \"bytes\": " << m.data.size() << ", \"fields\": [";
    for (size_t i = 0; i < m.fields.size(); ++i)
        o << (i ? ", " : "") << "[\"" << m.fields[i].name << "\", " << m.fields[i].offset << ", " << (int)m.fields[i].datatype << ", "
          << m.fields[i].count << "]";
    o << "]}";
    write_files(o.str(), m.data.data(), m.data.size());
}

void dump(const std::string& topic, const geometry_msgs::msg::PointStamped& m) {
    char buf[256];
    snprintf(buf, sizeof buf, "\"point\": [%.17g, %.17g, %.17g]", m.point.x, m.point.y, m.point.z);
    write_files("{\"topic\": \"" + topic + "\", \"type\": \"PointStamped\", " + header_json(m.header) + ", " + buf + "}", nullptr, 0);
}

void dump(const std::string& topic, const visualization_msgs::msg::MarkerArray& m) {
    // per marker: action, type, id, position xyz, scale xyz, colour rgba, lifetime seconds — 13 doubles; namespaces in the json
    std::vector<double> rows;
    std::ostringstream o;
    o << "{\"topic\": \"" << topic << "\", \"type\": \"MarkerArray\", \"markers\": " << m.markers.size() << ", \"namespaces\": [";
    std::string last = "\x01";
    bool first = true;
    for (const auto& k : m.markers) {
        if (k.ns != last) { o << (first ? "" : ", ") << "\"" << esc(k.ns) << "\""; last = k.ns; first = false; }
        const double row[14] = {(double)k.action, (double)k.type, (double)k.id, k.pose.position.x, k.pose.position.y, k.pose.position.z,
                                k.scale.x, k.scale.y, k.scale.z, k.color.r, k.color.g, k.color.b, k.color.a,
                                k.lifetime.sec + 1e-9 * k.lifetime.nanosec};
        rows.insert(rows.end(), row, row + 14);
    }
    o << "], \"frame_id\": \"" << (m.markers.size() > 1 ? esc(m.markers[1].header.frame_id) : "") << "\"}";
    write_files(o.str(), rows.data(), rows.size() * sizeof(double));
}

void dump(const std::string& topic, const nav_msgs::msg::OccupancyGrid& m) {
    char buf[512];
    snprintf(buf, sizeof buf, "\"resolution\": %.9g, \"width\": %u, \"height\": %u, \"origin\": [%.17g, %.17g, %.17g], \"origin_w\": %.17g",
             m.info.resolution, m.info.width, m.info.height, m.info.origin.position.x, m.info.origin.position.y,
             m.info.origin.position.z, m.info.origin.orientation.w);
    write_files("{\"topic\": \"" + topic + "\", \"type\": \"OccupancyGrid\", " + header_json(m.header) + ", " + buf + "}", m.data.data(),
                m.data.size());
}

void dump(const std::string& topic, const geometry_msgs::msg::TransformStamped& m) {
    char buf[512];
    snprintf(buf, sizeof buf, "\"child_frame_id\": \"%s\", \"translation\": [%.17g, %.17g, %.17g], \"rotation\": [%.17g, %.17g, %.17g, %.17g]",
             esc(m.child_frame_id).c_str(), m.transform.translation.x, m.transform.translation.y, m.transform.translation.z,
             m.transform.rotation.x, m.transform.rotation.y, m.transform.rotation.z, m.transform.rotation.w);
    write_files("{\"topic\": \"" + topic + "\", \"type\": \"TransformStamped\", " + header_json(m.header) + ", " + buf + "}", nullptr, 0);
}

template <class T>
static void deliver(const std::string& topic, const std::shared_ptr<T>& msg) {
    auto range = subs().equal_range(topic);
    for (auto it = range.first; it != range.second; ++it)
        if (it->second.type == std::type_index(typeid(T))) it->second.cb(msg);
}

void run_script(const std::function<void(const std::string&, const ParamValue&)>& set_param) {
    load();
    for (const std::string& line : script()) {
        std::istringstream in(line);
        std::string cmd;
        in >> cmd;
        if (cmd == "setparam") {
            std::string name;
            ParamValue v;
            in >> name;
            if (!parse_value(in, &v)) throw std::runtime_error("bad setparam line: " + line);
            set_param(name, v);
        } else if (cmd == "tf") {
            std::string target, source;
            geometry_msgs::msg::TransformStamped t;
            in >> target >> source >> t.transform.translation.x >> t.transform.translation.y >> t.transform.translation.z >>
                t.transform.rotation.x >> t.transform.rotation.y >> t.transform.rotation.z >> t.transform.rotation.w;
            t.header.frame_id = target;
            t.child_frame_id = source;
            tfs()[target + "<-" + source] = t;
        } else if (cmd == "notf") {
            std::string target, source;
            in >> target >> source;
            tfs().erase(target + "<-" + source);
        } else if (cmd == "timers") {
            for (auto& t : timers()) t();
        } else if (cmd == "cloud") {
            std::string topic, file, frame;
            uint64_t n;
            uint32_t step;
            auto msg = std::make_shared<sensor_msgs::msg::PointCloud2>();
            in >> topic >> file >> n >> step >> frame >> msg->header.stamp.sec >> msg->header.stamp.nanosec;
            msg->header.frame_id = frame;
            msg->height = 1;
            msg->width = static_cast<uint32_t>(n);
            msg->point_step = step;
            msg->row_step = step * msg->width;
            for (std::string f; in >> f;) {
                sensor_msgs::msg::PointField pf;
                pf.name = f.substr(0, f.find(':'));
                pf.offset = static_cast<uint32_t>(atoi(f.substr(f.find(':') + 1).c_str()));
                pf.datatype = sensor_msgs::msg::PointField::FLOAT32;
                pf.count = 1;
                msg->fields.push_back(pf);
            }
            msg->data.resize(static_cast<size_t>(n) * step);
            std::ifstream(file, std::ios::binary).read(reinterpret_cast<char*>(msg->data.data()), (std::streamsize)msg->data.size());
            deliver(topic, msg);
        } else if (cmd == "fix") {
            std::string topic;
            int status = 0;
            auto msg = std::make_shared<sensor_msgs::msg::NavSatFix>();
            in >> topic >> msg->latitude >> msg->longitude >> msg->altitude >> status;
            msg->status.status = static_cast<int8_t>(status);
            msg->header.stamp = rclcpp::Time(now_ns());
            deliver(topic, msg);
        } else if (cmd == "heading") {
            std::string topic;
            auto msg = std::make_shared<geometry_msgs::msg::QuaternionStamped>();
            in >> topic >> msg->quaternion.x >> msg->quaternion.y >> msg->quaternion.z >> msg->quaternion.w;
            msg->header.stamp = rclcpp::Time(now_ns());
            deliver(topic, msg);
        } else {
            throw std::runtime_error("unknown script command: " + line);
        }
    }
}

}  // namespace mock_ros
