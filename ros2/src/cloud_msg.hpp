// cloud_msg.hpp — sensor_msgs/PointCloud2 <-> plc_cloud_view, and the two wire layouts the nodes publish.
#pragma once
#include <cstring>
#include <string>

#include <sensor_msgs/msg/point_cloud2.hpp>
#include <sensor_msgs/msg/point_field.hpp>

#include "plc.h"

namespace plc_ros {

// Field lookup by name, the way pcl::fromROSMsg maps fields.  Returns false if x, y or z
// is missing or not FLOAT32 (upstream would warn "Failed to find match for field" and read zeros).
inline bool view_of(const sensor_msgs::msg::PointCloud2& msg, plc_cloud_view* out) {
    int off[3] = {-1, -1, -1};
    for (const auto& f : msg.fields) {
        if (f.datatype != sensor_msgs::msg::PointField::FLOAT32) continue;
        if (f.name == "x") off[0] = static_cast<int>(f.offset);
        else if (f.name == "y") off[1] = static_cast<int>(f.offset);
        else if (f.name == "z") off[2] = static_cast<int>(f.offset);
    }
    if (off[0] < 0 || off[1] < 0 || off[2] < 0 || msg.is_bigendian) return false;
    out->data = msg.data.data();
    out->n_points = static_cast<uint64_t>(msg.width) * msg.height;
    out->point_step = msg.point_step;
    out->off_x = static_cast<uint32_t>(off[0]);
    out->off_y = static_cast<uint32_t>(off[1]);
    out->off_z = static_cast<uint32_t>(off[2]);
    return true;
}

inline sensor_msgs::msg::PointField field(const char* name, uint32_t offset) {
    sensor_msgs::msg::PointField f;
    f.name = name;
    f.offset = offset;
    f.datatype = sensor_msgs::msg::PointField::FLOAT32;
    f.count = 1;
    return f;
}

// pcl::toROSMsg(pcl::PointCloud<pcl::PointXYZ>): x,y,z at 0/4/8, point_step 16.
// with_rgb: pcl::PointXYZRGB: additionally rgb (FLOAT32 bit-cast) at 16, point_step 32.
inline void shape_output(sensor_msgs::msg::PointCloud2& msg, uint64_t n_points, bool with_rgb) {
    msg.fields = {field("x", 0), field("y", 4), field("z", 8)};
    if (with_rgb) msg.fields.push_back(field("rgb", 16));
    msg.point_step = with_rgb ? 32 : 16;
    msg.height = 1;
    msg.width = static_cast<uint32_t>(n_points);
    msg.row_step = msg.point_step * msg.width;
    msg.is_bigendian = false;
    msg.is_dense = true;
    msg.data.resize(static_cast<size_t>(msg.row_step));
}

}  // namespace plc_ros
