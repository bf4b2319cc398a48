// generator.cpp — CPU restatement of the excavated_surface_generator node's data path.  TEST INFRASTRUCTURE
// (oracle.h): only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may use it.
//
// Follows /root/reference/src/excavated_surface_generator.cpp:
//   getExcavationBoxes        138-181
//   getTerrainHeight          183-225   (pcl::KdTreeFLANN radiusSearch + nearestKSearch: restated, SURVEY §A.3 item 2;
//                                        upstream rebuilds the tree in every call, here one index is shared)
//   isInsideAnyBox            228-236,  isOuterEdge 239-262
//   matchedCloudCallback      263-325   (centre and yaw from the TF: tf2::Transform / Matrix3x3::getRPY, restated)
//   isInsideExcavationArea    327-348
//   generateExcavationArea    350-455
//   processExcavation         457-481
//   generateExcavatedSurface  487-584
// excavated_surface_generator_L.cpp (not built upstream) is this algorithm with vertical wall points in the area
// cloud (its :391-407; plc_excavation_params::area_walls_vertical); the _square file is another formulation, not restated.
// Parity: the reference has no tests or fixtures for this node; tests/golden/refsrc_generator_*.npz hold what
// its own excavated_surface_generator.cpp (unmodified, stand-in headers, oracle/_ref) publishes, which this file reproduces.  KD-tree result order decides
// the order of the double sum in getTerrainHeight (restated as ascending distance, ties by index) and which of two
// equidistant points nearestKSearch returns (restated as the lower index).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "oracle.h"
#include "radius_index.hpp"

namespace {

struct Box {
    double center_x_local, center_y_local, length, width, min_x, max_x, min_y, max_y;
};

std::vector<Box> excavation_boxes(const plc_excavation_params& P) {          // :138-181
    std::vector<Box> boxes;
    if (P.l_shape_enabled) {
        Box arm1;
        arm1.center_x_local = 0.0;
        arm1.center_y_local = -P.arm1_length / 2.0;
        arm1.length = P.arm1_width;
        arm1.width = P.arm1_length;
        arm1.min_x = arm1.center_x_local - arm1.length / 2.0;
        arm1.max_x = arm1.center_x_local + arm1.length / 2.0;
        arm1.min_y = arm1.center_y_local - arm1.width / 2.0;
        arm1.max_y = arm1.center_y_local + arm1.width / 2.0;
        boxes.push_back(arm1);
        Box arm2;
        arm2.center_x_local = P.arm2_length / 2.0;
        arm2.center_y_local = -P.arm1_length + P.arm2_width / 2.0;
        arm2.length = P.arm2_length;
        arm2.width = P.arm2_width;
        arm2.min_x = arm2.center_x_local - arm2.length / 2.0;
        arm2.max_x = arm2.center_x_local + arm2.length / 2.0;
        arm2.min_y = arm2.center_y_local - arm2.width / 2.0;
        arm2.max_y = arm2.center_y_local + arm2.width / 2.0;
        boxes.push_back(arm2);
    } else {
        Box box;
        box.center_x_local = 0.0;
        box.center_y_local = 0.0;
        box.length = P.length;
        box.width = P.width;
        box.min_x = -P.length / 2.0;
        box.max_x = P.length / 2.0;
        box.min_y = -P.width / 2.0;
        box.max_y = P.width / 2.0;
        boxes.push_back(box);
    }
    return boxes;
}

bool inside_any_box(double x, double y, const std::vector<Box>& boxes) {      // :228-236
    for (const Box& b : boxes)
        if (x >= b.min_x && x <= b.max_x && y >= b.min_y && y <= b.max_y) return true;
    return false;
}

bool outer_edge(double x, double y, const std::vector<Box>& boxes, double tolerance) {   // :239-262
    if (!inside_any_box(x, y, boxes)) return false;
    const double delta = tolerance;
    bool has_outside_neighbor = false;
    if (!inside_any_box(x + delta, y, boxes)) has_outside_neighbor = true;
    if (!inside_any_box(x - delta, y, boxes)) has_outside_neighbor = true;
    if (!inside_any_box(x, y + delta, boxes)) has_outside_neighbor = true;
    if (!inside_any_box(x, y - delta, boxes)) has_outside_neighbor = true;
    return has_outside_neighbor;
}

struct Terrain {
    const float* xyz;                 // the cloud as stored (non-finite points included)
    uint64_t n;
    orc::RadiusIndex* index;          // over its finite points
    double radius;

    double height(double x, double y) const {                                // :183-225
        if (n == 0) return 0.0;
        const float qx = static_cast<float>(x), qy = static_cast<float>(y), qz = 0.0f;
        std::vector<int> idx;
        std::vector<float> d2;
        index->search(qx, qy, qz, orc::radius_to_r2f(radius), idx, d2, true);
        if (!idx.empty()) {
            // ascending distance, ties by index: the order the sum below runs in
            std::vector<std::pair<float, int>> order(idx.size());
            for (size_t k = 0; k < idx.size(); ++k) order[k] = {d2[k], idx[k]};
            std::sort(order.begin(), order.end());
            double sum_z = 0.0;
            int valid_count = 0;
            for (const auto& o : order) {
                const float* p = xyz + 3 * static_cast<uint64_t>(o.second);
                const double dx = p[0] - x, dy = p[1] - y;
                const double distance_2d = std::sqrt(dx * dx + dy * dy);
                if (distance_2d <= radius) {
                    sum_z += p[2];
                    valid_count++;
                }
            }
            if (valid_count > 0) return sum_z / valid_count;
        }
        // nearestKSearch(search_point, 1): FLANN L2_Simple in float32, lowest index among equals
        float best = std::numeric_limits<float>::infinity();
        int best_i = -1;
        for (size_t k = 0; k < index->size(); ++k) {
            const float d = orc::l2_simple(qx, qy, qz, index->x[k], index->y[k], index->z[k]);
            if (d < best) {
                best = d;
                best_i = index->src[k];
            }
        }
        if (best_i >= 0) return xyz[3 * static_cast<uint64_t>(best_i) + 2];
        return 0.0;
    }
};

struct Rec {                         // pcl::PointXYZRGB as pcl::toROSMsg writes it: 32 bytes
    float x, y, z, one;
    uint32_t rgba;
    uint32_t pad[3];
};
static_assert(sizeof(Rec) == 32, "PointXYZRGB wire record");

Rec make_rec(double x, double y, double z, unsigned r, unsigned g, unsigned b) {
    Rec p{};
    p.x = static_cast<float>(x);
    p.y = static_cast<float>(y);
    p.z = static_cast<float>(z);
    p.one = 1.0f;
    p.rgba = 0xff000000u | (r << 16) | (g << 8) | b;
    return p;
}

// generateExcavationArea, :350-455
void excavation_area(std::vector<Rec>& cloud, const double center[3], double yaw, const Terrain& T,
                     const plc_excavation_params& P) {
    const std::vector<Box> boxes = excavation_boxes(P);
    const double slope_offset = P.depth / std::tan(P.slope_angle_deg * M_PI / 180.0);
    double overall_min_x = std::numeric_limits<double>::max(), overall_max_x = std::numeric_limits<double>::lowest();
    double overall_min_y = std::numeric_limits<double>::max(), overall_max_y = std::numeric_limits<double>::lowest();
    for (const Box& b : boxes) {
        overall_min_x = std::min(overall_min_x, b.min_x);
        overall_max_x = std::max(overall_max_x, b.max_x);
        overall_min_y = std::min(overall_min_y, b.min_y);
        overall_max_y = std::max(overall_max_y, b.max_y);
    }
    const double pd = P.point_density;
    const int n_x = static_cast<int>((overall_max_x - overall_min_x) / pd) + 1;
    const int n_y = static_cast<int>((overall_max_y - overall_min_y) / pd) + 1;
    const int n_depth = static_cast<int>(P.depth / pd);
    for (int i = 0; i <= n_x; ++i)
        for (int j = 0; j <= n_y; ++j) {
            const double x_local = overall_min_x + i * pd;
            const double y_local = overall_min_y + j * pd;
            if (!inside_any_box(x_local, y_local, boxes)) continue;
            const double x_global = center[0] + x_local * std::cos(yaw) - y_local * std::sin(yaw);
            const double y_global = center[1] + x_local * std::sin(yaw) + y_local * std::cos(yaw);
            const double local_terrain_height = T.height(x_global, y_global);
            cloud.push_back(make_rec(x_global, y_global, local_terrain_height - P.depth, 255, 255, 0));
            if (outer_edge(x_local, y_local, boxes, pd)) {
                for (int k = 1; k < n_depth; ++k) {
                    const double z_ratio = static_cast<double>(k) / n_depth;
                    const double offset = slope_offset * z_ratio;
                    double offset_x = 0.0, offset_y = 0.0;
                    if (!inside_any_box(x_local + pd, y_local, boxes)) offset_x = offset;
                    else if (!inside_any_box(x_local - pd, y_local, boxes)) offset_x = -offset;
                    if (!inside_any_box(x_local, y_local + pd, boxes)) offset_y = offset;
                    else if (!inside_any_box(x_local, y_local - pd, boxes)) offset_y = -offset;
                    const double x_slope = x_local + offset_x, y_slope = y_local + offset_y;
                    const double xg = center[0] + x_slope * std::cos(yaw) - y_slope * std::sin(yaw);
                    const double yg = center[1] + x_slope * std::sin(yaw) + y_slope * std::cos(yaw);
                    const double z_local = local_terrain_height - P.depth + k * pd;
                    // excavated_surface_generator_L.cpp:399-400 keeps the wall points straight above the edge
                    if (P.area_walls_vertical) cloud.push_back(make_rec(x_global, y_global, z_local, 200, 200, 0));
                    else cloud.push_back(make_rec(xg, yg, z_local, 200, 200, 0));
                }
            }
        }
}

// isInsideExcavationArea, :327-348
bool inside_excavation(double x_local, double y_local, double z_relative, const std::vector<Box>& boxes,
                       const plc_excavation_params& P) {
    if (z_relative < -P.depth || z_relative > 0) return false;
    const double slope_offset = P.depth / std::tan(P.slope_angle_deg * M_PI / 180.0);
    const double slope_factor = (P.depth + z_relative) / P.depth;
    const double current_offset = slope_offset * slope_factor;
    for (const Box& b : boxes) {
        const double dx = x_local - b.center_x_local, dy = y_local - b.center_y_local;
        const double half_length = b.length / 2.0 + current_offset, half_width = b.width / 2.0 + current_offset;
        if (std::abs(dx) <= half_length && std::abs(dy) <= half_width) return true;
    }
    return false;
}

// generateExcavatedSurface, :487-584
void excavated_surface(std::vector<Rec>& cloud, const double center[3], double yaw, const Terrain& T,
                       const plc_excavation_params& P) {
    const std::vector<Box> boxes = excavation_boxes(P);
    const double slope_offset = P.depth / std::tan(P.slope_angle_deg * M_PI / 180.0);
    double overall_min_x = std::numeric_limits<double>::max(), overall_max_x = std::numeric_limits<double>::lowest();
    double overall_min_y = std::numeric_limits<double>::max(), overall_max_y = std::numeric_limits<double>::lowest();
    for (const Box& b : boxes) {
        overall_min_x = std::min(overall_min_x, b.min_x);
        overall_max_x = std::max(overall_max_x, b.max_x);
        overall_min_y = std::min(overall_min_y, b.min_y);
        overall_max_y = std::max(overall_max_y, b.max_y);
    }
    const double pd = P.point_density;
    const int n_x = static_cast<int>((overall_max_x - overall_min_x) / pd) + 1;
    const int n_y = static_cast<int>((overall_max_y - overall_min_y) / pd) + 1;
    for (int i = 0; i <= n_x; ++i)
        for (int j = 0; j <= n_y; ++j) {
            const double x_local = overall_min_x + i * pd, y_local = overall_min_y + j * pd;
            if (!inside_any_box(x_local, y_local, boxes)) continue;
            const double x_global = center[0] + x_local * std::cos(yaw) - y_local * std::sin(yaw);
            const double y_global = center[1] + x_local * std::sin(yaw) + y_local * std::cos(yaw);
            const double h = T.height(x_global, y_global);
            cloud.push_back(make_rec(x_global, y_global, h - P.depth, 0, 139, 0));
        }
    const int n_slope = static_cast<int>(slope_offset / pd) + 1;
    for (int i = 0; i <= n_x; ++i)
        for (int j = 0; j <= n_y; ++j) {
            const double x_local = overall_min_x + i * pd, y_local = overall_min_y + j * pd;
            if (!outer_edge(x_local, y_local, boxes, pd)) continue;
            for (int k = 0; k <= n_slope; ++k) {
                const double z_ratio = static_cast<double>(k) / n_slope;
                const double offset = slope_offset * z_ratio;
                double offset_x = 0.0, offset_y = 0.0;
                if (!inside_any_box(x_local + pd, y_local, boxes)) offset_x = offset;
                else if (!inside_any_box(x_local - pd, y_local, boxes)) offset_x = -offset;
                if (!inside_any_box(x_local, y_local + pd, boxes)) offset_y = offset;
                else if (!inside_any_box(x_local, y_local - pd, boxes)) offset_y = -offset;
                const double x_slope = x_local + offset_x, y_slope = y_local + offset_y;
                const double x_global = center[0] + x_slope * std::cos(yaw) - y_slope * std::sin(yaw);
                const double y_global = center[1] + x_slope * std::sin(yaw) + y_slope * std::cos(yaw);
                const double h = T.height(x_global, y_global);
                const double z_local = h - P.depth * (1.0 - z_ratio);
                cloud.push_back(make_rec(x_global, y_global, z_local, 144, 238, 144));
            }
        }
}

}  // namespace

// tf2::Transform from (t, q): Matrix3x3::setRotation, operator*(Vector3), Matrix3x3::getRPY's yaw (:280-302)
extern "C" void orc_excavation_frame(const double* t, const double* q, double offset_x, double offset_y, double* center_xy,
                                     double* yaw) {
    const double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    const double s = 2.0 / d;
    const double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
    const double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
    const double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
    const double yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
    const double m[3][3] = {{1.0 - (yy + zz), xy - wz, xz + wy}, {xy + wz, 1.0 - (xx + zz), yz - wx}, {xz - wy, yz + wx, 1.0 - (xx + yy)}};
    // basis * v + origin, v = (offset_x, offset_y, 0); Vector3::dot adds x, y, z terms left to right
    center_xy[0] = (m[0][0] * offset_x + m[0][1] * offset_y + m[0][2] * 0.0) + t[0];
    center_xy[1] = (m[1][0] * offset_x + m[1][1] * offset_y + m[1][2] * 0.0) + t[1];
    // getRPY -> getEulerYPR, solution 1
    if (std::fabs(m[2][0]) >= 1.0) {
        *yaw = 0.0;
    } else {
        const double pitch = -std::asin(m[2][0]);
        *yaw = std::atan2(m[1][0] / std::cos(pitch), m[0][0] / std::cos(pitch));
    }
}

// matchedCloudCallback's data path (:280-310).  xyz: n packed float triples; rgba: n packed colours as stored in
// the PointXYZRGB cloud.  out_terrain needs n + surface points, out_area the area points (orc_excavation_counts).
extern "C" void orc_excavate(const float* xyz, const uint32_t* rgba, uint64_t n, const double* base_t, const double* base_q,
                             const plc_excavation_params* P, void* out_terrain, uint64_t* n_terrain, void* out_area,
                             uint64_t* n_area, plc_excavation_info* info) {
    double center[3] = {0, 0, 0}, yaw = 0;
    orc_excavation_frame(base_t, base_q, P->offset_x, P->offset_y, center, &yaw);
    orc::RadiusIndex* ix = orc::make_index(xyz, n, ORC_INDEX_GRID);
    Terrain T{xyz, n, ix, P->terrain_search_radius};
    center[2] = T.height(center[0], center[1]);                              // :289-297
    std::vector<Rec> area, result;
    excavation_area(area, center, yaw, T, *P);                               // :305-306
    const std::vector<Box> boxes = excavation_boxes(*P);
    uint64_t kept = 0;
    for (uint64_t i = 0; i < n; ++i) {                                       // processExcavation, :457-472
        const float* p = xyz + 3 * i;
        const double dx = p[0] - center[0], dy = p[1] - center[1];
        const double x_local = dx * std::cos(-yaw) - dy * std::sin(-yaw);
        const double y_local = dx * std::sin(-yaw) + dy * std::cos(-yaw);
        // a non-finite x or y makes every comparison of the carve test false whatever the height is (upstream's
        // KD-tree query with such a point is undefined): the height is not asked for
        const double h = (std::isfinite(p[0]) && std::isfinite(p[1])) ? T.height(p[0], p[1]) : 0.0;
        const double z_rel = p[2] - h;
        if (!inside_excavation(x_local, y_local, z_rel, boxes, *P)) {
            Rec r{};
            r.x = p[0];
            r.y = p[1];
            r.z = p[2];
            r.one = 1.0f;
            r.rgba = rgba ? rgba[i] : 0xff000000u;
            result.push_back(r);
            ++kept;
        }
    }
    excavated_surface(result, center, yaw, T, *P);                           // :474
    delete ix;
    if (out_terrain && !result.empty()) std::memcpy(out_terrain, result.data(), result.size() * sizeof(Rec));
    if (out_area && !area.empty()) std::memcpy(out_area, area.data(), area.size() * sizeof(Rec));
    *n_terrain = result.size();
    *n_area = area.size();
    if (info) {
        info->center[0] = center[0];
        info->center[1] = center[1];
        info->center[2] = center[2];
        info->yaw = yaw;
        info->n_kept = kept;
        info->n_surface = result.size() - kept;
        info->n_area = area.size();
    }
}
