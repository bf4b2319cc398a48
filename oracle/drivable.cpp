// drivable.cpp — CPU restatement of calc_drivable_area's occupancy grid.  TEST INFRASTRUCTURE (oracle.h): only
// tests/, __graft_entry__.smoke() and bench.py's CPU baseline may use it.
//
// Follows /root/reference/src/calc_drivable_area.cpp:
//   grid size                          36-37
//   transform into the map frame       87-92   (tf2_sensor_msgs doTransform: float32 affine, oracle's orc_transform)
//   binning of z values                137-160
//   occupancy decision                 178-222
//   calculateGradient                  238-250
// Parity: the reference has no tests or fixtures for this node; tests/golden/refsrc_drivable.npz holds what
// its own calc_drivable_area.cpp (unmodified, stand-in PCL / tf2 / ROS2 headers, oracle/_ref) publishes, which this file reproduces.
#include <algorithm>
#include <cmath>
#include <vector>

#include "oracle.h"

extern "C" int64_t orc_drivable_area(const float* xyz, uint64_t n, const double* t, const double* q, double robot_x,
                                     double robot_y, double start_x, double start_y, double grid_resolution,
                                     double map_width, double map_height, double max_gradient, int min_points_per_cell,
                                     double start_clear_radius, int8_t* out, int32_t* width_out, int32_t* height_out) {
    const int grid_width = static_cast<int>(map_width / grid_resolution);      // :36
    const int grid_height = static_cast<int>(map_height / grid_resolution);    // :37
    if (width_out) *width_out = grid_width;
    if (height_out) *height_out = grid_height;
    if (n == 0) return -1;                                                     // empty cloud: nothing is published (:108-112)
    std::vector<float> map_xyz(3 * n);
    if (t && q) orc_transform(xyz, n, t, q, map_xyz.data());                   // :87-92
    else std::copy(xyz, xyz + 3 * n, map_xyz.begin());

    const double map_origin_x = robot_x - map_width / 2.0;                     // :137
    const double map_origin_y = robot_y - map_height / 2.0;                    // :138
    struct Cell {
        std::vector<float> z_values;
        bool has_points = false;
    };
    std::vector<Cell> grid_map(static_cast<size_t>(grid_width) * grid_height);
    for (uint64_t i = 0; i < n; ++i) {                                         // :142-160
        const float px = map_xyz[3 * i], py = map_xyz[3 * i + 1], pz = map_xyz[3 * i + 2];
        if (!std::isfinite(px) || !std::isfinite(py) || !std::isfinite(pz)) continue;
        const double fx = (px - map_origin_x) / grid_resolution, fy = (py - map_origin_y) / grid_resolution;
        // the C++ cast is undefined outside int's range; such points are outside every map
        if (!(std::abs(fx) < 2.0e9 && std::abs(fy) < 2.0e9)) continue;
        const int grid_x = static_cast<int>(fx), grid_y = static_cast<int>(fy);
        if (grid_x >= 0 && grid_x < grid_width && grid_y >= 0 && grid_y < grid_height) {
            Cell& c = grid_map[static_cast<size_t>(grid_y) * grid_width + grid_x];
            c.z_values.push_back(pz);
            c.has_points = true;
        }
    }
    for (int y = 0; y < grid_height; ++y)                                      // :178-222
        for (int x = 0; x < grid_width; ++x) {
            const size_t index = static_cast<size_t>(y) * grid_width + x;
            const double cell_x = map_origin_x + (x + 0.5) * grid_resolution;
            const double cell_y = map_origin_y + (y + 0.5) * grid_resolution;
            const double dist_to_start = std::sqrt(std::pow(cell_x - start_x, 2) + std::pow(cell_y - start_y, 2));
            const Cell& c = grid_map[index];
            if (dist_to_start <= start_clear_radius) {
                out[index] = 0;
            } else if (!c.has_points || c.z_values.size() < static_cast<size_t>(min_points_per_cell)) {
                out[index] = -1;
            } else {
                float gradient = 0.0f;                                         // calculateGradient, :238-250
                if (c.z_values.size() >= 2) {
                    const float min_z = *std::min_element(c.z_values.begin(), c.z_values.end());
                    const float max_z = *std::max_element(c.z_values.begin(), c.z_values.end());
                    gradient = (max_z - min_z) / grid_resolution;
                }
                out[index] = gradient > max_gradient ? 100 : 0;
            }
        }
    return static_cast<int64_t>(grid_width) * grid_height;
}
