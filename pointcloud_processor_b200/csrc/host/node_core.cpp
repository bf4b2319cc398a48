// node_core.cpp — ROS-free cores of the three GPU-backed nodes (include/plc_node.h).
//
// State and control flow of the reference's rclcpp::Node subclasses with every ROS/PCL type
// replaced by a POD; all point-cloud work goes through the C ABI of include/plc.h (CUDA) and the
// host setup steps of include/plc_host.h.  What each function mirrors:
//   plc_vlnode_*      SimplifiedDualLidarOptimizer, /root/reference/src/virtual_lidar.cpp:60-548
//   plc_filter_scan   SimplifiedScanMatcher::processCloudSimple, src/pointcloud_filter.cpp:64-139
//   plc_merge_clouds  GnssGicpMatcher::processPointClouds, src/pointcloud_merger.cpp:308-394
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>
#include <new>
#include <string>
#include <vector>

#include "plc_node.h"

namespace {

// What the node remembers of the last /excavated_terrain message.  The points themselves live in the context
// (plc_vl_set_terrain copied or read them during the callback; plc_cloud_view::data may be device memory, which
// the host must not touch): only the count and the reference's "cloud pointer is non-null" are kept.
struct StoredCloud {
    uint64_t n_points = 0;
    bool present = false;
    void store(const plc_cloud_view& v) {
        n_points = v.n_points;
        present = true;
    }
};

bool view_ok(const plc_cloud_view* v) {
    if (!v) return false;
    if (v->n_points == 0) return true;
    return v->data && v->point_step >= 12 && v->off_x + 4 <= v->point_step && v->off_y + 4 <= v->point_step &&
           v->off_z + 4 <= v->point_step;
}

// ZX120_OFFSET_*, ZX120_PITCH, ZX120_YAW, NORMAL_SEARCH_RADIUS (virtual_lidar.cpp:103-110)
constexpr double kFixedOffset[3] = {0.4, 0.5, 3.5};
constexpr double kFixedPitch = -M_PI / 6;
constexpr double kFixedYaw = 0.0;
constexpr double kNormalSearchRadius = 1.5;

double percent(int part, int total) { return total > 0 ? static_cast<double>(part) / total * 100.0 : 0.0; }

}  // namespace

struct plc_vlnode {
    plc_ctx* ctx = nullptr;
    plc_group* group = nullptr;      // set instead of ctx when the node drives several GPUs
    plc_ctx* setup_ctx() { return group ? plc_group_ctx(group, 0) : ctx; }
    plc_status set_terrain(const plc_cloud_view* c) { return group ? plc_group_set_terrain(group, c) : plc_vl_set_terrain(ctx, c); }
    plc_status set_terrain_unindexed() { return group ? plc_group_set_terrain_unindexed(group) : plc_vl_set_terrain_unindexed(ctx); }
    plc_status set_observed(const plc_cloud_view* c) { return group ? plc_group_set_observed(group, c) : plc_vl_set_observed(ctx, c); }
    plc_status set_cells(const plc_cell* c, uint64_t m) { return group ? plc_group_set_cells(group, c, m) : plc_vl_set_cells(ctx, c, m); }
    // declared parameters (:66-71) and the copies updateParameters() refreshes (:155-162)
    double p_grid_resolution = 0.1, p_sensor_height = 1.1, p_search_radius = 3.0, p_max_distance = 15.0;
    int64_t p_num_candidates = 100, p_vertical_layers = 10;
    double grid_resolution = 0.1, sensor_height = 1.1, search_radius = 3.0, max_distance = 15.0;
    int num_candidates = 100, vertical_layers = 10;
    // clouds
    StoredCloud terrain;            // terrain_cloud_
    bool terrain_indexed = false;   // terrain_kdtree_ exists
    uint64_t observed_points = 0;   // zx120_cloud_->size()
    // grid
    std::vector<plc_cell> cells;    // excavation_grid_3d_
    std::vector<uint8_t> flags;
    plc_grid_bounds bounds{};
    // TF
    bool have_tf = false;
    double base[3] = {0, 0, 0};
    // results
    std::vector<plc_pose> candidates;   // candidate_positions_
    std::vector<double> scores;
    plc_pose best{0.0, 0.0, 10.0, -M_PI / 2, 0.0};
    std::string log, log_out;

    void info(const char* fmt, ...) {
        char line[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(line, sizeof(line), fmt, ap);
        va_end(ap);
        log += line;
        log += '\n';
    }
    void update_parameters() {   // :155-162
        grid_resolution = p_grid_resolution;
        sensor_height = p_sensor_height;
        search_radius = p_search_radius;
        max_distance = p_max_distance;
        num_candidates = static_cast<int>(p_num_candidates);
        vertical_layers = static_cast<int>(p_vertical_layers);
    }
};

static plc_status vlnode_create(plc_ctx* ctx, plc_group* group, plc_vlnode** out) {
    plc_vlnode* n = new (std::nothrow) plc_vlnode();
    if (!n) return PLC_E_OOM;
    n->ctx = ctx;
    n->group = group;
    n->update_parameters();     // constructor, :73
    // a freshly constructed node has no clouds and no grid: drop whatever the context(s) held before
    plc_cloud_view none{};
    plc_status st = n->set_terrain(&none);
    if (st == PLC_OK) st = n->set_observed(&none);
    if (st == PLC_OK) st = n->set_cells(nullptr, 0);
    if (st != PLC_OK) {
        delete n;
        return st;
    }
    *out = n;
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_create(plc_ctx* ctx, plc_vlnode** out) {
    if (!ctx || !out) return PLC_E_INVALID_ARG;
    return vlnode_create(ctx, nullptr, out);
}

extern "C" plc_status plc_vlnode_create_on_group(plc_group* group, plc_vlnode** out) {
    if (!group || !out) return PLC_E_INVALID_ARG;
    return vlnode_create(nullptr, group, out);
}

extern "C" void plc_vlnode_destroy(plc_vlnode* node) { delete node; }

extern "C" plc_status plc_vlnode_set_parameter_double(plc_vlnode* n, const char* name, double v) {
    if (!n || !name) return PLC_E_INVALID_ARG;
    if (!strcmp(name, "grid_resolution")) n->p_grid_resolution = v;
    else if (!strcmp(name, "sensor_height")) n->p_sensor_height = v;
    else if (!strcmp(name, "search_radius")) n->p_search_radius = v;
    else if (!strcmp(name, "max_distance")) n->p_max_distance = v;
    else return PLC_E_INVALID_ARG;
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_set_parameter_int(plc_vlnode* n, const char* name, int64_t v) {
    if (!n || !name) return PLC_E_INVALID_ARG;
    if (!strcmp(name, "num_candidates")) n->p_num_candidates = v;
    else if (!strcmp(name, "vertical_layers")) n->p_vertical_layers = v;
    else return PLC_E_INVALID_ARG;
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_update_parameters(plc_vlnode* n) {
    if (!n) return PLC_E_INVALID_ARG;
    n->update_parameters();     // :73 — launch-time overrides hold from the first message on
    return PLC_OK;
}

// excavationAreaCallback (:164-178): normals, then the evaluation grid, with the parameter values of
// the last updateParameters().  An empty cloud leaves the previous grid in place (:168).
extern "C" plc_status plc_vlnode_excavation_area_callback(plc_vlnode* n, const plc_cloud_view* cloud) {
    if (!n || !view_ok(cloud)) return PLC_E_INVALID_ARG;
    if (cloud->n_points == 0) return PLC_OK;
    std::vector<float> normals(static_cast<size_t>(cloud->n_points) * 3);
    plc_ctx* dev = n->setup_ctx();            // both steps run on the device (plc_setup.cu)
    plc_status st = plc_vl_compute_normals(dev, cloud, kNormalSearchRadius, normals.data());       // :173
    if (st != PLC_OK) return st;
    uint64_t count = 0;
    plc_grid_bounds b{};
    // room for the last count (or 64 Ki cells) makes this one pass; a larger lattice reports its count and is redone
    std::vector<plc_cell> cells(std::max<size_t>(n->cells.size(), size_t(1) << 16));
    st = plc_vl_generate_cells(dev, cloud, normals.data(), n->grid_resolution, n->vertical_layers, // :174
                               kNormalSearchRadius, cells.data(), cells.size(), &count, &b);
    if (st == PLC_E_CAPACITY) {
        cells.resize(count);
        st = plc_vl_generate_cells(dev, cloud, normals.data(), n->grid_resolution, n->vertical_layers,
                                   kNormalSearchRadius, cells.data(), count, &count, &b);
    }
    if (st != PLC_OK) return st;
    cells.resize(count);
    n->cells.swap(cells);                     // excavation_grid_3d_.clear() + refill (:239, :278)
    n->flags.assign(n->cells.size(), 0);      // GridCell(): every flag false
    n->bounds = b;
    n->info("Generated 3D grid: %d valid cells across %d vertical layers", static_cast<int>(n->cells.size()),
            n->vertical_layers);              // :284-285
    return n->set_cells(n->cells.data(), n->cells.size());
}

// terrainCallback (:180-192): the cloud pointer is replaced even when empty, the index only when not.
extern "C" plc_status plc_vlnode_terrain_callback(plc_vlnode* n, const plc_cloud_view* cloud) {
    if (!n || !view_ok(cloud)) return PLC_E_INVALID_ARG;
    n->terrain.store(*cloud);
    // an empty message replaces the cloud but builds no tree: before any tree exists every ray is visible (:721, :727)
    if (cloud->n_points == 0) return n->terrain_indexed ? PLC_OK : n->set_terrain_unindexed();
    plc_status st = n->set_terrain(cloud);
    if (st == PLC_OK) n->terrain_indexed = true;
    return st;
}

// zx120PointsCallback (:194-207)
extern "C" plc_status plc_vlnode_zx120_points_callback(plc_vlnode* n, const plc_cloud_view* cloud) {
    if (!n || !view_ok(cloud)) return PLC_E_INVALID_ARG;
    n->observed_points = cloud->n_points;
    // an empty message empties zx120_cloud_, and checkVisibility then takes the ray path (:719) whatever
    // tree is left over: clearing the observed index has the same effect
    plc_status st = n->set_observed(cloud);
    if (st == PLC_OK && cloud->n_points) n->info("ZX120 point cloud updated: %zu points", static_cast<size_t>(cloud->n_points));   // :204
    return st;
}

extern "C" plc_status plc_vlnode_set_base_transform(plc_vlnode* n, int32_t available, double x, double y, double z) {
    if (!n) return PLC_E_INVALID_ARG;
    n->have_tf = available != 0;
    n->base[0] = x;
    n->base[1] = y;
    n->base[2] = z;
    return PLC_OK;
}

namespace {

plc_color_stats colour_stats(const std::vector<uint8_t>& flags, bool dual) {
    // :387-397 (fixed sensor only) and :487-501 (either sensor)
    plc_color_stats s{};
    for (uint8_t f : flags) {
        const bool range = (f & PLC_FLAG_RANGE_FIXED) || (dual && (f & PLC_FLAG_RANGE_MOBILE));
        const bool fov = (f & PLC_FLAG_FOV_FIXED) || (dual && (f & PLC_FLAG_FOV_MOBILE));
        const bool vis = (f & PLC_FLAG_VIS_FIXED) || (dual && (f & PLC_FLAG_VIS_MOBILE));
        ++s.total;
        if (!range) ++s.blue;
        else if (!fov) ++s.yellow;
        else if (!vis) ++s.red;
        else ++s.green;
    }
    return s;
}

void log_colour_block(plc_vlnode* n, const plc_color_stats& s) {
    const double ratio = s.green > 0 ? static_cast<double>(s.red) / s.green
                                     : (s.red > 0 ? std::numeric_limits<double>::infinity() : 0.0);
    const int unobservable = s.red + s.blue + s.yellow;
    n->info("  Total cells: %d", s.total);
    n->info("  Green (Observable): %d cells (%.1f%%)", s.green, percent(s.green, s.total));
    n->info("  Red (Occluded): %d cells (%.1f%%)", s.red, percent(s.red, s.total));
    n->info("  Blue (Out of range): %d cells (%.1f%%)", s.blue, percent(s.blue, s.total));
    n->info("  Yellow (Out of FOV): %d cells (%.1f%%)", s.yellow, percent(s.yellow, s.total));
    n->info("  ---");
    n->info("  Red/Green Ratio: %.3f", ratio);
    n->info("  Total Unobservable: %d cells (%.1f%%)", unobservable, percent(unobservable, s.total));
}

}  // namespace

// runOptimization (:454-548).  The fixed-sensor pass, the candidate loop and the argmax are one
// plc_vl_sweep; the candidate lattice is generated on the device as well.
extern "C" plc_status plc_vlnode_run_optimization(plc_vlnode* n, plc_vlnode_result* out) {
    if (!n || !out) return PLC_E_INVALID_ARG;
    memset(out, 0, sizeof(*out));
    out->best = n->best;
    out->best_score = -std::numeric_limits<double>::infinity();
    if (n->cells.empty() || !n->terrain.present || !n->have_tf) return PLC_OK;     // :455
    n->update_parameters();                                                         // :457

    plc_pose fixed{n->base[0] + kFixedOffset[0], n->base[1] + kFixedOffset[1], n->base[2] + kFixedOffset[2],
                   kFixedPitch, kFixedYaw};                                         // :347-351

    // generateCandidatePositions (:550-598): lattice side^2 is an upper bound of the count
    const int side = static_cast<int>(std::ceil(std::sqrt(static_cast<double>(n->num_candidates))));
    std::vector<plc_pose> cands(static_cast<size_t>(side) * side);
    uint64_t n_cands = 0;
    // on the device, against the terrain plc_vl_set_terrain left there; an empty cloud that arrived after it
    // replaced terrain_cloud_ but not the tree (:180-192), so getGroundHeight answers 0 everywhere (:601)
    const plc_cloud_view empty_cloud{};
    const plc_cloud_view* ground = n->terrain.n_points ? nullptr : &empty_cloud;
    plc_status st = plc_vl_generate_candidates(n->setup_ctx(), ground, &n->bounds, &fixed, n->num_candidates,
                                               n->search_radius, n->sensor_height, cands.data(), cands.size(),
                                               &n_cands);
    if (st != PLC_OK) return st;
    cands.resize(n_cands);

    plc_params prm;
    plc_params_default(&prm);
    prm.max_distance = n->max_distance;
    n->scores.assign(n_cands, 0.0);
    std::vector<double> fixed_scores(n->cells.size());
    plc_sweep_out so{};
    so.scores = n_cands ? n->scores.data() : nullptr;
    so.cell_flags = n->flags.data();
    so.fixed_scores = fixed_scores.data();
    st = n->group ? plc_group_sweep(n->group, &prm, &fixed, cands.data(), n_cands, &so)
                  : plc_vl_sweep(n->ctx, &prm, &fixed, cands.data(), n_cands, 0, n_cands, &so);
    if (st != PLC_OK) return st;

    // ---- evaluateZX120Only statistics (:360-452)
    out->ran = 1;
    out->fixed = fixed;
    out->fixed_total = so.fixed_total;
    out->n_cells = static_cast<int32_t>(n->cells.size());
    out->observed_points = n->observed_points;
    for (uint8_t f : n->flags) {
        out->range_ok += (f & PLC_FLAG_RANGE_FIXED) ? 1 : 0;
        out->fov_ok += (f & PLC_FLAG_FOV_FIXED) ? 1 : 0;
        out->visibility_ok += (f & PLC_FLAG_VIS_FIXED) ? 1 : 0;
    }
    out->zx120_only = colour_stats(n->flags, false);
    const int total = out->zx120_only.total;
    n->info("========================================");
    n->info("ZX120 LiDAR Only Evaluation");
    n->info("========================================");
    n->info("ZX120 Position: (%.2f, %.2f, %.2f)", fixed.x, fixed.y, fixed.z);
    n->info("Total Score (ZX120 only): %.2f", so.fixed_total);
    n->info("----------------------------------------");
    n->info("Debug Info:");
    n->info("  Cells in range: %d (%.1f%%)", out->range_ok, total > 0 ? out->range_ok * 100.0 / total : 0.0);
    n->info("  Cells in FOV: %d (%.1f%%)", out->fov_ok, total > 0 ? out->fov_ok * 100.0 / total : 0.0);
    n->info("  Cells visible: %d (%.1f%%)", out->visibility_ok, total > 0 ? out->visibility_ok * 100.0 / total : 0.0);
    n->info("  ZX120 point cloud size: %zu", static_cast<size_t>(n->observed_points));
    n->info("----------------------------------------");
    n->info("Color-based Area Analysis (ZX120 only):");
    log_colour_block(n, out->zx120_only);
    n->info("========================================");
    n->info("");

    // ---- best candidate (:464-478); none => LidarPosition() (Q5)
    plc_pose best{0.0, 0.0, 10.0, -M_PI / 2, 0.0};
    if (so.best_index >= 0) best = cands[static_cast<size_t>(so.best_index)];
    n->best = best;
    n->candidates.swap(cands);
    out->best = best;
    out->best_score = so.best_score;
    out->n_candidates = static_cast<int32_t>(n_cands);

    // ---- dual statistics (:480-543)
    out->dual = colour_stats(n->flags, true);
    n->info("========================================");
    n->info("Dual LiDAR Configuration (ZX120 + Mobile)");
    n->info("========================================");
    n->info("Best Mobile LiDAR Position: (%.2f, %.2f, %.2f)", best.x, best.y, best.z);
    n->info("Total Score: %.2f", so.best_score);
    n->info("========================================");
    n->info("Color-based Area Analysis:");
    log_colour_block(n, out->dual);
    n->info("========================================");
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_get_candidates(const plc_vlnode* n, plc_pose* out, uint64_t capacity, uint64_t* count) {
    if (!n || !count) return PLC_E_INVALID_ARG;
    *count = n->candidates.size();
    if (capacity < n->candidates.size()) return PLC_E_CAPACITY;
    if (!n->candidates.empty()) {
        if (!out) return PLC_E_INVALID_ARG;
        memcpy(out, n->candidates.data(), n->candidates.size() * sizeof(plc_pose));
    }
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_get_cells(const plc_vlnode* n, plc_cell* cells, uint8_t* flags, uint64_t capacity,
                                           uint64_t* count) {
    if (!n || !count) return PLC_E_INVALID_ARG;
    *count = n->cells.size();
    if (capacity < n->cells.size()) return PLC_E_CAPACITY;
    if (!n->cells.empty()) {
        if (cells) memcpy(cells, n->cells.data(), n->cells.size() * sizeof(plc_cell));
        if (flags) memcpy(flags, n->flags.data(), n->flags.size());
    }
    return PLC_OK;
}

extern "C" plc_status plc_vlnode_get_bounds(const plc_vlnode* n, plc_grid_bounds* out) {
    if (!n || !out) return PLC_E_INVALID_ARG;
    *out = n->bounds;
    return PLC_OK;
}

extern "C" const char* plc_vlnode_take_log(plc_vlnode* n) {
    if (!n) return "";
    n->log_out.swap(n->log);
    n->log.clear();
    return n->log_out.c_str();
}

// ------------------------------------------------------------------ pointcloud_filter -----------

extern "C" plc_status plc_filter_scan(plc_ctx* ctx, const plc_cloud_view* scan, double front_range, double side_range,
                                      double height_range, double voxel_leaf_size, void* out_xyz16,
                                      uint64_t capacity, uint64_t* n_out) {
    if (!ctx || !view_ok(scan) || !n_out) return PLC_E_INVALID_ARG;
    // cropFrontArea bounds (pointcloud_filter.cpp:111-113)
    plc_crop_box box{0.0, front_range, -side_range, side_range, -1.5, height_range};
    uint64_t n_cropped = 0;
    return plc_filter(ctx, scan, &box, voxel_leaf_size, out_xyz16, capacity, n_out, &n_cropped);
}

// ------------------------------------------------------------------ pointcloud_merger -----------

extern "C" plc_status plc_merge_clouds(plc_ctx* ctx, const plc_merger_input* robot, const plc_merger_input* backhoe,
                                       void* out_merged, uint64_t cap_merged, uint64_t* n_merged,
                                       void* out_robot, uint64_t cap_robot, uint64_t* n_robot,
                                       void* out_backhoe, uint64_t cap_backhoe, uint64_t* n_backhoe) {
    if (!ctx || !robot || !backhoe || !n_merged || !n_robot || !n_backhoe) return PLC_E_INVALID_ARG;
    const plc_merger_input* in[2] = {robot, backhoe};
    const uint8_t rgb[2][3] = {{255, 0, 0}, {0, 0, 255}};       // :316-325
    plc_merge_src srcs[2];
    uint64_t expect[2] = {0, 0};
    for (int i = 0; i < 2; ++i) {
        memset(&srcs[i], 0, sizeof(srcs[i]));
        const bool use = in[i]->has_cloud && in[i]->has_transform;      // :315/:321 and :389-393
        if (in[i]->has_cloud && !view_ok(&in[i]->cloud)) return PLC_E_INVALID_ARG;
        srcs[i].present = use ? 1 : 0;
        if (use) {
            srcs[i].cloud = in[i]->cloud;
            memcpy(srcs[i].t, in[i]->t, sizeof(srcs[i].t));
            memcpy(srcs[i].q, in[i]->q, sizeof(srcs[i].q));
            expect[i] = in[i]->cloud.n_points;
        }
        srcs[i].r = rgb[i][0];
        srcs[i].g = rgb[i][1];
        srcs[i].b = rgb[i][2];
    }
    if (expect[0] > cap_robot || expect[1] > cap_backhoe || expect[0] + expect[1] > cap_merged) return PLC_E_CAPACITY;
    void* each[2] = {out_robot, out_backhoe};
    uint64_t n_each[2] = {0, 0};
    *n_merged = *n_robot = *n_backhoe = 0;
    if (expect[0] + expect[1] == 0) return PLC_OK;               // nothing to publish (:327)
    plc_status st = plc_merge(ctx, srcs, 2, out_merged, cap_merged, n_merged, each, n_each);
    *n_robot = n_each[0];
    *n_backhoe = n_each[1];
    return st;
}
