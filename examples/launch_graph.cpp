// launch_graph.cpp — a C++ host for the whole dataflow of launch/start.launch.xml, written against the C ABI only
// (include/plc.h, include/plc_node.h; no CUDA headers, no Python): what a maintainer's composed-node container, or a
// rosbag replay tool, would do with libplc.so.
//
//   pointcloud_filter x2 -> pointcloud_merger -> excavated_surface_generator -> virtual_lidar (normals, evaluation
//   cells, candidate poses, sweep), calc_drivable_area beside it on the robot's filtered scan.
//
//   launch_graph [--robot scan.f32 --zx120 scan.f32] [--resident] [--candidates N] [--res R] [--layers L] [--reps K]
//
// scan.f32: raw little-endian float32 PointXYZ records (x y z pad), sensor frame.  Without files, two synthetic scans
// of a rolling ground are generated.  --resident keeps every intermediate cloud in device buffers (plc_dev_alloc).
// Prints one JSON object: sizes, the optimal pose, its score and the wall time of the graph.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "plc.h"
#include "plc_node.h"

namespace {

struct Tf {
    double t[3];
    double q[4];   // x y z w
};

// static_transform_publisher "x y z yaw pitch roll" (roll = 0) composed with a base pose (xy, yaw) in the map
Tf sensor_in_map(double bx, double by, double byaw, double sx, double sy, double sz, double pitch) {
    const double cy = std::cos(byaw / 2), sy2 = std::sin(byaw / 2), cp = std::cos(pitch / 2), sp = std::sin(pitch / 2);
    Tf tf;
    tf.q[0] = -sy2 * sp;
    tf.q[1] = cy * sp;
    tf.q[2] = sy2 * cp;
    tf.q[3] = cy * cp;
    tf.t[0] = bx + std::cos(byaw) * sx - std::sin(byaw) * sy;
    tf.t[1] = by + std::sin(byaw) * sx + std::cos(byaw) * sy;
    tf.t[2] = sz;
    return tf;
}

void rotation(const double q[4], double m[3][3]) {
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    m[0][0] = 1 - 2 * (y * y + z * z); m[0][1] = 2 * (x * y - z * w);     m[0][2] = 2 * (x * z + y * w);
    m[1][0] = 2 * (x * y + z * w);     m[1][1] = 1 - 2 * (x * x + z * z); m[1][2] = 2 * (y * z - x * w);
    m[2][0] = 2 * (x * z - y * w);     m[2][1] = 2 * (y * z + x * w);     m[2][2] = 1 - 2 * (x * x + y * y);
}

// ground points seen from a sensor at `tf`, expressed in the sensor frame (p_s = R^T (p_m - t))
std::vector<float> synthetic_scan(const Tf& tf, size_t n, unsigned seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> ux(tf.t[0] - 18.0, tf.t[0] + 18.0), uy(tf.t[1] - 18.0, tf.t[1] + 18.0), un(-0.01, 0.01);
    double m[3][3];
    rotation(tf.q, m);
    std::vector<float> out(n * 4);
    for (size_t i = 0; i < n; ++i) {
        const double x = ux(rng), y = uy(rng), z = 0.15 * std::sin(0.4 * x) * std::cos(0.3 * y) + un(rng);
        const double d[3] = {x - tf.t[0], y - tf.t[1], z - tf.t[2]};
        for (int k = 0; k < 3; ++k) out[4 * i + k] = static_cast<float>(m[0][k] * d[0] + m[1][k] * d[1] + m[2][k] * d[2]);
        out[4 * i + 3] = 1.0f;
    }
    return out;
}

bool read_scan(const char* path, std::vector<float>* out) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize(static_cast<size_t>(bytes) / 4);
    const size_t got = std::fread(out->data(), 4, out->size(), f);
    std::fclose(f);
    return got == out->size() && out->size() % 4 == 0;
}

#define CHECK(expr)                                                                                             \
    do {                                                                                                        \
        const plc_status st__ = (expr);                                                                         \
        if (st__ != PLC_OK) {                                                                                   \
            std::fprintf(stderr, "%s: %s (%s)\n", #expr, plc_strerror(st__), ctx ? plc_last_error(ctx) : "");   \
            return 1;                                                                                           \
        }                                                                                                       \
    } while (0)

// a cloud buffer that lives on the host or on the device
struct Cloud {
    plc_ctx* ctx = nullptr;
    bool on_device = false;
    void* dev = nullptr;
    std::vector<unsigned char> host;
    uint64_t capacity_bytes = 0;
    plc_status reserve(plc_ctx* c, bool device, uint64_t bytes) {
        ctx = c;
        on_device = device;
        if (bytes <= capacity_bytes) return PLC_OK;
        capacity_bytes = bytes + bytes / 4 + 256;
        if (!device) {
            host.resize(capacity_bytes);
            return PLC_OK;
        }
        if (dev) plc_dev_free(ctx, dev);
        return plc_dev_alloc(ctx, capacity_bytes, &dev);
    }
    void* data() { return on_device ? dev : static_cast<void*>(host.data()); }
    plc_cloud_view view(uint64_t n, uint32_t step) {
        plc_cloud_view v{};
        v.data = n ? data() : nullptr;
        v.n_points = n;
        v.point_step = step;
        v.off_x = 0;
        v.off_y = 4;
        v.off_z = 8;
        return v;
    }
    void release() {
        if (dev) plc_dev_free(ctx, dev);
        dev = nullptr;
        capacity_bytes = 0;
    }
    ~Cloud() { release(); }
};

}  // namespace

int main(int argc, char** argv) {
    const char *robot_path = nullptr, *zx120_path = nullptr;
    bool resident = false;
    int num_candidates = 100, layers = 10, reps = 1, device = 0;
    double res = 0.1;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--robot") robot_path = next();
        else if (a == "--zx120") zx120_path = next();
        else if (a == "--resident") resident = true;
        else if (a == "--candidates") num_candidates = std::atoi(next());
        else if (a == "--layers") layers = std::atoi(next());
        else if (a == "--res") res = std::atof(next());
        else if (a == "--reps") reps = std::atoi(next());
        else if (a == "--device") device = std::atoi(next());
        else {
            std::fprintf(stderr, "unknown argument %s\n", a.c_str());
            return 2;
        }
    }
    // zx120/base_link at the map origin (pointcloud_merger.cpp:218-228); static TFs of launch/tf_*.launch.xml:4
    const double robot_base[3] = {9.0, -3.0, 2.6};
    const Tf zx = sensor_in_map(0.0, 0.0, 0.0, 0.55, 0.4, 3.5, 0.4363);
    const Tf rb = sensor_in_map(robot_base[0], robot_base[1], robot_base[2], 0.25, 0.0, 2.0, 0.1245);
    std::vector<float> robot_scan, zx120_scan;
    if (robot_path && zx120_path) {
        if (!read_scan(robot_path, &robot_scan) || !read_scan(zx120_path, &zx120_scan)) {
            std::fprintf(stderr, "cannot read the scan files\n");
            return 2;
        }
    } else {
        robot_scan = synthetic_scan(rb, 400000, 4);
        zx120_scan = synthetic_scan(zx, 400000, 5);
    }

    plc_ctx* ctx = nullptr;
    CHECK(plc_ctx_create(device, nullptr, &ctx));       // no device: fails here, there is no CPU fallback
    plc_excavation_params eprm;
    plc_excavation_params_default(&eprm);
    uint64_t n_surface = 0, n_area_bound = 0;
    CHECK(plc_excavation_counts(&eprm, &n_surface, &n_area_bound));

    Cloud f_robot, f_zx, merged, own_r, own_b, terrain, area, normals;
    std::vector<plc_cell> cells;
    std::vector<plc_pose> cands;
    std::vector<double> scores;
    std::vector<int8_t> grid;
    uint64_t n_fr = 0, n_fz = 0, n_m = 0, n_or = 0, n_ob = 0, n_t = 0, n_a = 0, n_cells = 0, n_cands = 0;
    plc_excavation_info einfo{};
    plc_sweep_out so{};
    plc_drivable_info dinfo{};
    plc_grid_bounds bounds{};
    double wall_ms = 0.0;

    for (int rep = 0; rep < reps; ++rep) {
        const auto t0 = std::chrono::steady_clock::now();
        // pointcloud_filter (defaults 15 / 10 / 10, leaf 0.2; pointcloud_filter.cpp:30-39)
        plc_cloud_view vr{robot_scan.data(), robot_scan.size() / 4, 16, 0, 4, 8}, vz{zx120_scan.data(), zx120_scan.size() / 4, 16, 0, 4, 8};
        CHECK(f_robot.reserve(ctx, resident, vr.n_points * 16));
        CHECK(f_zx.reserve(ctx, resident, vz.n_points * 16));
        CHECK(plc_filter_scan(ctx, &vr, 15.0, 10.0, 10.0, 0.2, f_robot.data(), vr.n_points, &n_fr));
        CHECK(plc_filter_scan(ctx, &vz, 15.0, 10.0, 10.0, 0.2, f_zx.data(), vz.n_points, &n_fz));
        // pointcloud_merger
        plc_merger_input in_r{}, in_z{};
        in_r.cloud = f_robot.view(n_fr, 16);
        in_z.cloud = f_zx.view(n_fz, 16);
        in_r.has_cloud = in_r.has_transform = in_z.has_cloud = in_z.has_transform = 1;
        std::memcpy(in_r.t, rb.t, sizeof(rb.t));
        std::memcpy(in_r.q, rb.q, sizeof(rb.q));
        std::memcpy(in_z.t, zx.t, sizeof(zx.t));
        std::memcpy(in_z.q, zx.q, sizeof(zx.q));
        CHECK(merged.reserve(ctx, resident, (n_fr + n_fz) * 32));
        CHECK(own_r.reserve(ctx, resident, n_fr * 32));
        CHECK(own_b.reserve(ctx, resident, n_fz * 32));
        CHECK(plc_merge_clouds(ctx, &in_r, &in_z, merged.data(), n_fr + n_fz, &n_m, own_r.data(), n_fr, &n_or, own_b.data(), n_fz, &n_ob));
        // excavated_surface_generator
        const double base_t[3] = {0.0, 0.0, 0.0}, base_q[4] = {0.0, 0.0, 0.0, 1.0};
        CHECK(terrain.reserve(ctx, resident, (n_m + n_surface) * 32));
        CHECK(area.reserve(ctx, resident, n_area_bound * 32));
        plc_cloud_view vm = merged.view(n_m, 32);
        CHECK(plc_excavate(ctx, &vm, 16, base_t, base_q, &eprm, terrain.data(), n_m + n_surface, &n_t, area.data(), n_area_bound, &n_a, &einfo));
        // virtual_lidar: excavationAreaCallback, terrainCallback, zx120PointsCallback, runOptimization
        plc_cloud_view va = area.view(n_a, 32), vt = terrain.view(n_t, 32), vo = f_zx.view(n_fz, 16);
        CHECK(normals.reserve(ctx, resident, n_a * 12));
        CHECK(plc_vl_compute_normals(ctx, &va, 1.5, static_cast<float*>(normals.data())));
        // room for the last count (64 Ki cells on the first pass): one call when it fits, else the count comes back
        if (cells.size() < (size_t(1) << 16)) cells.resize(size_t(1) << 16);
        plc_status st = plc_vl_generate_cells(ctx, &va, static_cast<const float*>(normals.data()), res, layers, 1.5, cells.data(), cells.size(), &n_cells, &bounds);
        if (st == PLC_E_CAPACITY) {
            cells.resize(n_cells);
            st = plc_vl_generate_cells(ctx, &va, static_cast<const float*>(normals.data()), res, layers, 1.5, cells.data(), n_cells, &n_cells, &bounds);
        }
        CHECK(st);
        CHECK(plc_vl_set_terrain(ctx, &vt));
        CHECK(plc_vl_set_cells(ctx, cells.data(), n_cells));
        CHECK(plc_vl_set_observed(ctx, &vo));
        const plc_pose fixed{0.4, 0.5, 3.5, -M_PI / 6, 0.0};                       // virtual_lidar.cpp:103-107, :347-351
        const int side = static_cast<int>(std::ceil(std::sqrt(static_cast<double>(num_candidates))));
        cands.resize(static_cast<size_t>(side) * side);
        CHECK(plc_vl_generate_candidates(ctx, nullptr, &bounds, &fixed, num_candidates, 3.0, 1.1, cands.data(), cands.size(), &n_cands));
        plc_params prm;
        plc_params_default(&prm);
        scores.assign(n_cands, 0.0);
        so = plc_sweep_out{};
        so.scores = n_cands ? scores.data() : nullptr;
        CHECK(plc_vl_sweep(ctx, &prm, &fixed, cands.data(), n_cands, 0, n_cands, &so));
        // calc_drivable_area
        plc_drivable_params dprm;
        plc_drivable_params_default(&dprm);
        grid.assign(static_cast<size_t>(dprm.map_width / dprm.grid_resolution) * static_cast<size_t>(dprm.map_height / dprm.grid_resolution), 0);
        plc_cloud_view vfr = f_robot.view(n_fr, 16);
        CHECK(plc_drivable_area(ctx, &vfr, rb.t, rb.q, robot_base[0], robot_base[1], robot_base[0], robot_base[1], &dprm, grid.data(), grid.size(), &dinfo));
        wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    long occupied = 0, unknown = 0;
    for (int8_t v : grid) {
        occupied += v == 100;
        unknown += v == -1;
    }
    const plc_pose best = so.best_index >= 0 ? cands[static_cast<size_t>(so.best_index)] : plc_pose{0, 0, 10, -M_PI / 2, 0};
    std::printf("{\"resident\": %s, \"robot_filtered\": %llu, \"zx120_filtered\": %llu, \"merged\": %llu, \"terrain\": %llu, "
                "\"kept\": %llu, \"area\": %llu, \"cells\": %llu, \"candidates\": %llu, \"best_index\": %lld, "
                "\"best_pose\": [%.17g, %.17g, %.17g], \"best_score\": %.17g, \"fixed_total\": %.17g, \"rays_visible\": %llu, "
                "\"grid_occupied\": %ld, \"grid_unknown\": %ld, \"kernels_launched\": %llu, \"wall_ms\": %.3f}\n",
                resident ? "true" : "false", (unsigned long long)n_fr, (unsigned long long)n_fz, (unsigned long long)n_m,
                (unsigned long long)n_t, (unsigned long long)einfo.n_kept, (unsigned long long)n_a, (unsigned long long)n_cells,
                (unsigned long long)n_cands, (long long)so.best_index, best.x, best.y, best.z, so.best_score, so.fixed_total,
                (unsigned long long)so.counters.rays_visible, occupied, unknown, (unsigned long long)plc_launch_count(ctx), wall_ms);
    for (Cloud* c : {&f_robot, &f_zx, &merged, &own_r, &own_b, &terrain, &area, &normals}) c->release();   // before their context goes
    plc_ctx_destroy(ctx);
    return 0;
}
