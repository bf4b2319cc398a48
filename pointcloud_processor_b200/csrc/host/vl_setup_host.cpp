// vl_setup_host.cpp — host-side setup of the virtual_lidar node: evaluation-cell lattice, cell
// normals, candidate poses.  These run once per /excavation_area message or once per sweep on a
// few thousand items; they are the sweep's *inputs*, not the hot path.
//
// Reference behaviour reproduced (file:line in /root/reference/src/virtual_lidar.cpp):
//   generateExcavationGrid3D 236-287, isPointNearExcavation 289-299, computeCellSurfaceNormal 301-340,
//   computeTerrainNormals 209-234, generateCandidatePositions 550-598, getGroundHeight 600-625.
#include <cmath>
#include <limits>
#include <utility>
#include <vector>

#include "plc_host.h"
#include "point_buckets.hpp"

using plc_host::Buckets;
using plc_host::HostCloud;

namespace {

inline float radius_sq_f32(double radius) { return static_cast<float>(radius * radius); }

bool view_ok(const plc_cloud_view* v) {
    if (!v) return false;
    if (v->n_points == 0) return true;
    return v->data && v->point_step >= 12 && v->off_x + 4 <= v->point_step &&
           v->off_y + 4 <= v->point_step && v->off_z + 4 <= v->point_step;
}

// Eigenvector of the smallest eigenvalue of a symmetric 3x3 matrix: closed-form roots, then the
// best-conditioned cross product of two rows of (A - lambda I).
void min_eigenvector_sym3(const double m_in[6], double out[3]) {
    // m = [xx xy xz yy yz zz]
    double scale = 0.0;
    for (int i = 0; i < 6; ++i) scale = std::max(scale, std::abs(m_in[i]));
    if (scale <= 0.0) { out[0] = 0; out[1] = 0; out[2] = 1; return; }
    double a00 = m_in[0] / scale, a01 = m_in[1] / scale, a02 = m_in[2] / scale;
    double a11 = m_in[3] / scale, a12 = m_in[4] / scale, a22 = m_in[5] / scale;
    double q = (a00 + a11 + a22) / 3.0;
    double b00 = a00 - q, b11 = a11 - q, b22 = a22 - q;
    double p2 = (b00 * b00 + b11 * b11 + b22 * b22 + 2.0 * (a01 * a01 + a02 * a02 + a12 * a12)) / 6.0;
    double p = std::sqrt(p2);
    double lambda_min;
    if (p < 1e-300) {
        lambda_min = q;
    } else {
        double c00 = b00 / p, c11 = b11 / p, c22 = b22 / p, c01 = a01 / p, c02 = a02 / p, c12 = a12 / p;
        double half_det = 0.5 * (c00 * (c11 * c22 - c12 * c12) - c01 * (c01 * c22 - c12 * c02) +
                                 c02 * (c01 * c12 - c11 * c02));
        half_det = std::min(1.0, std::max(-1.0, half_det));
        double phi = std::acos(half_det) / 3.0;
        // eigenvalues q + 2p cos(phi + 2k pi/3); the smallest is k = 1
        lambda_min = q + 2.0 * p * std::cos(phi + 2.0 * M_PI / 3.0);
    }
    double r0[3] = {a00 - lambda_min, a01, a02};
    double r1[3] = {a01, a11 - lambda_min, a12};
    double r2[3] = {a02, a12, a22 - lambda_min};
    auto cross = [](const double* u, const double* v, double* w) {
        w[0] = u[1] * v[2] - u[2] * v[1];
        w[1] = u[2] * v[0] - u[0] * v[2];
        w[2] = u[0] * v[1] - u[1] * v[0];
    };
    double c[3][3];
    cross(r0, r1, c[0]);
    cross(r0, r2, c[1]);
    cross(r1, r2, c[2]);
    int best = 0;
    double best_n = -1.0;
    for (int i = 0; i < 3; ++i) {
        double n = c[i][0] * c[i][0] + c[i][1] * c[i][1] + c[i][2] * c[i][2];
        if (n > best_n) { best_n = n; best = i; }
    }
    if (best_n <= 1e-300) { out[0] = 0; out[1] = 0; out[2] = 1; return; }
    double inv = 1.0 / std::sqrt(best_n);
    for (int k = 0; k < 3; ++k) out[k] = c[best][k] * inv;
}

}  // namespace

extern "C" plc_status plc_host_compute_normals(const plc_cloud_view* area, double radius, float* out) {
    if (!view_ok(area) || !out || !(radius > 0)) return PLC_E_INVALID_ARG;
    HostCloud cloud;
    cloud.assign(*area);
    const float quiet = std::numeric_limits<float>::quiet_NaN();
    for (uint64_t i = 0; i < area->n_points; ++i) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = quiet;
    Buckets buckets;
    buckets.build(cloud, std::max(radius * 0.5, 1e-3));
    const float r2f = radius_sq_f32(radius);
    plc_host::parallel_for(static_cast<int64_t>(cloud.size()), [&](int64_t s, int64_t e) {
        std::vector<uint32_t> nb;
        for (int64_t i = s; i < e; ++i) {
            nb.clear();
            buckets.in_ball(cloud.x[i], cloud.y[i], cloud.z[i], r2f, [&](uint32_t j, float) {
                nb.push_back(j);
                return false;
            });
            if (nb.size() < 3) continue;     // PCL: fewer than 3 neighbours => NaN normal
            double mean[3] = {0, 0, 0};
            for (uint32_t j : nb) { mean[0] += cloud.x[j]; mean[1] += cloud.y[j]; mean[2] += cloud.z[j]; }
            for (double& m : mean) m /= static_cast<double>(nb.size());
            double cov[6] = {0, 0, 0, 0, 0, 0};
            for (uint32_t j : nb) {
                double dx = cloud.x[j] - mean[0], dy = cloud.y[j] - mean[1], dz = cloud.z[j] - mean[2];
                cov[0] += dx * dx; cov[1] += dx * dy; cov[2] += dx * dz;
                cov[3] += dy * dy; cov[4] += dy * dz; cov[5] += dz * dz;
            }
            double n[3];
            min_eigenvector_sym3(cov, n);
            // towards the viewpoint (0,0,0) as PCL does, then up as the node does (:223-229)
            double towards = -(cloud.x[i] * n[0] + cloud.y[i] * n[1] + cloud.z[i] * n[2]);
            if (towards < 0) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
            if (n[2] < 0) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
            float* o = out + 3 * static_cast<uint64_t>(cloud.src[i]);
            o[0] = static_cast<float>(n[0]); o[1] = static_cast<float>(n[1]); o[2] = static_cast<float>(n[2]);
        }
    });
    return PLC_OK;
}

extern "C" plc_status plc_host_generate_cells(const plc_cloud_view* area, const float* normals,
                                              double res, int32_t layers, double normal_radius,
                                              plc_cell* out, uint64_t capacity, uint64_t* n_out,
                                              plc_grid_bounds* bounds) {
    if (!view_ok(area) || !n_out || !(res > 0) || (capacity && !out)) return PLC_E_INVALID_ARG;
    *n_out = 0;
    if (area->n_points == 0) return PLC_OK;     // reference returns early on an empty area (:168, :237)

    // bounding box over the stored points (NaNs fall out of std::min/max the same way, :242-249)
    plc_grid_bounds bb;
    bb.min_x = bb.min_y = bb.min_z = std::numeric_limits<double>::max();
    bb.max_x = bb.max_y = bb.max_z = std::numeric_limits<double>::lowest();
    {
        const unsigned char* base = static_cast<const unsigned char*>(area->data);
        for (uint64_t i = 0; i < area->n_points; ++i) {
            const unsigned char* p = base + i * area->point_step;
            double px = plc_host::load_f32(p + area->off_x), py = plc_host::load_f32(p + area->off_y),
                   pz = plc_host::load_f32(p + area->off_z);
            bb.min_x = std::min(bb.min_x, px); bb.max_x = std::max(bb.max_x, px);
            bb.min_y = std::min(bb.min_y, py); bb.max_y = std::max(bb.max_y, py);
            bb.min_z = std::min(bb.min_z, pz); bb.max_z = std::max(bb.max_z, pz);
        }
    }
    bb.min_x -= res; bb.max_x += res;
    bb.min_y -= res; bb.max_y += res;
    bb.min_z -= res; bb.max_z += res;
    bb.width = static_cast<int>(std::ceil((bb.max_x - bb.min_x) / res)) + 1;
    bb.height = static_cast<int>(std::ceil((bb.max_y - bb.min_y) / res)) + 1;
    if (bounds) *bounds = bb;
    if (layers <= 0) return PLC_OK;             // the k-loop never runs (:271)

    HostCloud cloud;
    cloud.assign(*area);
    Buckets near_buckets, normal_buckets;
    const double near_radius = res * 1.5;
    near_buckets.build(cloud, std::max(near_radius, 1e-3));
    if (normals) normal_buckets.build(cloud, std::max(normal_radius * 0.5, 1e-3));
    const float near_r2 = radius_sq_f32(near_radius), normal_r2 = radius_sq_f32(normal_radius);
    const double z_step = (bb.max_z - bb.min_z) / std::max(1, layers);

    const int64_t columns = static_cast<int64_t>(bb.width) * bb.height;
    std::vector<std::vector<plc_cell>> per_column(static_cast<size_t>(columns));
    plc_host::parallel_for(columns, [&](int64_t s, int64_t e) {
        std::vector<std::pair<float, uint32_t>> found;
        for (int64_t col = s; col < e; ++col) {
            int i = static_cast<int>(col / bb.width), j = static_cast<int>(col % bb.width);
            double x = bb.min_x + j * res, y = bb.min_y + i * res;
            for (int k = 0; k < layers; ++k) {
                double z = bb.min_z + k * z_step + z_step / 2.0;
                float qx = static_cast<float>(x), qy = static_cast<float>(y), qz = static_cast<float>(z);
                if (!near_buckets.any_in_ball(qx, qy, qz, near_r2)) continue;
                plc_cell cell{x, y, z, 0.0f, 0.0f, 1.0f, 0u};
                if (normals) {
                    // the reference adds the neighbours' normals in KD-tree result order (nearest
                    // first); ties broken by point index to make the double sum reproducible
                    found.clear();
                    normal_buckets.in_ball(qx, qy, qz, normal_r2, [&](uint32_t p, float d2) {
                        found.emplace_back(d2, cloud.src[p]);
                        return false;
                    });
                    std::sort(found.begin(), found.end());
                    double sx = 0, sy = 0, sz = 0;
                    int used = 0;
                    for (const auto& f : found) {
                        const float* nv = normals + 3 * static_cast<uint64_t>(f.second);
                        if (std::isfinite(nv[0]) && std::isfinite(nv[1]) && std::isfinite(nv[2])) {
                            sx += nv[0]; sy += nv[1]; sz += nv[2];
                            ++used;
                        }
                    }
                    if (used > 0) {
                        double len = std::sqrt(sx * sx + sy * sy + sz * sz);
                        if (len > 1e-6) {
                            cell.nx = static_cast<float>(sx / len);
                            cell.ny = static_cast<float>(sy / len);
                            cell.nz = static_cast<float>(sz / len);
                        }
                    }
                }
                per_column[static_cast<size_t>(col)].push_back(cell);
            }
        }
    });
    uint64_t total = 0;
    for (const auto& v : per_column) total += v.size();
    *n_out = total;
    if (total > capacity) return PLC_E_CAPACITY;
    uint64_t w = 0;
    for (const auto& v : per_column)
        for (const plc_cell& c : v) out[w++] = c;
    return PLC_OK;
}

namespace {

// Max z of terrain points that are inside the 3-D ball of radius 2 around (x,y,0) [float32 test]
// AND closer than 1 m horizontally [double test]; 0 when there are none (:600-625).
// A 2-D column index is enough because both conditions bound the horizontal distance by 1 m.
class GroundColumns {
public:
    void build(const HostCloud& c) {
        cloud_ = &c;
        if (c.size() == 0) return;
        min_x_ = *std::min_element(c.x.begin(), c.x.end());
        min_y_ = *std::min_element(c.y.begin(), c.y.end());
        float max_x = *std::max_element(c.x.begin(), c.x.end());
        float max_y = *std::max_element(c.y.begin(), c.y.end());
        nx_ = static_cast<int64_t>(std::floor((static_cast<double>(max_x) - min_x_))) + 1;
        ny_ = static_cast<int64_t>(std::floor((static_cast<double>(max_y) - min_y_))) + 1;
        if (nx_ * ny_ > (int64_t{1} << 26)) { nx_ = ny_ = 0; return; }   // absurd extent: fall back to a scan
        start_.assign(static_cast<size_t>(nx_ * ny_) + 1, 0);
        std::vector<uint32_t> col(c.size());
        for (size_t i = 0; i < c.size(); ++i) {
            col[i] = static_cast<uint32_t>(column(c.x[i], c.y[i]));
            start_[col[i] + 1]++;
        }
        for (size_t k = 0; k + 1 < start_.size(); ++k) start_[k + 1] += start_[k];
        std::vector<uint32_t> cursor(start_.begin(), start_.end() - 1);
        members_.resize(c.size());
        for (size_t i = 0; i < c.size(); ++i) members_[cursor[col[i]]++] = static_cast<uint32_t>(i);
    }
    double height(double x, double y) const {
        const HostCloud& c = *cloud_;
        if (c.size() == 0) return 0.0;
        const float qx = static_cast<float>(x), qy = static_cast<float>(y), qz = 0.0f;
        const float r2f = radius_sq_f32(2.0);
        double top = std::numeric_limits<double>::lowest();
        auto consider = [&](uint32_t i) {
            if (plc_host::squared_distance_f32(qx, qy, qz, c.x[i], c.y[i], c.z[i]) < r2f) {
                double dx = c.x[i] - x, dy = c.y[i] - y;
                if (std::sqrt(dx * dx + dy * dy) < 1.0) top = std::max(top, static_cast<double>(c.z[i]));
            }
        };
        if (nx_ == 0) {
            for (size_t i = 0; i < c.size(); ++i) consider(static_cast<uint32_t>(i));
        } else {
            int64_t cx0 = static_cast<int64_t>(std::floor(x - 1.001 - min_x_)), cx1 = static_cast<int64_t>(std::floor(x + 1.001 - min_x_));
            int64_t cy0 = static_cast<int64_t>(std::floor(y - 1.001 - min_y_)), cy1 = static_cast<int64_t>(std::floor(y + 1.001 - min_y_));
            cx0 = std::max<int64_t>(cx0, 0); cy0 = std::max<int64_t>(cy0, 0);
            cx1 = std::min<int64_t>(cx1, nx_ - 1); cy1 = std::min<int64_t>(cy1, ny_ - 1);
            for (int64_t cy = cy0; cy <= cy1; ++cy)
                for (int64_t cx = cx0; cx <= cx1; ++cx)
                    for (uint32_t s = start_[cy * nx_ + cx]; s < start_[cy * nx_ + cx + 1]; ++s) consider(members_[s]);
        }
        return top != std::numeric_limits<double>::lowest() ? top : 0.0;
    }

private:
    int64_t column(float px, float py) const {
        int64_t cx = static_cast<int64_t>(std::floor(static_cast<double>(px) - min_x_));
        int64_t cy = static_cast<int64_t>(std::floor(static_cast<double>(py) - min_y_));
        cx = std::min(std::max<int64_t>(cx, 0), nx_ - 1);
        cy = std::min(std::max<int64_t>(cy, 0), ny_ - 1);
        return cy * nx_ + cx;
    }
    const HostCloud* cloud_ = nullptr;
    double min_x_ = 0, min_y_ = 0;
    int64_t nx_ = 0, ny_ = 0;
    std::vector<uint32_t> start_, members_;
};

}  // namespace

extern "C" plc_status plc_host_generate_candidates(const plc_cloud_view* terrain,
                                                   const plc_grid_bounds* b, const plc_pose* fixed,
                                                   int32_t num_candidates, double search_radius,
                                                   double sensor_height, plc_pose* out,
                                                   uint64_t capacity, uint64_t* n_out) {
    if (!view_ok(terrain) || !b || !fixed || !n_out || (capacity && !out)) return PLC_E_INVALID_ARG;
    *n_out = 0;
    if (num_candidates < 2) return PLC_E_INVALID_ARG;   // lattice side 1 divides by zero upstream (:563)
    const int side = static_cast<int>(std::ceil(std::sqrt(static_cast<double>(num_candidates))));
    HostCloud cloud;
    cloud.assign(*terrain);
    GroundColumns ground;
    ground.build(cloud);

    const double lo_x = b->min_x - search_radius, hi_x = b->max_x + search_radius;
    const double lo_y = b->min_y - search_radius, hi_y = b->max_y + search_radius;
    const double mid_x = (b->min_x + b->max_x) / 2.0, mid_y = (b->min_y + b->max_y) / 2.0,
                 mid_z = (b->min_z + b->max_z) / 2.0;
    const double step_x = (hi_x - lo_x) / (side - 1), step_y = (hi_y - lo_y) / (side - 1);
    const double elev_lo = -85.0 * M_PI / 180.0, elev_hi = 85.0 * M_PI / 180.0;

    struct Slot { bool keep; plc_pose pose; };
    std::vector<Slot> slots(static_cast<size_t>(side) * side);
    plc_host::parallel_for(static_cast<int64_t>(slots.size()), [&](int64_t s, int64_t e) {
        for (int64_t t = s; t < e; ++t) {
            Slot& slot = slots[static_cast<size_t>(t)];
            slot.keep = false;
            int i = static_cast<int>(t / side), j = static_cast<int>(t % side);
            double x = lo_x + i * step_x, y = lo_y + j * step_y;
            double ex = x - fixed->x, ey = y - fixed->y;
            if (std::sqrt(ex * ex + ey * ey) < 0.5) continue;                            // :571-573
            if (x >= b->min_x && x <= b->max_x && y >= b->min_y && y <= b->max_y) continue;   // :575
            double z = ground.height(x, y) + sensor_height;
            double dx = mid_x - x, dy = mid_y - y, dz = mid_z - z;
            double horizontal = std::sqrt(dx * dx + dy * dy);
            if (horizontal < 0.1) continue;                                              // :585
            double elevation = std::atan2(-dz, horizontal);
            if (elevation < elev_lo || elevation > elev_hi) continue;                    // :589
            slot.pose = plc_pose{x, y, z, -M_PI / 2 + elevation, std::atan2(dy, dx)};
            slot.keep = true;
        }
    });
    uint64_t total = 0;
    for (const Slot& s : slots) total += s.keep;
    *n_out = total;
    if (total > capacity) return PLC_E_CAPACITY;
    uint64_t w = 0;
    for (const Slot& s : slots)
        if (s.keep) out[w++] = s.pose;
    return PLC_OK;
}
