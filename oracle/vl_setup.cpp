// vl_setup.cpp — CPU restatement of the virtual_lidar setup steps that produce the sweep's inputs
// (evaluation cells, cell normals, candidate poses).  TEST INFRASTRUCTURE (oracle.h).
//
// Follows /root/reference/src/virtual_lidar.cpp:
//   computeTerrainNormals      209-234   (PCL NormalEstimation, restated from its published algorithm)
//   generateExcavationGrid3D   236-287
//   isPointNearExcavation      289-299
//   computeCellSurfaceNormal   301-340
//   generateCandidatePositions 550-598
//   getGroundHeight            600-625
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "oracle.h"
#include "radius_index.hpp"

using orc::RadiusIndex;

extern "C" void* orc_index_create(const float* xyz, uint64_t n, int kind) {
    return orc::make_index(xyz, n, kind);
}
extern "C" void orc_index_destroy(void* index) { delete static_cast<RadiusIndex*>(index); }
extern "C" uint64_t orc_index_size(const void* index) {
    return index ? static_cast<const RadiusIndex*>(index)->size() : 0;
}
extern "C" int64_t orc_index_radius_count(const void* index, float qx, float qy, float qz, double radius) {
    return static_cast<const RadiusIndex*>(index)->count(qx, qy, qz, orc::radius_to_r2f(radius));
}

extern "C" void orc_unpack_cloud(const plc_cloud_view* v, float* out) {
    const unsigned char* base = static_cast<const unsigned char*>(v->data);
    for (uint64_t i = 0; i < v->n_points; ++i) {
        const unsigned char* p = base + i * v->point_step;
        std::memcpy(&out[3 * i + 0], p + v->off_x, 4);
        std::memcpy(&out[3 * i + 1], p + v->off_y, 4);
        std::memcpy(&out[3 * i + 2], p + v->off_z, 4);
    }
}

// ---------------------------------------------------------------------------------------------
// 3x3 symmetric eigen-decomposition (Jacobi) — used for the PCA normal.
static void smallest_eigenvector(double a[3][3], double v_out[3]) {
    double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
        if (off < 1e-300) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (std::abs(a[p][q]) < 1e-300) continue;
                double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                double t = (theta >= 0 ? 1.0 : -1.0) / (std::abs(theta) + std::sqrt(theta * theta + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 3; ++k) {
                    double akp = a[k][p], akq = a[k][q];
                    a[k][p] = c * akp - s * akq;
                    a[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 3; ++k) {
                    double apk = a[p][k], aqk = a[q][k];
                    a[p][k] = c * apk - s * aqk;
                    a[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 3; ++k) {
                    double vkp = v[k][p], vkq = v[k][q];
                    v[k][p] = c * vkp - s * vkq;
                    v[k][q] = s * vkp + c * vkq;
                }
            }
    }
    int m = 0;
    if (a[1][1] < a[m][m]) m = 1;
    if (a[2][2] < a[m][m]) m = 2;
    for (int k = 0; k < 3; ++k) v_out[k] = v[k][m];
}

// computeTerrainNormals, :209-234.  PCL NormalEstimation per point: neighbours within `radius`
// (the point itself included); fewer than 3 => NaN normal; covariance about the centroid; normal =
// eigenvector of the smallest eigenvalue; flipped towards the viewpoint (0,0,0); then the node
// flips to n_z >= 0 (:223-229).  Not parity-gated (setup input shared by both sides).
extern "C" void orc_compute_normals(const float* xyz, uint64_t n, double radius, float* out) {
    RadiusIndex* ix = orc::make_index(xyz, n, ORC_INDEX_GRID);
    const float r2f = orc::radius_to_r2f(radius);
    const float nanf_ = std::numeric_limits<float>::quiet_NaN();
    std::vector<int> idx;
    std::vector<float> d2;
    for (uint64_t i = 0; i < n; ++i) {
        float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
        out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = nanf_;
        if (!std::isfinite(px) || !std::isfinite(py) || !std::isfinite(pz)) continue;
        idx.clear();
        d2.clear();
        ix->search(px, py, pz, r2f, idx, d2, false);
        if (idx.size() < 3) continue;
        double c[3] = {0, 0, 0};
        for (int j : idx)
            for (int k = 0; k < 3; ++k) c[k] += xyz[3 * j + k];
        for (int k = 0; k < 3; ++k) c[k] /= static_cast<double>(idx.size());
        double cov[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
        for (int j : idx) {
            double d[3] = {xyz[3 * j] - c[0], xyz[3 * j + 1] - c[1], xyz[3 * j + 2] - c[2]};
            for (int r = 0; r < 3; ++r)
                for (int s = 0; s < 3; ++s) cov[r][s] += d[r] * d[s];
        }
        double nv[3];
        smallest_eigenvector(cov, nv);
        // flipNormalTowardsViewpoint with viewpoint (0,0,0): (vp - p) . n < 0 => flip
        if ((-px) * nv[0] + (-py) * nv[1] + (-pz) * nv[2] < 0)
            for (double& k : nv) k = -k;
        if (nv[2] < 0)
            for (double& k : nv) k = -k;
        for (int k = 0; k < 3; ++k) out[3 * i + k] = static_cast<float>(nv[k]);
    }
    delete ix;
}

// generateExcavationGrid3D, :236-287
extern "C" int64_t orc_generate_cells(const float* area, uint64_t n_area, const float* normals,
                                      double res, int32_t layers, double normal_radius,
                                      plc_cell* out, uint64_t capacity, orc_grid_bounds* bounds) {
    if (!n_area) return 0;
    RadiusIndex* ix = orc::make_index(area, n_area, ORC_INDEX_GRID);
    double min_x, min_y, min_z, max_x, max_y, max_z;
    min_x = min_y = min_z = std::numeric_limits<double>::max();
    max_x = max_y = max_z = std::numeric_limits<double>::lowest();
    for (uint64_t i = 0; i < n_area; ++i) {                      // :242-249 (all points, as stored)
        min_x = std::min(min_x, static_cast<double>(area[3 * i]));
        max_x = std::max(max_x, static_cast<double>(area[3 * i]));
        min_y = std::min(min_y, static_cast<double>(area[3 * i + 1]));
        max_y = std::max(max_y, static_cast<double>(area[3 * i + 1]));
        min_z = std::min(min_z, static_cast<double>(area[3 * i + 2]));
        max_z = std::max(max_z, static_cast<double>(area[3 * i + 2]));
    }
    double margin = res;                                          // :251-254
    min_x -= margin; max_x += margin;
    min_y -= margin; max_y += margin;
    min_z -= margin; max_z += margin;
    int width = static_cast<int>(std::ceil((max_x - min_x) / res)) + 1;    // :256
    int height = static_cast<int>(std::ceil((max_y - min_y) / res)) + 1;   // :257
    if (bounds) *bounds = orc_grid_bounds{min_x, max_x, min_y, max_y, min_z, max_z, width, height};

    double z_range = max_z - min_z;
    double z_step = z_range / std::max(1, layers);                // :262
    const float r2_near = orc::radius_to_r2f(res * 1.5);         // :274
    const float r2_norm = orc::radius_to_r2f(normal_radius);     // :312
    std::vector<int> idx;
    std::vector<float> d2;
    int64_t count = 0;
    for (int i = 0; i < height; ++i)
        for (int j = 0; j < width; ++j) {
            double x = min_x + j * res;
            double y = min_y + i * res;
            for (int k = 0; k < layers; ++k) {
                double z = min_z + k * z_step + z_step / 2.0;
                float qx = static_cast<float>(x), qy = static_cast<float>(y), qz = static_cast<float>(z);
                if (!ix->any(qx, qy, qz, r2_near)) continue;      // isPointNearExcavation, :289-299
                plc_cell cell{x, y, z, 0.0f, 0.0f, 1.0f, 0u};     // GridCell ctor, :34-43
                if (normals) {                                    // computeCellSurfaceNormal, :301-340
                    idx.clear();
                    d2.clear();
                    ix->search(qx, qy, qz, r2_norm, idx, d2, true);
                    double sx = 0, sy = 0, sz = 0;
                    int valid = 0;
                    for (int p : idx) {
                        float a = normals[3 * p], b = normals[3 * p + 1], c = normals[3 * p + 2];
                        if (std::isfinite(a) && std::isfinite(b) && std::isfinite(c)) {
                            sx += a; sy += b; sz += c;
                            ++valid;
                        }
                    }
                    if (valid > 0) {
                        double norm = std::sqrt(sx * sx + sy * sy + sz * sz);
                        if (norm > 1e-6) {
                            cell.nx = static_cast<float>(sx / norm);
                            cell.ny = static_cast<float>(sy / norm);
                            cell.nz = static_cast<float>(sz / norm);
                        }
                    }
                }
                if (static_cast<uint64_t>(count) < capacity) out[count] = cell;
                ++count;
            }
        }
    delete ix;
    return count;
}

// getGroundHeight, :600-625
extern "C" double orc_ground_height(const void* terrain_index, const float* txyz, double x, double y) {
    const RadiusIndex* ix = static_cast<const RadiusIndex*>(terrain_index);
    if (!ix || ix->size() == 0) return 0.0;
    std::vector<int> idx;
    std::vector<float> d2;
    ix->search(static_cast<float>(x), static_cast<float>(y), 0.0f, orc::radius_to_r2f(2.0), idx, d2, true);
    if (!idx.empty()) {
        double max_z = std::numeric_limits<double>::lowest();
        for (int p : idx) {
            double dx = txyz[3 * p] - x;
            double dy = txyz[3 * p + 1] - y;
            if (std::sqrt(dx * dx + dy * dy) < 1.0) max_z = std::max(max_z, static_cast<double>(txyz[3 * p + 2]));
        }
        if (max_z != std::numeric_limits<double>::lowest()) return max_z;
    }
    return 0.0;
}

// generateCandidatePositions, :550-598
extern "C" int64_t orc_generate_candidates(const void* terrain_index, const float* txyz,
                                           const orc_grid_bounds* b, const plc_pose* fixed,
                                           int32_t num_candidates, double search_radius,
                                           double sensor_height, plc_pose* out, uint64_t capacity) {
    const double MIN_ELEVATION = -85.0 * M_PI / 180.0;   // :113
    const double MAX_ELEVATION = 85.0 * M_PI / 180.0;    // :114
    double ex_min_x = b->min_x - search_radius, ex_max_x = b->max_x + search_radius;
    double ex_min_y = b->min_y - search_radius, ex_max_y = b->max_y + search_radius;
    double center_x = (b->min_x + b->max_x) / 2.0;
    double center_y = (b->min_y + b->max_y) / 2.0;
    double center_z = (b->min_z + b->max_z) / 2.0;
    int g = static_cast<int>(std::ceil(std::sqrt(static_cast<double>(num_candidates))));
    double x_step = (ex_max_x - ex_min_x) / (g - 1);
    double y_step = (ex_max_y - ex_min_y) / (g - 1);
    int64_t count = 0;
    for (int i = 0; i < g; ++i)
        for (int j = 0; j < g; ++j) {
            double x = ex_min_x + i * x_step;
            double y = ex_min_y + j * y_step;
            double to_fixed = std::sqrt(std::pow(x - fixed->x, 2) + std::pow(y - fixed->y, 2));
            if (to_fixed < 0.5) continue;
            if (x >= b->min_x && x <= b->max_x && y >= b->min_y && y <= b->max_y) continue;
            double z = orc_ground_height(terrain_index, txyz, x, y) + sensor_height;
            double dx = center_x - x, dy = center_y - y, dz = center_z - z;
            double hd = std::sqrt(dx * dx + dy * dy);
            if (hd < 0.1) continue;
            double elev = std::atan2(-dz, hd);
            if (elev >= MIN_ELEVATION && elev <= MAX_ELEVATION) {
                plc_pose p{x, y, z, -M_PI / 2 + elev, std::atan2(dy, dx)};
                if (static_cast<uint64_t>(count) < capacity) out[count] = p;
                ++count;
            }
        }
    return count;
}
