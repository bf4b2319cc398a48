// vl_sweep.cpp — CPU restatement of the virtual_lidar sweep.  TEST INFRASTRUCTURE (oracle.h).
//
// Follows /root/reference/src/virtual_lidar.cpp:
//   runOptimization            454-475   driver: fixed-sensor pass, candidate loop, strict '>' argmax
//   evaluateZX120Only          360-398   fixed-sensor pass (scores + flags + total)
//   evaluatePosition           627-654   per-candidate sum over cells
//   evaluateCellScore          656-701   range gate, FOV gate, visibility, incidence + 1/L score
//   isInFieldOfView            703-714   elevation test (memo variant 684-697 adds azimuth)
//   checkVisibility            716-730   which visibility test applies to which sensor
//   checkVisibility...Relaxed  732-752   fixed sensor: observed-cloud lookup, else ray
//   checkVisibilityWithRaycasting 754-800 sampled-sphere occlusion test
// Numerics per SURVEY §A.3.  Build with -ffp-contract=off (no FMA, like an x86-64 distro build).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "oracle.h"
#include "radius_index.hpp"

namespace {

using orc::RadiusIndex;

struct Sweep {
    const RadiusIndex* terrain;
    const RadiusIndex* observed;
    const plc_params* prm;
    bool faithful;
};

struct Tally {
    uint64_t samples = 0;
};

// kdtree->radiusSearch(search_point, radius, point_indices, point_distances) > 0, then the loop
// over point_distances at :784-790.  PCL returns SQUARED distances; the reference compares them
// with VISIBILITY_RADIUS*0.5 as if they were metres — restated literally.
bool sample_blocked(const Sweep& s, float qx, float qy, float qz) {
    const double radius = s.prm->vis_radius * s.prm->block_radius_scale;   // :782
    const float r2f = orc::radius_to_r2f(radius);
    const double half_vis = s.prm->vis_radius * 0.5;                       // :786
    if (s.faithful) {
        std::vector<int> point_indices;                                     // :779
        std::vector<float> point_distances;                                 // :780
        s.terrain->search(qx, qy, qz, r2f, point_indices, point_distances, true);
        if (!point_indices.empty()) {
            for (size_t i = 0; i < point_indices.size(); ++i)
                if (point_distances[i] < half_vis) return true;             // float promoted to double
        }
        return false;
    }
    // Same predicate without materialising the neighbour list:  exists p with d2 < r2f and d2 < half_vis.
    // d2 (float) < half_vis (double)  <=>  d2 < the smallest float >= half_vis.
    float hv = static_cast<float>(half_vis);
    if (static_cast<double>(hv) < half_vis) hv = std::nextafterf(hv, std::numeric_limits<float>::infinity());
    return s.terrain->any(qx, qy, qz, std::min(r2f, hv));
}

// checkVisibilityWithRaycasting, :754-800
bool ray_clear(const Sweep& s, const plc_pose& sensor, const plc_cell& cell, Tally& t) {
    double dx = cell.x - sensor.x;
    double dy = cell.y - sensor.y;
    double dz = cell.z - sensor.z;
    double distance = std::sqrt(dx * dx + dy * dy + dz * dz);
    double ux = dx / distance, uy = dy / distance, uz = dz / distance;
    double step = s.prm->ray_start;                          // :765-766
    double end = distance - s.prm->vis_radius;               // :767
    while (step < end) {                                     // :769
        double cx = sensor.x + ux * step;
        double cy = sensor.y + uy * step;
        double cz = sensor.z + uz * step;
        float qx = static_cast<float>(cx), qy = static_cast<float>(cy), qz = static_cast<float>(cz);  // PointXYZRGB fields
        ++t.samples;
        if (sample_blocked(s, qx, qy, qz)) return false;
        step += s.prm->ray_step;                             // :796 (accumulates in double)
    }
    return true;
}

// isInFieldOfView, :703-714; azimuth branch from virtual_lidar_memo.cpp:684-697 when fov_horizontal >= 0
bool in_fov(const plc_params& p, const plc_pose& sensor, double dx, double dy, double dz) {
    double elevation = std::atan2(dz, std::sqrt(dx * dx + dy * dy));
    double elevation_diff = elevation - sensor.pitch;
    bool ok = std::abs(elevation_diff) <= p.fov_vertical / 2.0;
    if (p.fov_horizontal >= 0.0) {
        double azimuth = std::atan2(dy, dx);
        double azimuth_diff = std::fmod(azimuth - sensor.yaw + M_PI, 2 * M_PI) - M_PI;
        ok = (std::abs(azimuth_diff) <= p.fov_horizontal / 2.0) && ok;
    }
    return ok;
}

// checkVisibility, :716-730 (+ memo variant 699-726 when fixed_mode == 1)
bool visible(const Sweep& s, const plc_pose& sensor, const plc_cell& cell, bool is_fixed, Tally& t) {
    if (is_fixed) {
        bool have_observed = s.observed && s.observed->size() > 0;
        if (s.prm->fixed_mode == 1) {
            if (!have_observed) return false;
            float r2f = orc::radius_to_r2f(s.prm->vis_radius * s.prm->observed_radius_scale);
            return s.observed->any(static_cast<float>(cell.x), static_cast<float>(cell.y),
                                   static_cast<float>(cell.z), r2f);
        }
        if (!have_observed) {                                              // :719-723
            if (!s.terrain) return true;
            return ray_clear(s, sensor, cell, t);
        }
        // checkVisibilityWithPointCloudRelaxed, :732-752
        float r2f = orc::radius_to_r2f(s.prm->vis_radius * s.prm->observed_radius_scale);
        if (s.observed->any(static_cast<float>(cell.x), static_cast<float>(cell.y),
                            static_cast<float>(cell.z), r2f))
            return true;
        if (!s.terrain) return true;
        return ray_clear(s, sensor, cell, t);
    }
    if (!s.terrain) return true;                                           // :727
    return ray_clear(s, sensor, cell, t);
}

// evaluateCellScore, :656-701.  `flags` may be null; bits are written only where the reference
// assigns the corresponding bool (early returns leave later flags untouched: SURVEY Q3).
double cell_score(const Sweep& s, const plc_pose& sensor, const plc_cell& cell, bool is_fixed,
                  uint8_t* flags, Tally& t, bool* gated_in, bool* vis_out) {
    auto put = [&](unsigned bit, bool v) {
        if (!flags) return;
        if (v) *flags |= bit;
        else *flags &= static_cast<uint8_t>(~bit);
    };
    double dx = cell.x - sensor.x;
    double dy = cell.y - sensor.y;
    double dz = cell.z - sensor.z;
    double L = std::sqrt(dx * dx + dy * dy + dz * dz);
    if (gated_in) *gated_in = false;
    if (vis_out) *vis_out = false;

    bool in_range = (L >= s.prm->min_distance && L <= s.prm->max_distance);
    put(is_fixed ? PLC_FLAG_RANGE_FIXED : PLC_FLAG_RANGE_MOBILE, in_range);
    if (!in_range) return 0.0;

    bool fov = in_fov(*s.prm, sensor, dx, dy, dz);
    put(is_fixed ? PLC_FLAG_FOV_FIXED : PLC_FLAG_FOV_MOBILE, fov);
    if (!fov) return 0.0;
    if (gated_in) *gated_in = true;

    bool vis = visible(s, sensor, cell, is_fixed, t);
    put(is_fixed ? PLC_FLAG_VIS_FIXED : PLC_FLAG_VIS_MOBILE, vis);
    if (!vis) return 0.0;
    if (vis_out) *vis_out = true;

    double bx = dx / L, by = dy / L, bz = dz / L;
    double dot = bx * cell.nx + by * cell.ny + bz * cell.nz;               // float normal promoted
    double theta = std::acos(std::max(0.0, std::min(1.0, std::abs(dot))));
    double score = s.prm->alpha * std::sin(M_PI / 2 - theta) + s.prm->beta * (1.0 / L);
    return std::max(0.0, score);
}

}  // namespace

extern "C" int orc_vl_sweep(orc_sweep_args* a) {
    if (!a || !a->prm || !a->fixed || (!a->cells && a->n_cells) || (!a->cands && a->n_cands)) return -1;
    Sweep s{static_cast<const RadiusIndex*>(a->terrain), static_cast<const RadiusIndex*>(a->observed),
            a->prm, a->faithful != 0};
    if (s.terrain && s.terrain->size() == 0) s.terrain = nullptr;   // empty cloud never builds a tree (:184)
    const uint64_t M = a->n_cells, C = a->n_cands;
    const uint64_t W = (M + 31) / 32;
    const bool sequential = a->threads <= 1;
    std::memset(&a->counters, 0, sizeof(a->counters));

    // evaluateZX120Only, :360-398
    std::vector<double> fixed_score(M, 0.0);
    {
        Tally t;
        double total = 0.0;
        for (uint64_t j = 0; j < M; ++j) {
            uint8_t* fl = (sequential && a->cell_flags) ? &a->cell_flags[j] : nullptr;
            fixed_score[j] = cell_score(s, *a->fixed, a->cells[j], true, fl, t, nullptr, nullptr);
            if (fixed_score[j] > 0) total += fixed_score[j];
        }
        a->fixed_total = total;
        a->counters.fixed_samples = t.samples;
        if (a->fixed_scores) std::memcpy(a->fixed_scores, fixed_score.data(), M * sizeof(double));
    }

    std::vector<double> scores(C, 0.0);
    std::vector<int32_t> covered(C, 0);
    std::vector<uint64_t> c_samples(C, 0), c_gated(C, 0), c_vis(C, 0);
    if (a->mask) std::memset(a->mask, 0, C * W * sizeof(uint32_t));

    auto eval_candidate = [&](uint64_t ci, bool with_flags) {
        // evaluatePosition, :627-654
        Tally t, t_fixed;
        double total = 0.0;
        int32_t cov = 0;
        uint64_t gated = 0, nvis = 0;
        for (uint64_t j = 0; j < M; ++j) {
            uint8_t* fl = (with_flags && a->cell_flags) ? &a->cell_flags[j] : nullptr;
            double s_fixed = fixed_score[j];
            if (s.faithful)   // :637 — recomputed for every candidate upstream (F9); same value
                s_fixed = cell_score(s, *a->fixed, a->cells[j], true, fl, t_fixed, nullptr, nullptr);
            bool g = false, v = false;
            double s_mobile = cell_score(s, a->cands[ci], a->cells[j], false, fl, t, &g, &v);
            double combined = std::max(s_fixed, s_mobile);
            if (combined > 0) {
                ++cov;
                total += combined;
            }
            gated += g;
            nvis += v;
            if (v && a->mask) a->mask[ci * W + j / 32] |= (1u << (j % 32));
        }
        scores[ci] = total;
        covered[ci] = cov;
        c_samples[ci] = t.samples;
        c_gated[ci] = gated;
        c_vis[ci] = nvis;
    };

    if (sequential) {
        for (uint64_t ci = 0; ci < C; ++ci) eval_candidate(ci, true);
    } else {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(a->threads)
#endif
        for (int64_t ci = 0; ci < static_cast<int64_t>(C); ++ci) eval_candidate(static_cast<uint64_t>(ci), false);
    }

    // argmax, :464-475 — strict '>' keeps the first maximum
    double best = -std::numeric_limits<double>::infinity();
    int64_t best_i = -1;
    for (uint64_t ci = 0; ci < C; ++ci) {
        if (scores[ci] > best) {
            best = scores[ci];
            best_i = static_cast<int64_t>(ci);
        }
        a->counters.samples += c_samples[ci];
        a->counters.rays_gated_in += c_gated[ci];
        a->counters.rays_visible += c_vis[ci];
    }
    a->counters.rays = C * M;
    a->best_index = best_i;
    a->best_score = best;
    a->best_covered = best_i >= 0 ? covered[best_i] : 0;
    if (a->scores && C) std::memcpy(a->scores, scores.data(), C * sizeof(double));       // an empty vector's data() may be null
    if (a->covered && C) std::memcpy(a->covered, covered.data(), C * sizeof(int32_t));
    return 0;
}
