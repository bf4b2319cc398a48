// point_buckets.hpp — host-side spatial buckets answering PCL-style radius queries for the node's
// setup steps (cell lattice, cell normals, ground height).  Not on the hot path.
//
// The acceptance rule is the one PCL KdTreeFLANN::radiusSearch applies:
// float32 query, d2 = ((dx*dx)+(dy*dy))+(dz*dz) in float32 without FMA, accept iff
// d2 < (float)(radius*radius).  This translation unit is built with -ffp-contract=off.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#include "plc.h"

namespace plc_host {

inline float load_f32(const unsigned char* p) {
    float v;
    std::memcpy(&v, p, 4);
    return v;
}

// x/y/z of every point of a PointCloud2 view, kept with its original index (non-finite dropped).
struct HostCloud {
    std::vector<float> x, y, z;
    std::vector<uint32_t> src;
    uint64_t n_input = 0;
    void assign(const plc_cloud_view& v) {
        n_input = v.n_points;
        x.clear(); y.clear(); z.clear(); src.clear();
        x.reserve(v.n_points); y.reserve(v.n_points); z.reserve(v.n_points); src.reserve(v.n_points);
        const unsigned char* base = static_cast<const unsigned char*>(v.data);
        for (uint64_t i = 0; i < v.n_points; ++i) {
            const unsigned char* p = base + i * v.point_step;
            float a = load_f32(p + v.off_x), b = load_f32(p + v.off_y), c = load_f32(p + v.off_z);
            if (std::isfinite(a) && std::isfinite(b) && std::isfinite(c)) {
                x.push_back(a); y.push_back(b); z.push_back(c);
                src.push_back(static_cast<uint32_t>(i));
            }
        }
    }
    size_t size() const { return x.size(); }
};

inline float squared_distance_f32(float qx, float qy, float qz, float px, float py, float pz) {
    float ex = qx - px, ey = qy - py, ez = qz - pz;
    float acc = ex * ex;
    acc = acc + ey * ey;
    acc = acc + ez * ez;
    return acc;
}

// Sorted-key buckets: each point gets a 63-bit key from its integer bucket coordinates; points are
// sorted by key and a bucket is located by binary search.  Memory is O(N) whatever the extent.
class Buckets {
public:
    void build(const HostCloud& c, double edge) {
        cloud_ = &c;
        h_ = edge;
        keys_.clear();
        order_.clear();
        if (c.size() == 0) return;
        lo_[0] = *std::min_element(c.x.begin(), c.x.end());
        lo_[1] = *std::min_element(c.y.begin(), c.y.end());
        lo_[2] = *std::min_element(c.z.begin(), c.z.end());
        std::vector<std::pair<uint64_t, uint32_t>> kv(c.size());
        for (size_t i = 0; i < c.size(); ++i)
            kv[i] = {key(coord(c.x[i], 0), coord(c.y[i], 1), coord(c.z[i], 2)), static_cast<uint32_t>(i)};
        std::sort(kv.begin(), kv.end());
        keys_.resize(kv.size());
        order_.resize(kv.size());
        for (size_t i = 0; i < kv.size(); ++i) {
            keys_[i] = kv[i].first;
            order_[i] = kv[i].second;
        }
    }
    // Calls fn(local_index, d2) for every point accepted by the float32 radius test; fn returns
    // true to stop early.
    template <class F>
    void in_ball(float qx, float qy, float qz, float r2f, F&& fn) const {
        if (keys_.empty()) return;
        const HostCloud& c = *cloud_;
        // real-number radius that surely covers everything the float test accepts
        double reach = std::sqrt(static_cast<double>(r2f)) * 1.00001 + 1e-9;
        int64_t a[3], b[3];
        const double q[3] = {qx, qy, qz};
        for (int d = 0; d < 3; ++d) {
            a[d] = std::max<int64_t>(0, static_cast<int64_t>(std::floor((q[d] - reach - lo_[d]) / h_)));
            b[d] = static_cast<int64_t>(std::floor((q[d] + reach - lo_[d]) / h_));
            if (b[d] < 0) return;
            b[d] = std::min<int64_t>(b[d], kMaxCoord);
            if (a[d] > b[d]) return;
        }
        for (int64_t k = a[2]; k <= b[2]; ++k)
            for (int64_t j = a[1]; j <= b[1]; ++j) {
                // x is the low field of the key: one contiguous key interval per (j,k) row
                uint64_t k0 = key(a[0], j, k), k1 = key(b[0], j, k);
                auto it = std::lower_bound(keys_.begin(), keys_.end(), k0);
                for (size_t s = static_cast<size_t>(it - keys_.begin()); s < keys_.size() && keys_[s] <= k1; ++s) {
                    uint32_t i = order_[s];
                    float d2 = squared_distance_f32(qx, qy, qz, c.x[i], c.y[i], c.z[i]);
                    if (d2 < r2f)
                        if (fn(i, d2)) return;
                }
            }
    }
    bool any_in_ball(float qx, float qy, float qz, float r2f) const {
        bool hit = false;
        in_ball(qx, qy, qz, r2f, [&](uint32_t, float) { return hit = true; });
        return hit;
    }

private:
    static constexpr int64_t kMaxCoord = (1 << 21) - 1;
    int64_t coord(float v, int d) const {
        int64_t c = static_cast<int64_t>(std::floor((static_cast<double>(v) - lo_[d]) / h_));
        return std::min<int64_t>(std::max<int64_t>(c, 0), kMaxCoord);
    }
    static uint64_t key(int64_t i, int64_t j, int64_t k) {
        return (static_cast<uint64_t>(k) << 42) | (static_cast<uint64_t>(j) << 21) | static_cast<uint64_t>(i);
    }
    const HostCloud* cloud_ = nullptr;
    double h_ = 1.0, lo_[3] = {0, 0, 0};
    std::vector<uint64_t> keys_;
    std::vector<uint32_t> order_;
};

// Static-partition parallel loop over [0,n) on the host cores.
inline void parallel_for(int64_t n, const std::function<void(int64_t, int64_t)>& body) {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    int64_t workers = std::min<int64_t>(hw, std::max<int64_t>(1, n / 64));
    if (workers <= 1) {
        body(0, n);
        return;
    }
    std::vector<std::thread> pool;
    int64_t chunk = (n + workers - 1) / workers;
    for (int64_t w = 0; w < workers; ++w) {
        int64_t s = w * chunk, e = std::min(n, s + chunk);
        if (s >= e) break;
        pool.emplace_back([=, &body] { body(s, e); });
    }
    for (auto& t : pool) t.join();
}

}  // namespace plc_host
