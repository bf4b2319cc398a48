// pcl_standin.hpp — stand-ins for the few PCL 1.12 / FLANN 1.9 / pcl_conversions pieces that the reference's
// src/virtual_lidar.cpp uses, so that the reference's OWN, UNMODIFIED source file can be compiled and run here (no PCL,
// FLANN or ROS2 in this image).  Test infrastructure only (oracle/ rules apply): used by `make -C oracle ref`, which
// compiles /root/reference/src/virtual_lidar.cpp where it lies against these headers and tests/mock_ros2 into
// oracle/_ref/.  Everything the reference itself wrote — lattices, gates, the sampled ray, the fixed-sensor rule, the
// stale flags, the sequential sums, the strict-> argmax, the statistics — then comes from the reference's code.  What
// is restated here is third-party arithmetic only:
//   pcl::KdTreeFLANN::radiusSearch   FLANN L2_Simple<float>: ((dx*dx) + (dy*dy)) + (dz*dz) in float32, accepted when
//                                    strictly below (float)(radius * radius), results in ascending distance; the tree
//                                    holds the finite points only (a uniform grid answers here: same result set)
//   pcl::NormalEstimation            neighbours by the same search, covariance about their centroid, eigenvector of the
//                                    smallest eigenvalue, flipped towards the viewpoint (0, 0, 0); fewer than three
//                                    neighbours: NaN
//   pcl::fromROSMsg                  fields matched by name (x, y, z FLOAT32; rgb / rgba if present)
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "mock_ros2.hpp"

namespace pcl {

// PCL's 16-byte-aligned layouts: data[4] = {x, y, z, 1}; PointXYZRGB adds data_c[4] with the packed colour in [0]
struct PointXYZ { float x = 0.f, y = 0.f, z = 0.f, data3 = 1.f; };
struct PointXYZRGB {
    float x = 0.f, y = 0.f, z = 0.f, data3 = 1.f;
    union {
        float rgb;
        uint32_t rgba;
        struct { uint8_t b, g, r, a; };
    };
    float pad_c[3] = {0.f, 0.f, 0.f};
    PointXYZRGB() : rgba(0xff000000u) {}
};
static_assert(sizeof(PointXYZ) == 16 && sizeof(PointXYZRGB) == 32, "PCL point sizes");
struct Normal { float normal_x = 0.f, normal_y = 0.f, normal_z = 0.f, curvature = 0.f; };

template <class T>
class PointCloud {
public:
    using Ptr = std::shared_ptr<PointCloud<T>>;
    using ConstPtr = std::shared_ptr<const PointCloud<T>>;
    struct Header { uint32_t seq = 0; uint64_t stamp = 0; std::string frame_id; } header;     // pcl::PCLHeader
    std::vector<T> points;
    uint32_t width = 0, height = 0;
    bool is_dense = true;
    void reserve(size_t n) { points.reserve(n); }
    bool empty() const { return points.empty(); }
    size_t size() const { return points.size(); }
    void clear() { points.clear(); width = height = 0; }
    void push_back(const T& p) { points.push_back(p); width = static_cast<uint32_t>(points.size()); height = 1; }
    T& operator[](size_t i) { return points[i]; }
    const T& operator[](size_t i) const { return points[i]; }
    auto begin() { return points.begin(); }
    auto end() { return points.end(); }
    auto begin() const { return points.begin(); }
    auto end() const { return points.end(); }
};

namespace detail {
// The set FLANN's radius search returns, from a uniform grid over the finite points.
class RadiusIndex {
public:
    void build(const float* xyz, size_t stride_floats, size_t n) {
        pts_.clear(); ids_.clear(); cells_.clear();
        for (size_t i = 0; i < n; ++i) {
            const float* p = xyz + i * stride_floats;
            if (std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2])) {
                pts_.insert(pts_.end(), p, p + 3);
                ids_.push_back(static_cast<int>(i));
            }
        }
        for (size_t k = 0; k < ids_.size(); ++k) cells_[key(cell(pts_[3 * k]), cell(pts_[3 * k + 1]), cell(pts_[3 * k + 2]))].push_back(static_cast<int>(k));
    }
    // (squared distance, original index) of every point strictly within the float32 radius, ascending
    void search(float qx, float qy, float qz, double radius, std::vector<std::pair<float, int>>& out) const {
        out.clear();
        const float r2 = static_cast<float>(radius * radius);
        const double reach = radius * 1.0001 + 1e-4;
        const long x0 = cell(qx - reach), x1 = cell(qx + reach), y0 = cell(qy - reach), y1 = cell(qy + reach),
                   z0 = cell(qz - reach), z1 = cell(qz + reach);
        for (long cz = z0; cz <= z1; ++cz)
            for (long cy = y0; cy <= y1; ++cy)
                for (long cx = x0; cx <= x1; ++cx) {
                    auto it = cells_.find(key(cx, cy, cz));
                    if (it == cells_.end()) continue;
                    for (int k : it->second) {
                        const float dx = qx - pts_[3 * k], dy = qy - pts_[3 * k + 1], dz = qz - pts_[3 * k + 2];
                        float d = 0.f;                 // L2_Simple<float>: result += diff * diff, x then y then z
                        d += dx * dx;
                        d += dy * dy;
                        d += dz * dz;
                        if (d < r2) out.emplace_back(d, ids_[k]);
                    }
                }
        std::sort(out.begin(), out.end());
    }
    bool nearest(float qx, float qy, float qz, std::pair<float, int>* best) const {
        bool any = false;
        for (size_t k = 0; k < ids_.size(); ++k) {
            const float dx = qx - pts_[3 * k], dy = qy - pts_[3 * k + 1], dz = qz - pts_[3 * k + 2];
            float d = 0.f;
            d += dx * dx;
            d += dy * dy;
            d += dz * dz;
            if (!any || d < best->first) { *best = {d, ids_[k]}; any = true; }
        }
        return any;
    }
private:
    static constexpr double kCell = 0.25;
    static long cell(double v) { return static_cast<long>(std::floor(v / kCell)); }
    static uint64_t key(long x, long y, long z) {
        return (static_cast<uint64_t>(x & 0x1fffff) << 42) | (static_cast<uint64_t>(y & 0x1fffff) << 21) | static_cast<uint64_t>(z & 0x1fffff);
    }
    std::vector<float> pts_;
    std::vector<int> ids_;
    std::unordered_map<uint64_t, std::vector<int>> cells_;
};
}  // namespace detail

template <class T>
class KdTreeFLANN {
public:
    using Ptr = std::shared_ptr<KdTreeFLANN<T>>;
    void setInputCloud(const typename PointCloud<T>::ConstPtr& cloud) {
        cloud_ = cloud;
        static_assert(sizeof(T) % sizeof(float) == 0, "points are float records");
        index_.build(reinterpret_cast<const float*>(cloud->points.data()), sizeof(T) / sizeof(float), cloud->points.size());
    }
    int radiusSearch(const T& p, double radius, std::vector<int>& indices, std::vector<float>& sqr_distances, unsigned max_nn = 0) const {
        std::vector<std::pair<float, int>> hits;
        index_.search(p.x, p.y, p.z, radius, hits);
        if (max_nn && hits.size() > max_nn) hits.resize(max_nn);
        indices.resize(hits.size());
        sqr_distances.resize(hits.size());
        for (size_t i = 0; i < hits.size(); ++i) { indices[i] = hits[i].second; sqr_distances[i] = hits[i].first; }
        return static_cast<int>(hits.size());
    }
    int nearestKSearch(const T& p, int k, std::vector<int>& indices, std::vector<float>& sqr_distances) const {
        std::pair<float, int> best;
        if (k < 1 || !index_.nearest(p.x, p.y, p.z, &best)) { indices.clear(); sqr_distances.clear(); return 0; }
        indices.assign(1, best.second);
        sqr_distances.assign(1, best.first);
        return 1;
    }
    typename PointCloud<T>::ConstPtr getInputCloud() const { return cloud_; }
private:
    typename PointCloud<T>::ConstPtr cloud_;
    detail::RadiusIndex index_;
};

namespace search {
template <class T>
class KdTree : public pcl::KdTreeFLANN<T> {
public:
    using Ptr = std::shared_ptr<KdTree<T>>;
};
}  // namespace search

template <class PointInT, class PointOutT>
class NormalEstimation {
public:
    void setInputCloud(const typename PointCloud<PointInT>::ConstPtr& cloud) { cloud_ = cloud; }
    void setSearchMethod(const typename search::KdTree<PointInT>::Ptr& tree) { tree_ = tree; }
    void setRadiusSearch(double r) { radius_ = r; }
    void compute(PointCloud<PointOutT>& out) {
        const float nan = std::numeric_limits<float>::quiet_NaN();
        out.points.assign(cloud_->points.size(), PointOutT());
        out.width = static_cast<uint32_t>(out.points.size());
        out.height = 1;
        std::vector<int> idx;
        std::vector<float> d2;
        for (size_t i = 0; i < cloud_->points.size(); ++i) {
            const PointInT& p = cloud_->points[i];
            PointOutT& n = out.points[i];
            n.normal_x = n.normal_y = n.normal_z = n.curvature = nan;
            if (!(std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z))) continue;
            if (tree_->radiusSearch(p, radius_, idx, d2) < 3) continue;
            double m[3] = {0, 0, 0};
            for (int j : idx) { m[0] += cloud_->points[j].x; m[1] += cloud_->points[j].y; m[2] += cloud_->points[j].z; }
            for (double& v : m) v /= static_cast<double>(idx.size());
            double c[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
            for (int j : idx) {
                const double d[3] = {cloud_->points[j].x - m[0], cloud_->points[j].y - m[1], cloud_->points[j].z - m[2]};
                for (int a = 0; a < 3; ++a)
                    for (int b = 0; b < 3; ++b) c[a][b] += d[a] * d[b];
            }
            double v[3];
            smallest_eigenvector(c, v);
            if ((0.0 - p.x) * v[0] + (0.0 - p.y) * v[1] + (0.0 - p.z) * v[2] < 0) { v[0] = -v[0]; v[1] = -v[1]; v[2] = -v[2]; }   // flipNormalTowardsViewpoint
            n.normal_x = static_cast<float>(v[0]);
            n.normal_y = static_cast<float>(v[1]);
            n.normal_z = static_cast<float>(v[2]);
            n.curvature = 0.f;
        }
    }
private:
    // cyclic Jacobi on the symmetric 3x3 matrix; returns the unit eigenvector of the smallest eigenvalue
    static void smallest_eigenvector(double a[3][3], double out[3]) {
        double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
        for (int sweep = 0; sweep < 64; ++sweep) {
            const double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
            if (off < 1e-300) break;
            for (int p = 0; p < 2; ++p)
                for (int q = p + 1; q < 3; ++q) {
                    if (std::fabs(a[p][q]) < 1e-300) continue;
                    const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                    const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                    const double cs = 1.0 / std::sqrt(t * t + 1.0), sn = t * cs;
                    for (int k = 0; k < 3; ++k) {
                        const double akp = a[k][p], akq = a[k][q];
                        a[k][p] = cs * akp - sn * akq;
                        a[k][q] = sn * akp + cs * akq;
                    }
                    for (int k = 0; k < 3; ++k) {
                        const double apk = a[p][k], aqk = a[q][k];
                        a[p][k] = cs * apk - sn * aqk;
                        a[q][k] = sn * apk + cs * aqk;
                    }
                    for (int k = 0; k < 3; ++k) {
                        const double vkp = v[k][p], vkq = v[k][q];
                        v[k][p] = cs * vkp - sn * vkq;
                        v[k][q] = sn * vkp + cs * vkq;
                    }
                }
        }
        int m = 0;
        for (int k = 1; k < 3; ++k)
            if (a[k][k] < a[m][m]) m = k;
        const double len = std::sqrt(v[0][m] * v[0][m] + v[1][m] * v[1][m] + v[2][m] * v[2][m]);
        for (int k = 0; k < 3; ++k) out[k] = v[k][m] / len;
    }
    typename PointCloud<PointInT>::ConstPtr cloud_;
    typename search::KdTree<PointInT>::Ptr tree_;
    double radius_ = 0.0;
};

// pcl_conversions: fields matched by name; x, y, z must be FLOAT32 (anything else reads as zero, as upstream warns)
template <class T>
void fromROSMsg(const sensor_msgs::msg::PointCloud2& msg, PointCloud<T>& cloud) {
    int ox = -1, oy = -1, oz = -1, orgb = -1;
    for (const auto& f : msg.fields) {
        if (f.name == "x" && f.datatype == sensor_msgs::msg::PointField::FLOAT32) ox = static_cast<int>(f.offset);
        else if (f.name == "y" && f.datatype == sensor_msgs::msg::PointField::FLOAT32) oy = static_cast<int>(f.offset);
        else if (f.name == "z" && f.datatype == sensor_msgs::msg::PointField::FLOAT32) oz = static_cast<int>(f.offset);
        else if (f.name == "rgb" || f.name == "rgba") orgb = static_cast<int>(f.offset);
    }
    const size_t n = static_cast<size_t>(msg.width) * msg.height;
    cloud.points.assign(n, T());
    cloud.width = msg.width;
    cloud.height = msg.height;
    cloud.is_dense = msg.is_dense;
    for (size_t i = 0; i < n; ++i) {
        const uint8_t* p = msg.data.data() + i * msg.point_step;
        T& q = cloud.points[i];
        if (ox >= 0) std::memcpy(&q.x, p + ox, 4);
        if (oy >= 0) std::memcpy(&q.y, p + oy, 4);
        if (oz >= 0) std::memcpy(&q.z, p + oz, 4);
        if constexpr (sizeof(T) > 16) {
            if (orgb >= 0) std::memcpy(&q.rgba, p + orgb, 4);
        }
    }
}

// pcl::toROSMsg: the point struct's bytes with its registered fields (x, y, z at 0 / 4 / 8; rgb at 16 for PointXYZRGB)
template <class T>
void toROSMsg(const PointCloud<T>& cloud, sensor_msgs::msg::PointCloud2& msg) {
    auto field = [](const char* name, uint32_t offset) {
        sensor_msgs::msg::PointField f;
        f.name = name; f.offset = offset; f.datatype = sensor_msgs::msg::PointField::FLOAT32; f.count = 1;
        return f;
    };
    msg.fields = {field("x", 0), field("y", 4), field("z", 8)};
    if constexpr (sizeof(T) > 16) msg.fields.push_back(field("rgb", 16));
    // pcl::PointCloud::push_back keeps width = size, height = 1; a cloud never touched has width = height = 0 and
    // toPCLPointCloud2 then takes width = size, height = 1 as well
    msg.height = 1;
    msg.width = static_cast<uint32_t>(cloud.points.size());
    msg.point_step = sizeof(T);
    msg.row_step = msg.point_step * msg.width;
    msg.is_bigendian = false;
    msg.is_dense = cloud.is_dense;
    msg.data.resize(cloud.points.size() * sizeof(T));
    if (!cloud.points.empty()) std::memcpy(msg.data.data(), cloud.points.data(), msg.data.size());
}

// pcl::VoxelGrid<PointT>::applyFilter with the defaults the reference leaves (downsample_all_data, no field filter,
// min_points_per_voxel 0), SURVEY A.3 item 13: float inverse leaf, bounds over the finite points, index
// i + j dx + k dx dy from floor(p * inv) - min_b, one centroid per index in ascending order, float running sums in the
// order the (index, point) pairs come out of the sort — taken as ascending point index, the order the oracle fixes
// (the real sort is not stable: SURVEY flags this), divided by the count.
template <class T>
class VoxelGrid {
public:
    void setInputCloud(const typename PointCloud<T>::ConstPtr& cloud) { input_ = cloud; }
    void setLeafSize(float lx, float ly, float lz) { leaf_[0] = lx; leaf_[1] = ly; leaf_[2] = lz; }
    void filter(PointCloud<T>& out) {
        out.points.clear();
        out.height = 1;
        out.is_dense = true;
        const float inv[3] = {1.0f / leaf_[0], 1.0f / leaf_[1], 1.0f / leaf_[2]};
        float mn[3] = {std::numeric_limits<float>::max(), std::numeric_limits<float>::max(), std::numeric_limits<float>::max()};
        float mx[3] = {-mn[0], -mn[1], -mn[2]};
        for (const T& p : input_->points) {
            if (!(std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z))) continue;
            const float v[3] = {p.x, p.y, p.z};
            for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], v[a]); mx[a] = std::max(mx[a], v[a]); }
        }
        int64_t d64[3];
        for (int a = 0; a < 3; ++a) d64[a] = static_cast<int64_t>((mx[a] - mn[a]) * inv[a]) + 1;
        if (d64[0] * d64[1] * d64[2] > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) {
            out = *input_;                    // "Leaf size is too small for the input dataset"
            return;
        }
        int min_b[3], div_b[3];
        for (int a = 0; a < 3; ++a) {
            min_b[a] = static_cast<int>(std::floor(mn[a] * inv[a]));
            div_b[a] = static_cast<int>(std::floor(mx[a] * inv[a])) - min_b[a] + 1;
        }
        const int mul[3] = {1, div_b[0], div_b[0] * div_b[1]};
        std::vector<std::pair<unsigned, unsigned>> order;      // (voxel index, point index)
        for (size_t i = 0; i < input_->points.size(); ++i) {
            const T& p = input_->points[i];
            if (!(std::isfinite(p.x) && std::isfinite(p.y) && std::isfinite(p.z))) continue;
            const int ijk0 = static_cast<int>(std::floor(p.x * inv[0]) - static_cast<float>(min_b[0]));
            const int ijk1 = static_cast<int>(std::floor(p.y * inv[1]) - static_cast<float>(min_b[1]));
            const int ijk2 = static_cast<int>(std::floor(p.z * inv[2]) - static_cast<float>(min_b[2]));
            order.emplace_back(static_cast<unsigned>(ijk0 * mul[0] + ijk1 * mul[1] + ijk2 * mul[2]), static_cast<unsigned>(i));
        }
        std::sort(order.begin(), order.end());
        for (size_t a = 0; a < order.size();) {
            size_t b = a;
            float sx = 0.f, sy = 0.f, sz = 0.f;
            for (; b < order.size() && order[b].first == order[a].first; ++b) {
                const T& p = input_->points[order[b].second];
                sx += p.x; sy += p.y; sz += p.z;
            }
            T c;
            const float n = static_cast<float>(b - a);
            c.x = sx / n; c.y = sy / n; c.z = sz / n;
            out.points.push_back(c);
            a = b;
        }
        out.width = static_cast<uint32_t>(out.points.size());
    }
private:
    typename PointCloud<T>::ConstPtr input_;
    float leaf_[3] = {0.f, 0.f, 0.f};
};

}  // namespace pcl
