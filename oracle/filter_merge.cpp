// filter_merge.cpp — CPU restatement of the pointcloud_filter and pointcloud_merger data paths.
// TEST INFRASTRUCTURE (oracle.h).
//
// Follows /root/reference/src/pointcloud_filter.cpp:87-120 (cropFrontArea), :122-139
// (downsampleCloud -> pcl::VoxelGrid<PointXYZ>, third-party, restated from PCL 1.12's published
// applyFilter: SURVEY §A.3 item 13) and /root/reference/src/pointcloud_merger.cpp:354-387
// (tf2::doTransform float affine, third-party, §A.3 item 14).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

#include "oracle.h"

// cropFrontArea, pointcloud_filter.cpp:106-116 — float coordinates promoted to double, six strict tests.
extern "C" int64_t orc_crop(const float* xyz, uint64_t n, const plc_crop_box* box, float* out_xyz,
                            int64_t* out_index) {
    int64_t kept = 0;
    for (uint64_t i = 0; i < n; ++i) {
        float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
        if (px > box->x_min && px < box->x_max && py > box->y_min && py < box->y_max &&
            pz > box->z_min && pz < box->z_max) {
            if (out_xyz) {
                out_xyz[3 * kept] = px;
                out_xyz[3 * kept + 1] = py;
                out_xyz[3 * kept + 2] = pz;
            }
            if (out_index) out_index[kept] = static_cast<int64_t>(i);
            ++kept;
        }
    }
    return kept;
}

// pcl::VoxelGrid<pcl::PointXYZ>::applyFilter with the node's settings (leaf only).
extern "C" int64_t orc_voxel_grid(const float* xyz, uint64_t n, double leaf, float* out_xyz,
                                  int32_t* out_key, int32_t* out_count, int32_t* point_key) {
    if (n == 0) return 0;
    const float leaf_f = static_cast<float>(leaf);      // setLeafSize(float, float, float)
    const float inv = 1.0f / leaf_f;                    // inverse_leaf_size_
    float mn[3], mx[3];
    for (int d = 0; d < 3; ++d) {
        mn[d] = std::numeric_limits<float>::max();
        mx[d] = -std::numeric_limits<float>::max();
    }
    for (uint64_t i = 0; i < n; ++i)
        for (int d = 0; d < 3; ++d) {
            mn[d] = std::min(mn[d], xyz[3 * i + d]);
            mx[d] = std::max(mx[d], xyz[3 * i + d]);
        }
    int64_t ext[3];
    for (int d = 0; d < 3; ++d) ext[d] = static_cast<int64_t>((mx[d] - mn[d]) * inv) + 1;
    if (ext[0] * ext[1] * ext[2] > static_cast<int64_t>(std::numeric_limits<int32_t>::max())) return -1;
    int min_b[3], max_b[3], div_b[3];
    for (int d = 0; d < 3; ++d) {
        min_b[d] = static_cast<int>(std::floor(mn[d] * inv));
        max_b[d] = static_cast<int>(std::floor(mx[d] * inv));
        div_b[d] = max_b[d] - min_b[d] + 1;
    }
    const int mul[3] = {1, div_b[0], div_b[0] * div_b[1]};
    std::vector<std::pair<uint32_t, uint32_t>> pairs(n);   // (voxel idx, point index)
    for (uint64_t i = 0; i < n; ++i) {
        int ijk[3];
        for (int d = 0; d < 3; ++d)
            ijk[d] = static_cast<int>(std::floor(xyz[3 * i + d] * inv) - static_cast<float>(min_b[d]));
        int idx = ijk[0] * mul[0] + ijk[1] * mul[1] + ijk[2] * mul[2];
        pairs[i] = {static_cast<uint32_t>(idx), static_cast<uint32_t>(i)};
        if (point_key) point_key[i] = idx;
    }
    // PCL sorts by idx only (unstable); sorting (idx, point index) pairs fixes one of the orders
    // PCL may produce.  The member order only moves the float32 centroid in its last bits.
    std::sort(pairs.begin(), pairs.end());
    int64_t out = 0;
    uint64_t i = 0;
    while (i < n) {
        uint64_t j = i;
        float sx = 0.f, sy = 0.f, sz = 0.f;   // AccumulatorXYZ: Eigen::Vector3f running sum
        while (j < n && pairs[j].first == pairs[i].first) {
            uint32_t p = pairs[j].second;
            sx += xyz[3 * p];
            sy += xyz[3 * p + 1];
            sz += xyz[3 * p + 2];
            ++j;
        }
        float cnt = static_cast<float>(j - i);
        if (out_xyz) {
            out_xyz[3 * out] = sx / cnt;
            out_xyz[3 * out + 1] = sy / cnt;
            out_xyz[3 * out + 2] = sz / cnt;
        }
        if (out_key) out_key[out] = static_cast<int32_t>(pairs[i].first);
        if (out_count) out_count[out] = static_cast<int32_t>(j - i);
        ++out;
        i = j;
    }
    return out;
}

// tf2_sensor_msgs doTransform: Eigen::Translation3f * Eigen::Quaternionf applied per point in float32.
extern "C" void orc_transform(const float* xyz, uint64_t n, const double* t, const double* q,
                              float* out_xyz) {
    const float qx = static_cast<float>(q[0]), qy = static_cast<float>(q[1]),
                qz = static_cast<float>(q[2]), qw = static_cast<float>(q[3]);
    const float tr[3] = {static_cast<float>(t[0]), static_cast<float>(t[1]), static_cast<float>(t[2])};
    // Eigen QuaternionBase::toRotationMatrix
    const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
    const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    const float m[3][3] = {{1.0f - (tyy + tzz), txy - twz, txz + twy},
                           {txy + twz, 1.0f - (txx + tzz), tyz - twx},
                           {txz - twy, tyz + twx, 1.0f - (txx + tyy)}};
    for (uint64_t i = 0; i < n; ++i) {
        float v0 = xyz[3 * i], v1 = xyz[3 * i + 1], v2 = xyz[3 * i + 2];
        for (int r = 0; r < 3; ++r) {
            float acc = m[r][0] * v0;
            acc = acc + m[r][1] * v1;
            acc = acc + m[r][2] * v2;
            out_xyz[3 * i + r] = acc + tr[r];
        }
    }
}
