// radius_index.hpp — three interchangeable point indices answering the reference's only spatial
// query: "which indexed points lie within `radius` of q" with PCL/FLANN float32 semantics
// (SURVEY §A.3 item 2).  TEST INFRASTRUCTURE (see oracle.h).
//
//   Brute   O(N) scan — the definition.
//   Grid    uniform buckets — used by the fast oracle paths.
//   KdTree  FLANN KDTreeSingleIndex-shaped tree (leaf size 15 as PCL passes, points reordered,
//           results gathered then sorted by distance) — used as the *timed* CPU baseline because it
//           has the cost structure of the reference's KdTreeFLANN::radiusSearch calls
//           (virtual_lidar.cpp:779-783).
// All three must return identical neighbour sets; tests/test_oracle.py checks that bit for bit.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace orc {

// FLANN L2_Simple<float>::operator(): result starts at 0, adds diff*diff per dimension in x,y,z
// order, every operation rounded to float32 (build with -ffp-contract=off: no FMA, as distro FLANN).
static inline float l2_simple(float qx, float qy, float qz, float px, float py, float pz) {
    float result = 0.0f;
    float diff = qx - px;
    result += diff * diff;
    diff = qy - py;
    result += diff * diff;
    diff = qz - pz;
    result += diff * diff;
    return result;
}

// PCL: flann_index_->radiusSearch(query, ..., static_cast<float>(radius * radius), params);
// RadiusResultSet::addPoint accepts iff dist < radius (strict).
static inline float radius_to_r2f(double radius) { return static_cast<float>(radius * radius); }

class RadiusIndex {
public:
    virtual ~RadiusIndex() {}
    // Appends (index, d2) of every point with d2 < r2f.  `sorted` mimics PCL's default sorted_=true.
    virtual void search(float qx, float qy, float qz, float r2f, std::vector<int>& idx,
                        std::vector<float>& d2, bool sorted) const = 0;
    // Early-out existence test with the same acceptance rule.
    virtual bool any(float qx, float qy, float qz, float r2f) const = 0;
    virtual int64_t count(float qx, float qy, float qz, float r2f) const = 0;
    size_t size() const { return x.size(); }
    // finite points, in input order (original index in `src`)
    std::vector<float> x, y, z;
    std::vector<int> src;
};

RadiusIndex* make_index(const float* xyz, uint64_t n, int kind);

}  // namespace orc
