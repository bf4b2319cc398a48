// radius_index.cpp — see radius_index.hpp.  TEST INFRASTRUCTURE (oracle.h).
#include "radius_index.hpp"

#include <algorithm>
#include <limits>
#include <numeric>
#include <utility>

namespace orc {

namespace {

void sort_by_distance(std::vector<int>& idx, std::vector<float>& d2, size_t from) {
    // FLANN copies a RadiusResultSet out sorted by distance when SearchParams.sorted is set
    // (PCL's default).  Order never affects the sweep (existence test only).
    size_t n = idx.size() - from;
    if (n < 2) return;
    std::vector<std::pair<float, int>> tmp(n);
    for (size_t i = 0; i < n; ++i) tmp[i] = {d2[from + i], idx[from + i]};
    std::sort(tmp.begin(), tmp.end());
    for (size_t i = 0; i < n; ++i) {
        d2[from + i] = tmp[i].first;
        idx[from + i] = tmp[i].second;
    }
}

// ------------------------------------------------------------------------------- brute force
class Brute : public RadiusIndex {
public:
    void search(float qx, float qy, float qz, float r2f, std::vector<int>& idx,
                std::vector<float>& d2, bool sorted) const override {
        size_t from = idx.size();
        for (size_t i = 0; i < x.size(); ++i) {
            float d = l2_simple(qx, qy, qz, x[i], y[i], z[i]);
            if (d < r2f) {
                idx.push_back(src[i]);
                d2.push_back(d);
            }
        }
        if (sorted) sort_by_distance(idx, d2, from);
    }
    bool any(float qx, float qy, float qz, float r2f) const override {
        for (size_t i = 0; i < x.size(); ++i)
            if (l2_simple(qx, qy, qz, x[i], y[i], z[i]) < r2f) return true;
        return false;
    }
    int64_t count(float qx, float qy, float qz, float r2f) const override {
        int64_t c = 0;
        for (size_t i = 0; i < x.size(); ++i) c += l2_simple(qx, qy, qz, x[i], y[i], z[i]) < r2f;
        return c;
    }
};

// A real-number radius that certainly contains every point the float32 test accepts.
double safe_radius(float r2f) { return std::sqrt(static_cast<double>(r2f)) * (1.0 + 1e-5) + 1e-9; }

// ------------------------------------------------------------------------------- uniform grid
class Grid : public RadiusIndex {
public:
    void build() {
        if (x.empty()) return;
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (size_t i = 0; i < x.size(); ++i) {
            double p[3] = {x[i], y[i], z[i]};
            for (int d = 0; d < 3; ++d) {
                lo[d] = std::min(lo[d], p[d]);
                hi[d] = std::max(hi[d], p[d]);
            }
        }
        h = 0.25;
        for (;;) {
            double cells = 1.0;
            for (int d = 0; d < 3; ++d) {
                dim[d] = static_cast<int>(std::floor((hi[d] - lo[d]) / h)) + 1;
                cells *= dim[d];
            }
            if (cells <= 64.0 * 1024 * 1024) break;
            h *= 2.0;
        }
        for (int d = 0; d < 3; ++d) org[d] = lo[d];
        size_t ncell = static_cast<size_t>(dim[0]) * dim[1] * dim[2];
        start.assign(ncell + 1, 0);
        std::vector<uint32_t> cell_of(x.size());
        for (size_t i = 0; i < x.size(); ++i) {
            cell_of[i] = static_cast<uint32_t>(linear(bin(x[i], 0), bin(y[i], 1), bin(z[i], 2)));
            start[cell_of[i] + 1]++;
        }
        for (size_t c = 0; c < ncell; ++c) start[c + 1] += start[c];
        std::vector<uint32_t> fill(start.begin(), start.end() - 1);
        order.resize(x.size());
        for (size_t i = 0; i < x.size(); ++i) order[fill[cell_of[i]]++] = static_cast<uint32_t>(i);
    }
    template <class F>
    void visit(float qx, float qy, float qz, float r2f, F&& f) const {
        if (x.empty()) return;
        double R = safe_radius(r2f);
        double q[3] = {qx, qy, qz};
        int a[3], b[3];
        for (int d = 0; d < 3; ++d) {
            double l = std::floor((q[d] - R - org[d]) / h), u = std::floor((q[d] + R - org[d]) / h);
            if (u < 0 || l > dim[d] - 1) return;
            a[d] = static_cast<int>(std::max(l, 0.0));
            b[d] = static_cast<int>(std::min(u, static_cast<double>(dim[d] - 1)));
        }
        for (int k = a[2]; k <= b[2]; ++k)
            for (int j = a[1]; j <= b[1]; ++j) {
                // buckets adjacent in x are adjacent in `order`
                size_t c0 = linear(a[0], j, k), c1 = linear(b[0], j, k);
                for (uint32_t s = start[c0]; s < start[c1 + 1]; ++s) {
                    uint32_t i = order[s];
                    float d = l2_simple(qx, qy, qz, x[i], y[i], z[i]);
                    if (d < r2f)
                        if (f(i, d)) return;
                }
            }
    }
    void search(float qx, float qy, float qz, float r2f, std::vector<int>& idx,
                std::vector<float>& d2, bool sorted) const override {
        size_t from = idx.size();
        visit(qx, qy, qz, r2f, [&](uint32_t i, float d) {
            idx.push_back(src[i]);
            d2.push_back(d);
            return false;
        });
        if (sorted) sort_by_distance(idx, d2, from);
    }
    bool any(float qx, float qy, float qz, float r2f) const override {
        bool hit = false;
        visit(qx, qy, qz, r2f, [&](uint32_t, float) {
            hit = true;
            return true;
        });
        return hit;
    }
    int64_t count(float qx, float qy, float qz, float r2f) const override {
        int64_t c = 0;
        visit(qx, qy, qz, r2f, [&](uint32_t, float) {
            ++c;
            return false;
        });
        return c;
    }

private:
    int bin(float v, int d) const {
        int b = static_cast<int>(std::floor((static_cast<double>(v) - org[d]) / h));
        return std::min(std::max(b, 0), dim[d] - 1);
    }
    size_t linear(int i, int j, int k) const {
        return (static_cast<size_t>(k) * dim[1] + j) * dim[0] + i;
    }
    double h = 0.25, org[3] = {0, 0, 0};
    int dim[3] = {0, 0, 0};
    std::vector<uint32_t> start, order;
};

// ------------------------------------------------------------------------------- kd-tree
// Shape of FLANN's KDTreeSingleIndex as PCL configures it: bounding-box tree, split on the
// widest dimension at the (clamped) box midpoint, leaves of at most 15 points, dataset reordered
// into leaf order.  Pruning here is done in double with a safety factor so that the answer is
// always the brute-force answer (FLANN prunes in float; it could differ only for a point whose
// distance is within an ulp of the radius — inside the parity contract's grazing tolerance).
class KdTree : public RadiusIndex {
public:
    static constexpr int kLeaf = 15;
    struct Node {
        int left, right;        // leaf: point range [left,right) in reordered arrays
        int child0, child1;     // inner: children (child0 < 0 for leaf)
        int axis;
        float divlow, divhigh;  // max of left child / min of right child along axis
    };
    void build() {
        size_t n = x.size();
        if (!n) return;
        perm.resize(n);
        std::iota(perm.begin(), perm.end(), 0);
        float lo[3], hi[3];
        bounds(0, static_cast<int>(n), lo, hi);
        for (int d = 0; d < 3; ++d) {
            root_lo[d] = lo[d];
            root_hi[d] = hi[d];
        }
        nodes.reserve(2 * n / kLeaf + 16);
        divide(0, static_cast<int>(n), lo, hi);
        rx.resize(n);
        ry.resize(n);
        rz.resize(n);
        for (size_t i = 0; i < n; ++i) {
            rx[i] = x[perm[i]];
            ry[i] = y[perm[i]];
            rz[i] = z[perm[i]];
        }
    }
    template <class F>
    void visit(float qx, float qy, float qz, float r2f, F&& f) const {
        if (nodes.empty()) return;
        double q[3] = {qx, qy, qz};
        double dists[3], mind = 0.0;
        for (int d = 0; d < 3; ++d) {
            double e = 0.0;
            if (q[d] < root_lo[d]) e = root_lo[d] - q[d];
            if (q[d] > root_hi[d]) e = q[d] - root_hi[d];
            dists[d] = e * e;
            mind += dists[d];
        }
        double worst = static_cast<double>(r2f) * (1.0 + 1e-4) + 1e-12;
        bool stop = false;
        descend(0, q, qx, qy, qz, r2f, worst, mind, dists, f, stop);
    }
    void search(float qx, float qy, float qz, float r2f, std::vector<int>& idx,
                std::vector<float>& d2, bool sorted) const override {
        size_t from = idx.size();
        visit(qx, qy, qz, r2f, [&](int i, float d) {
            idx.push_back(src[perm[i]]);
            d2.push_back(d);
            return false;
        });
        if (sorted) sort_by_distance(idx, d2, from);
    }
    bool any(float qx, float qy, float qz, float r2f) const override {
        bool hit = false;
        visit(qx, qy, qz, r2f, [&](int, float) {
            hit = true;
            return true;
        });
        return hit;
    }
    int64_t count(float qx, float qy, float qz, float r2f) const override {
        int64_t c = 0;
        visit(qx, qy, qz, r2f, [&](int, float) {
            ++c;
            return false;
        });
        return c;
    }

private:
    void bounds(int l, int r, float* lo, float* hi) const {
        for (int d = 0; d < 3; ++d) {
            lo[d] = std::numeric_limits<float>::max();
            hi[d] = std::numeric_limits<float>::lowest();
        }
        for (int i = l; i < r; ++i) {
            float p[3] = {x[perm[i]], y[perm[i]], z[perm[i]]};
            for (int d = 0; d < 3; ++d) {
                lo[d] = std::min(lo[d], p[d]);
                hi[d] = std::max(hi[d], p[d]);
            }
        }
    }
    float coord(int pi, int axis) const { return axis == 0 ? x[pi] : axis == 1 ? y[pi] : z[pi]; }
    int divide(int l, int r, const float* lo, const float* hi) {
        int me = static_cast<int>(nodes.size());
        nodes.push_back(Node{l, r, -1, -1, 0, 0.f, 0.f});
        if (r - l <= kLeaf) return me;
        int axis = 0;
        float span = hi[0] - lo[0];
        for (int d = 1; d < 3; ++d)
            if (hi[d] - lo[d] > span) {
                span = hi[d] - lo[d];
                axis = d;
            }
        float cut = 0.5f * (lo[axis] + hi[axis]);
        // three-way partition around the cut, then take the split nearest to the median
        int lim1 = l, lim2;
        {
            int i = l, j = r - 1;
            while (i <= j) {
                if (coord(perm[i], axis) < cut) ++i;
                else std::swap(perm[i], perm[j--]);
            }
            lim1 = i;
            j = r - 1;
            while (i <= j) {
                if (coord(perm[i], axis) <= cut) ++i;
                else std::swap(perm[i], perm[j--]);
            }
            lim2 = i;
        }
        int half = l + (r - l) / 2, mid;
        if (lim1 > half) mid = lim1;
        else if (lim2 < half) mid = lim2;
        else mid = half;
        if (mid == l || mid == r) {  // all coordinates equal along the widest axis: split by count
            mid = half;
        }
        float llo[3], lhi[3], rlo[3], rhi[3];
        bounds(l, mid, llo, lhi);
        bounds(mid, r, rlo, rhi);
        int c0 = divide(l, mid, llo, lhi);
        int c1 = divide(mid, r, rlo, rhi);
        nodes[me].child0 = c0;
        nodes[me].child1 = c1;
        nodes[me].axis = axis;
        nodes[me].divlow = lhi[axis];
        nodes[me].divhigh = rlo[axis];
        return me;
    }
    template <class F>
    void descend(int ni, const double* q, float qx, float qy, float qz, float r2f, double worst,
                 double mind, double* dists, F& f, bool& stop) const {
        const Node& nd = nodes[ni];
        if (nd.child0 < 0) {
            for (int i = nd.left; i < nd.right && !stop; ++i) {
                float d = l2_simple(qx, qy, qz, rx[i], ry[i], rz[i]);
                if (d < r2f) stop = f(i, d);
            }
            return;
        }
        int ax = nd.axis;
        double val = q[ax];
        double diff1 = val - nd.divlow, diff2 = val - nd.divhigh;
        int best, other;
        double cut;
        if (diff1 + diff2 < 0) {
            best = nd.child0;
            other = nd.child1;
            cut = diff2 * diff2;
        } else {
            best = nd.child1;
            other = nd.child0;
            cut = diff1 * diff1;
        }
        descend(best, q, qx, qy, qz, r2f, worst, mind, dists, f, stop);
        if (stop) return;
        double saved = dists[ax];
        // boxes nest, so the separation along `ax` can only grow: replace the ancestor's share
        double m2 = mind + cut - saved;
        if (m2 <= worst) {
            dists[ax] = cut;
            descend(other, q, qx, qy, qz, r2f, worst, m2, dists, f, stop);
            dists[ax] = saved;
        }
    }
    std::vector<Node> nodes;
    std::vector<int> perm;
    std::vector<float> rx, ry, rz;
    double root_lo[3], root_hi[3];
};

}  // namespace

RadiusIndex* make_index(const float* xyz, uint64_t n, int kind) {
    RadiusIndex* ix;
    Grid* g = nullptr;
    KdTree* k = nullptr;
    if (kind == 1) ix = g = new Grid();
    else if (kind == 2) ix = k = new KdTree();
    else ix = new Brute();
    ix->x.reserve(n);
    ix->y.reserve(n);
    ix->z.reserve(n);
    ix->src.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        float a = xyz[3 * i], b = xyz[3 * i + 1], c = xyz[3 * i + 2];
        if (!std::isfinite(a) || !std::isfinite(b) || !std::isfinite(c)) continue;
        ix->x.push_back(a);
        ix->y.push_back(b);
        ix->z.push_back(c);
        ix->src.push_back(static_cast<int>(i));
    }
    if (g) g->build();
    if (k) k->build();
    return ix;
}

}  // namespace orc
