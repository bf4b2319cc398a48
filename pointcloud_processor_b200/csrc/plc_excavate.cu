// plc_excavate.cu — the excavated_surface_generator node's data path on the device: the
// producer of both inputs of the sweep, /excavated_terrain and /excavation_area.
//
// Replaces /root/reference/src/excavated_surface_generator.cpp:
//   getTerrainHeight          183-225  -> terrain_height()   (one warp per query; upstream rebuilds a KD-tree per call)
//   isInsideAnyBox, isOuterEdge 228-262 -> inside_any_box(), k_plan_lattice
//   isInsideExcavationArea    327-348  -> k_carve_prefilter (height-independent rejection) + k_carve_heights
//   processExcavation         457-481  -> k_carve_prefilter, k_carve_heights, scan, k_emit_kept
//   generateExcavationArea    350-455  -> k_plan_lattice, k_lattice_heights, k_emit_area
//   generateExcavatedSurface  487-584  -> k_plan_lattice, k_lattice_heights, k_emit_bottoms, k_emit_slopes
// (excavated_surface_generator_L.cpp, not built upstream, differs only in the area cloud's wall points: see
// plc_excavation_params::area_walls_vertical.)
//
// Parity against the CPU restatement: the kept set, every generated x/y and the output order are bit-identical;
// a height is a mean of hundreds of float z values summed in double — upstream in KD-tree result order, here in a
// fixed lane/tree order — so a generated z can differ in its last float32 bit, and a point whose carve test is
// decided inside that 1e-16 relative band could flip (none observed).
//
// HBM/L2-bound gathers: 16 B per neighbour tested; the carve needs heights only for the points inside the pit's
// outline widened by the slope (a height-independent bound), not for all N.
#include "plc_buckets.cuh"

namespace plc {
namespace {

struct Boxes {
    int n;
    double cx[2], cy[2], len[2], wid[2], min_x[2], max_x[2], min_y[2], max_y[2];
};

struct ExcavDev {
    Boxes B;
    double center_x, center_y;
    double cos_yaw, sin_yaw, cos_neg, sin_neg;     // libm values of cos(yaw), sin(yaw), cos(-yaw), sin(-yaw), as upstream forms them per point
    double depth, slope_offset, pd, radius;
    double overall_min_x, overall_min_y;
    double reach;                                  // widened search radius for the bucket walk
    int n_x, n_y, n_depth, n_slope;
    float r2f;
    int have_points;                               // the cloud is not empty (:185)
    int area_walls_vertical;                       // excavated_surface_generator_L.cpp: wall points of the area cloud without the slope offset
    unsigned n_finite;
};

__device__ __forceinline__ bool inside_any_box(const Boxes& B, double x, double y) {           // :228-236
    for (int b = 0; b < B.n; ++b)
        if (x >= B.min_x[b] && x <= B.max_x[b] && y >= B.min_y[b] && y <= B.max_y[b]) return true;
    return false;
}

// getTerrainHeight (:183-225), all lanes of the warp with the same (x, y); every lane returns the result
__device__ double terrain_height(const BucketGrid& G, const ExcavDev& E, double x, double y, int lane) {
    if (!E.have_points) return 0.0;                                                             // :185
    const unsigned FULL = 0xffffffffu;
    const float qx = (float)x, qy = (float)y;                                                   // :190-193
    double sum_z = 0.0;
    unsigned valid = 0;
    if (E.n_finite) {
        const RowRange R = rows_of_box(G, qx, qy, 0.0, E.reach, E.reach, E.reach);
        if (!R.empty())
            for (int bz = R.z0; bz <= R.z1; ++bz)
                for (int by = R.y0; by <= R.y1; ++by) {
                    const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                    const uint32_t e = __ldg(&G.start[row + R.x1 + 1]);
                    for (uint32_t p = __ldg(&G.start[row + R.x0]) + lane; p < e; p += 32) {
                        const float4 v = __ldg(&G.pts[p]);
                        if (!(dist2(qx, qy, 0.f, v.x, v.y, v.z) < E.r2f)) continue;               // :198-199
                        const double dx = __dsub_rn((double)v.x, x), dy = __dsub_rn((double)v.y, y);
                        if (__dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy))) <= E.radius) {   // :204-208
                            sum_z = __dadd_rn(sum_z, (double)v.z);
                            ++valid;
                        }
                    }
                }
    }
    valid = __reduce_add_sync(FULL, valid);
    if (valid > 0) return __ddiv_rn(warp_sum(sum_z), (double)(int)valid);                       // :213-215
    if (!E.n_finite) return 0.0;
    // nearestKSearch(search_point, 1) (:219-223): rings of buckets around the query's bucket; after ring k every
    // unvisited point is at least k bucket edges away, so the search ends once the best distance is inside that
    const double edge = 1.0 / G.inv_e;
    const int cbx = bucket_of((double)qx, G.ox, G.inv_e, G.nx), cby = bucket_of((double)qy, G.oy, G.inv_e, G.ny),
              cbz = bucket_of(0.0, G.oz, G.inv_e, G.nz);
    float best = INFINITY, best_z = 0.f;
    unsigned best_i = 0xffffffffu;
    const int k_max = max(G.nx, max(G.ny, G.nz));
    for (int k = 0; k <= k_max; ++k) {
        const int z0 = max(cbz - k, 0), z1 = min(cbz + k, G.nz - 1), y0 = max(cby - k, 0), y1 = min(cby + k, G.ny - 1);
        const int x0 = max(cbx - k, 0), x1 = min(cbx + k, G.nx - 1);
        for (int bz = z0; bz <= z1; ++bz)
            for (int by = y0; by <= y1; ++by) {
                const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                const bool shell = (bz == cbz - k) || (bz == cbz + k) || (by == cby - k) || (by == cby + k);
                // on the shell's faces the whole row segment is new; elsewhere only its two end buckets
                for (int part = 0; part < (shell ? 1 : 2); ++part) {
                    int sx0 = x0, sx1 = x1;
                    if (!shell) {
                        const int bx = part == 0 ? cbx - k : cbx + k;
                        if (bx < 0 || bx >= G.nx || (part == 1 && k == 0)) continue;
                        sx0 = sx1 = bx;
                    }
                    const uint32_t e = __ldg(&G.start[row + sx1 + 1]);
                    for (uint32_t p = __ldg(&G.start[row + sx0]) + lane; p < e; p += 32) {
                        const float4 v = __ldg(&G.pts[p]);
                        const float d = dist2(qx, qy, 0.f, v.x, v.y, v.z);
                        const unsigned id = __float_as_uint(v.w);
                        if (d < best || (d == best && id < best_i)) {       // lowest index among equidistant points
                            best = d;
                            best_i = id;
                            best_z = v.z;
                        }
                    }
                }
            }
        float wb = best, wz = best_z;
        unsigned wi = best_i;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(FULL, wb, o), oz = __shfl_xor_sync(FULL, wz, o);
            const unsigned oi = __shfl_xor_sync(FULL, wi, o);
            if (ob < wb || (ob == wb && oi < wi)) {
                wb = ob;
                wi = oi;
                wz = oz;
            }
        }
        const double safe = (double)k * edge;
        if ((wi != 0xffffffffu && (double)wb < safe * safe * (1.0 - 1e-5)) || k == k_max) {
            best_i = wi;
            best_z = wz;
            break;
        }
    }
    if (best_i == 0xffffffffu) return 0.0;
    return (double)best_z;                                                                      // :220-222
}

// ---- carve (:457-472) ---------------------------------------------------------------------------------
// One thread per input point.  keep[i] = 1 where the carve test cannot succeed whatever the terrain height is
// (non-finite coordinates, or outside every box widened by the full slope offset); the others are queued for
// k_carve_heights.
__global__ void __launch_bounds__(256)
k_carve_prefilter(const float4* __restrict__ pts, unsigned n, ExcavDev E, uint32_t* __restrict__ keep,
                  uint32_t* __restrict__ queue, unsigned* __restrict__ queue_n) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = pts[i];
    bool need = false;
    if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
        const double dx = __dsub_rn((double)p.x, E.center_x), dy = __dsub_rn((double)p.y, E.center_y);
        const double xl = __dsub_rn(__dmul_rn(dx, E.cos_neg), __dmul_rn(dy, E.sin_neg));           // :464
        const double yl = __dadd_rn(__dmul_rn(dx, E.sin_neg), __dmul_rn(dy, E.cos_neg));           // :465
        const double grow = fmax(E.slope_offset, 0.0) * (1.0 + 1e-9) + 1e-9;
        for (int b = 0; b < E.B.n; ++b)
            if (fabs(xl - E.B.cx[b]) <= E.B.len[b] / 2.0 + grow && fabs(yl - E.B.cy[b]) <= E.B.wid[b] / 2.0 + grow) need = true;
    }
    keep[i] = need ? 0u : 1u;
    if (need) queue[atomicAdd(queue_n, 1u)] = i;
}

// one warp per queued point: its terrain height, then isInsideExcavationArea (:327-348)
__global__ void __launch_bounds__(256)
k_carve_heights(BucketGrid G, ExcavDev E, const float4* __restrict__ pts, const uint32_t* __restrict__ queue, unsigned n_queue,
                uint32_t* __restrict__ keep) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_queue) return;
    const unsigned i = queue[w];
    const float4 p = pts[i];
    const double h = terrain_height(G, E, (double)p.x, (double)p.y, lane);                          // :467
    if (lane != 0) return;
    const double z_rel = __dsub_rn((double)p.z, h);                                                 // :468
    bool inside = false;
    if (!(z_rel < -E.depth || z_rel > 0)) {                                                         // :329
        const double dx0 = __dsub_rn((double)p.x, E.center_x), dy0 = __dsub_rn((double)p.y, E.center_y);
        const double xl = __dsub_rn(__dmul_rn(dx0, E.cos_neg), __dmul_rn(dy0, E.sin_neg));
        const double yl = __dadd_rn(__dmul_rn(dx0, E.sin_neg), __dmul_rn(dy0, E.cos_neg));
        const double factor = __ddiv_rn(__dadd_rn(E.depth, z_rel), E.depth);                        // :332
        const double cur = __dmul_rn(E.slope_offset, factor);                                       // :333
        for (int b = 0; b < E.B.n; ++b) {
            const double dx = __dsub_rn(xl, E.B.cx[b]), dy = __dsub_rn(yl, E.B.cy[b]);
            const double hl = __dadd_rn(__ddiv_rn(E.B.len[b], 2.0), cur), hw = __dadd_rn(__ddiv_rn(E.B.wid[b], 2.0), cur);
            if (fabs(dx) <= hl && fabs(dy) <= hw) inside = true;                                    // :342
        }
    }
    keep[i] = inside ? 0u : 1u;
}

struct Rec32 {                 // pcl::PointXYZRGB as pcl::toROSMsg writes it
    uint4 lo, hi;
};
__device__ __forceinline__ Rec32 make_rec(double x, double y, double z, unsigned rgba) {
    Rec32 r;
    r.lo = make_uint4(__float_as_uint((float)x), __float_as_uint((float)y), __float_as_uint((float)z), 0x3f800000u);
    r.hi = make_uint4(rgba, 0u, 0u, 0u);
    return r;
}

// kept input points, in input order, with their own colour (:470-472)
__global__ void __launch_bounds__(256)
k_emit_kept(const float4* __restrict__ pts, unsigned n, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ slot,
            const unsigned char* __restrict__ raw, unsigned step, unsigned off_rgb, Rec32* __restrict__ out) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    const float4 p = pts[i];
    unsigned rgba = 0xff000000u;              // PointXYZRGB's default when the message carries no rgb field
    if (off_rgb != 0xffffffffu) {
        const unsigned char* c = raw + (size_t)i * step + off_rgb;
        rgba = (unsigned)c[0] | ((unsigned)c[1] << 8) | ((unsigned)c[2] << 16) | ((unsigned)c[3] << 24);
    }
    Rec32 r;
    r.lo = make_uint4(__float_as_uint(p.x), __float_as_uint(p.y), __float_as_uint(p.z), 0x3f800000u);
    r.hi = make_uint4(rgba, 0u, 0u, 0u);
    out[slot[i]] = r;
}

// ---- the pit's lattice (:350-455, :487-584) -----------------------------------------------------------
// entry e = i * (n_y + 1) + j, i over x (outer), j over y, as both upstream loops run
struct LatticePlan {
    uint32_t* area_off;      // exclusive scans of the per-entry point counts (one extra element = totals)
    uint32_t* bottom_off;
    uint32_t* slope_off;
    signed char* dir;        // 2 per entry: outward direction on x and y (-1, 0, +1); dir[2e] = 2 marks "not inside"
    unsigned char* edge;     // isOuterEdge
};

__device__ __forceinline__ void lattice_local(const ExcavDev& E, unsigned e, double& xl, double& yl) {
    const unsigned i = e / (unsigned)(E.n_y + 1), j = e % (unsigned)(E.n_y + 1);
    xl = __dadd_rn(E.overall_min_x, __dmul_rn((double)(int)i, E.pd));           // :387, :520
    yl = __dadd_rn(E.overall_min_y, __dmul_rn((double)(int)j, E.pd));
}

__global__ void __launch_bounds__(256)
k_plan_lattice(ExcavDev E, unsigned n_entries, LatticePlan P) {
    const unsigned e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_entries) return;
    double xl, yl;
    lattice_local(E, e, xl, yl);
    const bool in = inside_any_box(E.B, xl, yl);
    const bool xp = !inside_any_box(E.B, __dadd_rn(xl, E.pd), yl), xm = !inside_any_box(E.B, __dsub_rn(xl, E.pd), yl);
    const bool yp = !inside_any_box(E.B, xl, __dadd_rn(yl, E.pd)), ym = !inside_any_box(E.B, xl, __dsub_rn(yl, E.pd));
    const bool edge = in && (xp || xm || yp || ym);                             // isOuterEdge with tolerance = point_density
    P.edge[e] = edge ? 1 : 0;
    P.dir[2 * e] = in ? (xp ? 1 : (xm ? -1 : 0)) : 2;                           // :420-424, :552-556
    P.dir[2 * e + 1] = yp ? 1 : (ym ? -1 : 0);                                  // :426-430, :558-562
    P.area_off[e] = in ? 1u + (edge ? (unsigned)max(E.n_depth - 1, 0) : 0u) : 0u;
    P.bottom_off[e] = in ? 1u : 0u;
    P.slope_off[e] = edge ? (unsigned)(E.n_slope + 1) : 0u;
}

__device__ __forceinline__ void to_global(const ExcavDev& E, double xl, double yl, double& xg, double& yg) {
    xg = __dsub_rn(__dadd_rn(E.center_x, __dmul_rn(xl, E.cos_yaw)), __dmul_rn(yl, E.sin_yaw));     // :394, :527
    yg = __dadd_rn(__dadd_rn(E.center_y, __dmul_rn(xl, E.sin_yaw)), __dmul_rn(yl, E.cos_yaw));     // :395, :528
}

// one warp per lattice entry inside the outline: the terrain height under it (shared by the area cloud's bottom and
// slope points and by the surface's bottom point, which all query the same (x, y))
__global__ void __launch_bounds__(256)
k_lattice_heights(BucketGrid G, ExcavDev E, unsigned n_entries, LatticePlan P, double* __restrict__ height) {
    const unsigned e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (e >= n_entries || P.dir[2 * e] == 2) return;
    double xl, yl, xg, yg;
    lattice_local(E, e, xl, yl);
    to_global(E, xl, yl, xg, yg);
    const double h = terrain_height(G, E, xg, yg, lane);
    if (lane == 0) height[e] = h;
}

__global__ void __launch_bounds__(256)
k_emit_area(ExcavDev E, unsigned n_entries, LatticePlan P, const double* __restrict__ height, Rec32* __restrict__ out) {
    const unsigned e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_entries || P.dir[2 * e] == 2) return;
    double xl, yl, xg, yg;
    lattice_local(E, e, xl, yl);
    to_global(E, xl, yl, xg, yg);
    const double h = height[e];
    Rec32* o = out + P.area_off[e];
    o[0] = make_rec(xg, yg, __dsub_rn(h, E.depth), 0xffffff00u);                                   // 255,255,0 (:400-405)
    if (!P.edge[e]) return;
    const double sx = (double)P.dir[2 * e], sy = (double)P.dir[2 * e + 1];
    for (int k = 1; k < E.n_depth; ++k) {                                                          // :410-444
        const double ratio = __ddiv_rn((double)k, (double)E.n_depth);
        const double offset = __dmul_rn(E.slope_offset, ratio);
        // offset_x = +offset, -offset or 0.0 (:417-430): the sign is exact
        const double xs = __dadd_rn(xl, sx * offset), ys = __dadd_rn(yl, sy * offset);
        double xgs = xg, ygs = yg;                                                                 // _L variant: straight above the edge
        if (!E.area_walls_vertical) to_global(E, xs, ys, xgs, ygs);
        const double z = __dadd_rn(__dsub_rn(h, E.depth), __dmul_rn((double)k, E.pd));             // :437
        o[k] = make_rec(xgs, ygs, z, 0xffc8c800u);                                                 // 200,200,0
    }
}

__global__ void __launch_bounds__(256)
k_emit_bottoms(ExcavDev E, unsigned n_entries, LatticePlan P, const double* __restrict__ height, Rec32* __restrict__ out) {
    const unsigned e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_entries || P.dir[2 * e] == 2) return;
    double xl, yl, xg, yg;
    lattice_local(E, e, xl, yl);
    to_global(E, xl, yl, xg, yg);
    out[P.bottom_off[e]] = make_rec(xg, yg, __dsub_rn(height[e], E.depth), 0xff008b00u);           // 0,139,0 (:532-537)
}

// one warp per (outer-edge entry, k): each slope point of the surface has its own height query (:569)
__global__ void __launch_bounds__(256)
k_emit_slopes(BucketGrid G, ExcavDev E, unsigned n_entries, LatticePlan P, Rec32* __restrict__ out) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const unsigned per = (unsigned)(E.n_slope + 1);
    const unsigned e = w / per, k = w % per;
    if (e >= n_entries || !P.edge[e]) return;
    double xl, yl;
    lattice_local(E, e, xl, yl);
    const double ratio = __ddiv_rn((double)(int)k, (double)E.n_slope);                             // :548
    const double offset = __dmul_rn(E.slope_offset, ratio);                                        // :549
    const double xs = __dadd_rn(xl, (double)P.dir[2 * e] * offset), ys = __dadd_rn(yl, (double)P.dir[2 * e + 1] * offset);
    double xg, yg;
    to_global(E, xs, ys, xg, yg);
    const double h = terrain_height(G, E, xg, yg, lane);
    if (lane != 0) return;
    const double z = __dsub_rn(h, __dmul_rn(E.depth, __dsub_rn(1.0, ratio)));                      // :570
    out[P.slope_off[e] + k] = make_rec(xg, yg, z, 0xff90ee90u);                                    // 144,238,144
}

__global__ void k_one_height(BucketGrid G, ExcavDev E, double x, double y, double* out) {
    const double h = terrain_height(G, E, x, y, threadIdx.x & 31);
    if (threadIdx.x == 0) *out = h;
}

// ---- host side ----------------------------------------------------------------------------------------
Boxes make_boxes(const plc_excavation_params& P) {                // getExcavationBoxes, :138-181
    Boxes B{};
    auto set = [&](int b, double cx, double cy, double len, double wid) {
        B.cx[b] = cx;
        B.cy[b] = cy;
        B.len[b] = len;
        B.wid[b] = wid;
        B.min_x[b] = cx - len / 2.0;
        B.max_x[b] = cx + len / 2.0;
        B.min_y[b] = cy - wid / 2.0;
        B.max_y[b] = cy + wid / 2.0;
    };
    if (P.l_shape_enabled) {
        B.n = 2;
        set(0, 0.0, -P.arm1_length / 2.0, P.arm1_width, P.arm1_length);
        set(1, P.arm2_length / 2.0, -P.arm1_length + P.arm2_width / 2.0, P.arm2_length, P.arm2_width);
    } else {
        B.n = 1;
        set(0, 0.0, 0.0, P.length, P.width);
        B.min_x[0] = -P.length / 2.0;
        B.max_x[0] = P.length / 2.0;
        B.min_y[0] = -P.width / 2.0;
        B.max_y[0] = P.width / 2.0;
    }
    return B;
}

bool params_ok(const plc_excavation_params* p) {
    if (!p) return false;
    const double v[] = {p->depth, p->slope_angle_deg, p->offset_x, p->offset_y, p->point_density, p->terrain_search_radius,
                        p->arm1_length, p->arm1_width, p->arm2_length, p->arm2_width, p->width, p->length};
    for (double d : v)
        if (!std::isfinite(d)) return false;
    return p->point_density > 0 && p->depth > 0 && p->terrain_search_radius > 0;
}

// the part of the device parameters that depends on the node parameters only
bool lattice_of(const plc_excavation_params& P, ExcavDev* E) {
    E->B = make_boxes(P);
    E->depth = P.depth;
    E->pd = P.point_density;
    E->radius = P.terrain_search_radius;
    E->slope_offset = P.depth / std::tan(P.slope_angle_deg * M_PI / 180.0);          // :353, :492
    double lo_x = std::numeric_limits<double>::max(), hi_x = std::numeric_limits<double>::lowest(), lo_y = lo_x, hi_y = hi_x;
    for (int b = 0; b < E->B.n; ++b) {
        lo_x = std::min(lo_x, E->B.min_x[b]);
        hi_x = std::max(hi_x, E->B.max_x[b]);
        lo_y = std::min(lo_y, E->B.min_y[b]);
        hi_y = std::max(hi_y, E->B.max_y[b]);
    }
    E->overall_min_x = lo_x;
    E->overall_min_y = lo_y;
    const double nx = (hi_x - lo_x) / P.point_density, ny = (hi_y - lo_y) / P.point_density;
    const double nd = P.depth / P.point_density, ns = E->slope_offset / P.point_density;
    if (!(nx >= 0 && ny >= 0 && (nx + 2) * (ny + 2) < 6.0e7 && nd < 1e6 && std::fabs(ns) < 1e6)) return false;
    // the per-entry counts are scanned in 32 bits and the slope kernel runs one warp per (entry, slope step)
    if (!((nx + 2) * (ny + 2) * (std::fabs(ns) + 2) < 2.0e9 && (nx + 2) * (ny + 2) * (nd + 1) < 2.0e9)) return false;
    E->n_x = (int)nx + 1;                                                             // :368-370, :507-508
    E->n_y = (int)ny + 1;
    E->n_depth = (int)nd;
    E->n_slope = (int)ns + 1;                                                         // :541
    return E->n_slope >= 1;       // a slope of 90 degrees or more has no synthetic wall upstream either (division by n_slope)
}

}  // namespace
}  // namespace plc

using namespace plc;

extern "C" void plc_excavation_params_default(plc_excavation_params* p) {
    if (!p) return;
    p->depth = 1.0;                    // excavated_surface_generator.cpp:28-47
    p->slope_angle_deg = 75.0;
    p->offset_x = 4.0;
    p->offset_y = 1.0;
    p->point_density = 0.05;
    p->terrain_search_radius = 0.5;
    p->arm1_length = 2.0;
    p->arm1_width = 1.2;
    p->arm2_length = 2.0;
    p->arm2_width = 1.2;
    p->width = 1.2;
    p->length = 1.8;
    p->l_shape_enabled = 1;
    p->area_walls_vertical = 0;
}

extern "C" plc_status plc_excavation_counts(const plc_excavation_params* prm, uint64_t* n_surface, uint64_t* n_area) {
    ExcavDev E{};
    if (!params_ok(prm) || !lattice_of(*prm, &E)) return PLC_E_INVALID_ARG;
    // the same classification k_plan_lattice makes, on the parameters alone (sizing, not data)
    auto inside = [&](double x, double y) {
        for (int b = 0; b < E.B.n; ++b)
            if (x >= E.B.min_x[b] && x <= E.B.max_x[b] && y >= E.B.min_y[b] && y <= E.B.max_y[b]) return true;
        return false;
    };
    uint64_t surface = 0, area = 0;
    for (int i = 0; i <= E.n_x; ++i)
        for (int j = 0; j <= E.n_y; ++j) {
            const double x = E.overall_min_x + i * E.pd, y = E.overall_min_y + j * E.pd;
            if (!inside(x, y)) continue;
            const bool edge = !inside(x + E.pd, y) || !inside(x - E.pd, y) || !inside(x, y + E.pd) || !inside(x, y - E.pd);
            surface += 1 + (edge ? (uint64_t)(E.n_slope + 1) : 0);
            area += 1 + (edge ? (uint64_t)std::max(E.n_depth - 1, 0) : 0);
        }
    if (n_surface) *n_surface = surface;
    if (n_area) *n_area = area;
    return PLC_OK;
}

extern "C" plc_status plc_excavate(plc_ctx* ctx, const plc_cloud_view* matched, uint32_t off_rgb, const double* base_t,
                                   const double* base_q, const plc_excavation_params* prm, void* out_terrain,
                                   uint64_t cap_terrain, uint64_t* n_terrain, void* out_area, uint64_t cap_area,
                                   uint64_t* n_area, plc_excavation_info* info) {
    PLC_RANGE();
    if (!ctx || !cloud_view_ok(matched) || !base_t || !base_q || !params_ok(prm) || !n_terrain || !n_area)
        return PLC_E_INVALID_ARG;
    if (matched->n_points > 0x7fffff00ull) return PLC_E_INVALID_ARG;
    if (off_rgb != 0xffffffffu && (uint64_t)off_rgb + 4 > matched->point_step) return PLC_E_INVALID_ARG;
    *n_terrain = *n_area = 0;
    ExcavDev E{};
    if (!lattice_of(*prm, &E)) return PLC_E_INVALID_ARG;
    const size_t n = matched->n_points;

    // excavation centre and yaw from the TF (:280-302): tf2::Transform * Vector3 and Matrix3x3::getRPY, in double
    double yaw = 0.0;
    {
        const double* q = base_q;
        const double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
        if (!(d > 0)) return PLC_E_INVALID_ARG;
        const double s = 2.0 / d;
        const double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
        const double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
        const double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs, yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
        const double r0[3] = {1.0 - (yy + zz), xy - wz, xz + wy}, r1[3] = {xy + wz, 1.0 - (xx + zz), yz - wx};
        const double m20 = xz - wy;
        E.center_x = (r0[0] * prm->offset_x + r0[1] * prm->offset_y + r0[2] * 0.0) + base_t[0];
        E.center_y = (r1[0] * prm->offset_x + r1[1] * prm->offset_y + r1[2] * 0.0) + base_t[1];
        if (std::fabs(m20) < 1.0) {
            const double pitch = -std::asin(m20);
            yaw = std::atan2(r1[0] / std::cos(pitch), r0[0] / std::cos(pitch));
        }
    }
    E.cos_yaw = std::cos(yaw);
    E.sin_yaw = std::sin(yaw);
    E.cos_neg = std::cos(-yaw);
    E.sin_neg = std::sin(-yaw);
    E.r2f = static_cast<float>(prm->terrain_search_radius * prm->terrain_search_radius);
    E.reach = widened(prm->terrain_search_radius);
    E.have_points = n ? 1 : 0;
    E.area_walls_vertical = prm->area_walls_vertical ? 1 : 0;

    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    ctx->timer.reset();
    BucketGrid G{};
    if (n) {
        Bounds b;
        PLC_TRY(load_cloud(ctx, matched, &b));
        PLC_TRY(build_buckets(ctx, ctx->su_pts.as<float4>(), n, b, std::max(prm->terrain_search_radius * 0.5, 1e-3), &G));
        PLC_TRY(finite_count(ctx, G, &E.n_finite));
    }
    const float4* pts = ctx->su_pts.as<float4>();

    // ---- plan of the pit lattice: counts, scans, totals
    const size_t n_entries = (size_t)(E.n_x + 1) * (size_t)(E.n_y + 1);
    const size_t plan_words = 3 * (n_entries + 1);
    PLC_CUDA_TRY(ctx, ctx->ex_plan.reserve(plan_words * sizeof(uint32_t) + n_entries * 3 + n_entries * sizeof(double) + 64));
    PLC_CUDA_TRY(ctx, ctx->su_tmp.reserve(scan_tmp_bytes(n_entries + 1) + 64));
    LatticePlan P;
    P.area_off = ctx->ex_plan.as<uint32_t>();
    P.bottom_off = P.area_off + (n_entries + 1);
    P.slope_off = P.bottom_off + (n_entries + 1);
    double* height = reinterpret_cast<double*>(P.slope_off + (n_entries + 1) + ((n_entries + 1) & 1));   // 8-byte aligned
    P.dir = reinterpret_cast<signed char*>(height + n_entries);
    P.edge = reinterpret_cast<unsigned char*>(P.dir + 2 * n_entries);
    ctx->timer.begin(s);
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(P.area_off, 0, plan_words * sizeof(uint32_t), s));
    k_plan_lattice<<<blocks_for(n_entries, 256), 256, 0, s>>>(E, (unsigned)n_entries, P);
    ctx->launches += 1 + scan_exclusive_u32(P.area_off, n_entries + 1, ctx->lb, s);
    ctx->launches += scan_exclusive_u32(P.bottom_off, n_entries + 1, ctx->lb, s);
    ctx->launches += scan_exclusive_u32(P.slope_off, n_entries + 1, ctx->lb, s);
    // ---- carve: height-independent rejection, then heights for the rest
    unsigned* queue_n = nullptr;
    uint32_t *keep = nullptr, *slot = nullptr, *queue = nullptr;
    if (n) {
        PLC_CUDA_TRY(ctx, ctx->su_flags.reserve((3 * (n + 1) + 4) * sizeof(uint32_t)));
        keep = ctx->su_flags.as<uint32_t>();
        slot = keep + (n + 1);
        queue = slot + (n + 1);
        queue_n = queue + (n + 1);
        PLC_CUDA_TRY(ctx, cudaMemsetAsync(queue_n, 0, sizeof(unsigned), s));
        k_carve_prefilter<<<blocks_for(n, 256), 256, 0, s>>>(pts, (unsigned)n, E, keep, queue, queue_n);
        ctx->launches += 1;
    }
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    uint32_t* host = static_cast<uint32_t*>(ctx->pinned_reserve(4 * sizeof(uint32_t)));
    if (!host) return PLC_E_OOM;
    host[3] = 0;
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host + 0, P.area_off + n_entries, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host + 1, P.bottom_off + n_entries, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host + 2, P.slope_off + n_entries, 4, cudaMemcpyDeviceToHost, s));
    if (n) PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host + 3, queue_n, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint64_t area_total = host[0], bottoms = host[1], slopes = host[2];
    const unsigned n_queue = host[3];

    uint64_t kept = 0;
    if (n) {
        ctx->timer.begin(s);
        if (n_queue) k_carve_heights<<<blocks_for((size_t)n_queue * 32, 256), 256, 0, s>>>(G, E, pts, queue, n_queue, keep);
        ctx->timer.end(s);
        ctx->launches += n_queue ? 1 : 0;
        PLC_TRY(scan_flags(ctx, keep, slot, n, &kept));
    }
    const uint64_t terrain_total = kept + bottoms + slopes;
    *n_terrain = terrain_total;
    *n_area = area_total;
    if (info) {
        memset(info, 0, sizeof(*info));
        info->center[0] = E.center_x;
        info->center[1] = E.center_y;
        info->yaw = yaw;
        info->n_kept = kept;
        info->n_surface = bottoms + slopes;
        info->n_area = area_total;
    }
    if (terrain_total > cap_terrain || area_total > cap_area || (terrain_total && !out_terrain) || (area_total && !out_area))
        return PLC_E_CAPACITY;

    PLC_CUDA_TRY(ctx, ctx->ex_terrain.reserve((terrain_total + 1) * sizeof(Rec32)));
    PLC_CUDA_TRY(ctx, ctx->ex_area.reserve((area_total + 1) * sizeof(Rec32) + sizeof(double)));
    Rec32* terrain = ctx->ex_terrain.as<Rec32>();
    Rec32* area = ctx->ex_area.as<Rec32>();
    double* centre_h = reinterpret_cast<double*>(area + area_total);
    ctx->timer.begin(s);
    if (n && kept)
        k_emit_kept<<<blocks_for(n, 256), 256, 0, s>>>(pts, (unsigned)n, keep, slot, ctx->su_raw,
                                                       matched->point_step, off_rgb, terrain);
    k_lattice_heights<<<blocks_for(n_entries * 32, 256), 256, 0, s>>>(G, E, (unsigned)n_entries, P, height);
    k_emit_area<<<blocks_for(n_entries, 256), 256, 0, s>>>(E, (unsigned)n_entries, P, height, area);
    k_emit_bottoms<<<blocks_for(n_entries, 256), 256, 0, s>>>(E, (unsigned)n_entries, P, height, terrain + kept);
    k_emit_slopes<<<blocks_for(n_entries * (size_t)(E.n_slope + 1) * 32, 256), 256, 0, s>>>(G, E, (unsigned)n_entries, P,
                                                                                          terrain + kept + bottoms);
    k_one_height<<<1, 32, 0, s>>>(G, E, E.center_x, E.center_y, centre_h);
    ctx->timer.end(s);
    ctx->launches += 5 + ((n && kept) ? 1 : 0);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    double* host_h = reinterpret_cast<double*>(static_cast<char*>(ctx->pinned) + 64);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host_h, centre_h, sizeof(double), cudaMemcpyDeviceToHost, s));
    if (terrain_total) PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_terrain, terrain, terrain_total * sizeof(Rec32), cudaMemcpyDefault, s));
    if (area_total) PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_area, area, area_total * sizeof(Rec32), cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    if (info) info->center[2] = *host_h;
    ctx->times.excavate_ms = ctx->timer.total_ms();
    return PLC_OK;
}
