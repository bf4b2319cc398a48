// plc_setup.cu — the steps of the virtual_lidar node that produce the sweep's inputs, on the device
// (SURVEY §8 row N1): evaluation-cell lattice + cell normals, per-point PCA normals, candidate poses.
//
// Replaces /root/reference/src/virtual_lidar.cpp:
//   computeTerrainNormals      209-234  -> k_pca_normals      (one warp per area point)
//   generateExcavationGrid3D   236-287  -> k_load_points (bounds), k_lattice_near, k_emit_cells
//   isPointNearExcavation      289-299  -> k_lattice_near     (one thread per lattice point)
//   computeCellSurfaceNormal   301-340  -> k_cell_normals     (one warp per kept cell)
//   generateCandidatePositions 550-598  -> k_candidate_slots, k_emit_poses
//   getGroundHeight            600-625  -> k_candidate_slots  (one warp per lattice slot)
// The reference answers every neighbourhood question with a pcl::KdTreeFLANN radius search; here the cloud is
// bucketed once (stable radix sort by bucket, so the order inside a bucket is the cloud's own) and a query
// walks the rows of buckets its ball can touch.  The acceptance rule is PCL's: float32
// query, d2 = ((dx*dx)+(dy*dy))+(dz*dz) in float32 without FMA (this file is built with -fmad=false), accept
// iff d2 < (float)(radius*radius).
//
// Parity against the CPU restatement: cell positions, the kept set and its order, candidate x/y/z, the kept
// candidates and their order are bit-identical (all decided by exact float32 tests and double arithmetic
// written in upstream's operation order).  Cell normals are sums of hundreds of float normals in double; the
// reference adds them in KD-tree result order, this kernel in a fixed lane/tree order, so a component can
// differ in the last float32 bit.  Pitch/yaw come from the device atan2 (2 ulp).  PCA normals are not pinned
// upstream (PCL's eigen solver) and are compared by angle.
//
// All streams of this file are HBM/L2-bound gathers: algorithmic bytes = 16 B per point tested.
#include <cmath>
#include <limits>

#include "plc_buckets.cuh"

namespace plc {
namespace {

// ---- evaluation-cell lattice (:236-299) -------------------------------------------------------
struct Lattice {
    double min_x, min_y, min_z, res, z_step;
    int width, height, layers;
};
__device__ __forceinline__ void lattice_point(const Lattice& L, unsigned id, double& x, double& y, double& z) {
    // loop order of :264-271: y index outer, x index, layer inner
    const unsigned k = id % (unsigned)L.layers, col = id / (unsigned)L.layers;
    const unsigned j = col % (unsigned)L.width, i = col / (unsigned)L.width;
    x = __dadd_rn(L.min_x, __dmul_rn((double)(int)j, L.res));                  // :268
    y = __dadd_rn(L.min_y, __dmul_rn((double)(int)i, L.res));                  // :269
    z = __dadd_rn(__dadd_rn(L.min_z, __dmul_rn((double)(int)k, L.z_step)), __ddiv_rn(L.z_step, 2.0));   // :272
}

// flag[id] = 1 iff an area point lies within res*1.5 of the lattice point (isPointNearExcavation)
__global__ void __launch_bounds__(256)
k_lattice_near(BucketGrid G, Lattice L, unsigned n_lattice, float r2f, double reach, uint32_t* __restrict__ flag) {
    unsigned id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_lattice) return;
    double x, y, z;
    lattice_point(L, id, x, y, z);
    const float qx = (float)x, qy = (float)y, qz = (float)z;
    const RowRange R = rows_of_box(G, qx, qy, qz, reach, reach, reach);
    uint32_t hit = 0;
    if (!R.empty())
        for (int bz = R.z0; bz <= R.z1 && !hit; ++bz)
            for (int by = R.y0; by <= R.y1 && !hit; ++by) {
                const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                const uint32_t e = __ldg(&G.start[row + R.x1 + 1]);
                for (uint32_t p = __ldg(&G.start[row + R.x0]); p < e; ++p) {
                    const float4 v = __ldg(&G.pts[p]);
                    if (dist2(qx, qy, qz, v.x, v.y, v.z) < r2f) {
                        hit = 1;
                        break;
                    }
                }
            }
    flag[id] = hit;
}

// kept lattice points -> cells in lattice order, default normal (GridCell ctor :34-43); `slot` is the
// exclusive scan of `flag`
__global__ void __launch_bounds__(256)
k_emit_cells(Lattice L, unsigned n_lattice, const uint32_t* __restrict__ flag, const uint32_t* __restrict__ slot,
             plc_cell* __restrict__ out) {
    unsigned id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_lattice || !flag[id]) return;
    plc_cell c;
    lattice_point(L, id, c.x, c.y, c.z);
    c.nx = 0.f;
    c.ny = 0.f;
    c.nz = 1.f;
    c.pad = 0u;
    out[slot[id]] = c;
}

// computeCellSurfaceNormal (:301-340): sum of the finite normals of the area points within normal_radius of the
// cell centre, normalised; one warp per cell, lanes stride over each row of buckets
// per-point normals brought into bucket order once (w = 1 where all three components are finite, :320-324), so that
// k_cell_normals reads them next to the points instead of gathering by index for every cell
__global__ void __launch_bounds__(256)
k_normals_in_bucket_order(BucketGrid G, const float* __restrict__ normals, unsigned n_finite, float4* __restrict__ out) {
    unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_finite) return;
    const float* nv = normals + 3 * (size_t)__float_as_uint(G.pts[p].w);
    const float a = nv[0], b = nv[1], c = nv[2];
    out[p] = make_float4(a, b, c, (isfinite(a) && isfinite(b) && isfinite(c)) ? 1.f : 0.f);
}

__global__ void __launch_bounds__(256)
k_cell_normals(BucketGrid G, const float4* __restrict__ normals, unsigned n_cells, float r2f, double reach,
               plc_cell* __restrict__ cells) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_cells) return;
    const float qx = (float)cells[w].x, qy = (float)cells[w].y, qz = (float)cells[w].z;
    const RowRange R = rows_of_box(G, qx, qy, qz, reach, reach, reach);
    double sx = 0.0, sy = 0.0, sz = 0.0;
    unsigned used = 0;
    if (!R.empty())
        for (int bz = R.z0; bz <= R.z1; ++bz)
            for (int by = R.y0; by <= R.y1; ++by) {
                const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                const uint32_t e = __ldg(&G.start[row + R.x1 + 1]);
                for (uint32_t p = __ldg(&G.start[row + R.x0]) + lane; p < e; p += 32) {
                    const float4 v = __ldg(&G.pts[p]);
                    if (dist2(qx, qy, qz, v.x, v.y, v.z) < r2f) {
                        const float4 nv = __ldg(&normals[p]);
                        if (nv.w != 0.f) {                                               // :320-324
                            sx = __dadd_rn(sx, (double)nv.x);
                            sy = __dadd_rn(sy, (double)nv.y);
                            sz = __dadd_rn(sz, (double)nv.z);
                            ++used;
                        }
                    }
                }
            }
    sx = warp_sum(sx);
    sy = warp_sum(sy);
    sz = warp_sum(sz);
    used = __reduce_add_sync(0xffffffffu, used);
    if (lane == 0 && used > 0) {
        const double len = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(sx, sx), __dmul_rn(sy, sy)), __dmul_rn(sz, sz)));
        if (len > 1e-6) {                                                                // :329-333
            cells[w].nx = (float)__ddiv_rn(sx, len);
            cells[w].ny = (float)__ddiv_rn(sy, len);
            cells[w].nz = (float)__ddiv_rn(sz, len);
        }
    }
}

// ---- per-point PCA normals (:209-234, PCL NormalEstimation restated) ----------------------------
// Unit eigenvector of the smallest eigenvalue of the symmetric matrix {xx xy xz yy yz zz}: trigonometric
// roots of the characteristic polynomial, then the longest cross product of two rows of (A - lambda I).
__device__ void smallest_eigenvector(const double m[6], double out[3]) {
    double scale = 0.0;
    for (int i = 0; i < 6; ++i) scale = fmax(scale, fabs(m[i]));
    out[0] = 0.0;
    out[1] = 0.0;
    out[2] = 1.0;
    if (!(scale > 0.0)) return;
    const double xx = m[0] / scale, xy = m[1] / scale, xz = m[2] / scale, yy = m[3] / scale, yz = m[4] / scale,
                 zz = m[5] / scale;
    const double q = (xx + yy + zz) / 3.0;
    const double dx = xx - q, dy = yy - q, dz = zz - q;
    const double p = sqrt((dx * dx + dy * dy + dz * dz + 2.0 * (xy * xy + xz * xz + yz * yz)) / 6.0);
    double lambda = q;
    if (p >= 1e-300) {
        const double bxx = dx / p, byy = dy / p, bzz = dz / p, bxy = xy / p, bxz = xz / p, byz = yz / p;
        double hd = 0.5 * (bxx * (byy * bzz - byz * byz) - bxy * (bxy * bzz - byz * bxz) + bxz * (bxy * byz - byy * bxz));
        hd = fmin(1.0, fmax(-1.0, hd));
        lambda = q + 2.0 * p * cos(acos(hd) / 3.0 + 2.0 * M_PI / 3.0);        // the smallest of the three roots
    }
    const double r[3][3] = {{xx - lambda, xy, xz}, {xy, yy - lambda, yz}, {xz, yz, zz - lambda}};
    const int pair[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    double best = -1.0, bv[3] = {0.0, 0.0, 1.0};
    for (int t = 0; t < 3; ++t) {
        const double* u = r[pair[t][0]];
        const double* v = r[pair[t][1]];
        const double c[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
        const double n2 = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
        if (n2 > best) {
            best = n2;
            bv[0] = c[0];
            bv[1] = c[1];
            bv[2] = c[2];
        }
    }
    if (!(best > 1e-300)) return;
    const double inv = 1.0 / sqrt(best);
    for (int k = 0; k < 3; ++k) out[k] = bv[k] * inv;
}

__global__ void __launch_bounds__(256) k_fill_nan(float* __restrict__ p, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = __int_as_float(0x7fc00000);
}

// one warp per bucketed (finite) point: neighbours within `radius` (itself included), mean, covariance about
// the mean, eigenvector, flipped towards the viewpoint (0,0,0) as PCL does and then to n_z >= 0 (:223-229)
__global__ void __launch_bounds__(256)
k_pca_normals(BucketGrid G, unsigned n_finite, float r2f, double reach, float* __restrict__ normals) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_finite) return;
    const float4 q = __ldg(&G.pts[w]);
    const RowRange R = rows_of_box(G, q.x, q.y, q.z, reach, reach, reach);
    double mean[3] = {0.0, 0.0, 0.0}, cov[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    unsigned count = 0;
    for (int pass = 0; pass < 2; ++pass) {
        for (int bz = R.z0; bz <= R.z1; ++bz)
            for (int by = R.y0; by <= R.y1; ++by) {
                const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                const uint32_t e = __ldg(&G.start[row + R.x1 + 1]);
                for (uint32_t p = __ldg(&G.start[row + R.x0]) + lane; p < e; p += 32) {
                    const float4 v = __ldg(&G.pts[p]);
                    if (!(dist2(q.x, q.y, q.z, v.x, v.y, v.z) < r2f)) continue;
                    if (pass == 0) {
                        mean[0] += (double)v.x;
                        mean[1] += (double)v.y;
                        mean[2] += (double)v.z;
                        ++count;
                    } else {
                        const double ex = (double)v.x - mean[0], ey = (double)v.y - mean[1], ez = (double)v.z - mean[2];
                        cov[0] += ex * ex;
                        cov[1] += ex * ey;
                        cov[2] += ex * ez;
                        cov[3] += ey * ey;
                        cov[4] += ey * ez;
                        cov[5] += ez * ez;
                    }
                }
            }
        if (pass == 0) {
            count = __reduce_add_sync(0xffffffffu, count);
            if (count < 3) return;                       // PCL: fewer than three neighbours => NaN normal
            for (int a = 0; a < 3; ++a) mean[a] = warp_sum(mean[a]) / (double)count;
        }
    }
    for (int a = 0; a < 6; ++a) cov[a] = warp_sum(cov[a]);
    if (lane != 0) return;
    double n[3];
    smallest_eigenvector(cov, n);
    if (-((double)q.x * n[0] + (double)q.y * n[1] + (double)q.z * n[2]) < 0.0) {
        n[0] = -n[0];
        n[1] = -n[1];
        n[2] = -n[2];
    }
    if (n[2] < 0.0) {
        n[0] = -n[0];
        n[1] = -n[1];
        n[2] = -n[2];
    }
    float* o = normals + 3 * (size_t)__float_as_uint(q.w);
    o[0] = (float)n[0];
    o[1] = (float)n[1];
    o[2] = (float)n[2];
}

// ---- candidate poses (:550-625) -----------------------------------------------------------------
struct CandidateRule {
    double lo_x, lo_y, step_x, step_y;            // lattice over the bounds widened by search_radius (:555-565)
    double min_x, max_x, min_y, max_y;            // grid bounds: lattice points inside are skipped (:575)
    double mid_x, mid_y, mid_z;                   // what the sensor looks at (:558-560)
    double fixed_x, fixed_y, sensor_height, elev_lo, elev_hi, neg_half_pi;
    int side;
};

// one warp per lattice slot (x index outer, y index inner, :567-568): the gates that need no terrain, then
// getGroundHeight — max z over terrain points inside the float32 ball of radius 2 around (x, y, 0) that are also
// closer than 1 m horizontally in double — then the elevation gate and the pose
__global__ void __launch_bounds__(256)
k_candidate_slots(BucketGrid G, int have_terrain, CandidateRule C, unsigned n_slots, uint32_t* __restrict__ keep,
                  plc_pose* __restrict__ slot_pose) {
    const unsigned t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= n_slots) return;
    if (lane == 0) keep[t] = 0;
    const int i = (int)(t / (unsigned)C.side), j = (int)(t % (unsigned)C.side);
    const double x = __dadd_rn(C.lo_x, __dmul_rn((double)i, C.step_x));              // :569
    const double y = __dadd_rn(C.lo_y, __dmul_rn((double)j, C.step_y));              // :570
    const double fx = __dsub_rn(x, C.fixed_x), fy = __dsub_rn(y, C.fixed_y);
    if (__dsqrt_rn(__dadd_rn(__dmul_rn(fx, fx), __dmul_rn(fy, fy))) < 0.5) return;   // :571-573
    if (x >= C.min_x && x <= C.max_x && y >= C.min_y && y <= C.max_y) return;        // :575
    double top = -INFINITY;
    if (have_terrain) {
        const float qx = (float)x, qy = (float)y;
        // both conditions bound the horizontal offset by 1 m, the ball bounds |z| by 2 m
        const RowRange R = rows_of_box(G, x, y, 0.0, 1.001, 1.001, widened(2.0));
        if (!R.empty())
            for (int bz = R.z0; bz <= R.z1; ++bz)
                for (int by = R.y0; by <= R.y1; ++by) {
                    const size_t row = ((size_t)bz * G.ny + by) * G.nx;
                    const uint32_t e = __ldg(&G.start[row + R.x1 + 1]);
                    for (uint32_t p = __ldg(&G.start[row + R.x0]) + lane; p < e; p += 32) {
                        const float4 v = __ldg(&G.pts[p]);
                        if (!(dist2(qx, qy, 0.f, v.x, v.y, v.z) < 4.0f)) continue;           // :611, radius 2.0
                        const double ex = __dsub_rn((double)v.x, x), ey = __dsub_rn((double)v.y, y);
                        if (__dsqrt_rn(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey))) < 1.0)   // :616-618
                            top = fmax(top, (double)v.z);
                    }
                }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) top = fmax(top, __shfl_xor_sync(0xffffffffu, top, o));
    }
    if (lane != 0) return;
    const double ground = top != -INFINITY ? top : 0.0;                              // :620-624
    const double z = __dadd_rn(ground, C.sensor_height);                             // :577
    const double dx = __dsub_rn(C.mid_x, x), dy = __dsub_rn(C.mid_y, y), dz = __dsub_rn(C.mid_z, z);
    const double hd = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    if (hd < 0.1) return;                                                            // :585
    const double elevation = atan2(-dz, hd);                                         // :587
    if (!(elevation >= C.elev_lo && elevation <= C.elev_hi)) return;                 // :589
    plc_pose ps;
    ps.x = x;
    ps.y = y;
    ps.z = z;
    ps.pitch = __dadd_rn(C.neg_half_pi, elevation);                                  // :590
    ps.yaw = atan2(dy, dx);                                                          // :591
    slot_pose[t] = ps;
    keep[t] = 1;
}

__global__ void __launch_bounds__(256)
k_emit_poses(unsigned n_slots, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ slot,
             const plc_pose* __restrict__ slot_pose, plc_pose* __restrict__ out) {
    unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n_slots && keep[t]) out[slot[t]] = slot_pose[t];
}

}  // namespace
}  // namespace plc

using namespace plc;

extern "C" plc_status plc_vl_compute_normals(plc_ctx* ctx, const plc_cloud_view* area, double radius, float* normals_out) {
    PLC_RANGE();
    if (!ctx || !cloud_view_ok(area) || !normals_out || !(radius > 0) || area->n_points > 0x7fffff00ull)
        return PLC_E_INVALID_ARG;
    const size_t n = area->n_points;
    if (n == 0) return PLC_OK;
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ctx->timer.reset();
    Bounds b;
    PLC_TRY(load_cloud(ctx, area, &b));
    BucketGrid G;
    PLC_TRY(build_buckets(ctx, ctx->su_pts.as<float4>(), n, b, std::max(radius * 0.25, 1e-3), &G));
    uint32_t n_finite = 0;
    PLC_TRY(finite_count(ctx, G, &n_finite));
    PLC_CUDA_TRY(ctx, ctx->su_normals.reserve(3 * n * sizeof(float)));
    cudaStream_t s = ctx->stream;
    ctx->timer.begin(s);
    k_fill_nan<<<blocks_for(3 * n, 256), 256, 0, s>>>(ctx->su_normals.as<float>(), 3 * n);
    if (n_finite)
        k_pca_normals<<<blocks_for((size_t)n_finite * 32, 256), 256, 0, s>>>(G, n_finite, radius_sq_f32(radius), widened(radius),
                                                                             ctx->su_normals.as<float>());
    ctx->timer.end(s);
    ctx->launches += n_finite ? 2 : 1;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(normals_out, ctx->su_normals.ptr, 3 * n * sizeof(float), cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->times.normals_ms = ctx->timer.total_ms();
    return PLC_OK;
}

extern "C" plc_status plc_vl_generate_cells(plc_ctx* ctx, const plc_cloud_view* area, const float* normals, double res,
                                            int32_t layers, double normal_radius, plc_cell* out, uint64_t capacity,
                                            uint64_t* n_out, plc_grid_bounds* bounds) {
    PLC_RANGE();
    if (!ctx || !cloud_view_ok(area) || !n_out || !(res > 0) || (capacity && !out) || area->n_points > 0x7fffff00ull)
        return PLC_E_INVALID_ARG;
    if (normals && !(normal_radius > 0)) return PLC_E_INVALID_ARG;
    *n_out = 0;
    const size_t n = area->n_points;
    if (n == 0) return PLC_OK;                    // the reference returns early on an empty area (:168, :237)
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ctx->timer.reset();
    Bounds b;
    PLC_TRY(load_cloud(ctx, area, &b));
    // :242-262 — bounds over the stored points, margin, lattice size
    plc_grid_bounds bb;
    const double dmax = std::numeric_limits<double>::max(), dlow = std::numeric_limits<double>::lowest();
    auto lo_of = [&](int a) { return b.any_lo[a] <= b.any_hi[a] ? (double)b.any_lo[a] : dmax; };   // all-NaN axis: untouched
    auto hi_of = [&](int a) { return b.any_lo[a] <= b.any_hi[a] ? (double)b.any_hi[a] : dlow; };
    bb.min_x = lo_of(0) - res;
    bb.max_x = hi_of(0) + res;
    bb.min_y = lo_of(1) - res;
    bb.max_y = hi_of(1) + res;
    bb.min_z = lo_of(2) - res;
    bb.max_z = hi_of(2) + res;
    const double wd = std::ceil((bb.max_x - bb.min_x) / res), hd = std::ceil((bb.max_y - bb.min_y) / res);
    // a lattice that is not finite or does not fit 31 bits cannot be walked (upstream would not return either)
    if (!(wd >= 0 && hd >= 0 && wd < 2.0e9 && hd < 2.0e9)) return PLC_E_INVALID_ARG;
    bb.width = (int)wd + 1;
    bb.height = (int)hd + 1;
    if (bounds) *bounds = bb;
    if (layers <= 0) return PLC_OK;               // the layer loop never runs (:271)
    const double n_lat_d = (double)bb.width * bb.height * layers;
    if (!(n_lat_d < 2.0e9)) return PLC_E_INVALID_ARG;
    const size_t n_lat = (size_t)bb.width * bb.height * layers;

    Lattice L{bb.min_x, bb.min_y, bb.min_z, res, (bb.max_z - bb.min_z) / std::max(1, layers), bb.width, bb.height, layers};
    const double near_radius = res * 1.5;         // :274
    BucketGrid G;
    PLC_TRY(build_buckets(ctx, ctx->su_pts.as<float4>(), n, b, std::max(near_radius, 1e-3), &G));
    PLC_CUDA_TRY(ctx, ctx->su_flags.reserve(2 * (n_lat + 1) * sizeof(uint32_t)));
    uint32_t* flag = ctx->su_flags.as<uint32_t>();
    uint32_t* slot = flag + (n_lat + 1);
    cudaStream_t s = ctx->stream;
    ctx->timer.begin(s);
    k_lattice_near<<<blocks_for(n_lat, 256), 256, 0, s>>>(G, L, (unsigned)n_lat, radius_sq_f32(near_radius), widened(near_radius), flag);
    ctx->timer.end(s);
    ctx->launches += 1;
    uint64_t total = 0;
    PLC_TRY(scan_flags(ctx, flag, slot, n_lat, &total));
    *n_out = total;
    if (total > capacity) return PLC_E_CAPACITY;
    if (total == 0) return PLC_OK;
    PLC_CUDA_TRY(ctx, ctx->su_out.reserve(total * sizeof(plc_cell)));
    plc_cell* cells = ctx->su_out.as<plc_cell>();
    ctx->timer.begin(s);
    k_emit_cells<<<blocks_for(n_lat, 256), 256, 0, s>>>(L, (unsigned)n_lat, flag, slot, cells);
    ctx->launches += 1;
    ctx->timer.end(s);
    if (normals) {
        // the near test wanted buckets of 1.5 res; a ball of normal_radius over those would walk thousands of empty
        // rows per cell, so the cloud is bucketed again at a quarter of the normal radius (the first grid is done with)
        PLC_TRY(build_buckets(ctx, ctx->su_pts.as<float4>(), n, b, std::max(normal_radius * 0.25, std::max(near_radius, 1e-3)), &G));
        uint32_t n_finite = 0;
        PLC_TRY(finite_count(ctx, G, &n_finite));
        PLC_CUDA_TRY(ctx, ctx->su_normals.reserve(3 * n * sizeof(float) + 16 + (size_t)(n + 1) * sizeof(float4)));
        float* raw_normals = ctx->su_normals.as<float>();
        float4* ordered = reinterpret_cast<float4*>(raw_normals + ((3 * n + 3) & ~(size_t)3));
        PLC_CUDA_TRY(ctx, cudaMemcpyAsync(raw_normals, normals, 3 * n * sizeof(float), cudaMemcpyDefault, s));
        ctx->timer.begin(s);
        if (n_finite) k_normals_in_bucket_order<<<blocks_for(n_finite, 256), 256, 0, s>>>(G, raw_normals, n_finite, ordered);
        k_cell_normals<<<blocks_for(total * 32, 256), 256, 0, s>>>(G, ordered, (unsigned)total, radius_sq_f32(normal_radius),
                                                                  widened(normal_radius), cells);
        ctx->timer.end(s);
        ctx->launches += n_finite ? 2 : 1;
    }
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out, cells, total * sizeof(plc_cell), cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->times.cells_ms = ctx->timer.total_ms();
    return PLC_OK;
}

extern "C" plc_status plc_vl_generate_candidates(plc_ctx* ctx, const plc_cloud_view* terrain, const plc_grid_bounds* gb,
                                                 const plc_pose* fixed, int32_t num_candidates, double search_radius,
                                                 double sensor_height, plc_pose* out, uint64_t capacity, uint64_t* n_out) {
    PLC_RANGE();
    if (!ctx || !gb || !fixed || !n_out || (capacity && !out)) return PLC_E_INVALID_ARG;
    if (terrain && (!cloud_view_ok(terrain) || terrain->n_points > 0x7fffff00ull)) return PLC_E_INVALID_ARG;
    *n_out = 0;
    if (num_candidates < 2) return PLC_E_INVALID_ARG;        // a lattice of side 1 divides by zero upstream (:563)
    const int side = (int)std::ceil(std::sqrt((double)num_candidates));
    const size_t n_slots = (size_t)side * side;
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ctx->timer.reset();
    cudaStream_t s = ctx->stream;

    // the terrain: the caller's cloud, or the points plc_vl_set_terrain keeps on the device
    const float4* pts = nullptr;
    size_t n = 0;
    Bounds b{};
    if (terrain) {
        n = terrain->n_points;
        if (n) {
            PLC_TRY(load_cloud(ctx, terrain, &b));
            pts = ctx->su_pts.as<float4>();
        }
    } else if (ctx->terrain.n_raw) {
        n = ctx->terrain.n_raw;
        pts = ctx->terrain.raw.as<float4>();
        PLC_CUDA_TRY(ctx, ctx->su_state.reserve(sizeof(BoundsState)));
        ctx->timer.begin(s);
        k_bounds_init<<<1, 1, 0, s>>>(ctx->su_state.as<BoundsState>());
        k_bounds_of<<<blocks_for(n, 256), 256, 0, s>>>(pts, (unsigned)n, ctx->su_state.as<BoundsState>());
        ctx->timer.end(s);
        ctx->launches += 2;
        PLC_TRY(read_bounds(ctx, ctx->su_state.as<BoundsState>(), &b));
    }
    BucketGrid G{};
    const int have_terrain = (n && b.any_finite) ? 1 : 0;
    if (have_terrain) PLC_TRY(build_buckets(ctx, pts, n, b, 0.5, &G));

    CandidateRule C{};
    C.lo_x = gb->min_x - search_radius;                      // :555-556
    C.lo_y = gb->min_y - search_radius;
    const double hi_x = gb->max_x + search_radius, hi_y = gb->max_y + search_radius;
    C.step_x = (hi_x - C.lo_x) / (side - 1);                 // :564-565
    C.step_y = (hi_y - C.lo_y) / (side - 1);
    C.min_x = gb->min_x;
    C.max_x = gb->max_x;
    C.min_y = gb->min_y;
    C.max_y = gb->max_y;
    C.mid_x = (gb->min_x + gb->max_x) / 2.0;                 // :558-560
    C.mid_y = (gb->min_y + gb->max_y) / 2.0;
    C.mid_z = (gb->min_z + gb->max_z) / 2.0;
    C.fixed_x = fixed->x;
    C.fixed_y = fixed->y;
    C.sensor_height = sensor_height;
    C.elev_lo = -85.0 * M_PI / 180.0;                        // :113-114
    C.elev_hi = 85.0 * M_PI / 180.0;
    C.neg_half_pi = -M_PI / 2;
    C.side = side;

    PLC_CUDA_TRY(ctx, ctx->su_flags.reserve(2 * (n_slots + 1) * sizeof(uint32_t)));
    PLC_CUDA_TRY(ctx, ctx->su_out.reserve(2 * n_slots * sizeof(plc_pose)));
    uint32_t* keep = ctx->su_flags.as<uint32_t>();
    uint32_t* slot = keep + (n_slots + 1);
    plc_pose* slot_pose = ctx->su_out.as<plc_pose>();
    plc_pose* packed = slot_pose + n_slots;
    ctx->timer.begin(s);
    k_candidate_slots<<<blocks_for(n_slots * 32, 256), 256, 0, s>>>(G, have_terrain, C, (unsigned)n_slots, keep, slot_pose);
    ctx->timer.end(s);
    ctx->launches += 1;
    uint64_t total = 0;
    PLC_TRY(scan_flags(ctx, keep, slot, n_slots, &total));
    *n_out = total;
    if (total > capacity) return PLC_E_CAPACITY;
    if (total == 0) return PLC_OK;
    ctx->timer.begin(s);
    k_emit_poses<<<blocks_for(n_slots, 256), 256, 0, s>>>((unsigned)n_slots, keep, slot, slot_pose, packed);
    ctx->timer.end(s);
    ctx->launches += 1;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out, packed, total * sizeof(plc_pose), cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->times.candidates_ms = ctx->timer.total_ms();
    return PLC_OK;
}
