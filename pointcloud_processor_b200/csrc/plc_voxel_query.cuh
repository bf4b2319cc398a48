// plc_voxel_query.cuh — device-side queries against a VoxelGeom (built by plc_voxel.cu).
//
// The query answered is the reference's only spatial primitive: does any
// indexed point p satisfy  d2_f32(q,p) < thr  with  d2 = ((dx*dx)+(dy*dy))+(dz*dz)  evaluated in
// float32 without FMA (FLANN L2_Simple<float>, strict compare as RadiusResultSet does).
//
// Two levels, both addressed from the fine-voxel coordinate of q (edge g, a power of two):
//   brick   two bits per fine voxel, 8x8x8 fine voxels per 128-byte brick, bricks stored densely:
//           clear / uncertain / certainly blocked (plc_internal.cuh)
//   cand    per mid voxel (4x4x4 fine voxels): the points within reach of its box, stored
//           contiguously in SoA groups of four (a point appears in up to 27 lists) — the exact
//           test scans just that list
//
// Every translation unit that includes this header is compiled with -fmad=false, so `a*b+c`
// below is two roundings, as on the reference's x86-64 build.
#pragma once
#include "plc_internal.cuh"

namespace plc {

// Read-only loads that do not allocate in L1: candidate lists and cell records are touched once per
// block, while the occupancy words they would evict are reused by neighbouring rays.
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long ldg_stream(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

#ifdef PLC_BOUNDS_CHECK
// Debug build only (tools/sanitize.py): out-of-range structure accesses are counted and skipped.
__device__ unsigned long long plc_oob_count = 0;
#define PLC_CHECK_OOB(cond, fallback)              \
    if (cond) {                                     \
        atomicAdd(&plc_oob_count, 1ull);            \
        return fallback;                            \
    }
#else
#define PLC_CHECK_OOB(cond, fallback)
#endif

// One 16-bit occupancy row (an L1::evict_last hint made no measurable difference).
__device__ __forceinline__ unsigned ldg_brick_row(const char* p) {
    unsigned short v;
    asm("ld.global.nc.u16 %0, [%1];" : "=h"(v) : "l"(p));
    return v;
}

__device__ __forceinline__ float dist2_f32(float qx, float qy, float qz, float px, float py, float pz) {
    float ex = __fsub_rn(qx, px), ey = __fsub_rn(qy, py), ez = __fsub_rn(qz, pz);
    float acc = __fmul_rn(ex, ex);
    acc = __fadd_rn(acc, __fmul_rn(ey, ey));
    acc = __fadd_rn(acc, __fmul_rn(ez, ez));
    return acc;
}

// Fine-voxel coordinate along one axis; monotone non-decreasing in v (float sub, float mul by a
// power of two, floor).  The builder's dilation slack covers its rounding error.
__device__ __forceinline__ float fine_coord_f(float v, float origin, float inv_g) {
    return floorf(__fmul_rn(__fsub_rn(v, origin), inv_g));
}

__device__ __forceinline__ int clamp_i(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Brick of coarse voxel (cx, cy, cz): bricks are stored densely, x fastest.
__device__ __forceinline__ uint32_t brick_of(const VoxelGeom& g, int cx, int cy, int cz) {
    return (uint32_t)((cz * g.dim_y + cy) * g.dim_x + cx);
}

// Index of the 16-bit occupancy row that holds fine voxels (8 cx .. 8 cx + 7, y, z).  PLC_LAYOUT picks how the rows
// are ordered in memory (the builder, the point queries and k_sweep's march all go through this one function):
//   0  bricks: 8x8x8 fine voxels per 128-byte line, bricks x-fastest           [cz][cy][cx][z&7][y&7]
//   1  rows in plain z, y, cx order (a 128-byte line is 512 voxels along x)    [z][y][cx]
//   2  flat bricks: 64x8x1 fine voxels per 128-byte line                        [z][cy][cx][y&7]
#ifndef PLC_LAYOUT
#define PLC_LAYOUT 0
#endif
__device__ __forceinline__ unsigned row_index(const VoxelGeom& g, int cx, int y, int z) {
#if PLC_LAYOUT == 0
    return (brick_of(g, cx, y >> kFineShift, z >> kFineShift) << 6) | (unsigned)((z & 7) << 3) | (unsigned)(y & 7);
#elif PLC_LAYOUT == 1
    return (unsigned)((z * (g.dim_y << kFineShift) + y) * g.dim_x + cx);
#else
    return ((unsigned)((z * g.dim_y + (y >> kFineShift)) * g.dim_x + cx) << 3) | (unsigned)(y & 7);
#endif
}

__device__ __forceinline__ unsigned ldg_row(const VoxelGeom& g, unsigned idx) {
    return ldg_brick_row(reinterpret_cast<const char*>(reinterpret_cast<const unsigned short*>(g.bricks) + idx));
}

// State of fine voxel (ix, iy, iz): 0 clear, 1 uncertain (run the exact test), 3 certainly blocked.  One load.
__device__ __forceinline__ unsigned voxel_state(const VoxelGeom& g, int ix, int iy, int iz) {
    return (ldg_row(g, row_index(g, ix >> kFineShift, iy, iz)) >> (2 * (ix & 7))) & 3u;
}

// The same for a position in 16.16 fixed-point fine-voxel units (k_sweep's march), split for software pipelining:
// the row of the voxel holding (fx, fy, fz), and the state from a row.  In the brick layout the row's byte offset is
// assembled from bit fields of the three coordinates.
#if defined(PLC_TEX_MARCH)
// experiment: the row through the texture unit, which does the 3-D address arithmetic (integer texel coordinates)
__device__ __forceinline__ unsigned brick_row_fixed(const VoxelGeom& g, int fx, int fy, int fz) {
    unsigned r0, r1, r2, r3;
    asm("tex.3d.v4.u32.s32 {%0,%1,%2,%3}, [%4, {%5,%6,%7,%8}];"
        : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
        : "l"(g.occ_tex), "r"(fx >> (16 + kFineShift)), "r"(fy >> 16), "r"(fz >> 16), "r"(0));
    return r0;
}
#else
__device__ __forceinline__ unsigned brick_row_fixed(const VoxelGeom& g, int fx, int fy, int fz) {
    PLC_CHECK_OOB((unsigned)(fx >> 16) >= (unsigned)(g.dim_x << kFineShift) || (unsigned)(fy >> 16) >= (unsigned)(g.dim_y << kFineShift) ||
                  (unsigned)(fz >> 16) >= (unsigned)(g.dim_z << kFineShift), 0u)
#if PLC_LAYOUT == 0
    const unsigned brick = brick_of(g, fx >> (16 + kFineShift), fy >> (16 + kFineShift), fz >> (16 + kFineShift));
#ifdef PLC_WIDE_BRICKS
    const size_t off = ((size_t)brick << 7) | (((unsigned)fz >> 12) & 0x70u) | (((unsigned)fy >> 15) & 0xeu);
#else
    const unsigned off = (brick << 7) | (((unsigned)fz >> 12) & 0x70u) | (((unsigned)fy >> 15) & 0xeu);
#endif
    return ldg_brick_row(reinterpret_cast<const char*>(g.bricks) + off);
#else
    return ldg_row(g, row_index(g, fx >> (16 + kFineShift), fy >> 16, fz >> 16));
#endif
}
#endif
__device__ __forceinline__ unsigned row_shift(int fx) { return ((unsigned)fx >> 15) & 0xeu; }

__device__ __forceinline__ uint32_t mid_voxel_of(const VoxelGeom& g, int ix, int iy, int iz) {
    return (uint32_t)(((size_t)(iz >> g.mid_shift) * g.mdim_y + (iy >> g.mid_shift)) * g.mdim_x + (ix >> g.mid_shift));
}

// Exact existence test over the candidate list of mid voxel m, four points per iteration.
__device__ __forceinline__ bool cand_any_m(const VoxelGeom& g, uint32_t m, float qx, float qy, float qz, float thr) {
    PLC_CHECK_OOB(m >= (uint32_t)g.mdim_x * (uint32_t)g.mdim_y * (uint32_t)g.mdim_z, false)
    uint32_t s = __ldg(&g.gstart[m]), e = __ldg(&g.gstart[m + 1]);
    const float4* __restrict__ grp = g.cgrp + (size_t)s * 3;
    for (; s < e; ++s, grp += 3) {
        float4 X = ldg_stream(grp), Y = ldg_stream(grp + 1), Z = ldg_stream(grp + 2);
        float d0 = dist2_f32(qx, qy, qz, X.x, Y.x, Z.x), d1 = dist2_f32(qx, qy, qz, X.y, Y.y, Z.y);
        float d2 = dist2_f32(qx, qy, qz, X.z, Y.z, Z.z), d3 = dist2_f32(qx, qy, qz, X.w, Y.w, Z.w);
        if (fminf(fminf(d0, d1), fminf(d2, d3)) < thr) return true;
    }
    return false;
}

__device__ __forceinline__ bool cand_any(const VoxelGeom& g, int ix, int iy, int iz, float qx, float qy,
                                         float qz, float thr) {
    return cand_any_m(g, mid_voxel_of(g, ix, iy, iz), qx, qy, qz, thr);
}

// Occupancy first; the exact test only in the uncertain shell.  Outside the grid nothing is
// within reach (the grid has a one-coarse-voxel margin around every point).  `thr` must be the
// threshold the index was built for (the "certain" state is only valid for that one).
__device__ __forceinline__ bool ball_any(const VoxelGeom& g, float qx, float qy, float qz, float thr,
                                         unsigned long long* exact_counter) {
    float fx = fine_coord_f(qx, g.ox, g.inv_g), fy = fine_coord_f(qy, g.oy, g.inv_g),
          fz = fine_coord_f(qz, g.oz, g.inv_g);
    // NaN fails every comparison => treated as outside
    if (!(fx >= 0.0f && fy >= 0.0f && fz >= 0.0f && fx < g.fdim_xf && fy < g.fdim_yf && fz < g.fdim_zf))
        return false;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    unsigned st = voxel_state(g, ix, iy, iz);
    if (st == 0) return false;
    if (st == 3) return true;
    if (exact_counter) ++*exact_counter;
    return cand_any(g, ix, iy, iz, qx, qy, qz, thr);
}

}  // namespace plc
