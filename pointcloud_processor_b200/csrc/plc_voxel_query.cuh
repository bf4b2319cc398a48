// plc_voxel_query.cuh — device-side queries against a VoxelGeom (built by plc_voxel.cu).
//
// The query answered is the reference's only spatial primitive: does any
// indexed point p satisfy  d2_f32(q,p) < thr  with  d2 = ((dx*dx)+(dy*dy))+(dz*dz)  evaluated in
// float32 without FMA (FLANN L2_Simple<float>, strict compare as RadiusResultSet does).
//
// Two levels, both addressed from the fine-voxel coordinate of q (edge g, a power of two):
//   mask   one bit per fine voxel: 0 = no point within reach of the voxel's box (certainly clear)
//   cand   per mid voxel (2x2x2 fine voxels): the points within reach of its box, stored
//          contiguously (a point appears in up to 27 lists) — the exact test scans just that list
//
// Every translation unit that includes this header is compiled with -fmad=false, so `a*b+c`
// below is two roundings, as on the reference's x86-64 build.
#pragma once
#include "plc_internal.cuh"

namespace plc {

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

// Exact existence test over the candidate list of the mid voxel (ix>>1, iy>>1, iz>>1).
__device__ __forceinline__ bool cand_any(const VoxelGeom& g, int ix, int iy, int iz, float qx, float qy,
                                         float qz, float thr) {
    size_t m = ((size_t)(iz >> 1) * g.mdim_y + (iy >> 1)) * g.mdim_x + (ix >> 1);
    uint32_t s = __ldg(&g.cstart[m]), e = __ldg(&g.cstart[m + 1]);
    const float4* __restrict__ pts = g.cpts;
    for (; s < e; ++s) {
        float4 p = __ldg(&pts[s]);
        if (dist2_f32(qx, qy, qz, p.x, p.y, p.z) < thr) return true;
    }
    return false;
}

// Mask first; the exact test only where the fine voxel is flagged.  Outside the grid nothing is
// within reach (the grid has a one-coarse-voxel margin around every point).
__device__ __forceinline__ bool ball_any(const VoxelGeom& g, float qx, float qy, float qz, float thr,
                                         unsigned long long* exact_counter) {
    float fx = fine_coord_f(qx, g.ox, g.inv_g), fy = fine_coord_f(qy, g.oy, g.inv_g),
          fz = fine_coord_f(qz, g.oz, g.inv_g);
    // NaN fails every comparison => treated as outside
    if (!(fx >= 0.0f && fy >= 0.0f && fz >= 0.0f && fx < g.fdim_xf && fy < g.fdim_yf && fz < g.fdim_zf))
        return false;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    size_t v = ((size_t)(iz >> 2) * g.dim_y + (iy >> 2)) * g.dim_x + (ix >> 2);
    uint32_t w = __ldg(&g.mask32[2 * v + ((iz >> 1) & 1)]);
    int bit = (ix & 3) | ((iy & 3) << 2) | ((iz & 1) << 4);
    if (!((w >> bit) & 1u)) return false;
    if (exact_counter) ++*exact_counter;
    return cand_any(g, ix, iy, iz, qx, qy, qz, thr);
}

}  // namespace plc
