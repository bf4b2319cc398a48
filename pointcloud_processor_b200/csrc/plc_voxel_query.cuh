// plc_voxel_query.cuh — device-side queries against a VoxelGeom (built by plc_voxel.cu).
//
// The query answered is the reference's only spatial primitive: does any
// indexed point p satisfy  d2_f32(q,p) < thr  with  d2 = ((dx*dx)+(dy*dy))+(dz*dz)  evaluated in
// float32 without FMA (FLANN L2_Simple<float>, strict compare as RadiusResultSet does).
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
// power of two, floor), which is what makes range lookups complete.
__device__ __forceinline__ float fine_coord_f(float v, float origin, float inv_g) {
    return floorf(__fmul_rn(__fsub_rn(v, origin), inv_g));
}

// 1 if the fine voxel holding q is flagged (a point may be within the query radius), 0 if q is
// certainly clear.  Outside the grid nothing is within reach (the grid has a one-voxel margin).
__device__ __forceinline__ bool mask_flagged(const VoxelGeom& g, float qx, float qy, float qz) {
    float fx = fine_coord_f(qx, g.ox, g.inv_g), fy = fine_coord_f(qy, g.oy, g.inv_g),
          fz = fine_coord_f(qz, g.oz, g.inv_g);
    // NaN fails every comparison => treated as outside
    if (!(fx >= 0.0f && fy >= 0.0f && fz >= 0.0f && fx < (float)g.fdim_x && fy < (float)g.fdim_y &&
          fz < (float)g.fdim_z))
        return false;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    size_t v = ((size_t)(iz >> 2) * g.dim_y + (iy >> 2)) * g.dim_x + (ix >> 2);
    uint32_t w = __ldg(&g.mask32[2 * v + ((iz >> 1) & 1)]);
    int bit = (ix & 3) | ((iy & 3) << 2) | ((iz & 1) << 4);
    return (w >> bit) & 1u;
}

__device__ __forceinline__ int clamp_i(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Exact existence test over the (at most 2x2x2) coarse voxels the query ball can touch.
__device__ __forceinline__ bool exact_any(const VoxelGeom& g, float qx, float qy, float qz, float thr) {
    float r = g.reach;
    // coarse index = fine index >> 2 (h = 4g, same origin, power-of-two scale => exact)
    float lx = fine_coord_f(__fsub_rn(qx, r), g.ox, g.inv_g), hx = fine_coord_f(__fadd_rn(qx, r), g.ox, g.inv_g);
    float ly = fine_coord_f(__fsub_rn(qy, r), g.oy, g.inv_g), hy = fine_coord_f(__fadd_rn(qy, r), g.oy, g.inv_g);
    float lz = fine_coord_f(__fsub_rn(qz, r), g.oz, g.inv_g), hz = fine_coord_f(__fadd_rn(qz, r), g.oz, g.inv_g);
    if (!(hx >= 0.0f && hy >= 0.0f && hz >= 0.0f && lx < (float)g.fdim_x && ly < (float)g.fdim_y &&
          lz < (float)g.fdim_z))
        return false;
    int x0 = clamp_i((int)fmaxf(lx, 0.0f) >> 2, 0, g.dim_x - 1), x1 = clamp_i((int)fminf(hx, (float)(g.fdim_x - 1)) >> 2, 0, g.dim_x - 1);
    int y0 = clamp_i((int)fmaxf(ly, 0.0f) >> 2, 0, g.dim_y - 1), y1 = clamp_i((int)fminf(hy, (float)(g.fdim_y - 1)) >> 2, 0, g.dim_y - 1);
    int z0 = clamp_i((int)fmaxf(lz, 0.0f) >> 2, 0, g.dim_z - 1), z1 = clamp_i((int)fminf(hz, (float)(g.fdim_z - 1)) >> 2, 0, g.dim_z - 1);
    for (int z = z0; z <= z1; ++z)
        for (int y = y0; y <= y1; ++y) {
            size_t row = ((size_t)z * g.dim_y + y) * g.dim_x;
            uint32_t s = __ldg(&g.vstart[row + x0]), e = __ldg(&g.vstart[row + x1 + 1]);
            for (; s < e; ++s) {
                float4 p = __ldg(&g.pts[s]);
                if (dist2_f32(qx, qy, qz, p.x, p.y, p.z) < thr) return true;
            }
        }
    return false;
}

// mask first, exact test only where flagged
__device__ __forceinline__ bool ball_any(const VoxelGeom& g, float qx, float qy, float qz, float thr,
                                         unsigned long long* exact_counter) {
    if (!mask_flagged(g, qx, qy, qz)) return false;
    if (exact_counter) ++*exact_counter;
    return exact_any(g, qx, qy, qz, thr);
}

}  // namespace plc
