// plc_voxel.cu — K1: voxelise a point cloud into the occupancy structure the sweep traverses.
//
// Replaces the KD-tree builds of terrainCallback / zx120PointsCallback
// (/root/reference/src/virtual_lidar.cpp:180-207).  Instead of a tree, points are counting-sorted
// into a uniform grid: every coarse voxel is a 128-byte brick of 8x8x8 fine voxels with two bits each — "a sample here may be within the radius of a point" (conservative dilation)
// and "a sample here certainly is" (conservative erosion) — and every mid voxel (4x4x4 fine voxels)
// owns the list of points within reach of its box.  Most samples of a ray are resolved by one or
// two word loads; only samples in the thin uncertain shell run the exact float32 test, over one list.
//
// Kernels (all HBM-bound, N = points, V = coarse voxels):
//   k_extract   raw PointCloud2 bytes -> float4, drops non-finite, bbox           N*(step) rd, N*16 wr
//   k_candidates<false>  per-mid-voxel candidate counts                            N*16 rd, <=27 atomics/pt
//   k_scan_*    exclusive scan of the counts (3 passes)                            ~3*8V*4
//   k_candidates<true>   points into their candidate lists (SoA groups of 4)       N*16 rd, <=27*12 wr/pt
//   k_dilate    2-bit fine-voxel states within reach of each point                 N*16 rd, ~100 atomicOr/pt
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "plc_voxel_query.cuh"
#include "plc_scan.cuh"

namespace plc {

namespace {

__device__ __forceinline__ int float_order_key(float f) {
    int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float float_from_order_key(int k) {
    int b = k >= 0 ? k : k ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

__device__ __forceinline__ float load_field(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    unsigned int b = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(b);
}

struct ExtractState {
    unsigned long long n_out;
    int bbox_key[6];     // min xyz, max xyz as order keys
};

__global__ void k_extract_init(ExtractState* st) {
    st->n_out = 0;
    for (int i = 0; i < 3; ++i) {
        st->bbox_key[i] = 0x7fffffff;
        st->bbox_key[3 + i] = (int)0x80000000;
    }
}

// One thread per input point.  Finite points are appended to `out` (one atomic per block: the order of the
// indexed points is irrelevant, only existence tests are made against them); the bounds go through a block
// reduction and touch the global words only when they improve them (same-address atomics from every warp cost
// 0.66 ms on a 4 M-point cloud, 0.07 ms this way).
__global__ void __launch_bounds__(256)
k_extract(const unsigned char* __restrict__ data, unsigned long long n, unsigned step, unsigned ox,
          unsigned oy, unsigned oz, bool aligned, float4* __restrict__ out, ExtractState* st) {
    __shared__ unsigned wcount[8];
    __shared__ unsigned long long s_base;
    __shared__ int red[6][8];
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool ok = false;
    float x = 0.f, y = 0.f, z = 0.f;
    if (i < n) {
        const unsigned char* p = data + i * step;
        x = load_field(p + ox, aligned);
        y = load_field(p + oy, aligned);
        z = load_field(p + oz, aligned);
        ok = isfinite(x) && isfinite(y) && isfinite(z);
    }
    const unsigned FULL = 0xffffffffu;
    const unsigned ballot = __ballot_sync(FULL, ok);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) wcount[warp] = __popc(ballot);
    const int lo_id = 0x7fffffff, hi_id = (int)0x80000000;
    int k[6];
    k[0] = __reduce_min_sync(FULL, ok ? float_order_key(x) : lo_id);
    k[1] = __reduce_min_sync(FULL, ok ? float_order_key(y) : lo_id);
    k[2] = __reduce_min_sync(FULL, ok ? float_order_key(z) : lo_id);
    k[3] = __reduce_max_sync(FULL, ok ? float_order_key(x) : hi_id);
    k[4] = __reduce_max_sync(FULL, ok ? float_order_key(y) : hi_id);
    k[5] = __reduce_max_sync(FULL, ok ? float_order_key(z) : hi_id);
    if (lane < 6) {
        int mine = k[0];
#pragma unroll
        for (int a = 1; a < 6; ++a) mine = lane == a ? k[a] : mine;
        red[lane][warp] = mine;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned total = 0;
        for (int w = 0; w < 8; ++w) {
            const unsigned c = wcount[w];
            wcount[w] = total;
            total += c;
        }
        s_base = total ? atomicAdd(&st->n_out, (unsigned long long)total) : 0ull;
    } else if (threadIdx.x >= 32 && threadIdx.x < 38) {
        const int a = threadIdx.x - 32;
        int v = red[a][0];
        for (int w = 1; w < 8; ++w) v = a < 3 ? min(v, red[a][w]) : max(v, red[a][w]);
        // a plain read first: after the first blocks almost no block still moves a bound
        if (a < 3) {
            if (v < *(volatile int*)&st->bbox_key[a]) atomicMin(&st->bbox_key[a], v);
        } else {
            if (v > *(volatile int*)&st->bbox_key[a]) atomicMax(&st->bbox_key[a], v);
        }
    }
    __syncthreads();
    if (ok) out[s_base + wcount[warp] + __popc(ballot & ((1u << lane) - 1u))] = make_float4(x, y, z, 0.f);
}

// Candidate lists: a point belongs to the list of every mid voxel (edge h/2) whose box lies within
// `reach` of it (double arithmetic, same slack as the occupancy bits).  FILL == false counts,
// FILL == true writes the point into its slot.  Up to 3x3x3 mid voxels per point.
// A list is stored as groups of four points in SoA form — x4, y4, z4 as three float4 — so the exact
// test handles four points per iteration from three independent 16-byte loads; unused slots of the
// last group hold +inf (never within any radius).
template <bool FILL>
__global__ void __launch_bounds__(256)
k_candidates(VoxelGeom g, const float4* __restrict__ raw, unsigned n, double reach, double msz,
             uint32_t* __restrict__ ccount, const uint32_t* __restrict__ gstart, float* __restrict__ cgrp) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = raw[i];
    double px = (double)p.x - (double)g.ox, py = (double)p.y - (double)g.oy, pz = (double)p.z - (double)g.oz;
    int x0 = max(0, (int)floor((px - reach) / msz)), x1 = min(g.mdim_x - 1, (int)floor((px + reach) / msz));
    int y0 = max(0, (int)floor((py - reach) / msz)), y1 = min(g.mdim_y - 1, (int)floor((py + reach) / msz));
    int z0 = max(0, (int)floor((pz - reach) / msz)), z1 = min(g.mdim_z - 1, (int)floor((pz + reach) / msz));
    double r2 = reach * reach;
    for (int z = z0; z <= z1; ++z) {
        double dz = fmax(0.0, fmax(z * msz - pz, pz - (z + 1) * msz));
        for (int y = y0; y <= y1; ++y) {
            double dy = fmax(0.0, fmax(y * msz - py, py - (y + 1) * msz));
            double dyz = dy * dy + dz * dz;
            if (dyz > r2) continue;
            size_t row = ((size_t)z * g.mdim_y + y) * g.mdim_x;
            for (int x = x0; x <= x1; ++x) {
                double dx = fmax(0.0, fmax(x * msz - px, px - (x + 1) * msz));
                if (dx * dx + dyz > r2) continue;
                uint32_t k = atomicAdd(&ccount[row + x], 1u);
                if (FILL) {
                    float* grp = cgrp + ((size_t)gstart[row + x] + (k >> 2)) * 12 + (k & 3);
                    grp[0] = p.x;
                    grp[4] = p.y;
                    grp[8] = p.z;
                }
            }
        }
    }
}

// point counts -> group counts (in place), ahead of the exclusive scan that yields gstart
__global__ void __launch_bounds__(256) k_counts_to_groups(uint32_t* __restrict__ c, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) c[i] = (c[i] + 3u) >> 2;
}

__global__ void __launch_bounds__(256) k_fill_f32(float* __restrict__ p, size_t n, float v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) p[i] = v;
}

// Two bits per fine voxel (`reach` and `inner` already contain the slack that absorbs the queries' float32
// voxel-coordinate rounding):
//   bit 0 where the voxel's box comes within `reach` of the point,
//   bit 1 where the whole box lies within `inner` of the point (inner <= 0: never).
// One thread per (point, fine z slice), so a warp shares the point and spreads over the slices.  The point is
// placed in its own fine voxel in double; from there on everything is float32 in voxel units relative to that
// voxel (values below ~10), with `reach` widened and `inner` narrowed by a thousandth of a voxel, far more than
// the float32 error, so bit 0 stays a superset and bit 1 a subset of the exact classification.  Each (y, z) row
// is two x intervals (maybe, certain), written as bit runs: at most three atomicOr per row.
__global__ void __launch_bounds__(256)
k_dilate(VoxelGeom g, const float4* __restrict__ raw, unsigned n, double reach, double inner, double gsz,
         int slices, uint32_t* __restrict__ bricks) {
    unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned i = (unsigned)(t / (unsigned)slices);
    int sl = (int)(t - (unsigned long long)i * (unsigned)slices);
    if (i >= n) return;
    float4 p = raw[i];
    const double ux = ((double)p.x - (double)g.ox) / gsz, uy = ((double)p.y - (double)g.oy) / gsz,
                 uz = ((double)p.z - (double)g.oz) / gsz;
    const double ru = reach / gsz;
    const int fx_max = (g.dim_x << kFineShift) - 1, fy_max = (g.dim_y << kFineShift) - 1,
              fz_max = (g.dim_z << kFineShift) - 1;
    const int z = max(0, (int)floor(uz - ru)) + sl;
    if (z > min(fz_max, (int)floor(uz + ru))) return;
    const int y0 = max(0, (int)floor(uy - ru)), y1 = min(fy_max, (int)floor(uy + ru));
    const int bx = (int)floor(ux), by = (int)floor(uy), bz = (int)floor(uz);
    const float lx = (float)(ux - bx), ly = (float)(uy - by), lz = (float)(uz - bz);       // in [0, 1]
    const float R = (float)ru * (1.0f + 1e-6f) + 1e-3f;
    const float RI = inner > 0.0 ? (float)(inner / gsz) * (1.0f - 1e-6f) - 1e-3f : -1.0f;
    const float R2 = R * R, RI2 = RI > 0.f ? RI * RI : -1.0f;
    const float zr = (float)(z - bz);
    const float dz_lo = zr - lz, dz_hi = lz - (zr + 1.0f);         // > 0 when the box is above / below the point
    const float dz = fmaxf(0.f, fmaxf(dz_lo, dz_hi));
    const float fz = fmaxf(fabsf(dz_lo), fabsf(dz_hi));            // farthest face along z
    for (int y = y0; y <= y1; ++y) {
        const float yr = (float)(y - by);
        const float dy_lo = yr - ly, dy_hi = ly - (yr + 1.0f);
        const float dy = fmaxf(0.f, fmaxf(dy_lo, dy_hi));
        const float rem = R2 - dy * dy - dz * dz;
        if (rem < 0.f) continue;
        const float w = sqrtf(rem);
        const int xa = max(0, bx + (int)floorf(lx - w)), xb = min(fx_max, bx + (int)floorf(lx + w));
        if (xa > xb) continue;
        int xc = 1, xd = 0;                                        // certain interval (empty)
        if (RI2 > 0.f) {
            const float fy = fmaxf(fabsf(dy_lo), fabsf(dy_hi));
            const float remi = RI2 - fy * fy - fz * fz;
            if (remi > 0.f) {
                const float wi = sqrtf(remi);
                xc = max(xa, bx + (int)floorf(lx - wi) + 1);
                xd = min(xb, bx + (int)ceilf(lx + wi) - 2);
            }
        }
        for (int v = xa >> kFineShift; v <= (xb >> kFineShift); ++v) {
            const int lo = v << kFineShift;
            const int a = max(xa, lo) - lo, b = min(xb, lo + 7) - lo;                    // fields a..b of this brick row
            uint32_t bits = (((1u << (2 * (b + 1))) - 1u) ^ ((1u << (2 * a)) - 1u)) & 0x5555u;
            const int c = max(xc, lo) - lo, d = min(xd, lo + 7) - lo;
            if (c <= d) bits |= (((1u << (2 * (d + 1))) - 1u) ^ ((1u << (2 * c)) - 1u)) & 0xAAAAu;
            const unsigned row = row_index(g, v, y, z);                                  // two 16-bit rows per word
            // Bits are only ever set, so a plain read that already shows them makes the atomic redundant (a stale
            // read merely repeats it).  Every voxel near a surface lies within reach of tens to hundreds of points:
            // nine atomics in ten go away (L2 read instead of L2 read-modify-write).
            uint32_t* wp = &bricks[row >> 1];
            const uint32_t want = bits << ((row & 1u) << 4);
            if ((__ldcg(wp) & want) != want) atomicOr(wp, want);
        }
    }
}

// Zeroes the bricks the previous build can have written: those of its non-empty list voxels (a fine voxel only gets
// bits within `reach` of a point, which puts that point into the list of the voxel's list voxel; the lists are
// built with a slightly larger reach than the dilation uses, see voxel_build).  One thread per 16 bytes of a list
// voxel's bricks would be finer; a thread per list voxel writing its (at most 8) bricks is enough at 20-150 MB.
__global__ void __launch_bounds__(256)
k_clear_touched(VoxelGeom g, size_t n_mid, const uint32_t* __restrict__ gstart, uint4* __restrict__ bricks) {
    const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_mid || gstart[m + 1] == gstart[m]) return;
    const int mx = (int)(m % (size_t)g.mdim_x), my = (int)((m / (size_t)g.mdim_x) % (size_t)g.mdim_y),
              mz = (int)(m / ((size_t)g.mdim_x * g.mdim_y));
    // list voxel edge = 2^mid_shift fine voxels; brick edge = 8 fine voxels
    const int fine = 1 << g.mid_shift;
    const int bx0 = (mx * fine) >> kFineShift, by0 = (my * fine) >> kFineShift, bz0 = (mz * fine) >> kFineShift;
    const int nb = fine > 8 ? fine >> kFineShift : 1;
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    for (int z = 0; z < nb; ++z)
        for (int y = 0; y < nb; ++y)
            for (int x = 0; x < nb; ++x) {
#if PLC_LAYOUT == 0
                uint4* b = bricks + (size_t)brick_of(g, bx0 + x, by0 + y, bz0 + z) * 8;
#pragma unroll
                for (int q = 0; q < 8; ++q) b[q] = zero;
#else
                for (int fz = 0; fz < 8; ++fz)
                    for (int fy = 0; fy < 8; ++fy)
                        reinterpret_cast<unsigned short*>(bricks)[row_index(g, bx0 + x, ((by0 + y) << 3) + fy, ((bz0 + z) << 3) + fz)] = 0;
#endif
            }
}

#if defined(PLC_TEX_MARCH)
// experiment: copies the bricks of the non-empty list voxels into the 3-D array the march reads through a texture
__global__ void __launch_bounds__(256)
k_rows_to_surface(VoxelGeom g, size_t n_mid, const uint32_t* __restrict__ gstart, const unsigned short* __restrict__ rows,
                  cudaSurfaceObject_t surf) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t m = t >> 6;
    const int r = (int)(t & 63);                 // row of the brick: (z & 7) * 8 + (y & 7)
    if (m >= n_mid || gstart[m + 1] == gstart[m]) return;
    const int mx = (int)(m % (size_t)g.mdim_x), my = (int)((m / (size_t)g.mdim_x) % (size_t)g.mdim_y),
              mz = (int)(m / ((size_t)g.mdim_x * g.mdim_y));
    const int fine = 1 << g.mid_shift;
    const int bx0 = (mx * fine) >> kFineShift, by0 = (my * fine) >> kFineShift, bz0 = (mz * fine) >> kFineShift;
    const int nb = fine > 8 ? fine >> kFineShift : 1;
    for (int z = 0; z < nb; ++z)
        for (int y = 0; y < nb; ++y)
            for (int x = 0; x < nb; ++x) {
                const int fy = ((by0 + y) << 3) + (r & 7), fz = ((bz0 + z) << 3) + (r >> 3);
                const unsigned short v = rows[row_index(g, bx0 + x, fy, fz)];
                surf3Dwrite(v, surf, (bx0 + x) * 2, fy, fz);
            }
}
__global__ void __launch_bounds__(256) k_surface_zero(cudaSurfaceObject_t surf, int dx, int dy, int dz) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int x = (int)(t % (size_t)dx), y = (int)((t / (size_t)dx) % (size_t)dy), z = (int)(t / ((size_t)dx * dy));
    if (z < dz) surf3Dwrite((unsigned short)0, surf, x * 2, y, z);
}
#endif

}  // namespace

// Uploads the cloud, drops non-finite points, records bbox.  The voxel index itself is (re)built
// by voxel_build for a given query radius.
plc_status voxel_set_cloud(plc_ctx* ctx, VoxelIndex& ix, const plc_cloud_view* cloud) {
    ix.n_raw = 0;
    ix.n_input = cloud->n_points;
    ix.built_radius = -1.0;
    ix.n_vox = 0;
    if (cloud->n_points == 0) return PLC_OK;
    if (cloud->n_points > 0xfffffff0ull) return PLC_E_INVALID_ARG;
    size_t bytes = (size_t)cloud->n_points * cloud->point_step;
    PLC_CUDA_TRY(ctx, ix.raw.reserve((size_t)cloud->n_points * sizeof(float4)));
    PLC_CUDA_TRY(ctx, ctx->scratch.reserve(4096));
    const unsigned char* raw = nullptr;
    PLC_TRY(stage_bytes(ctx, cloud->data, bytes, &raw));
    ExtractState* st = ctx->scratch.as<ExtractState>();
    ctx->T().begin(ctx->stream);
    k_extract_init<<<1, 1, 0, ctx->stream>>>(st);
    bool aligned = ((cloud->point_step | cloud->off_x | cloud->off_y | cloud->off_z) & 3u) == 0;
    unsigned blocks = (unsigned)((cloud->n_points + 255) / 256);
    k_extract<<<blocks, 256, 0, ctx->stream>>>(raw, cloud->n_points, cloud->point_step,
                                               cloud->off_x, cloud->off_y, cloud->off_z, aligned,
                                               ix.raw.as<float4>(), st);
    ctx->launches += 2;
    ctx->T().end(ctx->stream);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    ExtractState* host = static_cast<ExtractState*>(ctx->pinned);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, st, sizeof(ExtractState), cudaMemcpyDeviceToHost, ctx->stream));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ix.n_raw = host->n_out;
    for (int a = 0; a < 6; ++a) ix.bbox[a] = float_from_order_key(host->bbox_key[a]);
    return PLC_OK;
}

plc_status voxel_build(plc_ctx* ctx, VoxelIndex& ix, double radius, float thr, double march_err, double march_err_m,
                       int refine) {
    if (!ix.has_points()) return PLC_OK;
    if (ix.built_radius == radius && ix.built_thr == thr && march_err <= ix.built_march_err &&
        march_err_m <= ix.built_march_err_m && refine <= ix.asked_refine)
        return PLC_OK;
    ix.asked_refine = refine;
    refine = refine >= 4 ? 4 : (refine >= 2 ? 2 : 1);
    if (!(radius > 0) || !std::isfinite(radius) || !(thr >= 0) || !(march_err >= 0) || !(march_err_m >= 0))
        return PLC_E_INVALID_ARG;

    // real-number reach that covers everything the float32 test can accept, plus the slack that
    // absorbs float rounding of the voxel coordinate and the error of the fixed-point march
    const float r2f = static_cast<float>(radius * radius);
    double max_abs = 0.0;
    for (int a = 0; a < 6; ++a) max_abs = std::max(max_abs, (double)std::fabs(ix.bbox[a]));
    double extent = 0.0;
    for (int a = 0; a < 3; ++a) extent = std::max(extent, (double)ix.bbox[3 + a] - (double)ix.bbox[a]);
    // float32 rounding of a query coordinate (absolute) and of a grid-relative march position (extent)
    const double slack0 = 1e-4 + max_abs * 4.8e-7 + (extent + 2.0) * 6e-7;
    const double r_out = std::sqrt((double)r2f) * (1.0 + 1e-6);
    // h >= 2^-6 keeps the clip margins (h/16, h/32) far above float rounding at scene scale
    double h = std::max(std::exp2(std::ceil(std::log2(2.0 * (r_out + slack0)))), 1.0 / 64.0);
    double org[3], reach = 0.0, slack = 0.0;
    long long dim[3];
    int F = refine;      // bricks per list-voxel edge: the occupancy is F times finer than the candidate lists
    for (;;) {
        slack = slack0 + (h / (8.0 * F)) * march_err + march_err_m;
        reach = r_out + slack;
        if (h < 2.0 * reach) {
            h *= 2.0;
            continue;
        }
        double cells = 1.0;
        for (int a = 0; a < 3; ++a) {
            org[a] = (std::floor((double)ix.bbox[a] / h) - 1.0) * h;
            dim[a] = (long long)std::floor(((double)ix.bbox[3 + a] - org[a]) / h) + 2;
            cells *= (double)dim[a];
        }
        // fine coordinates must fit the march's signed 16.16 fixed point with room to spare
        static const double cap = getenv("PLC_DEV_BRICK_CAP") ? atof(getenv("PLC_DEV_BRICK_CAP")) : 1.6e7;
        bool fits = cells * F * F * F <= cap && dim[0] * F <= 3600 && dim[1] * F <= 3600 && dim[2] * F <= 3600;
        if (fits) break;
        if (F > 1) {
            F /= 2;  // the refinement is optional: give it up before coarsening the lists
            continue;
        }
        h *= 2.0;    // coarser voxels only make the lists longer, never incomplete
    }
    // "certain" bit: the float32 test accepts every sample of the fine voxel against this point
    const double inner = std::sqrt((double)std::min(thr, r2f)) * (1.0 - 1e-6) - slack;
    const size_t n_mid = (size_t)dim[0] * dim[1] * dim[2] * 8;      // list voxels: edge h / 2
    const size_t n_vox = (size_t)dim[0] * dim[1] * dim[2] * F * F * F;   // bricks: edge h / F, 8 fine voxels per edge
    const unsigned n = (unsigned)ix.n_raw;
    const double gsz = h / (8.0 * F);

    // Before anything of the previous build is freed or overwritten: un-set the bricks it wrote
    bool bricks_zero = false;
    if (ix.bricks_sparse_dirty && ix.bricks.ptr && ix.gstart.ptr && ix.n_mid) {
        ctx->T().begin(ctx->stream);
        k_clear_touched<<<(unsigned)((ix.n_mid + 255) / 256), 256, 0, ctx->stream>>>(ix.geom, ix.n_mid, ix.gstart.as<uint32_t>(),
                                                                                      ix.bricks.as<uint4>());
        ctx->T().end(ctx->stream);
        ctx->launches += 1;
        bricks_zero = true;
    }
    ix.bricks_sparse_dirty = false;
    PLC_CUDA_TRY(ctx, ix.gstart.reserve((n_mid + 1) * 4));
    PLC_CUDA_TRY(ctx, ix.cfill.reserve(n_mid * 4));
    PLC_CUDA_TRY(ctx, ix.scan_tmp.reserve(scan_tmp_bytes(n_mid + 1)));

    VoxelGeom g{};
    g.ox = (float)org[0];
    g.oy = (float)org[1];
    g.oz = (float)org[2];
    g.inv_g = (float)(8.0 * F / h);
    g.reach = (float)reach;
    g.dim_x = (int)dim[0] * F;
    g.dim_y = (int)dim[1] * F;
    g.dim_z = (int)dim[2] * F;
    g.mdim_x = (int)dim[0] * 2;
    g.mdim_y = (int)dim[1] * 2;
    g.mdim_z = (int)dim[2] * 2;
    g.mid_shift = F == 4 ? 4 : (F == 2 ? 3 : 2);
    g.fdim_xf = (float)(g.dim_x * 8);
    g.fdim_yf = (float)(g.dim_y * 8);
    g.fdim_zf = (float)(g.dim_z * 8);
    for (int a = 0; a < 3; ++a) {
        g.clip_lo[a] = (float)((double)ix.bbox[a] - reach - h / 16.0);
        g.clip_hi[a] = (float)((double)ix.bbox[3 + a] + reach + h / 16.0);
    }
    g.clip_t = (float)(h / 32.0);
    g.gstart = ix.gstart.as<uint32_t>();
    g.bricks = nullptr;
    g.cgrp = nullptr;

    cudaStream_t s = ctx->stream;
    unsigned pb = (n + 255) / 256;
    // The lists take every point within reach_l, a hair beyond the dilation's float32 reach (k_dilate widens its
    // radius by a thousandth of a fine voxel): every brick the dilation writes then belongs to a non-empty list
    // voxel, which is what lets the next build clear the occupancy sparsely.  Longer lists are harmless.
    const double reach_l = reach * (1.0 + 4e-6) + 4e-3 * gsz;
    // ---- candidate lists: counts -> offsets
    ctx->T().begin(s);
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.gstart.ptr, 0, (n_mid + 1) * 4, s));
    k_candidates<false><<<pb, 256, 0, s>>>(g, ix.raw.as<float4>(), n, reach_l, h / 2.0, ix.gstart.as<uint32_t>(),
                                           nullptr, nullptr);
    ctx->launches += 1;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    uint32_t* totals = static_cast<uint32_t*>(ctx->pinned);
    k_counts_to_groups<<<(unsigned)((n_mid + 255) / 256), 256, 0, s>>>(ix.gstart.as<uint32_t>(), n_mid);
    ctx->launches += 1 + scan_exclusive_u32(ix.gstart.as<uint32_t>(), n_mid + 1, ctx->lb, s);
    ctx->T().end(s);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(totals, ix.gstart.as<uint32_t>() + n_mid, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const size_t n_groups = totals[0];
    {
        const void* before = ix.bricks.ptr;
        PLC_CUDA_TRY(ctx, ix.bricks.reserve(n_vox * 128));
        if (ix.bricks.ptr != before) bricks_zero = false;            // a new allocation holds anything
    }
    PLC_CUDA_TRY(ctx, ix.cgrp.reserve(std::max<size_t>(n_groups, 1) * 48));
    g.bricks = ix.bricks.as<uint32_t>();
    g.cgrp = ix.cgrp.as<float4>();
    ctx->T().begin(s);
    // The occupancy dilation (issue-bound) and the list fill (bound by its scattered stores) touch different
    // structures: the fill runs on a side stream beside the dilation and is joined back before anything reads it.
    cudaStream_t side = s;
    if (ctx->side_stream || (cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking) == cudaSuccess &&
                             cudaEventCreateWithFlags(&ctx->side_fork, cudaEventDisableTiming) == cudaSuccess &&
                             cudaEventCreateWithFlags(&ctx->side_join, cudaEventDisableTiming) == cudaSuccess)) {
        if (ctx->side_fork && ctx->side_join && getenv("PLC_NO_SIDE_STREAM") == nullptr) side = ctx->side_stream;
    }
    if (side != s) {
        PLC_CUDA_TRY(ctx, cudaEventRecord(ctx->side_fork, s));
        PLC_CUDA_TRY(ctx, cudaStreamWaitEvent(side, ctx->side_fork, 0));
    }
    k_fill_f32<<<148 * 8, 256, 0, side>>>(ix.cgrp.as<float>(), n_groups * 12, INFINITY);
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.cfill.ptr, 0, n_mid * 4, side));
    k_candidates<true><<<pb, 256, 0, side>>>(g, ix.raw.as<float4>(), n, reach_l, h / 2.0, ix.cfill.as<uint32_t>(),
                                             ix.gstart.as<uint32_t>(), ix.cgrp.as<float>());
    if (side != s) PLC_CUDA_TRY(ctx, cudaEventRecord(ctx->side_join, side));
    if (!bricks_zero) PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.bricks.ptr, 0, n_vox * 128, s));
    const int slices = (int)std::floor(2.0 * reach / gsz) + 2;
    const unsigned long long dthreads = (unsigned long long)n * (unsigned)slices;
    k_dilate<<<(unsigned)((dthreads + 255) / 256), 256, 0, s>>>(g, ix.raw.as<float4>(), n, reach, inner, gsz, slices,
                                                                ix.bricks.as<uint32_t>());
    if (side != s) PLC_CUDA_TRY(ctx, cudaStreamWaitEvent(s, ctx->side_join, 0));
    ctx->launches += 3;
#if defined(PLC_TEX_MARCH)
    {
        const int dx = g.dim_x, dy = g.dim_y << kFineShift, dz = g.dim_z << kFineShift;
        if (ix.occ_dims[0] != dx || ix.occ_dims[1] != dy || ix.occ_dims[2] != dz) {
            if (ix.occ_tex) cudaDestroyTextureObject(ix.occ_tex);
            if (ix.occ_surf) cudaDestroySurfaceObject(ix.occ_surf);
            if (ix.occ_array) cudaFreeArray(ix.occ_array);
            ix.occ_tex = 0; ix.occ_surf = 0; ix.occ_array = nullptr;
            cudaChannelFormatDesc cd = cudaCreateChannelDesc(16, 0, 0, 0, cudaChannelFormatKindUnsigned);
            PLC_CUDA_TRY(ctx, cudaMalloc3DArray(&ix.occ_array, &cd, make_cudaExtent(dx, dy, dz), cudaArraySurfaceLoadStore));
            cudaResourceDesc rd{};
            rd.resType = cudaResourceTypeArray;
            rd.res.array.array = ix.occ_array;
            cudaTextureDesc td{};
            td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
            td.filterMode = cudaFilterModePoint;
            td.readMode = cudaReadModeElementType;
            td.normalizedCoords = 0;
            PLC_CUDA_TRY(ctx, cudaCreateTextureObject(&ix.occ_tex, &rd, &td, nullptr));
            PLC_CUDA_TRY(ctx, cudaCreateSurfaceObject(&ix.occ_surf, &rd));
            ix.occ_dims[0] = dx; ix.occ_dims[1] = dy; ix.occ_dims[2] = dz;
        }
        const size_t total = (size_t)dx * dy * dz;
        k_surface_zero<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(ix.occ_surf, dx, dy, dz);
        k_rows_to_surface<<<(unsigned)((n_mid * 64 + 255) / 256), 256, 0, s>>>(g, n_mid, ix.gstart.as<uint32_t>(),
                                                                                ix.bricks.as<unsigned short>(), ix.occ_surf);
        g.occ_tex = ix.occ_tex;
    }
#endif
    ctx->T().end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    ix.geom = g;
    ix.h = h;
    ix.n_vox = n_vox;
    ix.n_groups = n_groups;
    ix.built_radius = radius;
    ix.built_thr = thr;
    ix.built_march_err = march_err;
    ix.built_march_err_m = march_err_m;
    ix.built_refine = F;
    ix.n_mid = n_mid;
    ix.bricks_sparse_dirty = true;
    return PLC_OK;
}

}  // namespace plc
