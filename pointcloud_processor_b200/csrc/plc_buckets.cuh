// plc_buckets.cuh — a point cloud bucketed on the device, shared by the node-side steps that ask neighbourhood
// questions (plc_setup.cu, plc_excavate.cu).  The reference answers them with pcl::KdTreeFLANN radius searches;
// here the cloud is bucketed once (stable radix sort by bucket, so the order inside a bucket is the cloud's own)
// and a query walks the rows of buckets its ball can touch.  Acceptance is PCL's rule:
// float32 query, d2 = ((dx*dx)+(dy*dy))+(dz*dz) in float32 without FMA, accept iff d2 < (float)(radius*radius).
#pragma once
#include <cmath>
#include <limits>

#include "plc_internal.cuh"
#include "plc_sort.cuh"

namespace plc {
namespace {

// ---- bucketed cloud ---------------------------------------------------------------------------
struct BucketGrid {
    double ox, oy, oz, inv_e;      // bucket coordinate of p on an axis: floor((p - o) * inv_e), clamped
    int nx, ny, nz;
    const uint32_t* __restrict__ start;   // [nx*ny*nz + 1], x fastest
    const float4* __restrict__ pts;       // bucket order; w = index of the point in the caller's cloud
};

__device__ __forceinline__ int order_key(float f) {
    int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
inline float from_order_key(int k) {
    int b = k >= 0 ? k : k ^ 0x7fffffff;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

__device__ __forceinline__ float field_f32(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    unsigned int b = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(b);
}

// keys: [0..2] min, [3..5] max over finite points (what gets bucketed);
//       [6..8] min, [9..11] max over every coordinate that is not NaN — the std::min/std::max chain of :242-249
struct BoundsState {
    int key[12];
};
__global__ void k_bounds_init(BoundsState* st) {
    for (int a = 0; a < 3; ++a) {
        st->key[a] = st->key[6 + a] = 0x7fffffff;
        st->key[3 + a] = st->key[9 + a] = (int)0x80000000;
    }
}

// PointCloud2 bytes -> float4 {x, y, z, bit pattern of the point's index}; every point is kept in place
__global__ void __launch_bounds__(256)
k_load_points(const unsigned char* __restrict__ data, unsigned n, unsigned step, unsigned ox, unsigned oy, unsigned oz,
              bool aligned, float4* __restrict__ out, BoundsState* st) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    int k[12];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        k[a] = k[6 + a] = 0x7fffffff;
        k[3 + a] = k[9 + a] = (int)0x80000000;
    }
    if (i < n) {
        const unsigned char* p = data + (size_t)i * step;
        const float v[3] = {field_f32(p + ox, aligned), field_f32(p + oy, aligned), field_f32(p + oz, aligned)};
        out[i] = make_float4(v[0], v[1], v[2], __uint_as_float(i));
        const bool fin = isfinite(v[0]) && isfinite(v[1]) && isfinite(v[2]);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (fin) k[a] = k[3 + a] = order_key(v[a]);
            if (!isnan(v[a])) k[6 + a] = k[9 + a] = order_key(v[a]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            k[a] = min(k[a], __shfl_xor_sync(0xffffffffu, k[a], o));
            k[3 + a] = max(k[3 + a], __shfl_xor_sync(0xffffffffu, k[3 + a], o));
            k[6 + a] = min(k[6 + a], __shfl_xor_sync(0xffffffffu, k[6 + a], o));
            k[9 + a] = max(k[9 + a], __shfl_xor_sync(0xffffffffu, k[9 + a], o));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&st->key[a], k[a]);
            atomicMax(&st->key[3 + a], k[3 + a]);
            atomicMin(&st->key[6 + a], k[6 + a]);
            atomicMax(&st->key[9 + a], k[9 + a]);
        }
}

// the same bounds for points that are already float4 on the device (the terrain kept by plc_vl_set_terrain)
__global__ void __launch_bounds__(256)
k_bounds_of(const float4* __restrict__ pts, unsigned n, BoundsState* st) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    int k[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    if (i < n) {
        const float4 p = pts[i];
        if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
            k[0] = k[3] = order_key(p.x);
            k[1] = k[4] = order_key(p.y);
            k[2] = k[5] = order_key(p.z);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            k[a] = min(k[a], __shfl_xor_sync(0xffffffffu, k[a], o));
            k[3 + a] = max(k[3 + a], __shfl_xor_sync(0xffffffffu, k[3 + a], o));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&st->key[a], k[a]);
            atomicMax(&st->key[3 + a], k[3 + a]);
        }
}

__device__ __forceinline__ int bucket_of(double p, double o, double inv_e, int n) {
    const double b = floor((p - o) * inv_e);
    return b < 0.0 ? 0 : (b > (double)(n - 1) ? n - 1 : (int)b);
}

// key = bucket, or the sentinel n_buckets for a non-finite point (sorted to the end, never visited)
__global__ void __launch_bounds__(256)
k_bucket_keys(const float4* __restrict__ pts, unsigned n, BucketGrid G, uint32_t* __restrict__ keys,
              uint32_t* __restrict__ vals, uint32_t* __restrict__ counts) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = pts[i];
    uint32_t key = (uint32_t)G.nx * G.ny * G.nz;
    if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
        const int bx = bucket_of(p.x, G.ox, G.inv_e, G.nx), by = bucket_of(p.y, G.oy, G.inv_e, G.ny),
                  bz = bucket_of(p.z, G.oz, G.inv_e, G.nz);
        key = ((uint32_t)bz * G.ny + by) * G.nx + bx;
        atomicAdd(&counts[key], 1u);
    }
    keys[i] = key;
    vals[i] = i;
}

__global__ void __launch_bounds__(256)
k_gather_points(const float4* __restrict__ pts, const uint32_t* __restrict__ order, unsigned n, float4* __restrict__ out) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = pts[order[i]];
}

// Rows of buckets a query box can touch.  The box is the query's ball (or the tighter box a caller knows)
// widened by a margin far above the float32 error of the acceptance test, so every accepted point is visited.
struct RowRange {
    int x0, x1, y0, y1, z0, z1;
    __device__ __forceinline__ bool empty() const { return x0 > x1 || y0 > y1 || z0 > z1; }
};
__device__ __forceinline__ void axis_range(double lo, double hi, double o, double inv_e, int n, int& b0, int& b1) {
    const double f0 = floor((lo - o) * inv_e), f1 = floor((hi - o) * inv_e);
    // NaN bounds (a NaN query accepts nothing) compare false everywhere and leave the empty range 0 > -1
    b0 = 0;
    b1 = -1;
    if (f0 <= f1) {
        b0 = f0 < 0.0 ? 0 : (f0 > (double)n ? n : (int)f0);
        b1 = f1 < 0.0 ? -1 : (f1 > (double)(n - 1) ? n - 1 : (int)f1);
    }
}
__device__ __forceinline__ RowRange rows_of_box(const BucketGrid& G, double cx, double cy, double cz, double rx, double ry,
                                                double rz) {
    RowRange r;
    axis_range(cx - rx, cx + rx, G.ox, G.inv_e, G.nx, r.x0, r.x1);
    axis_range(cy - ry, cy + ry, G.oy, G.inv_e, G.ny, r.y0, r.y1);
    axis_range(cz - rz, cz + rz, G.oz, G.inv_e, G.nz, r.z0, r.z1);
    return r;
}
__host__ __device__ __forceinline__ double widened(double radius) { return radius * (1.0 + 1e-5) + 1e-6; }

__device__ __forceinline__ float dist2(float qx, float qy, float qz, float px, float py, float pz) {
    const float ex = qx - px, ey = qy - py, ez = qz - pz;
    return __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// keep flags are scanned in place; the last flag is needed afterwards for the total
__global__ void k_copy_u32(const uint32_t* __restrict__ src, size_t n, uint32_t* __restrict__ dst) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

// ---- host side --------------------------------------------------------------------------------
inline unsigned blocks_for(size_t n, unsigned per_block) { return (unsigned)((n + per_block - 1) / per_block); }

struct Bounds {
    float fin_lo[3], fin_hi[3];       // finite points
    float any_lo[3], any_hi[3];       // NaN-skipping min/max chain of :242-249
    bool any_finite;
};

plc_status read_bounds(plc_ctx* ctx, const BoundsState* dev, Bounds* out) {
    BoundsState* host = static_cast<BoundsState*>(ctx->pinned_reserve(sizeof(BoundsState)));
    if (!host) return PLC_E_OOM;
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, dev, sizeof(BoundsState), cudaMemcpyDeviceToHost, ctx->stream));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < 3; ++a) {
        out->fin_lo[a] = from_order_key(host->key[a]);
        out->fin_hi[a] = from_order_key(host->key[3 + a]);
        out->any_lo[a] = from_order_key(host->key[6 + a]);
        out->any_hi[a] = from_order_key(host->key[9 + a]);
    }
    out->any_finite = host->key[0] <= host->key[3];
    return PLC_OK;
}

// The PointCloud2 bytes (copied in, or read in place from a device buffer) and their conversion to float4 {x, y, z, index}; bounds come back to the host
plc_status load_cloud(plc_ctx* ctx, const plc_cloud_view* cloud, Bounds* bounds) {
    const size_t n = cloud->n_points;
    const size_t bytes = n * (size_t)cloud->point_step;
    PLC_CUDA_TRY(ctx, ctx->su_pts.reserve(n * sizeof(float4)));
    PLC_CUDA_TRY(ctx, ctx->su_state.reserve(sizeof(BoundsState)));
    PLC_TRY(stage_bytes(ctx, cloud->data, bytes, &ctx->su_raw));
    const bool aligned = cloud->point_step % 4 == 0 && cloud->off_x % 4 == 0 && cloud->off_y % 4 == 0 && cloud->off_z % 4 == 0;
    ctx->timer.begin(ctx->stream);
    k_bounds_init<<<1, 1, 0, ctx->stream>>>(ctx->su_state.as<BoundsState>());
    k_load_points<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(ctx->su_raw, (unsigned)n, cloud->point_step,
                                                               cloud->off_x, cloud->off_y, cloud->off_z, aligned,
                                                               ctx->su_pts.as<float4>(), ctx->su_state.as<BoundsState>());
    ctx->timer.end(ctx->stream);
    ctx->launches += 2;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    return read_bounds(ctx, ctx->su_state.as<BoundsState>(), bounds);
}

// Buckets `pts[0..n)` (device, float4 {x, y, z, index}) with edge >= `edge`, doubled until the table is small.
plc_status build_buckets(plc_ctx* ctx, const float4* pts, size_t n, const Bounds& b, double edge, BucketGrid* out) {
    BucketGrid G{};
    G.nx = G.ny = G.nz = 1;
    G.inv_e = 1.0;
    if (b.any_finite) {
        G.ox = b.fin_lo[0];
        G.oy = b.fin_lo[1];
        G.oz = b.fin_lo[2];
        const double ext[3] = {(double)b.fin_hi[0] - b.fin_lo[0], (double)b.fin_hi[1] - b.fin_lo[1],
                               (double)b.fin_hi[2] - b.fin_lo[2]};
        for (;; edge *= 2.0) {
            const double dims[3] = {std::floor(ext[0] / edge) + 1.0, std::floor(ext[1] / edge) + 1.0,
                                    std::floor(ext[2] / edge) + 1.0};
            if (dims[0] * dims[1] * dims[2] <= (double)(1u << 22)) {
                G.nx = (int)dims[0];
                G.ny = (int)dims[1];
                G.nz = (int)dims[2];
                break;
            }
        }
        G.inv_e = 1.0 / edge;
    }
    const size_t n_buckets = (size_t)G.nx * G.ny * G.nz;
    int key_bits = 1;
    while ((1ull << key_bits) <= n_buckets) ++key_bits;        // the sentinel n_buckets must fit too
    const size_t table = sort_table_entries(n);
    PLC_CUDA_TRY(ctx, ctx->su_start.reserve((n_buckets + 2) * sizeof(uint32_t)));
    PLC_CUDA_TRY(ctx, ctx->su_sorted.reserve((n + 1) * sizeof(float4)));
    PLC_CUDA_TRY(ctx, ctx->su_keys.reserve(4 * (n + 1) * sizeof(uint32_t)));
    PLC_CUDA_TRY(ctx, ctx->su_table.reserve((table + 1) * sizeof(uint32_t)));
    PLC_CUDA_TRY(ctx, ctx->su_tmp.reserve(scan_tmp_bytes(std::max(table, n_buckets + 1)) + 64));
    uint32_t* start = ctx->su_start.as<uint32_t>();
    uint32_t* keys_a = ctx->su_keys.as<uint32_t>();
    uint32_t *vals_a = keys_a + (n + 1), *keys_b = vals_a + (n + 1), *vals_b = keys_b + (n + 1);
    cudaStream_t s = ctx->stream;
    ctx->timer.begin(s);
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(start, 0, (n_buckets + 1) * sizeof(uint32_t), s));
    if (n) {
        k_bucket_keys<<<blocks_for(n, 256), 256, 0, s>>>(pts, (unsigned)n, G, keys_a, vals_a, start);
        ctx->launches += scan_exclusive_u32(start, n_buckets + 1, ctx->lb, s);
        const int where = radix_sort_pairs(keys_a, vals_a, keys_b, vals_b, (unsigned)n, key_bits, ctx->su_table.as<uint32_t>(),
                                           ctx->lb, s, &ctx->launches);
        k_gather_points<<<blocks_for(n, 256), 256, 0, s>>>(pts, where ? vals_b : vals_a, (unsigned)n, ctx->su_sorted.as<float4>());
        ctx->launches += 2;
    }
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    G.start = start;
    G.pts = ctx->su_sorted.as<float4>();
    *out = G;
    return PLC_OK;
}

plc_status finite_count(plc_ctx* ctx, const BucketGrid& G, uint32_t* n_finite) {
    // start[n_buckets] = number of finite points (the sentinel bucket begins there)
    uint32_t* host = static_cast<uint32_t*>(ctx->pinned_reserve(sizeof(uint32_t)));
    if (!host) return PLC_E_OOM;
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, G.start + (size_t)G.nx * G.ny * G.nz, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    *n_finite = *host;
    return PLC_OK;
}

inline float radius_sq_f32(double radius) { return static_cast<float>(radius * radius); }

// flags -> exclusive scan (in `slot`) and total
plc_status scan_flags(plc_ctx* ctx, const uint32_t* flags, uint32_t* slot, size_t n, uint64_t* total) {
    cudaStream_t s = ctx->stream;
    PLC_CUDA_TRY(ctx, ctx->su_tmp.reserve(scan_tmp_bytes(n + 1) + 64));
    k_copy_u32<<<blocks_for(n, 256), 256, 0, s>>>(flags, n, slot);
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(slot + n, 0, sizeof(uint32_t), s));
    ctx->launches += 1 + scan_exclusive_u32(slot, n + 1, ctx->lb, s);
    uint32_t* host = static_cast<uint32_t*>(ctx->pinned_reserve(sizeof(uint32_t)));
    if (!host) return PLC_E_OOM;
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, slot + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    *total = *host;
    return PLC_OK;
}

}  // namespace
}  // namespace plc
