// plc_voxel.cu — K1: voxelise a point cloud into the occupancy structure the sweep traverses.
//
// Replaces the KD-tree builds of terrainCallback / zx120PointsCallback
// (/root/reference/src/virtual_lidar.cpp:180-207).  Instead of a tree, points are counting-sorted
// into a dense uniform grid: a 4x4x4-per-coarse-voxel bitmask of fine voxels records where a sample
// could possibly be within the radius of a point (conservative dilation), and every mid voxel
// (2x2x2 fine voxels) owns the list of points within reach of its box.  Most samples of a ray are
// resolved by one bit test; only flagged samples run the exact float32 test, over one short list.
//
// Kernels (all HBM-bound, N = points, V = coarse voxels):
//   k_extract   raw PointCloud2 bytes -> float4, drops non-finite, bbox           N*(step) rd, N*16 wr
//   k_candidates<false>  per-mid-voxel candidate counts                            N*16 rd, <=27 atomics/pt
//   k_scan_*    exclusive scan of the counts (3 passes)                            ~3*8V*4
//   k_candidates<true>   points into their candidate lists                         N*16 rd, <=27*16 wr/pt
//   k_dilate    fine-voxel mask bits within reach of each point                    N*16 rd, <=125 atomicOr/pt
#include <algorithm>
#include <cmath>

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

// One thread per input point.  Finite points are appended (warp-aggregated) to `out`.
__global__ void __launch_bounds__(256)
k_extract(const unsigned char* __restrict__ data, unsigned long long n, unsigned step, unsigned ox,
          unsigned oy, unsigned oz, bool aligned, float4* __restrict__ out, ExtractState* st) {
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
    unsigned ballot = __ballot_sync(0xffffffffu, ok);
    int lane = threadIdx.x & 31;
    unsigned long long base = 0;
    if (lane == 0 && ballot) base = atomicAdd(&st->n_out, (unsigned long long)__popc(ballot));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (ok) out[base + __popc(ballot & ((1u << lane) - 1))] = make_float4(x, y, z, 0.f);

    // bbox: warp reduce then one atomic per warp and axis
    int k[6];
    k[0] = ok ? float_order_key(x) : 0x7fffffff;
    k[1] = ok ? float_order_key(y) : 0x7fffffff;
    k[2] = ok ? float_order_key(z) : 0x7fffffff;
    k[3] = ok ? float_order_key(x) : (int)0x80000000;
    k[4] = ok ? float_order_key(y) : (int)0x80000000;
    k[5] = ok ? float_order_key(z) : (int)0x80000000;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            k[a] = min(k[a], __shfl_xor_sync(0xffffffffu, k[a], o));
            k[3 + a] = max(k[3 + a], __shfl_xor_sync(0xffffffffu, k[3 + a], o));
        }
    }
    if (lane == 0 && ballot) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&st->bbox_key[a], k[a]);
            atomicMax(&st->bbox_key[3 + a], k[3 + a]);
        }
    }
}

// Candidate lists: a point belongs to the list of every mid voxel (edge 2g) whose box lies within
// `reach` of it (double arithmetic, same slack as the mask).  FILL == false counts, FILL == true
// writes the point into its slots.  Up to 3x3x3 mid voxels per point.
template <bool FILL>
__global__ void __launch_bounds__(256)
k_candidates(VoxelGeom g, const float4* __restrict__ raw, unsigned n, double reach, double msz,
             uint32_t* __restrict__ ccount, const uint32_t* __restrict__ cstart, float4* __restrict__ cpts) {
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
                if (FILL) cpts[cstart[row + x] + k] = p;
            }
        }
    }
}

// Sets the bit of every fine voxel whose box lies within `reach` of the point (double arithmetic;
// `reach` already contains the slack that covers float rounding in fine_coord_f).
__global__ void __launch_bounds__(128)
k_dilate(VoxelGeom g, const float4* __restrict__ raw, unsigned n, double reach, double gsz,
         uint32_t* __restrict__ mask32) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = raw[i];
    double px = (double)p.x - (double)g.ox, py = (double)p.y - (double)g.oy, pz = (double)p.z - (double)g.oz;
    int x0 = max(0, (int)floor((px - reach) / gsz)), x1 = min(g.dim_x * 4 - 1, (int)floor((px + reach) / gsz));
    int y0 = max(0, (int)floor((py - reach) / gsz)), y1 = min(g.dim_y * 4 - 1, (int)floor((py + reach) / gsz));
    int z0 = max(0, (int)floor((pz - reach) / gsz)), z1 = min(g.dim_z * 4 - 1, (int)floor((pz + reach) / gsz));
    double r2 = reach * reach;
    for (int z = z0; z <= z1; ++z) {
        double dz = fmax(0.0, fmax(z * gsz - pz, pz - (z + 1) * gsz));
        for (int y = y0; y <= y1; ++y) {
            double dy = fmax(0.0, fmax(y * gsz - py, py - (y + 1) * gsz));
            double dyz = dy * dy + dz * dz;
            if (dyz > r2) continue;
            // bits of this (y,z) row grouped per coarse voxel word
            size_t row = ((size_t)(z >> 2) * g.dim_y + (y >> 2)) * g.dim_x;
            int sub = ((z >> 1) & 1);
            int bit_yz = ((y & 3) << 2) | ((z & 1) << 4);
            uint32_t acc = 0;
            int acc_v = -1;
            for (int x = x0; x <= x1; ++x) {
                double dx = fmax(0.0, fmax(x * gsz - px, px - (x + 1) * gsz));
                if (dx * dx + dyz > r2) continue;
                int v = x >> 2;
                if (v != acc_v) {
                    if (acc) atomicOr(&mask32[2 * (row + acc_v) + sub], acc);
                    acc = 0;
                    acc_v = v;
                }
                acc |= 1u << ((x & 3) | bit_yz);
            }
            if (acc) atomicOr(&mask32[2 * (row + acc_v) + sub], acc);
        }
    }
}

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
    PLC_CUDA_TRY(ctx, ctx->upload.reserve(bytes));
    PLC_CUDA_TRY(ctx, ix.raw.reserve((size_t)cloud->n_points * sizeof(float4)));
    PLC_CUDA_TRY(ctx, ctx->scratch.reserve(4096));
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->upload.ptr, cloud->data, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ExtractState* st = ctx->scratch.as<ExtractState>();
    k_extract_init<<<1, 1, 0, ctx->stream>>>(st);
    bool aligned = ((cloud->point_step | cloud->off_x | cloud->off_y | cloud->off_z) & 3u) == 0;
    unsigned blocks = (unsigned)((cloud->n_points + 255) / 256);
    k_extract<<<blocks, 256, 0, ctx->stream>>>(ctx->upload.as<unsigned char>(), cloud->n_points, cloud->point_step,
                                               cloud->off_x, cloud->off_y, cloud->off_z, aligned,
                                               ix.raw.as<float4>(), st);
    ctx->launches += 2;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    ExtractState* host = static_cast<ExtractState*>(ctx->pinned);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, st, sizeof(ExtractState), cudaMemcpyDeviceToHost, ctx->stream));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ix.n_raw = host->n_out;
    for (int a = 0; a < 6; ++a) ix.bbox[a] = float_from_order_key(host->bbox_key[a]);
    return PLC_OK;
}

plc_status voxel_build(plc_ctx* ctx, VoxelIndex& ix, double radius) {
    if (!ix.has_points()) return PLC_OK;
    if (ix.built_radius == radius) return PLC_OK;
    if (!(radius > 0) || !std::isfinite(radius)) return PLC_E_INVALID_ARG;

    // real-number reach that covers everything the float32 test can accept, plus float slack
    const float r2f = static_cast<float>(radius * radius);
    double max_abs = 0.0;
    for (int a = 0; a < 6; ++a) max_abs = std::max(max_abs, (double)std::fabs(ix.bbox[a]));
    const double slack = 1e-4 + max_abs * 4.8e-7;
    const double reach = std::sqrt((double)r2f) * (1.0 + 1e-6) + slack;
    double h = std::exp2(std::ceil(std::log2(2.0 * reach)));
    if (h < 2.0 * reach) h *= 2.0;
    double org[3];
    long long dim[3];
    for (;;) {
        double cells = 1.0;
        for (int a = 0; a < 3; ++a) {
            org[a] = (std::floor((double)ix.bbox[a] / h) - 1.0) * h;
            dim[a] = (long long)std::floor(((double)ix.bbox[3 + a] - org[a]) / h) + 2;
            cells *= (double)dim[a];
        }
        bool fits = cells <= 2.5e8 && dim[0] * 4 < (1 << 23) && dim[1] * 4 < (1 << 23) && dim[2] * 4 < (1 << 23);
        if (fits) break;
        h *= 2.0;    // coarser voxels only make the lists longer, never incomplete
    }
    const size_t n_vox = (size_t)dim[0] * dim[1] * dim[2];
    const size_t n_mid = n_vox * 8;
    const unsigned n = (unsigned)ix.n_raw;

    PLC_CUDA_TRY(ctx, ix.mask32.reserve(n_vox * 8));
    PLC_CUDA_TRY(ctx, ix.cstart.reserve((n_mid + 1) * 4));
    PLC_CUDA_TRY(ctx, ix.cfill.reserve(n_mid * 4));
    PLC_CUDA_TRY(ctx, ix.scan_tmp.reserve(scan_tmp_bytes(n_mid + 1)));

    VoxelGeom g{};
    g.ox = (float)org[0];
    g.oy = (float)org[1];
    g.oz = (float)org[2];
    g.inv_g = (float)(4.0 / h);
    g.reach = (float)reach;
    g.dim_x = (int)dim[0];
    g.dim_y = (int)dim[1];
    g.dim_z = (int)dim[2];
    g.mdim_x = g.dim_x * 2;
    g.mdim_y = g.dim_y * 2;
    g.mdim_z = g.dim_z * 2;
    g.fdim_xf = (float)(g.dim_x * 4);
    g.fdim_yf = (float)(g.dim_y * 4);
    g.fdim_zf = (float)(g.dim_z * 4);
    g.mask32 = ix.mask32.as<uint32_t>();
    g.cstart = ix.cstart.as<uint32_t>();
    g.cpts = nullptr;

    cudaStream_t s = ctx->stream;
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.mask32.ptr, 0, n_vox * 8, s));
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.cstart.ptr, 0, (n_mid + 1) * 4, s));
    PLC_CUDA_TRY(ctx, cudaMemsetAsync(ix.cfill.ptr, 0, n_mid * 4, s));
    unsigned pb = (n + 255) / 256;
    k_candidates<false><<<pb, 256, 0, s>>>(g, ix.raw.as<float4>(), n, reach, h / 2.0, ix.cstart.as<uint32_t>(),
                                           nullptr, nullptr);
    scan_exclusive_u32(ix.cstart.as<uint32_t>(), n_mid + 1, ix.scan_tmp.as<uint32_t>(), s);
    k_dilate<<<(n + 127) / 128, 128, 0, s>>>(g, ix.raw.as<float4>(), n, reach, h / 4.0, ix.mask32.as<uint32_t>());
    ctx->launches += 5;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    // total number of list entries decides the size of the candidate array
    uint32_t* total = static_cast<uint32_t*>(ctx->pinned);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(total, ix.cstart.as<uint32_t>() + n_mid, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const size_t n_cand = *total;
    PLC_CUDA_TRY(ctx, ix.cpts.reserve(std::max<size_t>(n_cand, 1) * sizeof(float4)));
    g.cpts = ix.cpts.as<float4>();
    k_candidates<true><<<pb, 256, 0, s>>>(g, ix.raw.as<float4>(), n, reach, h / 2.0, ix.cfill.as<uint32_t>(),
                                          ix.cstart.as<uint32_t>(), ix.cpts.as<float4>());
    ctx->launches += 1;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    ix.geom = g;
    ix.h = h;
    ix.n_vox = n_vox;
    ix.n_cand = n_cand;
    ix.built_radius = radius;
    return PLC_OK;
}

}  // namespace plc
