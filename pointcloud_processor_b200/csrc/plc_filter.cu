// plc_filter.cu — K6 + K7: crop box and voxel-grid centroid downsample of one LiDAR scan.
//
// Replaces processCloudSimple's two stages in /root/reference/src/pointcloud_filter.cpp:
//   cropFrontArea   87-120   strict six-sided box test, input order kept        -> k_crop_flag + scan + k_crop_compact
//   downsampleCloud 122-139  pcl::VoxelGrid<PointXYZ> (leaf from the parameter) -> k_voxel_keys + radix sort +
//                                                                                  k_heads + scan + k_centroid
// PCL semantics restated in SURVEY §A.3 item 13: float32 inverse leaf, min/max over the cloud,
// idx = i + j*dx + k*dx*dy, output ascending in idx, centroid = float32 running sum / count.
// All kernels are bandwidth-bound streams over N (crop) or N' (downsample) points.
#include <cmath>
#include <limits>

#include "plc_internal.cuh"
#include "plc_sort.cuh"

namespace plc {
namespace {

__device__ __forceinline__ float ld_f32(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    unsigned b = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(b);
}

struct CropDev {
    double x_min, x_max, y_min, y_max, z_min, z_max;
};

// pointcloud_filter.cpp:111-113 — float promoted to double, all comparisons strict (NaN fails)
__global__ void __launch_bounds__(256)
k_crop_flag(const unsigned char* __restrict__ data, unsigned n, unsigned step, unsigned ox, unsigned oy,
            unsigned oz, bool aligned, CropDev b, uint32_t* __restrict__ flag) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    if (i == n) {
        flag[n] = 0;
        return;
    }
    const unsigned char* p = data + (size_t)i * step;
    double x = ld_f32(p + ox, aligned), y = ld_f32(p + oy, aligned), z = ld_f32(p + oz, aligned);
    flag[i] = (x > b.x_min && x < b.x_max && y > b.y_min && y < b.y_max && z > b.z_min && z < b.z_max) ? 1u : 0u;
}

__device__ __forceinline__ int order_key(float f) {
    int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}

struct FilterState {
    int bbox_key[6];
    unsigned n_voxels;
    unsigned pad;
};

__global__ void k_filter_state_init(FilterState* st) {
    for (int a = 0; a < 3; ++a) {
        st->bbox_key[a] = 0x7fffffff;
        st->bbox_key[3 + a] = (int)0x80000000;
    }
    st->n_voxels = 0;
}

// order-preserving compaction; also accumulates the min/max that VoxelGrid needs (getMinMax3D)
__global__ void __launch_bounds__(256)
k_crop_compact(const unsigned char* __restrict__ data, unsigned n, unsigned step, unsigned ox, unsigned oy,
               unsigned oz, bool aligned, const uint32_t* __restrict__ pos, float4* __restrict__ out,
               FilterState* st) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    bool keep = false;
    float x = 0, y = 0, z = 0;
    if (i < n) {
        keep = pos[i + 1] != pos[i];
        if (keep) {
            const unsigned char* p = data + (size_t)i * step;
            x = ld_f32(p + ox, aligned);
            y = ld_f32(p + oy, aligned);
            z = ld_f32(p + oz, aligned);
            out[pos[i]] = make_float4(x, y, z, 1.0f);   // pcl::PointXYZ keeps 1.0f in data[3]
        }
    }
    // block-wide min/max of the order-preserving keys, then one set of atomics per block
    __shared__ int red[6][8];
    const unsigned FULL = 0xffffffffu;
    const int lo_id = 0x7fffffff, hi_id = (int)0x80000000;
    int k[6];
    k[0] = __reduce_min_sync(FULL, keep ? order_key(x) : lo_id);
    k[1] = __reduce_min_sync(FULL, keep ? order_key(y) : lo_id);
    k[2] = __reduce_min_sync(FULL, keep ? order_key(z) : lo_id);
    k[3] = __reduce_max_sync(FULL, keep ? order_key(x) : hi_id);
    k[4] = __reduce_max_sync(FULL, keep ? order_key(y) : hi_id);
    k[5] = __reduce_max_sync(FULL, keep ? order_key(z) : hi_id);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < 6) {
        int mine = k[0];
#pragma unroll
        for (int a = 1; a < 6; ++a) mine = lane == a ? k[a] : mine;
        red[lane][warp] = mine;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int a = threadIdx.x;
        int v = red[a][0];
        for (int w = 1; w < 8; ++w) v = a < 3 ? min(v, red[a][w]) : max(v, red[a][w]);
        if (a < 3) {
            if (v != lo_id) atomicMin(&st->bbox_key[a], v);
        } else {
            if (v != hi_id) atomicMax(&st->bbox_key[a], v);
        }
    }
}

struct VoxDev {
    float inv;
    int min_b[3];
    int mul_y, mul_z;
};

__global__ void __launch_bounds__(256)
k_voxel_keys(const float4* __restrict__ pts, unsigned n, VoxDev v, uint32_t* __restrict__ keys,
             uint32_t* __restrict__ vals) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = pts[i];
    int i0 = (int)__fsub_rn(floorf(__fmul_rn(p.x, v.inv)), (float)v.min_b[0]);
    int i1 = (int)__fsub_rn(floorf(__fmul_rn(p.y, v.inv)), (float)v.min_b[1]);
    int i2 = (int)__fsub_rn(floorf(__fmul_rn(p.z, v.inv)), (float)v.min_b[2]);
    keys[i] = (uint32_t)(i0 + i1 * v.mul_y + i2 * v.mul_z);
    vals[i] = i;
}

__global__ void __launch_bounds__(256)
k_heads(const uint32_t* __restrict__ keys, unsigned n, uint32_t* __restrict__ head) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    head[i] = (i < n && (i == 0 || keys[i] != keys[i - 1])) ? 1u : 0u;
}

// position of every segment head, in output order (slot = exclusive scan of the head flags)
__global__ void __launch_bounds__(256)
k_segment_starts(const uint32_t* __restrict__ slot, unsigned n, uint32_t* __restrict__ seg_start) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    if (i == n) {
        seg_start[slot[n]] = n;       // end sentinel
        return;
    }
    if (slot[i + 1] != slot[i]) seg_start[slot[i]] = i;
}

// One warp per voxel.  PCL accumulates a voxel's points one after another in float32 (sorted order, which
// the oracle fixes to ascending point index), so the adds are serial by definition; the warp gathers 32
// points at a time in parallel and replays them in order through shuffles, every lane holding the same sums.
__global__ void __launch_bounds__(256)
k_centroid(const uint32_t* __restrict__ vals, const uint32_t* __restrict__ seg_start, unsigned n_seg,
           const float4* __restrict__ pts, float4* __restrict__ out) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_seg) return;
    const int lane = threadIdx.x & 31;
    const unsigned FULL = 0xffffffffu;
    const unsigned begin = seg_start[w], end = seg_start[w + 1];
    float sx = 0.f, sy = 0.f, sz = 0.f;
    float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
    if (begin + lane < end) nxt = pts[vals[begin + lane]];
    for (unsigned base = begin; base < end; base += 32) {
        const float4 cur = nxt;
        if (base + 32 + lane < end) nxt = pts[vals[base + 32 + lane]];     // in flight while this batch is summed
        const int cnt = (int)min(32u, end - base);
        for (int i = 0; i < cnt; ++i) {
            sx = __fadd_rn(sx, __shfl_sync(FULL, cur.x, i));
            sy = __fadd_rn(sy, __shfl_sync(FULL, cur.y, i));
            sz = __fadd_rn(sz, __shfl_sync(FULL, cur.z, i));
        }
    }
    if (lane == 0) {
        const float cnt = (float)(end - begin);
        out[w] = make_float4(__fdiv_rn(sx, cnt), __fdiv_rn(sy, cnt), __fdiv_rn(sz, cnt), 1.0f);
    }
}

float key_to_float(int k) {
    int b = k >= 0 ? k : k ^ 0x7fffffff;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

}  // namespace
}  // namespace plc

using namespace plc;

extern "C" plc_status plc_filter(plc_ctx* ctx, const plc_cloud_view* in, const plc_crop_box* box, double leaf,
                                 void* out_xyz16, uint64_t capacity, uint64_t* n_out, uint64_t* n_cropped) {
    PLC_RANGE();
    if (!ctx || !cloud_view_ok(in) || !box || !n_out || (capacity && !out_xyz16)) return PLC_E_INVALID_ARG;
    if (in->n_points > 0x7ffffff0ull) return PLC_E_INVALID_ARG;
    *n_out = 0;
    if (n_cropped) *n_cropped = 0;
    const unsigned n = (unsigned)in->n_points;
    if (n == 0) return PLC_OK;
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const size_t bytes = (size_t)n * in->point_step;
    PLC_CUDA_TRY(ctx, ctx->f_flags.reserve(((size_t)n + 1) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_tmp.reserve(scan_tmp_bytes((size_t)n + 1) + scan_tmp_bytes(sort_table_entries(n)) + 64));
    PLC_CUDA_TRY(ctx, ctx->f_out.reserve((size_t)n * sizeof(float4) * 2));   // [cropped | centroids]
    PLC_CUDA_TRY(ctx, ctx->scratch.reserve(4096));
    const unsigned char* raw = nullptr;
    PLC_TRY(stage_bytes(ctx, in->data, bytes, &raw));

    const bool aligned = ((in->point_step | in->off_x | in->off_y | in->off_z) & 3u) == 0;
    const CropDev cb{box->x_min, box->x_max, box->y_min, box->y_max, box->z_min, box->z_max};
    uint32_t* flag = ctx->f_flags.as<uint32_t>();
    uint32_t* scan_tmp = ctx->f_tmp.as<uint32_t>();
    float4* cropped = ctx->f_out.as<float4>();
    float4* centroids = cropped + n;
    FilterState* st = ctx->scratch.as<FilterState>();
    const unsigned blocks1 = (n + 1 + 255) / 256;

    ctx->timer.reset();
    ctx->times.filter_ms = 0.f;
    ctx->timer.begin(s);
    k_filter_state_init<<<1, 1, 0, s>>>(st);
    k_crop_flag<<<blocks1, 256, 0, s>>>(raw, n, in->point_step, in->off_x, in->off_y,
                                        in->off_z, aligned, cb, flag);
    ctx->launches += scan_exclusive_u32(flag, (size_t)n + 1, scan_tmp, s);
    k_crop_compact<<<(n + 255) / 256, 256, 0, s>>>(raw, n, in->point_step, in->off_x,
                                                   in->off_y, in->off_z, aligned, flag, cropped, st);
    ctx->launches += 3;
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());

    struct HostState { FilterState st; uint32_t total; };
    HostState* host = static_cast<HostState*>(ctx->pinned);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(&host->st, st, sizeof(FilterState), cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(&host->total, flag + n, 4, cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const unsigned m = host->total;
    ctx->times.filter_ms = ctx->timer.total_ms();
    if (n_cropped) *n_cropped = m;
    if (m == 0) return PLC_OK;                      // downsampleCloud returns the empty input (:125-127, Q7)

    auto copy_out = [&](const float4* src, unsigned count) -> plc_status {
        if (count > capacity) return PLC_E_CAPACITY;
        PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_xyz16, src, (size_t)count * sizeof(float4), cudaMemcpyDefault, s));
        PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
        *n_out = count;
        return PLC_OK;
    };
    if (!(leaf > 0)) return copy_out(cropped, m);

    // VoxelGrid geometry on the host, float32 exactly as PCL forms it
    const float leaf_f = static_cast<float>(leaf);
    const float inv = 1.0f / leaf_f;
    float mn[3], mx[3];
    for (int a = 0; a < 3; ++a) {
        mn[a] = key_to_float(host->st.bbox_key[a]);
        mx[a] = key_to_float(host->st.bbox_key[3 + a]);
    }
    int64_t ext[3];
    for (int a = 0; a < 3; ++a) ext[a] = static_cast<int64_t>((mx[a] - mn[a]) * inv) + 1;
    if (ext[0] * ext[1] * ext[2] > static_cast<int64_t>(std::numeric_limits<int32_t>::max()))
        return copy_out(cropped, m);                // PCL warns "leaf size too small" and returns the input
    VoxDev vd{};
    vd.inv = inv;
    int div_b[3];
    for (int a = 0; a < 3; ++a) {
        vd.min_b[a] = static_cast<int>(std::floor(mn[a] * inv));
        int max_b = static_cast<int>(std::floor(mx[a] * inv));
        div_b[a] = max_b - vd.min_b[a] + 1;
    }
    vd.mul_y = div_b[0];
    vd.mul_z = div_b[0] * div_b[1];
    uint64_t n_keys = (uint64_t)div_b[0] * div_b[1] * div_b[2];
    int key_bits = 1;
    while (key_bits < 32 && (1ull << key_bits) < n_keys) ++key_bits;

    PLC_CUDA_TRY(ctx, ctx->f_keys.reserve(((size_t)m + 1) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_vals.reserve((size_t)m * 4));
    PLC_CUDA_TRY(ctx, ctx->f_keys2.reserve(((size_t)m + 1) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_vals2.reserve((size_t)m * 4 + sort_table_entries(m) * 4));
    uint32_t* table = ctx->f_vals2.as<uint32_t>() + m;
    const unsigned mb = (m + 255) / 256;
    ctx->timer.begin(s);
    k_voxel_keys<<<mb, 256, 0, s>>>(cropped, m, vd, ctx->f_keys.as<uint32_t>(), ctx->f_vals.as<uint32_t>());
    ctx->launches += 1;
    int where = radix_sort_pairs(ctx->f_keys.as<uint32_t>(), ctx->f_vals.as<uint32_t>(), ctx->f_keys2.as<uint32_t>(),
                                 ctx->f_vals2.as<uint32_t>(), m, key_bits, table, scan_tmp, s, &ctx->launches);
    const uint32_t* skeys = where ? ctx->f_keys2.as<uint32_t>() : ctx->f_keys.as<uint32_t>();
    const uint32_t* svals = where ? ctx->f_vals2.as<uint32_t>() : ctx->f_vals.as<uint32_t>();
    k_heads<<<(m + 1 + 255) / 256, 256, 0, s>>>(skeys, m, flag);
    ctx->launches += scan_exclusive_u32(flag, (size_t)m + 1, scan_tmp, s);
    // the unsorted key buffer is free again: it holds the segment starts (at most m + 1 entries)
    uint32_t* seg_start = where ? ctx->f_keys.as<uint32_t>() : ctx->f_keys2.as<uint32_t>();
    k_segment_starts<<<(m + 1 + 255) / 256, 256, 0, s>>>(flag, m, seg_start);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(&host->total, flag + m, 4, cudaMemcpyDeviceToHost, s));
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const unsigned n_seg = host->total;
    ctx->timer.begin(s);
    k_centroid<<<(n_seg * 32 + 255) / 256, 256, 0, s>>>(svals, seg_start, n_seg, cropped, centroids);
    ctx->launches += 3;
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    plc_status rc = copy_out(centroids, n_seg);           // synchronises the stream
    ctx->times.filter_ms = ctx->timer.total_ms();
    return rc;
}
