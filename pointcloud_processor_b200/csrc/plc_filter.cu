// plc_filter.cu — K6 + K7: crop box and voxel-grid centroid downsample of one LiDAR scan.
//
// Replaces processCloudSimple's two stages in /root/reference/src/pointcloud_filter.cpp:
//   cropFrontArea   87-120   strict six-sided box test, input order kept        -> k_crop_onepass (test, ordered
//                                                                                  compaction, bounds, voxel keys)
//   downsampleCloud 122-139  pcl::VoxelGrid<PointXYZ> (leaf from the parameter) -> radix sort + k_heads_onepass +
//                                                                                  k_centroid
// The whole call is one stream of launches whose counts (cropped points, voxels) stay on the device; the host
// reads them once at the end.
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

__device__ __forceinline__ int order_key(float f) {
    int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}

// Counts and bounds the pipeline leaves on the device; the host reads them once, at the end.
struct FilterState {
    int bbox_key[6];       // min/max of the cropped points as order-preserving keys (VoxelGrid's getMinMax3D)
    unsigned m;            // points that passed the crop
    unsigned n_seg;        // voxels
};

__global__ void k_filter_state_init(FilterState* st) {
    for (int a = 0; a < 3; ++a) {
        st->bbox_key[a] = 0x7fffffff;
        st->bbox_key[3 + a] = (int)0x80000000;
    }
    st->m = 0;
    st->n_seg = 0;
}

// Voxel key of a point.  PCL's index  (i - min_b.x) + (j - min_b.y) dx + (k - min_b.z) dx dy  needs the minimum of
// the whole cropped cloud; but its ORDER is the lexicographic order of (k, j, i), which any fixed bias gives as
// well.  With the bias and extents taken from the crop box the key is known while cropping (`fixed` = 1); a box too
// large for 31 bits of key falls back to PCL's own formula after the bounds are known (k_voxel_keys).
struct VoxDev {
    float inv;
    int bias[3];
    int mul_y, mul_z;
    int fixed;
};

__device__ __forceinline__ uint32_t voxel_key(const VoxDev& v, float x, float y, float z) {
    // floor(p * inverse_leaf) - min_b in float32, then truncated, exactly as VoxelGrid::applyFilter forms it
    const int i0 = (int)__fsub_rn(floorf(__fmul_rn(x, v.inv)), (float)v.bias[0]);
    const int i1 = (int)__fsub_rn(floorf(__fmul_rn(y, v.inv)), (float)v.bias[1]);
    const int i2 = (int)__fsub_rn(floorf(__fmul_rn(z, v.inv)), (float)v.bias[2]);
    return (uint32_t)(i0 + i1 * v.mul_y + i2 * v.mul_z);
}

#ifndef PLC_CROP_ROUNDS
#define PLC_CROP_ROUNDS 16
#endif
constexpr int kCropRounds = PLC_CROP_ROUNDS;      // points per thread of k_crop_onepass (a tile is 256 x this many points)
constexpr int kCropTile = 256 * kCropRounds;
constexpr int kHeadRounds = 8;
constexpr int kHeadTile = 256 * kHeadRounds;

// K6 in one pass (pointcloud_filter.cpp:106-116): strict six-sided test on the float promoted to double (NaN fails),
// order-preserving compaction through a decoupled look-back over the tiles' counts, the cropped cloud's bounds, and
// — when the key is fixed by the crop box — the voxel key of every kept point.  N*12 read, M'*(16 + 8) written.
__global__ void __launch_bounds__(256, 8)
k_crop_onepass(const unsigned char* __restrict__ data, unsigned n, unsigned n_tiles, unsigned step, unsigned ox,
               unsigned oy, unsigned oz, bool aligned, bool vec16, CropDev b, VoxDev vox, float4* __restrict__ out,
               uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, FilterState* __restrict__ st, LookbackArgs lb) {
    __shared__ unsigned wcount[kCropRounds * 8];      // kept points per (round, warp), then their exclusive prefix
    __shared__ unsigned wkeep[kCropRounds * 8];       // which lanes of (round, warp) kept their point
    __shared__ unsigned s_excl, s_total;
    __shared__ int red[6][8];
    const unsigned tile = lookback_tile(lb);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int lo_id = 0x7fffffff, hi_id = (int)0x80000000;
    int k[6] = {lo_id, lo_id, lo_id, hi_id, hi_id, hi_id};
    // One record's coordinates.  The tile is read twice — here for the test, after the look-back for the copy (an
    // L2 hit by then) — so that no thread holds eight points in registers across the two barriers: half the
    // registers, twice the resident blocks, and it is the resident blocks that keep HBM busy while others wait.
    auto load_xyz = [&](unsigned i, float& x, float& y, float& z) {
        const unsigned char* p = data + (size_t)i * step;
        if (vec16) {              // x, y, z in one 16-byte word of the record (PointXYZ, PointXYZRGB): one load
            const float4 v = *reinterpret_cast<const float4*>(p + ox);
            x = v.x;
            y = v.y;
            z = v.z;
        } else {
            x = ld_f32(p + ox, aligned);
            y = ld_f32(p + oy, aligned);
            z = ld_f32(p + oz, aligned);
        }
    };
#pragma unroll
    for (int r = 0; r < kCropRounds; ++r) {
        const unsigned i = tile * kCropTile + r * 256 + threadIdx.x;
        bool keep = false;
        float x = 0.f, y = 0.f, z = 0.f;
        if (i < n) {
            load_xyz(i, x, y, z);
            const double dx = x, dy = y, dz = z;
            keep = dx > b.x_min && dx < b.x_max && dy > b.y_min && dy < b.y_max && dz > b.z_min && dz < b.z_max;
        }
        const unsigned bal = __ballot_sync(FULL, keep);
        if (lane == 0) {
            wcount[r * 8 + warp] = __popc(bal);
            wkeep[r * 8 + warp] = bal;
        }
        if (keep) {
            k[0] = min(k[0], order_key(x));
            k[1] = min(k[1], order_key(y));
            k[2] = min(k[2], order_key(z));
            k[3] = max(k[3], order_key(x));
            k[4] = max(k[4], order_key(y));
            k[5] = max(k[5], order_key(z));
        }
    }
    __syncthreads();
    if (warp == 0) {
        // exclusive prefix over the (round, warp) counts, kPer consecutive ones per lane, then the tile's place
        // among the tiles
        constexpr int kPer = kCropRounds * 8 / 32;
        unsigned c[kPer], sum = 0;
#pragma unroll
        for (int q = 0; q < kPer; ++q) {
            c[q] = wcount[kPer * lane + q];
            sum += c[q];
        }
        unsigned inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += t;
        }
        const unsigned total = __shfl_sync(FULL, inc, 31);
        unsigned run = inc - sum;
#pragma unroll
        for (int q = 0; q < kPer; ++q) {
            wcount[kPer * lane + q] = run;
            run += c[q];
        }
        const unsigned e = lookback_exclusive(lb, tile, total);
        if (lane == 0) {
            s_excl = e;
            s_total = total;
        }
    }
    __syncthreads();
    const unsigned base = s_excl;
#pragma unroll
    for (int r = 0; r < kCropRounds; ++r) {
        const unsigned bal = wkeep[r * 8 + warp];
        if (!((bal >> lane) & 1u)) continue;
        float x, y, z;
        load_xyz(tile * kCropTile + r * 256 + threadIdx.x, x, y, z);
        const unsigned pos = base + wcount[r * 8 + warp] + __popc(bal & lt);
        out[pos] = make_float4(x, y, z, 1.0f);          // pcl::PointXYZ keeps 1.0f in data[3]
        if (vox.fixed) {
            keys[pos] = voxel_key(vox, x, y, z);
            vals[pos] = pos;
        }
    }
    // bounds: warp, block, then one set of atomics per block that kept anything
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        k[a] = __reduce_min_sync(FULL, k[a]);
        k[3 + a] = __reduce_max_sync(FULL, k[3 + a]);
    }
    if (lane < 6) {
        int mine = k[0];
#pragma unroll
        for (int a = 1; a < 6; ++a) mine = lane == a ? k[a] : mine;
        red[lane][warp] = mine;
    }
    __syncthreads();
    if (threadIdx.x < 6 && s_total) {
        const int a = threadIdx.x;
        int v = red[a][0];
        for (int w = 1; w < 8; ++w) v = a < 3 ? min(v, red[a][w]) : max(v, red[a][w]);
        if (a < 3) atomicMin(&st->bbox_key[a], v);
        else atomicMax(&st->bbox_key[a], v);
    }
    if (threadIdx.x == 0 && tile == n_tiles - 1) st->m = base + s_total;
}

// PCL's own key, once the cropped cloud's minimum is known (crop boxes too large for a fixed key)
__global__ void __launch_bounds__(256)
k_voxel_keys(const float4* __restrict__ pts, unsigned n, VoxDev v, uint32_t* __restrict__ keys,
             uint32_t* __restrict__ vals) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = pts[i];
    keys[i] = voxel_key(v, p.x, p.y, p.z);
    vals[i] = i;
}

// First element of every run of equal keys, in order: seg_start[j] = index of the j-th voxel's first point,
// seg_start[n_seg] = m.  One pass: head flags, look-back over the tiles' head counts, scatter.
__global__ void __launch_bounds__(256, 8)
k_heads_onepass(const uint32_t* __restrict__ keys, unsigned n_max, unsigned n_tiles, uint32_t* __restrict__ seg_start,
                FilterState* __restrict__ st, LookbackArgs lb) {
    __shared__ unsigned wcount[kHeadRounds * 8];
    __shared__ unsigned s_excl, s_total;
    const unsigned tile = lookback_tile(lb);
    const unsigned m = min(n_max, st->m);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    unsigned head_bits = 0, rank[kHeadRounds];
#pragma unroll
    for (int r = 0; r < kHeadRounds; ++r) {
        const unsigned i = tile * kHeadTile + r * 256 + threadIdx.x;
        const bool head = i < m && (i == 0 || keys[i] != keys[i - 1]);
        const unsigned bal = __ballot_sync(FULL, head);
        rank[r] = __popc(bal & lt);
        if (lane == 0) wcount[r * 8 + warp] = __popc(bal);
        if (head) head_bits |= 1u << r;
    }
    __syncthreads();
    if (warp == 0) {
        const unsigned a0 = wcount[2 * lane], a1 = wcount[2 * lane + 1];
        unsigned inc = a0 + a1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += t;
        }
        const unsigned total = __shfl_sync(FULL, inc, 31);
        wcount[2 * lane] = inc - a0 - a1;
        wcount[2 * lane + 1] = inc - a1;
        const unsigned e = lookback_exclusive(lb, tile, total);
        if (lane == 0) {
            s_excl = e;
            s_total = total;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kHeadRounds; ++r)
        if ((head_bits >> r) & 1u) seg_start[s_excl + wcount[r * 8 + warp] + rank[r]] = tile * kHeadTile + r * 256 + threadIdx.x;
    if (threadIdx.x == 0 && tile == n_tiles - 1) {
        st->n_seg = s_excl + s_total;
        seg_start[s_excl + s_total] = m;          // end sentinel
    }
}

// One warp per voxel (warps stride over the voxels: their number is only known on the device).  PCL accumulates a
// voxel's points one after another in float32 (sorted order, which the oracle fixes to ascending point index), so
// the adds are serial by definition; the warp gathers 32 points at a time in parallel and replays them in order
// through shuffles, every lane holding the same sums.
__global__ void __launch_bounds__(256)
k_centroid(const uint32_t* __restrict__ vals, const uint32_t* __restrict__ seg_start,
           const FilterState* __restrict__ st, const float4* __restrict__ pts, float4* __restrict__ out) {
    const unsigned n_seg = st->n_seg;
    const unsigned n_warps = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    const unsigned FULL = 0xffffffffu;
    for (unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_seg; w += n_warps) {
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
    PLC_CUDA_TRY(ctx, ctx->f_out.reserve((size_t)n * sizeof(float4) * 2));   // [cropped | centroids]
    PLC_CUDA_TRY(ctx, ctx->f_keys.reserve(((size_t)n + 1) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_vals.reserve((size_t)n * 4));
    PLC_CUDA_TRY(ctx, ctx->f_keys2.reserve(((size_t)n + 1) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_vals2.reserve((size_t)n * 4 + sort_table_entries(n) * 4));
    PLC_CUDA_TRY(ctx, ctx->f_flags.reserve(((size_t)n + 2) * 4));            // segment starts
    PLC_CUDA_TRY(ctx, ctx->scratch.reserve(4096));
    const unsigned char* raw = nullptr;
    PLC_TRY(stage_bytes(ctx, in->data, bytes, &raw));

    const bool aligned = ((in->point_step | in->off_x | in->off_y | in->off_z) & 3u) == 0;
    const CropDev cb{box->x_min, box->x_max, box->y_min, box->y_max, box->z_min, box->z_max};
    float4* cropped = ctx->f_out.as<float4>();
    float4* centroids = cropped + n;
    uint32_t* seg_start = ctx->f_flags.as<uint32_t>();
    uint32_t* table = ctx->f_vals2.as<uint32_t>() + n;
    FilterState* st = ctx->scratch.as<FilterState>();
    const unsigned n_tiles = (n + kCropTile - 1) / kCropTile;

    // VoxelGrid geometry that does not depend on the data: float32 inverse leaf as PCL forms it, and — if the crop
    // box is small enough — a fixed key layout from the box (bias two voxels below its minimum)
    const bool downsample = leaf > 0;
    const float leaf_f = static_cast<float>(leaf);
    const float inv = downsample ? 1.0f / leaf_f : 0.0f;
    VoxDev vd{};
    vd.inv = inv;
    int key_bits = 0;
    if (downsample) {
        const double lo[3] = {box->x_min, box->y_min, box->z_min}, hi[3] = {box->x_max, box->y_max, box->z_max};
        double ext[3];
        bool ok = std::isfinite((double)inv);
        for (int a = 0; a < 3 && ok; ++a) {
            const double l = std::floor(lo[a] * (double)inv) - 2.0, h = std::floor(hi[a] * (double)inv) + 2.0;
            ok = std::isfinite(l) && std::isfinite(h) && std::fabs(l) < 8.0e6 && std::fabs(h) < 8.0e6 && h >= l;
            vd.bias[a] = ok ? (int)l : 0;
            ext[a] = h - l + 1.0;
        }
        if (ok && ext[0] * ext[1] * ext[2] < 2147483648.0) {
            vd.mul_y = (int)ext[0];
            vd.mul_z = (int)(ext[0] * ext[1]);
            vd.fixed = 1;
            const uint64_t n_keys = (uint64_t)(ext[0] * ext[1] * ext[2]);
            key_bits = 1;
            while (key_bits < 32 && (1ull << key_bits) < n_keys) ++key_bits;
        }
    }

    auto sort_and_reduce = [&](unsigned count_max, const unsigned* count_dev) -> plc_status {
        int where = radix_sort_pairs(ctx->f_keys.as<uint32_t>(), ctx->f_vals.as<uint32_t>(), ctx->f_keys2.as<uint32_t>(),
                                     ctx->f_vals2.as<uint32_t>(), count_max, key_bits, table, ctx->lb, s, &ctx->launches,
                                     count_dev);
        const uint32_t* skeys = where ? ctx->f_keys2.as<uint32_t>() : ctx->f_keys.as<uint32_t>();
        const uint32_t* svals = where ? ctx->f_vals2.as<uint32_t>() : ctx->f_vals.as<uint32_t>();
        const unsigned tiles = (count_max + kHeadTile - 1) / kHeadTile;
        LookbackArgs la{ctx->lb.status, ctx->lb.ticket, 0, 0};
        ctx->lb.begin(tiles, &la.ticket_base, &la.epoch);
        k_heads_onepass<<<tiles, 256, 0, s>>>(skeys, count_max, tiles, seg_start, st, la);
        k_centroid<<<148 * 4, 256, 0, s>>>(svals, seg_start, st, cropped, centroids);
        ctx->launches += 2;
        return PLC_OK;
    };

    ctx->timer.reset();
    ctx->times.filter_ms = 0.f;
    ctx->timer.begin(s);
    k_filter_state_init<<<1, 1, 0, s>>>(st);
    {
        LookbackArgs la{ctx->lb.status, ctx->lb.ticket, 0, 0};
        ctx->lb.begin(n_tiles, &la.ticket_base, &la.epoch);
        const bool vec16 = in->point_step % 16 == 0 && in->off_x % 16 == 0 && in->off_y == in->off_x + 4 &&
                           in->off_z == in->off_x + 8 && reinterpret_cast<uintptr_t>(raw) % 16 == 0;
        k_crop_onepass<<<n_tiles, 256, 0, s>>>(raw, n, n_tiles, in->point_step, in->off_x, in->off_y, in->off_z, aligned,
                                               vec16, cb, vd, cropped, ctx->f_keys.as<uint32_t>(), ctx->f_vals.as<uint32_t>(),
                                               st, la);
    }
    ctx->launches += 2;
    if (vd.fixed) PLC_TRY(sort_and_reduce(n, &st->m));       // the whole pipeline without coming back to the host
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());

    FilterState* host = static_cast<FilterState*>(ctx->pinned);
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, st, sizeof(FilterState), cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const unsigned m = host->m;
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
    if (!downsample) return copy_out(cropped, m);

    // PCL's own geometry from the cropped cloud's bounds, float32 exactly as VoxelGrid forms it
    float mn[3], mx[3];
    for (int a = 0; a < 3; ++a) {
        mn[a] = key_to_float(host->bbox_key[a]);
        mx[a] = key_to_float(host->bbox_key[3 + a]);
    }
    int64_t ext[3];
    for (int a = 0; a < 3; ++a) ext[a] = static_cast<int64_t>((mx[a] - mn[a]) * inv) + 1;
    if (ext[0] * ext[1] * ext[2] > static_cast<int64_t>(std::numeric_limits<int32_t>::max()))
        return copy_out(cropped, m);                // PCL warns "leaf size too small" and returns the input
    if (vd.fixed) return copy_out(centroids, host->n_seg);

    // crop box too large for a fixed key: PCL's formula from the bounds just read, then the same sort and reduction
    int div_b[3];
    for (int a = 0; a < 3; ++a) {
        vd.bias[a] = static_cast<int>(std::floor(mn[a] * inv));
        int max_b = static_cast<int>(std::floor(mx[a] * inv));
        div_b[a] = max_b - vd.bias[a] + 1;
    }
    vd.mul_y = div_b[0];
    vd.mul_z = div_b[0] * div_b[1];
    const uint64_t n_keys = (uint64_t)div_b[0] * div_b[1] * div_b[2];
    key_bits = 1;
    while (key_bits < 32 && (1ull << key_bits) < n_keys) ++key_bits;
    ctx->timer.begin(s);
    k_voxel_keys<<<(m + 255) / 256, 256, 0, s>>>(cropped, m, vd, ctx->f_keys.as<uint32_t>(), ctx->f_vals.as<uint32_t>());
    ctx->launches += 1;
    PLC_TRY(sort_and_reduce(m, nullptr));
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(host, st, sizeof(FilterState), cudaMemcpyDeviceToHost, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    plc_status rc = copy_out(centroids, host->n_seg);
    ctx->times.filter_ms = ctx->timer.total_ms();
    return rc;
}
