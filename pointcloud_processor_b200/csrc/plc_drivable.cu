// plc_drivable.cu — the occupancy grid of calc_drivable_area: the adjacent consumer of the
// filtered robot scan, after the same float32 rigid transform the merger applies.
//
// Replaces /root/reference/src/calc_drivable_area.cpp:
//   tf2::doTransform into the map frame          :87-92    -> k_bin_heights (fused: the transformed cloud is never stored)
//   binning of the z values per grid cell        :141-160  -> k_bin_heights (count, min z, max z per cell by atomics)
//   calculateGradient + the occupancy decision   :178-232  -> k_classify
// The reference keeps every z of a cell in a std::vector and takes min/max at the end; min and max are
// order-independent, so integer atomics on order-preserving keys give the same floats.  HBM-bound streams:
// N*point_step bytes read once, W*H*(12 + 1) bytes of cell state.
#include "plc_internal.cuh"

#include <mutex>

// Scratch of plc_drivable_area_static (bottom of this file): the per-cell (count, min z, max z) state lives in this
// array instead of a cudaMalloc'd buffer.  Every GPU that loads the library holds these 12 MiB until exit.  In host
// code the name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, per device.
static __device__ __align__(256) unsigned char g_drivable_static_scratch[(size_t)PLC_DRIVABLE_STATIC_MAX_CELLS * 12u];

namespace plc {
namespace {

struct Affine3 {
    float m[3][3];
    float t[3];
};

__device__ __forceinline__ float ld_f32d(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    unsigned b = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(b);
}
__device__ __forceinline__ int z_key(float f) {
    int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float z_from_key(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

struct CellState {
    unsigned count;
    int zmin, zmax;      // order keys
};

__global__ void __launch_bounds__(256) k_cells_reset(CellState* __restrict__ cells, unsigned n) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cells[i] = CellState{0u, 0x7fffffff, (int)0x80000000};
}

// static_cast<int>(double) of the reference truncates towards zero (so (-1, 0) lands in cell 0, :151-152);
// cvt.rzi saturates where the C++ cast is undefined, which the range test then rejects
__global__ void __launch_bounds__(256)
k_bin_heights(const unsigned char* __restrict__ data, unsigned n, unsigned step, unsigned ox, unsigned oy, unsigned oz,
              bool aligned, Affine3 a, int have_transform, double origin_x, double origin_y, double res, int width, int height,
              CellState* __restrict__ cells) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned char* p = data + (size_t)i * step;
    const float v0 = ld_f32d(p + ox, aligned), v1 = ld_f32d(p + oy, aligned), v2 = ld_f32d(p + oz, aligned);
    float r[3] = {v0, v1, v2};
    if (have_transform) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float acc = __fmul_rn(a.m[k][0], v0);
            acc = __fadd_rn(acc, __fmul_rn(a.m[k][1], v1));
            acc = __fadd_rn(acc, __fmul_rn(a.m[k][2], v2));
            r[k] = __fadd_rn(acc, a.t[k]);
        }
    }
    if (!(isfinite(r[0]) && isfinite(r[1]) && isfinite(r[2]))) return;                   // :145-146
    const int gx = __double2int_rz(__ddiv_rn(__dsub_rn((double)r[0], origin_x), res));   // :151
    const int gy = __double2int_rz(__ddiv_rn(__dsub_rn((double)r[1], origin_y), res));   // :152
    if (gx < 0 || gx >= width || gy < 0 || gy >= height) return;                         // :155
    CellState* c = cells + (size_t)gy * width + gx;
    atomicAdd(&c->count, 1u);
    const int key = z_key(r[2]);
    atomicMin(&c->zmin, key);
    atomicMax(&c->zmax, key);
}

__global__ void __launch_bounds__(256)
k_classify(const CellState* __restrict__ cells, int width, int height, double origin_x, double origin_y, double res,
           double start_x, double start_y, double clear_radius, double max_gradient, int min_points,
           signed char* __restrict__ grid) {
    const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (unsigned)width * (unsigned)height) return;
    const int y = (int)(idx / (unsigned)width), x = (int)(idx % (unsigned)width);
    const double cx = __dadd_rn(origin_x, __dmul_rn(__dadd_rn((double)x, 0.5), res));     // :185
    const double cy = __dadd_rn(origin_y, __dmul_rn(__dadd_rn((double)y, 0.5), res));     // :186
    const double ex = __dsub_rn(cx, start_x), ey = __dsub_rn(cy, start_y);
    const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)));      // :189-191
    const CellState c = cells[idx];
    signed char v;
    if (dist <= clear_radius) {
        v = 0;                                                                            // :194-197
    } else if (c.count == 0 || (unsigned long long)c.count < (unsigned long long)(long long)min_points) {
        // size() < static_cast<size_t>(min_points): a negative parameter wraps to a huge count (:198-199)
        v = -1;
    } else {
        float gradient = 0.0f;                                                            // calculateGradient, :238-250
        if (c.count >= 2)
            gradient = (float)__ddiv_rn((double)__fsub_rn(z_from_key(c.zmax), z_from_key(c.zmin)), res);
        v = ((double)gradient > max_gradient) ? 100 : 0;                                  // :209-218
    }
    grid[idx] = v;
}

// Eigen::Quaternionf::toRotationMatrix in float32, translation narrowed to float (tf2_sensor_msgs doTransform)
Affine3 affine_of(const double* t, const double* q) {
    const float qx = (float)q[0], qy = (float)q[1], qz = (float)q[2], qw = (float)q[3];
    const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
    const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    Affine3 a;
    a.m[0][0] = 1.0f - (tyy + tzz); a.m[0][1] = txy - twz;          a.m[0][2] = txz + twy;
    a.m[1][0] = txy + twz;          a.m[1][1] = 1.0f - (txx + tzz); a.m[1][2] = tyz - twx;
    a.m[2][0] = txz - twy;          a.m[2][1] = tyz + twx;          a.m[2][2] = 1.0f - (txx + tyy);
    a.t[0] = (float)t[0]; a.t[1] = (float)t[1]; a.t[2] = (float)t[2];
    return a;
}

}  // namespace
}  // namespace plc

using namespace plc;

extern "C" void plc_drivable_params_default(plc_drivable_params* p) {
    if (!p) return;
    p->grid_resolution = 1.0;          // calc_drivable_area.cpp:21-26
    p->map_width = 100.0;
    p->map_height = 100.0;
    p->max_gradient = 0.3;
    p->start_clear_radius = 3.0;
    p->min_points_per_cell = 10;
    p->reserved = 0;
}

extern "C" plc_status plc_drivable_area(plc_ctx* ctx, const plc_cloud_view* scan, const double* t, const double* q,
                                        double robot_x, double robot_y, double start_x, double start_y,
                                        const plc_drivable_params* prm, int8_t* out_grid, uint64_t capacity,
                                        plc_drivable_info* info) {
    PLC_RANGE();
    if (!ctx || !cloud_view_ok(scan) || !prm || !info || (t == nullptr) != (q == nullptr)) return PLC_E_INVALID_ARG;
    if (!(prm->grid_resolution > 0) || !(prm->map_width >= 0) || !(prm->map_height >= 0)) return PLC_E_INVALID_ARG;
    if (scan->n_points > 0x7fffff00ull) return PLC_E_INVALID_ARG;
    const double wd = prm->map_width / prm->grid_resolution, hd = prm->map_height / prm->grid_resolution;
    if (!(wd < 46000.0 && hd < 46000.0)) return PLC_E_INVALID_ARG;      // width*height must fit 31 bits
    memset(info, 0, sizeof(*info));
    info->width = (int32_t)wd;                                          // :36-37
    info->height = (int32_t)hd;
    info->origin_x = robot_x - prm->map_width / 2.0;                    // :137-138
    info->origin_y = robot_y - prm->map_height / 2.0;
    info->resolution = prm->grid_resolution;
    const uint64_t n_cells = (uint64_t)info->width * (uint64_t)info->height;
    if (scan->n_points == 0) return PLC_OK;                             // "Received empty point cloud": nothing published (:108-112)
    if (n_cells > capacity || (n_cells && !out_grid)) {
        info->n_cells = n_cells;
        return PLC_E_CAPACITY;
    }
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const size_t n = scan->n_points, bytes = n * (size_t)scan->point_step;
    PLC_CUDA_TRY(ctx, ctx->su_flags.reserve(n_cells * sizeof(CellState) + 16));
    PLC_CUDA_TRY(ctx, ctx->su_out.reserve(n_cells + 16));
    const unsigned char* raw = nullptr;
    PLC_TRY(stage_bytes(ctx, scan->data, bytes, &raw));
    CellState* cells = ctx->su_flags.as<CellState>();
    signed char* grid = ctx->su_out.as<signed char>();
    const bool aligned = ((scan->point_step | scan->off_x | scan->off_y | scan->off_z) & 3u) == 0;
    Affine3 a{};
    if (t) a = affine_of(t, q);
    ctx->timer.reset();
    ctx->timer.begin(s);
    if (n_cells) {
        k_cells_reset<<<(unsigned)((n_cells + 255) / 256), 256, 0, s>>>(cells, (unsigned)n_cells);
        k_bin_heights<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(raw, (unsigned)n, scan->point_step,
                                                                   scan->off_x, scan->off_y, scan->off_z, aligned, a, t ? 1 : 0,
                                                                   info->origin_x, info->origin_y, prm->grid_resolution,
                                                                   info->width, info->height, cells);
        k_classify<<<(unsigned)((n_cells + 255) / 256), 256, 0, s>>>(cells, info->width, info->height, info->origin_x,
                                                                    info->origin_y, prm->grid_resolution, start_x, start_y,
                                                                    prm->start_clear_radius, prm->max_gradient,
                                                                    prm->min_points_per_cell, grid);
        ctx->launches += 3;
    }
    ctx->timer.end(s);
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    if (n_cells) PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_grid, grid, n_cells, cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->times.drivable_ms = ctx->timer.total_ms();
    info->n_cells = n_cells;
    info->published = 1;
    return PLC_OK;
}

// ---- the allocation-free variant ------------------------------------------------------------------------------------
namespace plc {
namespace {

static_assert(sizeof(CellState) == 12, "g_drivable_static_scratch is sized for 12-byte cell states");
constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mu;
CellState* g_static_cells[kMaxStaticDevices] = {};       // device address of the scratch array, per device

// Device address of the static scratch on the context's device (which must be current): resolved once per device.
plc_status static_cells(plc_ctx* ctx, CellState** out) {
    if (ctx->device < 0 || ctx->device >= kMaxStaticDevices) {
        ctx->set_error("plc_drivable_area_static", "device ordinal beyond the per-device address table (64)");
        return PLC_E_INVALID_ARG;
    }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_cells[ctx->device]) {
        void* p = nullptr;
        PLC_CUDA_TRY(ctx, cudaGetSymbolAddress(&p, g_drivable_static_scratch));
        g_static_cells[ctx->device] = static_cast<CellState*>(p);
    }
    *out = g_static_cells[ctx->device];
    return PLC_OK;
}

// true when `p` is memory kernels on `device` may dereference without any staging copy
bool device_resident(const void* p, int device) {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    if (attr.type == cudaMemoryTypeManaged) return true;
    return attr.type == cudaMemoryTypeDevice && attr.device == device;
}

}  // namespace
}  // namespace plc

extern "C" plc_status plc_drivable_area_static(plc_ctx* ctx, const plc_cloud_view* scan, const double* t, const double* q,
                                               double robot_x, double robot_y, double start_x, double start_y,
                                               const plc_drivable_params* prm, int8_t* out_grid, uint64_t capacity,
                                               plc_drivable_info* info) {
    if (!ctx || !cloud_view_ok(scan) || !prm || !info || (t == nullptr) != (q == nullptr)) return PLC_E_INVALID_ARG;
    if (!(prm->grid_resolution > 0) || !(prm->map_width >= 0) || !(prm->map_height >= 0)) return PLC_E_INVALID_ARG;
    if (scan->n_points > 0x7fffff00ull) return PLC_E_INVALID_ARG;
    const double wd = prm->map_width / prm->grid_resolution, hd = prm->map_height / prm->grid_resolution;
    if (!(wd < 46000.0 && hd < 46000.0)) return PLC_E_INVALID_ARG;
    memset(info, 0, sizeof(*info));
    info->width = (int32_t)wd;
    info->height = (int32_t)hd;
    info->origin_x = robot_x - prm->map_width / 2.0;
    info->origin_y = robot_y - prm->map_height / 2.0;
    info->resolution = prm->grid_resolution;
    const uint64_t n_cells = (uint64_t)info->width * (uint64_t)info->height;
    if (scan->n_points == 0) return PLC_OK;
    info->n_cells = n_cells;
    if (n_cells > (uint64_t)PLC_DRIVABLE_STATIC_MAX_CELLS) {
        ctx->set_error("plc_drivable_area_static",
                       ("grid of " + std::to_string(n_cells) + " cells exceeds the static scratch of " +
                        std::to_string((uint64_t)PLC_DRIVABLE_STATIC_MAX_CELLS) +
                        " cells (PLC_DRIVABLE_STATIC_MAX_CELLS); use plc_drivable_area").c_str());
        return PLC_E_CAPACITY;
    }
    if (n_cells > capacity || (n_cells && !out_grid)) return PLC_E_CAPACITY;
    if (n_cells == 0) {
        info->published = 1;
        return PLC_OK;
    }
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (!device_resident(scan->data, ctx->device) || !device_resident(out_grid, ctx->device)) {
        ctx->set_error("plc_drivable_area_static",
                       "scan->data and out_grid must be device memory of the context's device (no staging buffer is allocated)");
        return PLC_E_INVALID_ARG;
    }
    CellState* cells = nullptr;
    PLC_TRY(static_cells(ctx, &cells));
    cudaStream_t s = ctx->stream;
    const size_t n = scan->n_points;
    const bool aligned = ((scan->point_step | scan->off_x | scan->off_y | scan->off_z) & 3u) == 0 &&
                         (reinterpret_cast<uintptr_t>(scan->data) & 3u) == 0;
    Affine3 a{};
    if (t) a = affine_of(t, q);
    // the array keeps the previous call's cell states: every cell this call reads is reset first
    k_cells_reset<<<(unsigned)((n_cells + 255) / 256), 256, 0, s>>>(cells, (unsigned)n_cells);
    k_bin_heights<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(static_cast<const unsigned char*>(scan->data), (unsigned)n,
                                                               scan->point_step, scan->off_x, scan->off_y, scan->off_z,
                                                               aligned, a, t ? 1 : 0, info->origin_x, info->origin_y,
                                                               prm->grid_resolution, info->width, info->height, cells);
    k_classify<<<(unsigned)((n_cells + 255) / 256), 256, 0, s>>>(cells, info->width, info->height, info->origin_x,
                                                                info->origin_y, prm->grid_resolution, start_x, start_y,
                                                                prm->start_clear_radius, prm->max_gradient,
                                                                prm->min_points_per_cell,
                                                                reinterpret_cast<signed char*>(out_grid));
    ctx->launches += 3;
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    info->published = 1;
    return PLC_OK;
}
