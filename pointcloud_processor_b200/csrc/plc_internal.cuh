// plc_internal.cuh — context, device buffers and the voxel index shared by the plc_* kernels.
// sm_100a only; there is deliberately no host fallback anywhere behind this header.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "plc.h"

// NVTX ranges around the C-ABI entry points (SURVEY §5 tracing row): header-only NVTX 3, a no-op unless a
// profiler injects its library, so timelines show which plc_* call the kernels below belong to.
#include <nvtx3/nvToolsExt.h>
namespace plc {
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
}  // namespace plc
#define PLC_RANGE() plc::NvtxRange plc_nvtx_range_(__func__)
#include "plc_scan.cuh"

#define PLC_CUDA_TRY(ctx, expr)                                                              \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            (ctx)->set_error(#expr, cudaGetErrorString(e__));                                \
            return e__ == cudaErrorMemoryAllocation ? PLC_E_OOM : PLC_E_CUDA;                \
        }                                                                                    \
    } while (0)

#define PLC_TRY(expr)                          \
    do {                                       \
        plc_status s__ = (expr);               \
        if (s__ != PLC_OK) return s__;         \
    } while (0)

namespace plc {

// Grow-only device allocation.
struct DevBuf {
    void* ptr = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&ptr, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const { return static_cast<T*>(ptr); }
};

// Geometry of a voxel index, passed to kernels by value.
//   coarse voxel edge h (a power of two >= 2*reach), mid voxel h/2, fine voxel g = h/8, shared origin.
//   bricks are stored densely, brick b = (cz*dim_y + cy)*dim_x + cx: one 128-byte line per coarse voxel, so a
//   sample's state is ONE load and the samples of neighbouring rays share lines.
//   bricks[32*b + (fz&7)*4 + ((fy&7)>>1)], 2 bits per fine voxel at bit 2*(((fy&1)<<3) | (fx&7)):
//              bit 0  a sample in the fine voxel MAY be within the radius of a point
//                     (conservative dilation of the points by `reach`)
//              bit 1  every sample in the fine voxel IS within the radius of one and the same point
//                     (conservative erosion by `inner`), so no exact test is needed
//   gstart[m] .. gstart[m+1]: candidate list of mid voxel m (x fastest), in groups of four points
//   stored as x4,y4,z4 (cgrp[3*group + 0..2]); it holds every point within reach of the mid voxel's
//   box, so a sample in mid voxel m can only be blocked by those points.  Pad slots are +inf.
constexpr int kFineShift = 3;            // fine voxels per coarse voxel edge = 8
constexpr int kFixedBits = 16;           // fractional bits of the fixed-point fine coordinate of the march

struct VoxelGeom {
    float ox, oy, oz;        // origin (multiple of h, exactly representable)
    float inv_g;             // 1/g (power of two)
    float reach;             // real-number radius + slack used by the builder
    int dim_x, dim_y, dim_z; // coarse dims
    int mdim_x, mdim_y, mdim_z;   // mid-voxel dims (candidate lists)
    int mid_shift;                // fine voxels per mid-voxel edge = 1 << mid_shift (4 at refinement 1, 8 at 2)
    float fdim_xf, fdim_yf, fdim_zf;   // fine dims = 8*coarse, as float for the bounds test
    float clip_lo[3], clip_hi[3];      // box outside which no position is within reach of a point
    float clip_t;                      // slack (m along the ray) the clipped sample range may overshoot by
    const uint32_t* __restrict__ bricks;
    const uint32_t* __restrict__ gstart;
    const float4* __restrict__ cgrp;
    unsigned long long occ_tex;        // PLC_TEX_MARCH experiment: the occupancy rows as a 3-D texture (x/8, y, z) of 16-bit texels
};

struct VoxelIndex {
    // raw finite points (unsorted) kept so the index can be rebuilt for another radius
    DevBuf raw;          // float4[n_raw]
    uint64_t n_raw = 0;
    uint64_t n_input = 0;
    float bbox[6] = {0, 0, 0, 0, 0, 0};   // min xyz, max xyz of finite points
    // built index
    double built_radius = -1.0;   // query radius (m) the structure was built for
    float built_thr = 0.0f;       // float32 squared-distance threshold of the "certain" bit
    double built_march_err = 0.0; // position error of the caller's march covered by the slack: in fine voxels ...
    double built_march_err_m = 0.0;   // ... plus in metres
    int built_refine = 0;         // occupancy refinement the structure was built with (1, 2 or 4) ...
    int asked_refine = 0;         // ... and the one asked for (the size caps may have lowered it)
    double h = 0.0;
    DevBuf bricks, gstart, cfill, cgrp, scan_tmp;
    VoxelGeom geom{};
    uint64_t n_vox = 0;       // bricks
    uint64_t n_mid = 0;       // list voxels
    uint64_t n_groups = 0;    // 4-point groups in cgrp
    // `bricks` holds zeros everywhere except in the bricks of the list voxels that (geom, gstart, n_mid) of the
    // last build describe as non-empty: the next build clears just those (k_clear_touched) instead of the whole array
    bool bricks_sparse_dirty = false;
    // PLC_TEX_MARCH experiment: a 3-D array copy of the occupancy rows for k_sweep's march
    cudaArray_t occ_array = nullptr;
    cudaTextureObject_t occ_tex = 0;
    cudaSurfaceObject_t occ_surf = 0;
    int occ_dims[3] = {0, 0, 0};
    bool has_points() const { return n_raw > 0; }
    // the "G" term of the algorithmic bytes: the indexed points, once (a lower bound of what the sweep must read)
    uint64_t bytes() const { return n_raw * 16; }
    // what is actually resident (candidate lists replicate a point into every mid voxel in reach)
    uint64_t resident_bytes() const { return n_vox * 128 + (n_mid + 1) * 4 + n_groups * 48; }
    void release() {
        raw.release(); bricks.release(); gstart.release(); cfill.release(); cgrp.release();
        scan_tmp.release();
        bricks_sparse_dirty = false;
    }
};

// Device time of one C-ABI call's kernels: event pairs on the launching stream around each stretch of kernels
// (host<->device copies and host synchronisation points stay outside), summed once the call has synchronised.
struct DevTimer {
    static constexpr int kPairs = 8;
    cudaEvent_t a[kPairs] = {}, b[kPairs] = {};
    int n = 0;
    bool open = false;
    void reset() { n = 0; open = false; }
    void begin(cudaStream_t s) {
        if (n >= kPairs || open) return;
        if (!a[n] && (cudaEventCreate(&a[n]) != cudaSuccess || cudaEventCreate(&b[n]) != cudaSuccess)) return;
        cudaEventRecord(a[n], s);
        open = true;
    }
    void end(cudaStream_t s) {
        if (!open) return;
        cudaEventRecord(b[n], s);
        open = false;
        ++n;
    }
    // the stream must have been synchronised after the last end()
    float total_ms() const {
        float t = 0.f;
        for (int i = 0; i < n; ++i) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, a[i], b[i]) == cudaSuccess) t += ms;
        }
        return t;
    }
    void destroy() {
        for (int i = 0; i < kPairs; ++i) {
            if (a[i]) cudaEventDestroy(a[i]);
            if (b[i]) cudaEventDestroy(b[i]);
            a[i] = b[i] = nullptr;
        }
    }
};

}  // namespace plc

struct plc_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint64_t launches = 0;
    std::string last_error;

    // staging
    plc::DevBuf upload;          // raw PointCloud2 bytes
    plc::DevBuf scratch;         // small reductions
    void* pinned = nullptr;      // pinned host scratch for small results
    size_t pinned_cap = 0;
    void* stage_pinned = nullptr;        // page-locked staging of large pageable inputs (plc::upload_pageable)
    size_t stage_cap = 0;
    cudaEvent_t stage_done = nullptr;    // the last staged copy has left the staging buffer
    bool stage_pending = false;

    // virtual_lidar state
    plc::VoxelIndex terrain, observed;
    plc::DevBuf cells;           // plc_cell[n_cells]
    uint64_t n_cells = 0;
    plc::DevBuf cells_f;         // float4[n_cells]: position relative to cell_ref, narrowed (k_sweep's float-first path)
    plc::DevBuf tile_boxes;      // per k_sweep tile: bounding box of its cells (plc_sweep.cu TileBox)
    double cell_ref[3] = {0, 0, 0};
    float cells_radius = 0.f;    // half diagonal of the cells' bounding box (inf if a finite box does not exist)
    plc::DevBuf cell_flags;      // uint8[n_cells], persists across sweeps (Q3)
    plc::DevBuf fixed_scores;    // double[n_cells]
    plc::DevBuf cands;           // plc_pose[n_cands] (full list)
    uint64_t n_cands = 0;
    uint64_t cand_begin = 0, cand_stride = 1, n_local = 0;   // this context evaluates cands[cand_begin + i*cand_stride], i < n_local
    plc_params prm{};
    plc_pose fixed{};
    bool have_job = false;
    bool ran = false, ran_mask = false, ran_flags = false;
    plc::DevBuf partial_score, partial_cov;   // per (candidate, tile)
    plc::DevBuf scores, covered, mask;
    plc::DevBuf pose_aux;        // per local candidate: FOV gate constants (plc_sweep.cu PoseAux)
    plc::DevBuf s_tab;           // double[n_tab]: sample parameters s_k of the upstream ray loop
    int n_tab = 0;
    double tab_start = 0, tab_step = 0, tab_max = 0;
    plc::DevBuf result;          // SweepResult
    uint64_t mask_words = 0;
    int tiles = 0;
    cudaEvent_t ev_sweep_begin = nullptr, ev_sweep_end = nullptr;   // around k_sweep
    // occupancy refinement policy (plc_sweep.cu pick_refine): rays swept over the current terrain so far, and the
    // size of the last sweep as the forecast for a terrain that has just arrived
    uint64_t terrain_rays_accum = 0, last_sweep_rays = 0;
    bool terrain_unindexed = false;   // plc_vl_set_terrain_unindexed: a terrain cloud exists, no index does (every ray visible)
    bool sweep_attr_set = false;      // k_sweep's shared-memory carve-out has been set on this context's device
    cudaEvent_t ev_run_begin = nullptr, ev_run_end = nullptr;       // around all kernels of plc_vl_run_async
    bool sweep_timed = false;
    bool count_samples = false;  // k_sweep in counting mode (plc_vl_count_samples) instead of any-hit mode
    // index build: the candidate-list fill runs on this stream beside the dilation (fork / join by events)
    cudaStream_t side_stream = nullptr;
    cudaEvent_t side_fork = nullptr, side_join = nullptr;
    // the instantiated CUDA graph of the last repeated sweep, the launch signature it was recorded for, the signature
    // of the previous sweep (a graph is recorded when a signature repeats), kernels per replay, replays so far
    cudaGraphExec_t sweep_graph_exec = nullptr;
    std::vector<unsigned char> sweep_graph_sig, sweep_graph_candidate;
    unsigned sweep_graph_launches = 0;
    uint64_t sweep_graph_replays = 0;
    plc::DevTimer timer;         // kernels of the call in progress
    // plc_vl_set_terrain returns with the tail of its index build still in flight (the caller's next host work —
    // staging the cells, the pose list — overlaps it), so its kernels are timed by a timer of their own that
    // plc_last_device_times reads after synchronising
    plc::DevTimer terrain_timer;
    plc::DevTimer* timer_override = nullptr;
    bool terrain_time_pending = false;
    plc::DevTimer& T() { return timer_override ? *timer_override : timer; }
    plc_device_times times{};    // last completed call of each kind (plc_last_device_times)

    // filter / merge scratch
    plc::DevBuf f_keys, f_vals, f_keys2, f_vals2, f_tmp, f_out, f_flags;
    plc::LookbackWs lb;          // status words and ticket counter of the single-pass scans (plc_scan.cuh)
    plc::DevBuf m_out;

    // node setup on the device (plc_setup.cu): loaded points, bucket order, keys/values of the sort, bucket
    // starts, sort table, scan scratch, flags + slots, packed outputs, per-point normals, bounds block
    plc::DevBuf su_pts, su_sorted, su_keys, su_start, su_table, su_tmp, su_flags, su_out, su_normals, su_state;
    plc::DevBuf ex_plan, ex_terrain, ex_area;     // plc_excavate.cu: lattice plan, output records
    const unsigned char* su_raw = nullptr;         // where load_cloud read the caller's records from (upload or the caller's device buffer)

    void set_error(const char* what, const char* why) {
        last_error = std::string(what) + ": " + why;
    }
    void* pinned_reserve(size_t bytes) {
        if (bytes <= pinned_cap) return pinned;
        if (pinned) cudaFreeHost(pinned);
        pinned = nullptr;
        pinned_cap = 0;
        if (cudaMallocHost(&pinned, bytes + 4096) != cudaSuccess) return nullptr;
        pinned_cap = bytes + 4096;
        return pinned;
    }
};

namespace plc {

// Device-side result block of a sweep (first 16 bytes are what the multi-GPU exchange gathers).
struct SweepResult {
    double best_score;
    long long best_index;       // global candidate index
    int best_covered;
    int pad;
    double fixed_total;
    unsigned long long rays_gated_in, rays_visible, samples, fixed_samples, exact_tests;
};

plc_status voxel_set_cloud(plc_ctx* ctx, VoxelIndex& ix, const plc_cloud_view* cloud);
plc_status fetch_strided(plc_ctx* ctx, plc_sweep_out* out, uint64_t dst_stride);
// Builds (or keeps) the index for `radius`.  thr: float32 squared-distance threshold of the exact test
// (normally (float)(radius*radius)); march_err + march_err_m: position error of the caller's march, in
// fine-voxel units, folded into the conservative slack.
plc_status voxel_build(plc_ctx* ctx, VoxelIndex& ix, double radius, float thr, double march_err, double march_err_m,
                       int refine = 1);

// Where a caller's cloud bytes are read from on the device.  `data` may be host memory (copied into ctx->upload) or
// device memory of this context's device — a buffer from plc_dev_alloc, typically the output of the previous stage —
// which is then read in place (clouds stay in HBM between the nodes' stages).
// Large pageable host buffer -> device, as a pipeline: a few host threads copy 1 MB pieces into page-locked staging
// while the pieces already there go over the bus (plc_ctx.cu).  False: not applicable, copy the plain way.
bool upload_pageable(plc_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes, plc_status* st);

inline plc_status stage_bytes(plc_ctx* ctx, const void* data, size_t bytes, const unsigned char** dev) {
    *dev = nullptr;
    if (bytes == 0) return PLC_OK;
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, data) == cudaSuccess &&
        (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged)) {
        if (attr.type == cudaMemoryTypeDevice && attr.device != ctx->device) {
            // another GPU's buffer (plc_group_* hand the same view to every context): kernels here cannot
            // dereference it without peer access, so it is copied device to device into this context's staging
            PLC_CUDA_TRY(ctx, ctx->upload.reserve(bytes));
            PLC_CUDA_TRY(ctx, cudaMemcpyPeerAsync(ctx->upload.ptr, ctx->device, data, attr.device, bytes, ctx->stream));
            *dev = ctx->upload.as<unsigned char>();
            return PLC_OK;
        }
        *dev = static_cast<const unsigned char*>(data);
        return PLC_OK;
    }
    const bool pageable = attr.type == cudaMemoryTypeUnregistered;
    cudaGetLastError();          // an unregistered host pointer is not an error
    PLC_CUDA_TRY(ctx, ctx->upload.reserve(bytes));
    plc_status st = PLC_OK;
    if (pageable && upload_pageable(ctx, ctx->upload.ptr, data, bytes, &st)) {
        if (st != PLC_OK) return st;
    } else {
        PLC_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->upload.ptr, data, bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    *dev = ctx->upload.as<unsigned char>();
    return PLC_OK;
}

inline bool cloud_view_ok(const plc_cloud_view* v) {
    if (!v) return false;
    if (v->n_points == 0) return true;
    return v->data && v->point_step >= 12 && v->off_x + 4 <= v->point_step &&
           v->off_y + 4 <= v->point_step && v->off_z + 4 <= v->point_step;
}

}  // namespace plc
