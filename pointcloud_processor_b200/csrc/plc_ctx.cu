// plc_ctx.cu — context lifetime, status text, parameter presets.
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <memory>
#include <new>
#include <thread>
#include <vector>

#include "plc_internal.cuh"

extern "C" int plc_abi_version(void) { return PLC_ABI_VERSION; }

extern "C" const char* plc_strerror(int status) {
    switch (status) {
        case PLC_OK: return "ok";
        case PLC_E_INVALID_ARG: return "invalid argument";
        case PLC_E_NO_TERRAIN: return "no terrain cloud has been set";
        case PLC_E_NO_CELLS: return "no evaluation cells have been set";
        case PLC_E_CAPACITY: return "output buffer too small";
        case PLC_E_CUDA: return "CUDA error (see plc_last_error)";
        case PLC_E_NCCL: return "NCCL error (see plc_last_error)";
        case PLC_E_OOM: return "out of device memory";
        case PLC_E_NO_CANDIDATES: return "no candidates have been uploaded";
        default: return "unknown status";
    }
}

// Built node: virtual_lidar.cpp:66-71 (parameters) and :100-114 (constants).
extern "C" void plc_params_default(plc_params* p) {
    if (!p) return;
    p->max_distance = 15.0;
    p->min_distance = 0.5;
    p->ray_step = 0.3;
    p->ray_start = 0.5;
    p->vis_radius = 0.08;
    p->block_radius_scale = 0.7;
    p->observed_radius_scale = 3.0;
    p->fov_vertical = 180.0 * M_PI / 180.0;   // formed exactly as :711
    p->fov_horizontal = -1.0;
    p->alpha = 1.0;
    p->beta = 1.0;
    p->fixed_mode = 0;
    p->reserved = 0;
}

// virtual_lidar_memo.cpp:102-113 (constants), :684-697 (FOV), :699-726 (fixed sensor: observed cloud only).
extern "C" void plc_params_memo(plc_params* p) {
    if (!p) return;
    plc_params_default(p);
    p->min_distance = 1.0;
    p->ray_step = 0.2;
    p->vis_radius = 0.05;
    p->observed_radius_scale = 1.0;
    p->fov_vertical = 90.0 * M_PI / 180.0;
    p->fov_horizontal = 180.0 * M_PI / 180.0;
    p->fixed_mode = 1;
}

extern "C" plc_status plc_ctx_create(int device, void* stream, plc_ctx** out) {
    PLC_RANGE();
    if (!out) return PLC_E_INVALID_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return PLC_E_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return PLC_E_CUDA;
    plc_ctx* c = new (std::nothrow) plc_ctx();
    if (!c) return PLC_E_OOM;
    c->device = device;
    if (stream) {
        c->stream = static_cast<cudaStream_t>(stream);
    } else {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete c;
            return PLC_E_CUDA;
        }
        c->own_stream = true;
    }
    if (c->result.reserve(sizeof(plc::SweepResult)) != cudaSuccess || !c->pinned_reserve(1 << 16) ||
        cudaEventCreate(&c->ev_sweep_begin) != cudaSuccess || cudaEventCreate(&c->ev_sweep_end) != cudaSuccess ||
        cudaEventCreate(&c->ev_run_begin) != cudaSuccess || cudaEventCreate(&c->ev_run_end) != cudaSuccess ||
        c->lb.reserve(size_t(1) << 20) != cudaSuccess) {     // status words for scans of up to 2^31 items (8 MB)
        plc_ctx_destroy(c);
        return PLC_E_OOM;
    }
    *out = c;
    return PLC_OK;
}

extern "C" void plc_ctx_destroy(plc_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->lb.release();
    c->terrain.release();
    c->observed.release();
    plc::DevBuf* bufs[] = {&c->upload, &c->scratch, &c->cells, &c->cells_f, &c->tile_boxes, &c->cell_flags, &c->fixed_scores, &c->cands,
                           &c->partial_score, &c->partial_cov, &c->scores, &c->covered, &c->mask, &c->result, &c->pose_aux, &c->s_tab,
                           &c->f_keys, &c->f_vals, &c->f_keys2, &c->f_vals2, &c->f_tmp, &c->f_out, &c->f_flags,
                           &c->m_out, &c->su_pts, &c->su_sorted, &c->su_keys, &c->su_start, &c->su_table, &c->su_tmp,
                           &c->su_flags, &c->su_out, &c->su_normals, &c->su_state, &c->ex_plan, &c->ex_terrain, &c->ex_area};
    for (plc::DevBuf* b : bufs) b->release();
    if (c->pinned) cudaFreeHost(c->pinned);
    if (c->stage_pinned) cudaFreeHost(c->stage_pinned);
    if (c->stage_done) cudaEventDestroy(c->stage_done);
    if (c->sweep_graph_exec) cudaGraphExecDestroy(c->sweep_graph_exec);
    if (c->side_fork) cudaEventDestroy(c->side_fork);
    if (c->side_join) cudaEventDestroy(c->side_join);
    if (c->side_stream) cudaStreamDestroy(c->side_stream);
    c->timer.destroy();
    c->terrain_timer.destroy();
    if (c->ev_sweep_begin) cudaEventDestroy(c->ev_sweep_begin);
    if (c->ev_sweep_end) cudaEventDestroy(c->ev_sweep_end);
    if (c->ev_run_begin) cudaEventDestroy(c->ev_run_begin);
    if (c->ev_run_end) cudaEventDestroy(c->ev_run_end);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" const char* plc_last_error(const plc_ctx* c) { return c ? c->last_error.c_str() : ""; }
extern "C" uint64_t plc_launch_count(const plc_ctx* c) { return c ? c->launches : 0; }

extern "C" plc_status plc_ctx_sync(plc_ctx* c) {
    if (!c) return PLC_E_INVALID_ARG;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    PLC_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return PLC_OK;
}

// ---- large pageable inputs (a ROS message's byte vector is pageable memory).  cudaMemcpyAsync from pageable memory
// is staged by the driver on the calling thread (14 GB/s at 64 MB); here a few threads fill page-locked staging piece by piece
// and each piece goes over the bus as soon as it is there.  PLC_STAGE_THREADS=0 turns it off.
namespace plc {
namespace {
// measured on the B200 box (tools/dev/stage_probe.py): 16 MB 0.84 ms plain / 1.07 ms staged (starting the helper threads
// is the fixed cost), 64 MB 4.75 ms plain / 2.66 ms staged — so only from 32 MB up
constexpr size_t kStagePiece = size_t(1) << 20, kStageMin = size_t(32) << 20, kStageMax = size_t(1) << 30;

int stage_threads() {
    static const int n = [] {
        if (const char* e = std::getenv("PLC_STAGE_THREADS")) return std::max(0, std::min(atoi(e), 16));
        const unsigned hw = std::thread::hardware_concurrency();
        return hw >= 8 ? 4 : hw >= 4 ? 2 : 0;
    }();
    return n;
}
}  // namespace

bool upload_pageable(plc_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes, plc_status* st) {
    *st = PLC_OK;
    const int n_threads = stage_threads();
    if (n_threads == 0 || bytes < kStageMin || bytes > kStageMax) return false;
    if (ctx->stage_pending) {                         // the previous staged copy may still be reading the buffer
        if (cudaEventSynchronize(ctx->stage_done) != cudaSuccess) return false;
        ctx->stage_pending = false;
    }
    if (bytes > ctx->stage_cap) {
        if (ctx->stage_pinned) cudaFreeHost(ctx->stage_pinned);
        ctx->stage_pinned = nullptr;
        ctx->stage_cap = 0;
        const size_t want = (bytes + kStagePiece - 1) / kStagePiece * kStagePiece;
        if (cudaMallocHost(&ctx->stage_pinned, want) != cudaSuccess) {
            cudaGetLastError();
            ctx->stage_pinned = nullptr;
            return false;
        }
        ctx->stage_cap = want;
    }
    if (!ctx->stage_done && cudaEventCreateWithFlags(&ctx->stage_done, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        ctx->stage_done = nullptr;
        return false;
    }
    const size_t n_pieces = (bytes + kStagePiece - 1) / kStagePiece;
    std::unique_ptr<std::atomic<unsigned char>[]> ready(new (std::nothrow) std::atomic<unsigned char>[n_pieces]);
    if (!ready) return false;
    for (size_t i = 0; i < n_pieces; ++i) ready[i].store(0, std::memory_order_relaxed);
    std::atomic<size_t> next{0};
    const unsigned char* src = static_cast<const unsigned char*>(host_src);
    unsigned char* stage = static_cast<unsigned char*>(ctx->stage_pinned);
    auto fill = [&] {
        for (;;) {
            const size_t i = next.fetch_add(1, std::memory_order_relaxed);
            if (i >= n_pieces) return;
            const size_t off = i * kStagePiece;
            memcpy(stage + off, src + off, std::min(kStagePiece, bytes - off));
            ready[i].store(1, std::memory_order_release);
        }
    };
    std::vector<std::thread> helpers;
    try {
        for (int t = 0; t < n_threads; ++t) helpers.emplace_back(fill);
    } catch (...) {
        // fewer helpers than wanted (or none): this thread fills what is left below
    }
    cudaError_t err = cudaSuccess;
    for (size_t i = 0; i < n_pieces; ++i) {
        while (!ready[i].load(std::memory_order_acquire)) {
            if (helpers.empty()) fill();              // no helper could be started: copy here
            else std::this_thread::yield();
        }
        const size_t off = i * kStagePiece;
        if (err == cudaSuccess)
            err = cudaMemcpyAsync(static_cast<unsigned char*>(dev_dst) + off, stage + off, std::min(kStagePiece, bytes - off),
                                  cudaMemcpyHostToDevice, ctx->stream);
    }
    for (std::thread& t : helpers) t.join();
    if (err == cudaSuccess) err = cudaEventRecord(ctx->stage_done, ctx->stream);
    if (err != cudaSuccess) {
        ctx->set_error("upload_pageable", cudaGetErrorString(err));
        cudaGetLastError();
        *st = PLC_E_CUDA;
        return true;
    }
    ctx->stage_pending = true;
    return true;
}
}  // namespace plc

// ---- device buffers handed between stages (plc.h)
extern "C" plc_status plc_dev_alloc(plc_ctx* c, uint64_t bytes, void** dev_ptr) {
    if (!c || !dev_ptr) return PLC_E_INVALID_ARG;
    *dev_ptr = nullptr;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    PLC_CUDA_TRY(c, cudaMalloc(dev_ptr, bytes ? bytes : 1));
    return PLC_OK;
}

extern "C" void plc_dev_free(plc_ctx* c, void* dev_ptr) {
    if (!c || !dev_ptr) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(dev_ptr);
}

extern "C" plc_status plc_dev_upload(plc_ctx* c, void* dev_dst, const void* host_src, uint64_t bytes) {
    PLC_RANGE();
    if (!c || (bytes && (!dev_dst || !host_src))) return PLC_E_INVALID_ARG;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    cudaPointerAttributes attr{};
    const bool pageable = cudaPointerGetAttributes(&attr, host_src) != cudaSuccess || attr.type == cudaMemoryTypeUnregistered;
    cudaGetLastError();
    plc_status st = PLC_OK;
    if (pageable && plc::upload_pageable(c, dev_dst, host_src, bytes, &st)) {
        if (st != PLC_OK) return st;
    } else {
        PLC_CUDA_TRY(c, cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, c->stream));
    }
    PLC_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return PLC_OK;
}

extern "C" plc_status plc_dev_download(plc_ctx* c, void* host_dst, const void* dev_src, uint64_t bytes) {
    PLC_RANGE();
    if (!c || (bytes && (!host_dst || !dev_src))) return PLC_E_INVALID_ARG;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    PLC_CUDA_TRY(c, cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, c->stream));
    PLC_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return PLC_OK;
}

// ---- the same buffers across processes (CUDA IPC; one GPU, one handle per plc_dev_alloc allocation)
static_assert(sizeof(cudaIpcMemHandle_t) == sizeof(plc_ipc_handle), "plc_ipc_handle carries a cudaIpcMemHandle_t");

extern "C" plc_status plc_dev_export(plc_ctx* c, void* dev_ptr, plc_ipc_handle* handle) {
    if (!c || !dev_ptr || !handle) return PLC_E_INVALID_ARG;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    PLC_CUDA_TRY(c, cudaStreamSynchronize(c->stream));          // whatever wrote the buffer has finished
    cudaIpcMemHandle_t h;
    PLC_CUDA_TRY(c, cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle->opaque, &h, sizeof(h));
    return PLC_OK;
}

extern "C" plc_status plc_dev_import(plc_ctx* c, const plc_ipc_handle* handle, void** dev_ptr) {
    if (!c || !handle || !dev_ptr) return PLC_E_INVALID_ARG;
    *dev_ptr = nullptr;
    PLC_CUDA_TRY(c, cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle->opaque, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaErrorInvalidValue || e == cudaErrorInvalidResourceHandle || e == cudaErrorDeviceUninitialized) {
        cudaGetLastError();                                     // a handle of this very process, or not a handle
        *dev_ptr = nullptr;
        return PLC_E_INVALID_ARG;
    }
    PLC_CUDA_TRY(c, e);
    return PLC_OK;
}

extern "C" void plc_dev_release(plc_ctx* c, void* dev_ptr) {
    if (!c || !dev_ptr) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaIpcCloseMemHandle(dev_ptr);
}
