// plc_group.cu — the sweep on several GPUs of one box from ONE host process (include/plc.h, plc_group_*).
//
// The reference's virtual_lidar is a single rclcpp process (virtual_lidar.cpp:965-969), so a drop-in that
// wants the whole box drives every device itself: one plc_ctx per GPU, the terrain index, cells and the
// hoisted fixed-sensor pass replicated, candidate poses dealt out round-robin (device g of G evaluates candidates
// g, g + G, g + 2G, ...: neighbouring poses cost about the same, contiguous blocks of the lattice do not).  The only exchange is 16 bytes per device,
// {f64 best_score, i64 best_global_index}, all-gathered over NCCL (NVLink) straight from each context's
// device result block; every device then holds all pairs and the first maximum is picked with the
// reference's strict-> rule (lowest global index among equals, :464-475; each pair carries its global index, so
// the rule does not depend on how the list was dealt out).
//
// NCCL is bound at run time (dlopen of libnccl.so.2): a process that already loaded an NCCL — PyTorch's
// bundled one in bench.py — gets that copy instead of a second one with clashing symbols.  A group of one
// device never touches NCCL.  One host thread per device is used for the replicated set_* calls, which
// synchronise; the sweep itself is launched asynchronously on every device from the calling thread.
#include <dlfcn.h>

#include <cmath>
#include <string>
#include <thread>
#include <vector>

#include "plc_internal.cuh"

namespace {

typedef void* nccl_comm_t;
struct NcclApi {
    int (*CommInitAll)(nccl_comm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    void* lib = nullptr;
    bool load(std::string& err) {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) {
            err = std::string("dlopen(libnccl.so.2): ") + dlerror();
            return false;
        }
        auto sym = [&](const char* n) { return dlsym(lib, n); };
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
        AllGather = reinterpret_cast<decltype(AllGather)>(sym("ncclAllGather"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        if (!CommInitAll || !CommDestroy || !GroupStart || !GroupEnd || !AllGather || !GetErrorString) {
            err = "libnccl is missing a required symbol";
            return false;
        }
        return true;
    }
};
constexpr int kNcclInt8 = 0;     // ncclInt8 / ncclChar

struct Pair {
    double score;
    long long index;
};

}  // namespace

struct plc_group {
    std::vector<plc_ctx*> ctx;
    std::vector<int> dev;
    NcclApi nccl;
    std::vector<nccl_comm_t> comms;
    std::vector<void*> gathered;     // per device: Pair[n] written by the all-gather
    std::string last_error;
};

namespace {

// runs f(i) for every device on its own host thread and returns the first failure
template <class F>
plc_status on_all(plc_group* g, F f) {
    const int n = (int)g->ctx.size();
    std::vector<plc_status> st(n, PLC_OK);
    if (n == 1) {
        st[0] = f(0);
    } else {
        std::vector<std::thread> th;
        th.reserve(n);
        for (int i = 0; i < n; ++i) th.emplace_back([&, i] { st[i] = f(i); });
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < n; ++i)
        if (st[i] != PLC_OK) {
            g->last_error = "device " + std::to_string(g->dev[i]) + ": " + plc_last_error(g->ctx[i]);
            return st[i];
        }
    return PLC_OK;
}

plc_status nccl_fail(plc_group* g, const char* what, int rc) {
    g->last_error = std::string(what) + ": " + (g->nccl.GetErrorString ? g->nccl.GetErrorString(rc) : "?");
    return PLC_E_NCCL;
}

}  // namespace

extern "C" plc_status plc_group_create(const int* devices, int n_devices, plc_group** out) {
    PLC_RANGE();
    if (!out || n_devices < 1 || n_devices > 64 || !devices) return PLC_E_INVALID_ARG;
    *out = nullptr;
    for (int i = 0; i < n_devices; ++i)
        for (int k = 0; k < i; ++k)
            if (devices[i] == devices[k]) return PLC_E_INVALID_ARG;
    plc_group* g = new (std::nothrow) plc_group();
    if (!g) return PLC_E_OOM;
    plc_status st = PLC_OK;
    for (int i = 0; i < n_devices && st == PLC_OK; ++i) {
        plc_ctx* c = nullptr;
        st = plc_ctx_create(devices[i], nullptr, &c);
        if (st == PLC_OK) {
            g->ctx.push_back(c);
            g->dev.push_back(devices[i]);
        }
    }
    if (st == PLC_OK && n_devices > 1) {
        if (!g->nccl.load(g->last_error)) st = PLC_E_NCCL;
        if (st == PLC_OK) {
            g->comms.assign(n_devices, nullptr);
            int rc = g->nccl.CommInitAll(g->comms.data(), n_devices, devices);
            if (rc != 0) {
                g->comms.clear();
                st = nccl_fail(g, "ncclCommInitAll", rc);
            }
        }
        for (int i = 0; i < n_devices && st == PLC_OK; ++i) {
            void* p = nullptr;
            if (cudaSetDevice(devices[i]) != cudaSuccess || cudaMalloc(&p, sizeof(Pair) * n_devices) != cudaSuccess) st = PLC_E_OOM;
            g->gathered.push_back(p);
        }
    }
    if (st != PLC_OK) {
        plc_group_destroy(g);
        return st;
    }
    *out = g;
    return PLC_OK;
}

extern "C" void plc_group_destroy(plc_group* g) {
    if (!g) return;
    for (size_t i = 0; i < g->gathered.size(); ++i)
        if (g->gathered[i]) {
            cudaSetDevice(g->dev[i]);
            cudaFree(g->gathered[i]);
        }
    for (nccl_comm_t c : g->comms)
        if (c) g->nccl.CommDestroy(c);
    for (plc_ctx* c : g->ctx) plc_ctx_destroy(c);
    delete g;
}

extern "C" int plc_group_size(const plc_group* g) { return g ? (int)g->ctx.size() : 0; }
extern "C" plc_ctx* plc_group_ctx(plc_group* g, int i) {
    return (g && i >= 0 && i < (int)g->ctx.size()) ? g->ctx[i] : nullptr;
}
extern "C" const char* plc_group_last_error(const plc_group* g) { return g ? g->last_error.c_str() : ""; }

extern "C" plc_status plc_group_set_terrain(plc_group* g, const plc_cloud_view* cloud) {
    PLC_RANGE();
    if (!g) return PLC_E_INVALID_ARG;
    return on_all(g, [&](int i) { return plc_vl_set_terrain(g->ctx[i], cloud); });
}
extern "C" plc_status plc_group_set_terrain_unindexed(plc_group* g) {
    if (!g) return PLC_E_INVALID_ARG;
    return on_all(g, [&](int i) { return plc_vl_set_terrain_unindexed(g->ctx[i]); });
}
extern "C" plc_status plc_group_set_observed(plc_group* g, const plc_cloud_view* cloud) {
    PLC_RANGE();
    if (!g) return PLC_E_INVALID_ARG;
    return on_all(g, [&](int i) { return plc_vl_set_observed(g->ctx[i], cloud); });
}
extern "C" plc_status plc_group_set_cells(plc_group* g, const plc_cell* cells, uint64_t n_cells) {
    PLC_RANGE();
    if (!g) return PLC_E_INVALID_ARG;
    return on_all(g, [&](int i) { return plc_vl_set_cells(g->ctx[i], cells, n_cells); });
}

extern "C" plc_status plc_group_sweep(plc_group* g, const plc_params* prm, const plc_pose* fixed, const plc_pose* cands,
                                      uint64_t n_cands, plc_sweep_out* out) {
    PLC_RANGE();
    if (!g || !out) return PLC_E_INVALID_ARG;
    const int n = (int)g->ctx.size();
    int launched = 0;           // devices whose sweep is in flight
    // a failure on one device must not leave the others running a sweep nobody collects: wait for them first
    auto fail = [&](int i, plc_status st) {
        g->last_error = "device " + std::to_string(g->dev[i]) + ": " + plc_last_error(g->ctx[i]);
        for (int k = 0; k < launched; ++k) plc_ctx_sync(g->ctx[k]);
        return st;
    };
    // upload + launch on every device without waiting for any of them
    for (int i = 0; i < n; ++i) {
        plc_status st = plc_vl_upload_candidates_strided(g->ctx[i], prm, fixed, cands, n_cands, (uint64_t)i, (uint64_t)n);
        if (st != PLC_OK) return fail(i, st);
        // the stale-flag bytes depend on the whole candidate list, which every context holds: device 0 makes them
        st = plc_vl_run_async(g->ctx[i], out->mask != nullptr, i == 0 && out->cell_flags != nullptr);
        if (st != PLC_OK) return fail(i, st);
        launched = i + 1;
    }
    // the one exchange: all-gather of {best_score, best_global_index}, ordered after the sweep on each stream
    if (n > 1) {
        // everything that can fail without NCCL is done before the NCCL group opens, and the group is closed on
        // every path: an open group would leave the communicators unusable
        std::vector<void*> pair(n, nullptr);
        for (int i = 0; i < n; ++i) {
            plc_status st = plc_vl_result_device_ptr(g->ctx[i], &pair[i]);
            if (st != PLC_OK) return fail(i, st);
        }
        int rc = g->nccl.GroupStart();
        if (rc != 0) {
            for (int k = 0; k < launched; ++k) plc_ctx_sync(g->ctx[k]);
            return nccl_fail(g, "ncclGroupStart", rc);
        }
        int rc_gather = 0;
        for (int i = 0; i < n && rc_gather == 0; ++i)
            rc_gather = g->nccl.AllGather(pair[i], g->gathered[i], sizeof(Pair), kNcclInt8, g->comms[i], g->ctx[i]->stream);
        rc = g->nccl.GroupEnd();
        if (rc_gather != 0 || rc != 0) {
            for (int k = 0; k < launched; ++k) plc_ctx_sync(g->ctx[k]);
            return rc_gather != 0 ? nccl_fail(g, "ncclAllGather", rc_gather) : nccl_fail(g, "ncclGroupEnd", rc);
        }
    }
    // results: each shard's arrays are interleaved straight into the caller's global arrays
    plc_sweep_out total = *out;
    total.counters = plc_counters{};
    std::vector<Pair> pairs(n);
    for (int i = 0; i < n; ++i) {
        plc_sweep_out part{};
        const uint64_t words = (g->ctx[i]->n_cells + 31) / 32;
        part.scores = out->scores ? out->scores + i : nullptr;
        part.covered = out->covered ? out->covered + i : nullptr;
        part.mask = out->mask ? out->mask + (uint64_t)i * words : nullptr;
        part.cell_flags = i == 0 ? out->cell_flags : nullptr;
        part.fixed_scores = i == 0 ? out->fixed_scores : nullptr;
        plc_status st = plc::fetch_strided(g->ctx[i], &part, (uint64_t)n);
        if (st != PLC_OK) return fail(i, st);
        pairs[i] = Pair{part.best_score, (long long)part.best_index};
        total.counters.rays += part.counters.rays;
        total.counters.rays_gated_in += part.counters.rays_gated_in;
        total.counters.rays_visible += part.counters.rays_visible;
        total.counters.samples += part.counters.samples;
        total.counters.exact_tests += part.counters.exact_tests;
        if (i == 0) {
            total.fixed_total = part.fixed_total;
            total.n_cells = part.n_cells;
            total.mask_words = part.mask_words;
            total.counters.fixed_samples = part.counters.fixed_samples;
        }
    }
    if (n > 1) {
        // every device holds the same gathered pairs; read device 0's copy (fetch has synchronised its stream)
        if (cudaSetDevice(g->dev[0]) != cudaSuccess ||
            cudaMemcpy(pairs.data(), g->gathered[0], sizeof(Pair) * n, cudaMemcpyDeviceToHost) != cudaSuccess) {
            g->last_error = "reading the gathered result pairs failed";
            return PLC_E_CUDA;
        }
    }
    // first maximum in candidate order: strictly greater score, or an equal score at a lower global index
    int best = -1;
    for (int i = 0; i < n; ++i) {
        if (pairs[i].index < 0) continue;
        if (best < 0 || pairs[i].score > pairs[best].score ||
            (pairs[i].score == pairs[best].score && pairs[i].index < pairs[best].index))
            best = i;
    }
    total.best_index = best >= 0 ? pairs[best].index : -1;
    total.best_score = best >= 0 ? pairs[best].score : -INFINITY;
    total.best_covered = 0;
    if (best >= 0) {
        plc_sweep_out part{};
        plc_status st = plc_vl_fetch(g->ctx[best], &part);      // scalars only
        if (st != PLC_OK) return fail(best, st);
        total.best_covered = part.best_covered;
    }
    *out = total;
    return PLC_OK;
}
