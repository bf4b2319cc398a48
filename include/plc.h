/*
 * plc.h — C ABI of the B200-native point-cloud path ("plc") that replaces the data-parallel
 * bodies of three reference ROS2 nodes.  POD only, no exceptions, no C++/torch types.
 *
 * The reference (YamaguchiAtsushi/pointcloud_processor) has no FFI or plugin interface: its hot
 * path sits in private member functions of rclcpp::Node subclasses.  Each entry point below names
 * the reference lines whose body it replaces; INTEGRATION.md shows the call sites a maintainer
 * adds to the three nodes.
 *
 * Conventions
 *   - every function returns plc_status (0 = OK, negative = error); plc_strerror() names it.
 *   - the caller owns every host pointer; the library copies in during the call and never keeps
 *     caller memory.  Output buffers are caller-allocated with a capacity; too small a capacity
 *     is PLC_E_CAPACITY and nothing is written.
 *   - a plc_ctx is bound to one CUDA device and one stream, is NOT thread-safe, and every call is
 *     synchronous unless its name ends in _async.
 *   - there is no CPU fallback: without a usable device plc_ctx_create fails with PLC_E_CUDA.
 *   - point clouds cross the boundary in sensor_msgs/PointCloud2 wire layout: a byte buffer,
 *     a point_step, and the byte offsets of the FLOAT32 x/y/z fields (plc_cloud_view).
 */
#ifndef PLC_H_
#define PLC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLC_ABI_VERSION 1

typedef enum plc_status {
    PLC_OK = 0,
    PLC_E_INVALID_ARG = -1,
    PLC_E_NO_TERRAIN = -2,   /* sweep asked for before plc_vl_set_terrain (reference: runOptimization guard, virtual_lidar.cpp:455) */
    PLC_E_NO_CELLS = -3,     /* sweep asked for before plc_vl_set_cells (virtual_lidar.cpp:455) */
    PLC_E_CAPACITY = -4,
    PLC_E_CUDA = -5,
    PLC_E_NCCL = -6,
    PLC_E_OOM = -7,
    PLC_E_NO_CANDIDATES = -8
} plc_status;

typedef struct plc_ctx plc_ctx;

/* ---------------------------------------------------------------- PODs ---------------------- */

/* A view of PointCloud2 `data`.  Offsets are byte offsets of FLOAT32 fields. */
typedef struct plc_cloud_view {
    const void* data;
    uint64_t n_points;      /* width*height */
    uint32_t point_step;
    uint32_t off_x, off_y, off_z;
} plc_cloud_view;

/* Evaluation target; mirrors the fields of `struct GridCell` the sweep reads
 * (virtual_lidar.cpp:20-44): double centre, float32 unit normal (pcl::Normal). 40 bytes. */
typedef struct plc_cell {
    double x, y, z;
    float nx, ny, nz;
    uint32_t pad;
} plc_cell;

/* Sensor pose; mirrors `struct LidarPosition` (virtual_lidar.cpp:46-51). 40 bytes. */
typedef struct plc_pose {
    double x, y, z, pitch, yaw;
} plc_pose;

/* Every ROS parameter and every `static constexpr` of the sweep (virtual_lidar.cpp:66-71,100-114)
 * as a runtime field, so the un-built `virtual_lidar_memo.cpp` parameterisation is testable too. */
typedef struct plc_params {
    double max_distance;          /* param "max_distance" = 15.0            (:69, :662) */
    double min_distance;          /* MIN_DISTANCE = 0.5                      (:102, :662) */
    double ray_step;              /* RAY_STEP_SIZE = 0.3                     (:111, :796) */
    double ray_start;             /* start_offset = 0.5                      (:765)       */
    double vis_radius;            /* VISIBILITY_RADIUS = 0.08; end margin    (:112, :767) */
    double block_radius_scale;    /* 0.7: search radius = vis_radius*0.7     (:782)       */
    double observed_radius_scale; /* 3.0: fixed-sensor lookup radius scale   (:743)       */
    double fov_vertical;          /* 180.0*M_PI/180.0 (full angle)           (:711-713)   */
    double fov_horizontal;        /* < 0: no azimuth test (built node); memo variant: 180.0*M_PI/180.0 */
    double alpha, beta;           /* ALPHA = BETA = 1.0                      (:100-101, :698) */
    int32_t fixed_mode;           /* 0: built node (:716-752)  1: memo variant (virtual_lidar_memo.cpp:699-726) */
    int32_t reserved;
} plc_params;

/* Defaults of the built node. */
void plc_params_default(plc_params* p);
/* Defaults of virtual_lidar_memo.cpp (constants :102-113, FOV :684-697). */
void plc_params_memo(plc_params* p);

/* Per-cell flag bits, same meaning as the GridCell bools (virtual_lidar.cpp:26-31). */
enum {
    PLC_FLAG_RANGE_FIXED = 1u << 0,
    PLC_FLAG_FOV_FIXED = 1u << 1,
    PLC_FLAG_VIS_FIXED = 1u << 2,
    PLC_FLAG_RANGE_MOBILE = 1u << 3,
    PLC_FLAG_FOV_MOBILE = 1u << 4,
    PLC_FLAG_VIS_MOBILE = 1u << 5
};

/* Counters echoed by the sweep so that both sides compute the same algorithmic bytes
 * (SURVEY §8 D4). All refer to mobile-sensor rays of the candidates this ctx evaluated. */
typedef struct plc_counters {
    uint64_t rays;            /* candidates evaluated * cells */
    uint64_t rays_gated_in;   /* rays that passed range and FOV */
    uint64_t rays_visible;    /* rays with the visibility bit set */
    uint64_t samples;         /* sample positions the reference would evaluate (up to and incl. first blocked);
                                 0 unless plc_vl_count_samples(ctx, 1) */
    uint64_t fixed_samples;   /* same, for the hoisted fixed-sensor pass (M rays) */
    uint64_t exact_tests;     /* samples that needed the exact point test (implementation detail, informational) */
} plc_counters;

typedef struct plc_sweep_out {
    /* scalars, always written */
    int64_t best_index;       /* global candidate index of the first maximum; -1 if no candidates (Q5) */
    double best_score;        /* -inf if no candidates */
    int32_t best_covered;     /* covered_cells of the best candidate */
    int32_t n_cells;
    double fixed_total;       /* "Total Score (ZX120 only)" (virtual_lidar.cpp:384-386) */
    plc_counters counters;
    /* optional arrays, may be NULL; indexed by LOCAL candidate (cand_begin..cand_end) */
    double* scores;           /* [n_local] total_score per candidate (:647) */
    int32_t* covered;         /* [n_local] covered_cells (:648) */
    uint32_t* mask;           /* [n_local * mask_words] bit j of word w = mobile ray to cell 32w+j passed range, FOV and occlusion */
    uint64_t mask_words;      /* ceil(n_cells/32), written by the library */
    uint8_t* cell_flags;      /* [n_cells] PLC_FLAG_* with the reference's carry-over rule */
    double* fixed_scores;     /* [n_cells] score_zx120 per cell (:377) */
} plc_sweep_out;

/* ---------------------------------------------------------------- context ------------------- */

/* Binds a context to `device`.  `stream` is a cudaStream_t (as void*) the library launches on, or
 * NULL to let the context create its own non-blocking stream. */
plc_status plc_ctx_create(int device, void* stream, plc_ctx** out);
void plc_ctx_destroy(plc_ctx* ctx);
const char* plc_strerror(int status);
/* Last CUDA/NCCL error text seen by this context ("" if none). */
const char* plc_last_error(const plc_ctx* ctx);
/* Number of kernels this context has launched since creation (bench.py "gpu_launches"). */
uint64_t plc_launch_count(const plc_ctx* ctx);
int plc_abi_version(void);
/* Blocks until everything queued on the context's stream has finished. */
plc_status plc_ctx_sync(plc_ctx* ctx);

/* ---------------------------------------------------------------- clouds that stay on the device -- */

/* Device buffers a caller can hand from one stage to the next without a host round trip.  Wherever an entry point
 * takes a point cloud (plc_cloud_view::data of plc_vl_set_terrain / plc_vl_set_observed / plc_filter / plc_merge /
 * plc_excavate / plc_drivable_area / plc_vl_compute_normals / plc_vl_generate_cells / plc_vl_generate_candidates),
 * the per-point normals of plc_vl_generate_cells, or writes a cloud, a normal array, cells or poses (out_* of the
 * same calls), the pointer may be host memory or memory of the context's device: device inputs are read in place,
 * device outputs are written with a device-to-device copy.  Counts and info structs always come back to the host.
 * (plc_vl_set_cells, the candidate list and the result arrays of the sweep are host memory: the node needs them
 * there for its markers and logs.) */
plc_status plc_dev_alloc(plc_ctx* ctx, uint64_t bytes, void** dev_ptr);
/* Waits for the context's stream, then frees.  Call it before plc_ctx_destroy(ctx). */
void plc_dev_free(plc_ctx* ctx, void* dev_ptr);
plc_status plc_dev_upload(plc_ctx* ctx, void* dev_dst, const void* host_src, uint64_t bytes);
plc_status plc_dev_download(plc_ctx* ctx, void* host_dst, const void* dev_src, uint64_t bytes);

/* The same hand-over between PROCESSES on one GPU (upstream's stages are separate ROS nodes: pointcloud_filter →
 * pointcloud_merger → excavated_surface_generator → virtual_lidar each publish a PointCloud2 the next one copies):
 * the producer exports a buffer it got from plc_dev_alloc, sends the 64 opaque bytes over any channel it has (a topic,
 * a pipe), the consumer imports them and passes the pointer as plc_cloud_view::data like any device pointer.
 * The producer keeps the buffer alive (and does not rewrite it) until the consumer has released it; every library call
 * that wrote the buffer has returned, hence finished on the device, before the handle is sent. */
typedef struct plc_ipc_handle {
    unsigned char opaque[64];
} plc_ipc_handle;
/* dev_ptr: a pointer plc_dev_alloc returned (the start of the allocation), in this process. */
plc_status plc_dev_export(plc_ctx* ctx, void* dev_ptr, plc_ipc_handle* handle);
/* Maps a buffer another process exported on the context's device.  PLC_E_INVALID_ARG in the exporting process itself. */
plc_status plc_dev_import(plc_ctx* ctx, const plc_ipc_handle* handle, void** dev_ptr);
/* Unmaps an imported buffer (not plc_dev_free: the memory belongs to the exporter). */
void plc_dev_release(plc_ctx* ctx, void* dev_ptr);

/* ---------------------------------------------------------------- virtual_lidar ------------- */

/* Replaces the KD-tree build of terrainCallback (virtual_lidar.cpp:180-192): voxelises the
 * /excavated_terrain cloud into the occupancy grid the sweep traverses.  Non-finite points are
 * dropped, as KdTreeFLANN::setInputCloud does.  n_points == 0 leaves "no terrain" (every ray
 * visible, :727), matching the reference's empty-cloud guard (:184).
 * The caller's buffer has been consumed when the call returns, but the last kernels of the index build may still be
 * running on the context's stream (later calls on the context are ordered behind them; plc_ctx_sync or
 * plc_last_device_times waits for them): a device fault in them is reported by the next call that synchronises. */
plc_status plc_vl_set_terrain(plc_ctx* ctx, const plc_cloud_view* cloud);
/* Replaces zx120PointsCallback (virtual_lidar.cpp:194-207): the cloud observed by the fixed sensor. */
/* terrainCallback with an EMPTY message (virtual_lidar.cpp:180-192): terrain_cloud_ becomes non-null, so
 * runOptimization's guard (:455) passes, but no KD-tree exists — checkVisibility then answers `true` for every
 * gated ray (:721, :727).  If no terrain has been indexed yet, sweeps run in that state instead of returning
 * PLC_E_NO_TERRAIN; an index that exists is kept, as upstream keeps its tree. */
plc_status plc_vl_set_terrain_unindexed(plc_ctx* ctx);
plc_status plc_vl_set_observed(plc_ctx* ctx, const plc_cloud_view* cloud);
/* Hands over excavation_grid_3d_ (built by generateExcavationGrid3D, :236-287). Resets cell flags. */
plc_status plc_vl_set_cells(plc_ctx* ctx, const plc_cell* cells, uint64_t n_cells);

/* Replaces runOptimization lines 460-475: evaluateZX120Only + the candidate loop + argmax.
 * `cands[0..n_cands)` is the full candidate list; this context evaluates [cand_begin, cand_end)
 * (pass 0, n_cands for a single GPU).  best_index is a GLOBAL index. */
plc_status plc_vl_sweep(plc_ctx* ctx, const plc_params* prm, const plc_pose* fixed,
                        const plc_pose* cands, uint64_t n_cands,
                        uint64_t cand_begin, uint64_t cand_end, plc_sweep_out* out);

/* The same sweep split so that a caller can keep inputs resident in HBM and time the device part:
 *   upload (H2D of params + poses)  ->  run_async (kernels only)  ->  fetch (D2H of results). */
plc_status plc_vl_upload_candidates(plc_ctx* ctx, const plc_params* prm, const plc_pose* fixed,
                                    const plc_pose* cands, uint64_t n_cands,
                                    uint64_t cand_begin, uint64_t cand_end);
/* The same with a strided share of the list: this context evaluates candidates first, first + stride,
 * first + 2 stride, ... (< n_cands).  Dealing the list out round-robin (first = rank, stride = ranks) balances
 * the devices better than contiguous blocks, because neighbouring poses cost about the same.  The optional
 * arrays of plc_sweep_out are indexed by LOCAL candidate i (global index first + i * stride); best_index is
 * global as always. */
plc_status plc_vl_upload_candidates_strided(plc_ctx* ctx, const plc_params* prm, const plc_pose* fixed,
                                            const plc_pose* cands, uint64_t n_cands,
                                            uint64_t first, uint64_t stride);
plc_status plc_vl_sweep_strided(plc_ctx* ctx, const plc_params* prm, const plc_pose* fixed,
                                const plc_pose* cands, uint64_t n_cands,
                                uint64_t first, uint64_t stride, plc_sweep_out* out);
plc_status plc_vl_run_async(plc_ctx* ctx, int want_mask, int want_flags);
plc_status plc_vl_fetch(plc_ctx* ctx, plc_sweep_out* out);
/* How the traversal treats a ray once one of its samples is known to be blocked.  The reference walks a ray from the
 * sensor outwards and returns at the first blocked sample (virtual_lidar.cpp:769-797); only "is any sample blocked"
 * reaches its outputs (visibility, scores, flags, optimum).
 *   on == 0 (default)  any-hit: samples are visited from the target end, and nothing of a ray is tested again once
 *                      a blocked sample is known.  Same masks, scores, flags and optimum, bit for bit;
 *                      plc_sweep_out::samples is reported as 0.
 *   on != 0            counting: the first blocked sample index of every ray is established, and
 *                      plc_sweep_out::samples is the number of sample positions the reference's loop evaluates
 *                      (what SURVEY D4's algorithmic-bytes figure is made of).  3-17 % slower.
 * No counterpart in the reference (it has no counters). */
plc_status plc_vl_count_samples(plc_ctx* ctx, int on);
/* A sweep that repeats the previous one's launch signature (same device buffers, sizes, parameters and outputs asked
 * for: the timer ticks of the node between two terrain messages, a benchmark's resident loop) is recorded into a CUDA
 * graph, and every further repeat replays it as one launch (the context owns its graphs; any
 * set_* / upload that moves a buffer or changes a size makes the next sweep record anew).  *n = replays so far.
 * PLC_NO_GRAPH in the environment turns the mechanism off.  No counterpart in the reference. */
plc_status plc_vl_graph_replays(plc_ctx* ctx, uint64_t* n);
/* Device address of this context's {double best_score; int64 best_global_index} pair (16 bytes),
 * valid after plc_vl_run_async; the multi-GPU exchange all-gathers it. */
plc_status plc_vl_result_device_ptr(plc_ctx* ctx, void** ptr);

/* Device time (ms, CUDA events on the context's stream) of the dominant kernel of the last
 * plc_vl_run_async — the per-(pose, cell) traversal kernel k_sweep — for roofline accounting.
 * Synchronises the stream.  0 if no candidates were evaluated. */
plc_status plc_vl_last_sweep_kernel_ms(plc_ctx* ctx, float* ms);

/* ---------------------------------------------------------------- virtual_lidar setup -- */

/* grid_min_x_ ... excavation_max_z_, grid_width_, grid_height_ (virtual_lidar.cpp:144-146) */
typedef struct plc_grid_bounds {
    double min_x, max_x, min_y, max_y, min_z, max_z;
    int32_t width, height;
} plc_grid_bounds;

/* Replaces computeTerrainNormals (virtual_lidar.cpp:209-234; pcl::NormalEstimation, radius search 1.5 m):
 * one PCA normal per /excavation_area point, flipped to n_z >= 0; NaN where a point has fewer than three
 * neighbours or is not finite.  normals_out: 3 floats per input point, in input order. */
plc_status plc_vl_compute_normals(plc_ctx* ctx, const plc_cloud_view* area, double radius, float* normals_out);

/* Replaces generateExcavationGrid3D + isPointNearExcavation + computeCellSurfaceNormal
 * (virtual_lidar.cpp:236-340): the lattice over the area cloud's bounds (margin = resolution), the lattice
 * points within 1.5 resolution of an area point kept in lattice order (y index, x index, layer), each with the
 * normalised sum of the normals of the area points within normal_search_radius.
 * normals: one float triple per area point (NaN allowed), or NULL => every cell keeps (0,0,1).
 * *n_out receives the number of kept cells even when it exceeds capacity (then PLC_E_CAPACITY). */
plc_status plc_vl_generate_cells(plc_ctx* ctx, const plc_cloud_view* area, const float* normals,
                                 double grid_resolution, int32_t vertical_layers, double normal_search_radius,
                                 plc_cell* out, uint64_t capacity, uint64_t* n_out, plc_grid_bounds* bounds);

/* Replaces generateCandidatePositions + getGroundHeight (virtual_lidar.cpp:550-625): the pose lattice around
 * the grid bounds, each kept pose standing sensor_height above the highest terrain point within 1 m.
 * terrain == NULL: the cloud last given to plc_vl_set_terrain, which is already on the device. */
plc_status plc_vl_generate_candidates(plc_ctx* ctx, const plc_cloud_view* terrain, const plc_grid_bounds* bounds,
                                      const plc_pose* fixed, int32_t num_candidates, double search_radius,
                                      double sensor_height, plc_pose* out, uint64_t capacity, uint64_t* n_out);

/* Grid facts for roofline accounting: bytes of the occupancy structure the sweep reads. */
/* ---------------------------------------------------------------- several GPUs, one process -- */

/* The reference node is one process (virtual_lidar.cpp:965-969).  A group owns one context per listed CUDA
 * device and runs the sweep over all of them: terrain, observed cloud and cells replicated, candidate poses
 * dealt out round-robin (device g of G evaluates g, g + G, ...), one NCCL all-gather of 16 bytes per device
 * ({f64 best_score, i64 best_global_index}), first maximum = lowest global index.  NCCL is
 * loaded at run time (libnccl.so.2) and only for groups of more than one device; failure is PLC_E_NCCL.
 * Same threading contract as a context: calls on one group must not overlap. */
typedef struct plc_group plc_group;
plc_status plc_group_create(const int* devices, int n_devices, plc_group** out);
void plc_group_destroy(plc_group* group);
int plc_group_size(const plc_group* group);
/* Context of the i-th device, for the per-context queries (plc_vl_grid_info, plc_last_device_times, ...). */
plc_ctx* plc_group_ctx(plc_group* group, int i);
const char* plc_group_last_error(const plc_group* group);
/* plc_vl_set_* on every device (one host thread each). */
plc_status plc_group_set_terrain(plc_group* group, const plc_cloud_view* cloud);
plc_status plc_group_set_terrain_unindexed(plc_group* g);   /* plc_vl_set_terrain_unindexed on every device */
plc_status plc_group_set_observed(plc_group* group, const plc_cloud_view* cloud);
plc_status plc_group_set_cells(plc_group* group, const plc_cell* cells, uint64_t n_cells);
/* plc_vl_sweep over the whole candidate list; the optional arrays of `out` are indexed by GLOBAL candidate. */
plc_status plc_group_sweep(plc_group* group, const plc_params* prm, const plc_pose* fixed,
                           const plc_pose* cands, uint64_t n_cands, plc_sweep_out* out);

/* Device time (ms, CUDA events on the context's stream) spent in the kernels of the most recent call of each
 * kind; host<->device copies and host-side waits are excluded.  0 until that call has run.  For benchmarks and
 * profiles: the reference has no counterpart. */
typedef struct plc_device_times {
    float set_terrain_ms;   /* plc_vl_set_terrain: extraction + occupancy/candidate-list build (K1) */
    float set_observed_ms;  /* plc_vl_set_observed */
    float filter_ms;        /* plc_filter: crop (K6) + voxel downsample (K7) */
    float merge_ms;         /* plc_merge: transform + colour + concatenate (K8) */
    float sweep_kernel_ms;  /* k_sweep of the last plc_vl_sweep / plc_vl_run_async */
    float normals_ms;       /* plc_vl_compute_normals */
    float cells_ms;         /* plc_vl_generate_cells */
    float candidates_ms;    /* plc_vl_generate_candidates */
    float drivable_ms;      /* plc_drivable_area */
    float excavate_ms;      /* plc_excavate */
    float sweep_fixed_ms;   /* last plc_vl_run_async: everything before k_sweep (reset, fixed-sensor pass k_fixed, k_pose_prep) */
    float sweep_reduce_ms;  /* last plc_vl_run_async: everything after k_sweep (k_finalize, k_argmax, k_flags if asked for) */
} plc_device_times;
plc_status plc_last_device_times(plc_ctx* ctx, plc_device_times* out);

typedef struct plc_grid_info {
    uint64_t n_points;        /* finite points indexed */
    uint64_t n_voxels;        /* coarse voxels */
    uint64_t grid_bytes;      /* mask bricks + voxel ranges + sorted float4 points */
    int32_t dim[3];
    double origin[3];
    double voxel;             /* coarse voxel edge (m) */
} plc_grid_info;
plc_status plc_vl_grid_info(const plc_ctx* ctx, plc_grid_info* terrain, plc_grid_info* observed);

/* ---------------------------------------------------------------- pointcloud_filter --------- */

/* cropFrontArea bounds (pointcloud_filter.cpp:111-113); all six comparisons strict. */
typedef struct plc_crop_box {
    double x_min, x_max;   /* 0.0, front_range  */
    double y_min, y_max;   /* -side, +side      */
    double z_min, z_max;   /* -1.5, height      */
} plc_crop_box;

/* Replaces processCloudSimple's crop + VoxelGrid (pointcloud_filter.cpp:64-85, 87-139).
 * Output: PointXYZ wire layout, 16 bytes per point: x,y,z and the 1.0f that pcl::PointXYZ keeps in
 * data[3] and pcl::toROSMsg copies to the wire.
 * leaf <= 0 skips the downsample (crop only).  An empty crop returns n_out = 0 (Q7). */
plc_status plc_filter(plc_ctx* ctx, const plc_cloud_view* in, const plc_crop_box* box, double leaf,
                      void* out_xyz16, uint64_t capacity, uint64_t* n_out, uint64_t* n_cropped);

/* ---------------------------------------------------------------- pointcloud_merger --------- */

typedef struct plc_merge_src {
    plc_cloud_view cloud;
    double t[3];           /* transform translation map <- velodyne_link */
    double q[4];           /* transform rotation quaternion x,y,z,w */
    uint8_t r, g, b, present;   /* present == 0: source skipped (TF failure / no cloud, :389-393) */
} plc_merge_src;

/* Replaces processRobotCloud x2 + concatenation (pointcloud_merger.cpp:308-394).
 * out_merged: PointXYZRGB wire layout, 32 bytes per point, sources in the order given.
 * out_each[i] (may be NULL): the i-th source's coloured cloud, same layout. */
plc_status plc_merge(plc_ctx* ctx, const plc_merge_src* srcs, int n_src,
                     void* out_merged, uint64_t capacity, uint64_t* n_out,
                     void* const* out_each, uint64_t* n_each);

/* ---------------------------------------------------------------- excavated_surface_generator -- */

/* Parameters of the node (excavated_surface_generator.cpp:28-47).  Of the two sources the reference keeps beside it
 * without building them (CMakeLists.txt:99 compiles only excavated_surface_generator.cpp):
 * excavated_surface_generator_L.cpp is this algorithm except that the wall points of /excavation_area stand
 * vertically above the edge (no slope offset, its :399-400) — area_walls_vertical = 1 — and its default offsets are
 * (5.0, 0.0); excavated_surface_genaerator_square.cpp is an earlier rectangle-only formulation with another lattice
 * enumeration and is not restated. */
typedef struct plc_excavation_params {
    double depth;                  /* excavation.depth = 1.0 */
    double slope_angle_deg;        /* excavation.slope_angle = 75.0 */
    double offset_x, offset_y;     /* excavation.offset_x = 4.0, offset_y = 1.0 (zx120/base_link frame) */
    double point_density;          /* excavation.point_density = 0.05 */
    double terrain_search_radius;  /* excavation.terrain_search_radius = 0.5 */
    double arm1_length, arm1_width, arm2_length, arm2_width;   /* 2.0, 1.2, 2.0, 1.2 */
    double width, length;          /* rectangle mode: 1.2, 1.8 */
    int32_t l_shape_enabled;       /* excavation.l_shape_enabled = true */
    int32_t area_walls_vertical;   /* 0: the built node (:410-444); 1: excavated_surface_generator_L.cpp:391-407 */
} plc_excavation_params;
void plc_excavation_params_default(plc_excavation_params* p);

typedef struct plc_excavation_info {
    double center[3];              /* excavation centre in the map frame; z = terrain height there (:289-297) */
    double yaw;                    /* yaw of zx120/base_link in the map frame (:299-302) */
    uint64_t n_kept;               /* input points that survive the carve (:470-472) */
    uint64_t n_surface;            /* generated bottom + slope points appended to them (:487-584) */
    uint64_t n_area;               /* points of the /excavation_area cloud (:350-455) */
} plc_excavation_info;

/* Number of points generateExcavatedSurface and generateExcavationArea produce for these parameters (they depend
 * on the parameters only), for sizing the output buffers: n_terrain <= matched points + *n_surface. */
plc_status plc_excavation_counts(const plc_excavation_params* prm, uint64_t* n_surface, uint64_t* n_area);

/* Replaces matchedCloudCallback's data path (excavated_surface_generator.cpp:263-325): getTerrainHeight per
 * point (mean z of the points within terrain_search_radius horizontally, nearest point otherwise, :183-225),
 * the carve test (:327-348, :457-472), the synthetic pit surface (:487-584) and the /excavation_area cloud
 * (:350-455).  Upstream rebuilds a KD-tree inside every getTerrainHeight call; here the cloud is bucketed once.
 * matched: the /matched_point_cloud message (PointXYZRGB); off_rgb: byte offset of its packed rgb field, or
 * UINT32_MAX when the message has none (pcl::fromROSMsg then leaves r = g = b = 0, a = 255).
 * base_t, base_q: the transform map <- zx120/base_link (translation, quaternion x y z w) from TF.
 * out_terrain / out_area: PointXYZRGB wire layout, 32 bytes per point. */
plc_status plc_excavate(plc_ctx* ctx, const plc_cloud_view* matched, uint32_t off_rgb,
                        const double* base_t, const double* base_q, const plc_excavation_params* prm,
                        void* out_terrain, uint64_t cap_terrain, uint64_t* n_terrain,
                        void* out_area, uint64_t cap_area, uint64_t* n_area, plc_excavation_info* info);

/* ---------------------------------------------------------------- calc_drivable_area -- */

/* Parameters of the node (calc_drivable_area.cpp:21-26). */
typedef struct plc_drivable_params {
    double grid_resolution;       /* 1.0   */
    double map_width, map_height; /* 100.0 */
    double max_gradient;          /* 0.3   */
    double start_clear_radius;    /* 3.0   */
    int32_t min_points_per_cell;  /* 10    */
    int32_t reserved;
} plc_drivable_params;
void plc_drivable_params_default(plc_drivable_params* p);

/* nav_msgs/OccupancyGrid info of the result (:167-176). */
typedef struct plc_drivable_info {
    int32_t width, height;        /* (int)(map_width / resolution), (int)(map_height / resolution)  (:36-37) */
    double origin_x, origin_y;    /* robot position minus half the map (:137-138) */
    double resolution;
    uint64_t n_cells;             /* width * height */
    int32_t published;            /* 0: empty scan, the reference returns without publishing (:108-112) */
    int32_t reserved;
} plc_drivable_info;

/* Replaces robotCloudCallback's data path (calc_drivable_area.cpp:87-232): the scan is transformed into the map
 * frame (t, q = map <- scan frame as tf2::doTransform applies it in float32; both NULL: the scan is in the map
 * frame already), its z values are binned into the robot-centred grid, and each cell becomes 0 (free: within
 * start_clear_radius of the start position, or gradient (max z - min z) / resolution <= max_gradient),
 * 100 (gradient above the limit) or -1 (fewer than min_points_per_cell points).
 * out_grid: row-major int8[height * width] (OccupancyGrid.data). */
plc_status plc_drivable_area(plc_ctx* ctx, const plc_cloud_view* scan, const double* t, const double* q,
                             double robot_x, double robot_y, double start_x, double start_y,
                             const plc_drivable_params* prm, int8_t* out_grid, uint64_t capacity,
                             plc_drivable_info* info);

/* plc_drivable_area without any allocation.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments, same grid as plc_drivable_area, with these differences:
 *   - the per-cell (count, min z, max z) scratch is a static __device__ array inside the library, sized for
 *     PLC_DRIVABLE_STATIC_MAX_CELLS cells (12 MiB that every GPU loading the library holds until exit); a grid
 *     with more cells is refused with PLC_E_CAPACITY (plc_last_error names the limit) and nothing is written;
 *   - scan->data and out_grid must be device memory of the context's device (plc_dev_alloc or the caller's own):
 *     no staging buffer exists, so host pointers are PLC_E_INVALID_ARG;
 *   - the three kernels are enqueued on the context's stream and the call returns without synchronising:
 *     out_grid is complete after plc_ctx_sync (or after whatever the caller orders behind that stream);
 *   - no cudaMalloc / cudaFree / allocating routine is called, no events are recorded (plc_last_device_times
 *     is not updated).  The first call on a device resolves the array's device address (cudaGetSymbolAddress).
 * Every call on a device shares the one array, which keeps the last call's contents (each call resets the cells
 * it uses before binning): calls that overlap on the GPU — two contexts or streams on one device — race; order
 * them on one stream. */
#define PLC_DRIVABLE_STATIC_MAX_CELLS 1048576u
plc_status plc_drivable_area_static(plc_ctx* ctx, const plc_cloud_view* scan, const double* t, const double* q,
                                    double robot_x, double robot_y, double start_x, double start_y,
                                    const plc_drivable_params* prm, int8_t* out_grid, uint64_t capacity,
                                    plc_drivable_info* info);

#ifdef __cplusplus
}
#endif
#endif /* PLC_H_ */
