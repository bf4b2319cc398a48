/*
 * plc_node.h — ROS-free cores of the three reference nodes whose data-parallel bodies run on the GPU.
 *
 * The reference's only stable interface is its ROS2 node surface (executables, topics, parameters,
 * TF frames).  Each core below holds the state and the callbacks of one rclcpp::Node subclass,
 * with every rclcpp/tf2/PCL type replaced by a POD: a PointCloud2 message is its `data` buffer
 * (plc_cloud_view), a TF lookup is its result, a log block is text.  The rclcpp wrappers in ros2/
 * only move messages in and out of these calls, so the node logic is testable without ROS.
 *
 *   plc_vlnode_*   class SimplifiedDualLidarOptimizer   src/virtual_lidar.cpp:60-970
 *   plc_filter_scan                                     SimplifiedScanMatcher::processCloudSimple, src/pointcloud_filter.cpp:64-85
 *   plc_merge_clouds                                    GnssGicpMatcher::processPointClouds, src/pointcloud_merger.cpp:308-352
 *
 * Same conventions as plc.h: status codes, caller-owned buffers, not thread-safe, no CPU fallback.
 */
#ifndef PLC_NODE_H_
#define PLC_NODE_H_

#include "plc.h"
#include "plc_host.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ virtual_lidar node -------- */

typedef struct plc_vlnode plc_vlnode;

/* Cell colours of the statistics blocks (virtual_lidar.cpp:387-397, :487-501). */
typedef struct plc_color_stats {
    int32_t total, green, red, blue, yellow;
} plc_color_stats;

typedef struct plc_vlnode_result {
    int32_t ran;                 /* 0: the guard at :455 returned early (no grid / no terrain / no TF) */
    int32_t n_candidates;        /* candidates.size() after the exclusions of :570-590 */
    plc_pose best;               /* best_mobile_position_; LidarPosition() = (0,0,10,-pi/2,0) if none (Q5) */
    double best_score;           /* -inf if no candidate */
    plc_pose fixed;              /* zx120_lidar_position_ (:342-352) */
    double fixed_total;          /* "Total Score (ZX120 only)" */
    int32_t range_ok, fov_ok, visibility_ok;   /* debug counters of evaluateZX120Only (:379-382) */
    int32_t n_cells;
    plc_color_stats zx120_only;  /* :387-397 */
    plc_color_stats dual;        /* :487-501, from the flags the candidate loop leaves behind (Q3) */
    uint64_t observed_points;    /* zx120_cloud_->size() (:433) */
} plc_vlnode_result;

/* The node's constructor (:62-97) minus rclcpp: parameters at their declared defaults. */
plc_status plc_vlnode_create(plc_ctx* ctx, plc_vlnode** out);
/* The same node over a multi-GPU group (plc.h, plc_group_*): every sweep is sharded over the group's devices. */
plc_status plc_vlnode_create_on_group(plc_group* group, plc_vlnode** out);
void plc_vlnode_destroy(plc_vlnode* node);

/* declare_parameter/get_parameter surface (:66-71): grid_resolution, sensor_height, search_radius,
 * max_distance (double); num_candidates, vertical_layers (int).  Values take effect the way they
 * do upstream: at the next runOptimization (updateParameters, :457).  Unknown name: INVALID_ARG. */
plc_status plc_vlnode_set_parameter_double(plc_vlnode* node, const char* name, double value);
plc_status plc_vlnode_set_parameter_int(plc_vlnode* node, const char* name, int64_t value);
/* updateParameters() (:154-162): makes the values set so far the ones the callbacks use.  The reference calls it at the
 * end of its constructor (:73), so that parameter overrides given at launch hold from the first message on, and at
 * the start of every runOptimization (:457), which plc_vlnode_run_optimization does itself.  A wrapper calls this once,
 * after handing over the declared parameter values. */
plc_status plc_vlnode_update_parameters(plc_vlnode* node);

/* Subscription callbacks (:164-207).  `cloud` is the message's data buffer with the offsets of its
 * x/y/z fields (pcl::fromROSMsg matches fields by name). */
plc_status plc_vlnode_excavation_area_callback(plc_vlnode* node, const plc_cloud_view* cloud);
plc_status plc_vlnode_terrain_callback(plc_vlnode* node, const plc_cloud_view* cloud);
plc_status plc_vlnode_zx120_points_callback(plc_vlnode* node, const plc_cloud_view* cloud);

/* Result of tf_buffer_->lookupTransform("map", "zx120/base_link") (:345): translation, or
 * available == 0 for a tf2::TransformException (the cycle is then skipped, :353-356). */
plc_status plc_vlnode_set_base_transform(plc_vlnode* node, int32_t available, double x, double y, double z);

/* Timer callback runOptimization (:454-548). */
plc_status plc_vlnode_run_optimization(plc_vlnode* node, plc_vlnode_result* out);

/* What the publishers read after a run: candidate_positions_ (:478) and the grid cells with their
 * flag bytes (PLC_FLAG_*), for the marker arrays (:813-962).  *n receives the count even when it
 * exceeds capacity (then PLC_E_CAPACITY). */
plc_status plc_vlnode_get_candidates(const plc_vlnode* node, plc_pose* out, uint64_t capacity, uint64_t* n);
plc_status plc_vlnode_get_cells(const plc_vlnode* node, plc_cell* cells, uint8_t* flags, uint64_t capacity, uint64_t* n);
plc_status plc_vlnode_get_bounds(const plc_vlnode* node, plc_grid_bounds* out);

/* The RCLCPP_INFO text produced since the last call (one line per message, '\n' separated):
 * ":284-285" grid size, ":204" ZX120 cloud size, ":419-451" and ":522-543" statistics blocks.
 * Returns a pointer owned by the node, valid until the next call on it. */
const char* plc_vlnode_take_log(plc_vlnode* node);

/* ------------------------------------------------------------------ pointcloud_filter node ---- */

/* processCloudSimple (pointcloud_filter.cpp:64-85) for one scan: crop with
 * (0, front) x (-side, side) x (-1.5, height), then VoxelGrid with `voxel_leaf_size`.
 * out_xyz16: PointXYZ wire layout (point_step 16).  An empty input or an empty crop publishes an
 * empty cloud (*n_out = 0), as upstream does (:125-127, Q7). */
plc_status plc_filter_scan(plc_ctx* ctx, const plc_cloud_view* scan, double front_range, double side_range,
                           double height_range, double voxel_leaf_size,
                           void* out_xyz16, uint64_t capacity, uint64_t* n_out);

/* ------------------------------------------------------------------ pointcloud_merger node ---- */

/* One source of processPointClouds (pointcloud_merger.cpp:308-352): the latest filtered cloud and
 * the result of lookupTransform("map", "<name>/velodyne_link") (:362-366). */
typedef struct plc_merger_input {
    plc_cloud_view cloud;     /* n_points == 0 with has_cloud == 0: no message received yet */
    int32_t has_cloud;
    int32_t has_transform;    /* 0: tf2::TransformException => source skipped with a warning (:389-393) */
    double t[3];
    double q[4];              /* x, y, z, w */
} plc_merger_input;

/* robot is coloured red (255,0,0) and comes first, backhoe blue (0,0,255) second (:316-325).
 * out_* are PointXYZRGB wire layout (point_step 32).  Returns PLC_OK with *n_merged = 0 when
 * neither source contributes (upstream publishes nothing in that case, :327). */
plc_status plc_merge_clouds(plc_ctx* ctx, const plc_merger_input* robot, const plc_merger_input* backhoe,
                            void* out_merged, uint64_t cap_merged, uint64_t* n_merged,
                            void* out_robot, uint64_t cap_robot, uint64_t* n_robot,
                            void* out_backhoe, uint64_t cap_backhoe, uint64_t* n_backhoe);

#ifdef __cplusplus
}
#endif
#endif /* PLC_NODE_H_ */
