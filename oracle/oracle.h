/*
 * oracle.h — CPU restatement of the reference's algorithm for the virtual_lidar sweep and its
 * filter/merge feeders.  TEST INFRASTRUCTURE ONLY: nothing under oracle/ is linked into, loaded by
 * or called from the product library (pointcloud_processor_b200/).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * PARITY: pinned against the reference's own sources, unpinned on its third-party binaries.  The reference ships no
 * tests, fixtures or golden vectors and its build cannot run here (rclcpp, tf2, PCL, FLANN, Eigen,
 * GeographicLib are absent), but five of its nodes compile unmodified against the stand-in headers of
 * oracle/ref_shim/ and tests/mock_ros2/ (`make -C oracle ref` -> oracle/_ref/), and tests/golden/refsrc_*.npz hold
 * what they publish on seeded inputs (tools/make_refsrc_golden.py); tests/test_reference_source.py requires this
 * restatement to reproduce them.  The arithmetic that lives in third-party code outside /root/reference stays
 * restated, in the stand-ins as here:
 *   PCL 1.12.1  KdTreeFLANN<PointXYZRGB>::radiusSearch   (call sites virtual_lidar.cpp:298,312,611,745,782)
 *   FLANN 1.9.1 KDTreeSingleIndex + L2_Simple<float>
 *   PCL 1.12.1  NormalEstimation, VoxelGrid<PointXYZ>::applyFilter   (pointcloud_filter.cpp:129-136)
 *   tf2_sensor_msgs 0.25.x doTransform<PointCloud2>       (pointcloud_merger.cpp:370, calc_drivable_area.cpp:91)
 * (versions inferred from "ROS2 Humble", not pinned by the reference); it is also checked by self-consistency
 * (brute-force == grid == kd-tree, bit for bit), hand-checkable fixtures and independent SciPy / NumPy
 * restatements (tests/test_oracle_independent*.py).
 *
 * POD types are shared with include/plc.h (the boundary declaration) so that the same bytes feed
 * both sides.
 */
#ifndef ORACLE_H_
#define ORACLE_H_

#include <stddef.h>
#include <stdint.h>
#include "../include/plc.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- radius search -------------------------------------------------- */
enum { ORC_INDEX_BRUTE = 0, ORC_INDEX_GRID = 1, ORC_INDEX_KDTREE = 2 };

/* xyz: n packed float triples.  Non-finite points are dropped (KdTreeFLANN::setInputCloud). */
void* orc_index_create(const float* xyz, uint64_t n, int kind);
void orc_index_destroy(void* index);
uint64_t orc_index_size(const void* index);
/* Number of indexed points p with  d2_f32(q,p) < (float)(radius*radius). */
int64_t orc_index_radius_count(const void* index, float qx, float qy, float qz, double radius);

/* Unpack a PointCloud2 byte buffer into packed float xyz (pcl::fromROSMsg field mapping, §A.5). */
void orc_unpack_cloud(const plc_cloud_view* v, float* xyz_out);

/* ---- sweep (virtual_lidar.cpp:360-398, 454-475, 627-800) --------------------------------- */
typedef struct orc_sweep_args {
    const void* terrain;        /* radius index or NULL: "no terrain tree" (:721,:727) */
    const void* observed;       /* radius index or NULL: "no zx120 cloud" (:719) */
    const plc_cell* cells;
    uint64_t n_cells;
    const plc_params* prm;
    const plc_pose* fixed;
    const plc_pose* cands;
    uint64_t n_cands;
    /* faithful != 0: keep the reference's loop structure for timing — the fixed sensor is
     * re-evaluated inside every candidate (F9, :637) and every radius search materialises the
     * index/distance vectors (:779-783).  Results are identical either way. */
    int32_t faithful;
    int32_t threads;            /* 1 = sequential (flags carry-over defined); >1 = OpenMP over candidates */
    /* outputs (nullable unless noted) */
    double* scores;             /* [n_cands] */
    int32_t* covered;           /* [n_cands] */
    uint32_t* mask;             /* [n_cands * ceil(n_cells/32)] */
    uint8_t* cell_flags;        /* [n_cells] in/out, PLC_FLAG_*; only touched when threads == 1 */
    double* fixed_scores;       /* [n_cells] */
    int64_t best_index;         /* out */
    double best_score;          /* out */
    int32_t best_covered;       /* out */
    double fixed_total;         /* out */
    plc_counters counters;      /* out */
} orc_sweep_args;

int orc_vl_sweep(orc_sweep_args* a);

/* ---- setup restatements (shared inputs) ---------------------------- */
typedef struct orc_grid_bounds {
    double min_x, max_x, min_y, max_y, min_z, max_z;   /* after the +-res margin (:251-254) */
    int32_t width, height;
} orc_grid_bounds;

/* generateExcavationGrid3D (:236-287) + computeCellSurfaceNormal (:301-340).
 * normals: n_area packed float triples (per-point normals) or NULL (cells keep (0,0,1)).
 * Returns the number of valid cells; writes at most `capacity`. */
int64_t orc_generate_cells(const float* area_xyz, uint64_t n_area, const float* normals,
                           double res, int32_t layers, double normal_radius,
                           plc_cell* out, uint64_t capacity, orc_grid_bounds* bounds);

/* computeTerrainNormals (:209-234): PCA normals, radius search, viewpoint (0,0,0), then n_z >= 0. */
void orc_compute_normals(const float* area_xyz, uint64_t n_area, double radius, float* normals_out);

/* getGroundHeight (:600-625) */
double orc_ground_height(const void* terrain_index, const float* terrain_xyz, double x, double y);

/* generateCandidatePositions (:550-598).  terrain_xyz must be the array the index was built from
 * with non-finite points already removed.  Returns count; writes at most capacity. */
int64_t orc_generate_candidates(const void* terrain_index, const float* terrain_xyz,
                                const orc_grid_bounds* b, const plc_pose* fixed,
                                int32_t num_candidates, double search_radius, double sensor_height,
                                plc_pose* out, uint64_t capacity);

/* ---- filter / merge (pointcloud_filter.cpp:87-139, pointcloud_merger.cpp:354-387) ---------- */
/* cropFrontArea: returns kept count; out_xyz packed triples in input order; out_index (nullable) source indices. */
int64_t orc_crop(const float* xyz, uint64_t n, const plc_crop_box* box, float* out_xyz, int64_t* out_index);
/* pcl::VoxelGrid<PointXYZ> (§A.3 item 13).  Returns number of output points, or -1 when the
 * voxel count overflows int32 (PCL warns and returns the input unchanged).
 * out_xyz: centroids in ascending voxel-index order.  out_key (nullable): the voxel index of each
 * output.  out_count (nullable): member count.  point_key (nullable, n entries): voxel index per input. */
int64_t orc_voxel_grid(const float* xyz, uint64_t n, double leaf, float* out_xyz,
                       int32_t* out_key, int32_t* out_count, int32_t* point_key);
/* tf2::doTransform float affine (§A.3 item 14): t (3 doubles) and quaternion x,y,z,w narrowed to float. */
void orc_transform(const float* xyz, uint64_t n, const double* t, const double* q, float* out_xyz);

/* excavated_surface_generator.cpp:263-584: the generator node's data path.  xyz / rgba: the /matched_point_cloud
 * as a PointXYZRGB cloud (rgba NULL: a = 255, r = g = b = 0).  Outputs are 32-byte PointXYZRGB wire records;
 * out_terrain needs n + n_surface records. */
void orc_excavation_frame(const double* t, const double* q, double offset_x, double offset_y, double* center_xy, double* yaw);
void orc_excavate(const float* xyz, const uint32_t* rgba, uint64_t n, const double* base_t, const double* base_q,
                  const plc_excavation_params* prm, void* out_terrain, uint64_t* n_terrain, void* out_area,
                  uint64_t* n_area, plc_excavation_info* info);

/* calc_drivable_area.cpp:36-37, 87-232: the occupancy grid (row-major int8, width x height).  t, q: map <- scan
 * frame (NULL: none).  Returns width*height, or -1 for an empty scan (nothing is published, :108-112). */
int64_t orc_drivable_area(const float* xyz, uint64_t n, const double* t, const double* q, double robot_x,
                          double robot_y, double start_x, double start_y, double grid_resolution, double map_width,
                          double map_height, double max_gradient, int min_points_per_cell, double start_clear_radius,
                          int8_t* out, int32_t* width_out, int32_t* height_out);

#ifdef __cplusplus
}
#endif
#endif
