/*
 * plc_host.h — host-side (CPU, non-hot) steps of the virtual_lidar node that produce the sweep's
 * inputs, exported so that headless drivers (bench.py, pytest) and the ROS2 wrapper share one
 * implementation.  These are the reference's setup callbacks, not the hot path:
 *   plc_host_generate_cells       generateExcavationGrid3D + computeCellSurfaceNormal (virtual_lidar.cpp:236-340)
 *   plc_host_compute_normals      computeTerrainNormals (virtual_lidar.cpp:209-234)
 *   plc_host_generate_candidates  generateCandidatePositions + getGroundHeight (virtual_lidar.cpp:550-625)
 * Same conventions as plc.h (status codes, caller-owned buffers, capacity checks).
 * The node runs the device versions of the same three steps (plc_vl_generate_cells, plc_vl_compute_normals,
 * plc_vl_generate_candidates in plc.h); these host versions remain for callers without a context (scene
 * generation in tests and tools) and as a second implementation the device versions are checked against.
 */
#ifndef PLC_HOST_H_
#define PLC_HOST_H_

#include "plc.h"

#ifdef __cplusplus
extern "C" {
#endif

/* plc_grid_bounds is declared in plc.h (the device versions plc_vl_generate_* share it). */

/* normals: one float triple per area point (NaN allowed) or NULL => every cell keeps (0,0,1).
 * *n_out receives the number of valid cells even when it exceeds capacity (then PLC_E_CAPACITY). */
plc_status plc_host_generate_cells(const plc_cloud_view* area, const float* normals,
                                   double grid_resolution, int32_t vertical_layers,
                                   double normal_search_radius,
                                   plc_cell* out, uint64_t capacity, uint64_t* n_out,
                                   plc_grid_bounds* bounds);

plc_status plc_host_compute_normals(const plc_cloud_view* area, double radius, float* normals_out);

plc_status plc_host_generate_candidates(const plc_cloud_view* terrain, const plc_grid_bounds* bounds,
                                        const plc_pose* fixed, int32_t num_candidates,
                                        double search_radius, double sensor_height,
                                        plc_pose* out, uint64_t capacity, uint64_t* n_out);

#ifdef __cplusplus
}
#endif
#endif
