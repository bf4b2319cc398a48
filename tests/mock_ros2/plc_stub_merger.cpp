// plc_stub_merger.cpp — test infrastructure: stands in for the three libplc.so entry points ros2/src/pointcloud_merger_node.cpp
// calls, so that the wrapper's HOST half (GNSS origin, map -> base_link broadcasts, tick gating, headers, stamps, log text)
// can be played on a machine without a GPU.  It computes nothing: the published clouds carry the right sizes and zero bytes.
// The data path itself is checked on the GPU (tests/test_reference_source.py, -m gpu) with the real library.
#include "plc_node.h"

extern "C" plc_status plc_ctx_create(int, void*, plc_ctx** out) {
    *out = reinterpret_cast<plc_ctx*>(1);
    return PLC_OK;
}
extern "C" void plc_ctx_destroy(plc_ctx*) {}
extern "C" plc_status plc_merge_clouds(plc_ctx*, const plc_merger_input* robot, const plc_merger_input* backhoe, void*, uint64_t,
                                       uint64_t* n_merged, void*, uint64_t, uint64_t* n_robot, void*, uint64_t, uint64_t* n_backhoe) {
    *n_robot = robot->has_cloud && robot->has_transform ? robot->cloud.n_points : 0;
    *n_backhoe = backhoe->has_cloud && backhoe->has_transform ? backhoe->cloud.n_points : 0;
    *n_merged = *n_robot + *n_backhoe;
    return PLC_OK;
}
