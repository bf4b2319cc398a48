"""Runs the stages around the sweep once more after a warm-up so that an ncu launch list shows their kernels:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/feeders_profile.py [sweep] [graph]
Default: config 4's filter -> merge -> sweep (nodes.run_pipeline).  With `graph`: the whole launch graph
(nodes.run_launch_graph: filter -> merger -> terrain generator -> normals / cells / candidates -> sweep, drivable grid)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_processor_b200 import Context, scenes  # noqa: E402
from pointcloud_processor_b200.nodes import run_launch_graph, run_pipeline  # noqa: E402

inp = scenes.pipeline_inputs(n_scan=1_000_000, sweep=sys.argv[1] if len(sys.argv) > 1 else "cfg1")
graph = len(sys.argv) > 2 and sys.argv[2] == "graph"
with Context(0) as ctx:
    for _ in range(2):
        got = run_launch_graph(ctx, inp, num_candidates=32 * 32, grid_resolution=0.05, vertical_layers=20) if graph else run_pipeline(ctx, inp)
    print(got["device_ms"], {k: v.shape for k, v in got.items() if hasattr(v, "shape")})
