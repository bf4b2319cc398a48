"""plc_filter / plc_merge / plc_vl_set_terrain at sizes where launch latency amortises (bench.py feeders_large), alone:
    python tools/dev/filter_large.py            (for an ncu launch list)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
from pointcloud_processor_b200 import Context
with Context(0) as ctx:
    print(json.dumps(bench.measure_feeders_large(ctx, bench.measured_peaks()[0], reps=3)))
