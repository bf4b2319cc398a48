"""Wall time of plc_dev_upload for a pageable buffer (PLC_STAGE_THREADS from the environment)."""
import os
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from pointcloud_processor_b200.context import Context, DevBuffer  # noqa: E402

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 16
data = np.random.default_rng(0).integers(0, 256, mb << 20, dtype=np.uint8)
with Context(0) as ctx:
    buf = DevBuffer(ctx, data.nbytes)
    ts = []
    for _ in range(12):
        t0 = time.perf_counter()
        buf.upload(data)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(os.environ.get("PLC_STAGE_THREADS", "default"), mb, "MB", "median ms", round(float(np.median(ts[2:])), 3),
          "GB/s", round(data.nbytes / np.median(ts[2:]) / 1e6, 1), "cpus", os.cpu_count(), len(os.sched_getaffinity(0)))
    buf.free()
