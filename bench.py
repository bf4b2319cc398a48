#!/usr/bin/env python
"""bench.py — the virtual_lidar visibility sweep on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W                  # our arm
    python -m torch.distributed.run ... bench.py --gpus N ...      # N > 1, launched by the driver
    python bench.py --impl reference --gpus N --steps K --warmup W # the reference's CPU path (restated)

One step = one full sweep of the workload: the fixed-sensor pass, every (candidate pose, target
cell) ray of this rank's candidate shard, per-candidate scores, argmax and — for N > 1 — the
16-byte-per-rank all-gather.  Metric: rays/s over the whole job (BASELINE.json).  Prints ONE JSON
line on rank 0.  See DESIGN.md §6 for every field.

Workload: N = 1 — BASELINE.json configs[1] (cfg2: rectangular pit, 200 k points, 32x32x2 poses, 32x1800 rays per
pose), plus a `north_star_cfg3` block.  N > 1 — configs[2] (cfg3: L-shape, 1 M points, 64x64x4 poses, 32x1800 rays),
the configuration the north star's 50 ms target is quoted on, strong-scaled over the N ranks ("scaling": "strong");
extra blocks: `breakdown` (k_sweep / fixed pass / finalize+argmax / all-gather / rank imbalance), `parity` (sharded
against unsharded on rank 0), `group` (plc_group_* from one process against the rank-sharded result),
`weak_cfg2_shard` (round 1's weak-scaling line) and, at N = 8, `cfg5` (configs[4]).
`e2e.value` is measured with PAGEABLE host buffers and the node core's outputs (flags + fixed scores);
`e2e.variants` has the pinned and scores-only figures beside it.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, os.fspath(ROOT))

METRIC = "candidate_beam_rays_per_sec"
UNIT = "rays/s"
WORKLOADS = {
    # BASELINE.json configs -> scenes.CONFIGS
    "cfg1": "L-shaped excavation, 50k pts, 16x16x1 poses, 16x900 rays/pose",
    "cfg2": "rectangular excavation, 200k pts, 32x32x2 poses, 32x1800 rays/pose",
    "cfg3": "L-shaped excavation, 1M pts, 64x64x4 poses, 32x1800 rays/pose",
    "cfg5": "L-shaped excavation, 4M pts, 128x128x8 poses, 64x2048 rays/pose",
}


def default_workload(n_gpus: int) -> str:
    """N = 1: BASELINE.json configs[1] ("... on 1xB200").  N > 1: configs[2] ("... sharded across 2/4/8 B200"), the
    configuration the north star's 50 ms target is quoted on, strong-scaled."""
    return "cfg2" if n_gpus <= 1 else "cfg3"


def build_scene(workload: str, world: int, weak: bool = False, setup=None):
    """`weak`: configs[1] with a fixed shard per rank (the pose lattice side grows with sqrt(N))."""
    from pointcloud_processor_b200 import scenes
    kw = {}
    if setup is not None:
        kw.update(setup=setup, layers=scenes.LAYERS_FOUND[workload])
    if weak and world > 1:
        kw["px"] = int(round(32 * np.sqrt(world)))
    return scenes.build(workload, **kw)


def workload_text(workload: str, world: int, weak: bool = False) -> str:
    text = WORKLOADS.get(workload, workload)
    if weak and world > 1:
        side = int(round(32 * np.sqrt(world)))
        text = text.replace("32x32x2 poses", f"{side}x{side}x2 poses (32x32x2 per GPU x {world})")
    return f"{workload}: {text}"


class ClockSampler:
    """Samples SM clocks and throttle reasons through NVML (every few ms, in a thread) while the timed
    region runs; falls back to an `nvidia-smi -lms` child process if NVML cannot be loaded."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index: int):
        self.sm, self.mx, self.reasons = [], [], set()
        self.proc = None
        self.thread = None
        self.stop = threading.Event()
        self.gpu = gpu_index
        self.nvml = None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return self
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read_smi, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.mx.append(float(n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)))
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        mask = int(get(self.handle))
        for name, bit in self.BITS.items():
            if mask & bit:
                self.reasons.add(name)

    def _poll_nvml(self):
        while not self.stop.is_set():
            try:
                self._sample_nvml()
            except Exception:
                return
            self.stop.wait(0.004)

    def _read_smi(self):
        for line in self.proc.stdout:
            r = [c.strip() for c in line.split(",")]
            try:
                self.sm.append(float(r[0]))
                self.mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(self.NAMES, r[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)

    def __exit__(self, *exc):
        if self.nvml:
            try:
                self._sample_nvml()          # at least one sample while the last step's kernels are still recent
            except Exception:
                pass
            self.stop.set()
            self.thread.join(timeout=2)
        elif self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self) -> dict:
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": float(max(self.mx)),
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nvml else "nvidia-smi"}


def measured_peaks() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except (KeyError, ValueError):
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def algorithmic_bytes(rays: int, samples: int, n_cands: int, grid_bytes: int) -> int:
    """32 B per ray (target record) + 4 B per evaluated sample (one occupancy word)
    + 48 B read and 16 B written per candidate + the grid structure once."""
    return 32 * rays + 4 * samples + 64 * n_cands + grid_bytes


def oracle_setup():
    """(generate_cells, generate_candidates) of oracle/vl_setup.cpp with the signatures scenes.build expects:
    the reference arm builds its scene with these, so it never loads the product library."""
    from oracle import binding as O
    index = {}

    def gen_cells(area, normals, res, layers, normal_radius):
        return O.generate_cells(area, normals, res, layers, normal_radius)

    def gen_cands(terrain, bounds, fixed, num_candidates, search_radius, sensor_height):
        if id(terrain) not in index:
            index.clear()
            index[id(terrain)] = O.Index(terrain, O.GRID)
        return O.generate_candidates(index[id(terrain)], bounds, fixed, num_candidates, search_radius, sensor_height)

    return gen_cells, gen_cands


def cpu_reference_sample(scene, budget_s: float, threads: int, seed: int = 1234):
    """Times the CPU restatement of the reference's sweep (KD-tree index, reference loop structure
    incl. the per-candidate fixed-sensor re-evaluation) on a seeded subset of candidates sized to
    take about `budget_s`.  Returns (rays_per_s, n_candidates_timed, seconds)."""
    from oracle import binding as O
    ix = O.Index(scene.terrain, O.KDTREE)
    rng = np.random.default_rng(seed)
    order = rng.permutation(scene.candidates.size)
    M = scene.cells.size
    # probe with a few candidates to size the sample
    n_probe = min(max(threads, 2), order.size)
    t0 = time.perf_counter()
    O.sweep(ix, None, scene.cells, scene.params, scene.fixed, scene.candidates[order[:n_probe]], faithful=True,
            threads=threads, want_mask=False)
    t_probe = time.perf_counter() - t0
    per_cand = t_probe / n_probe
    n = int(min(order.size, max(n_probe, budget_s / max(per_cand, 1e-9))))
    t0 = time.perf_counter()
    O.sweep(ix, None, scene.cells, scene.params, scene.fixed, scene.candidates[order[:n]], faithful=True,
            threads=threads, want_mask=False)
    dt = time.perf_counter() - t0
    return n * M / dt, n, dt


def reference_source_sample(scene, res: float, layers: int, k_small: int = 9, k_large: int = 49) -> dict:
    """Times the reference's OWN virtual_lidar.cpp — oracle/_ref/virtual_lidar_ref, compiled unmodified against the
    stand-in PCL / ROS2 headers where /root/reference exists (`make -C oracle ref`; the binary travels, its sources do
    not) — on this workload's terrain and area clouds: the node builds its own lattices (grid_resolution, vertical_layers
    as the scene's) and runs one optimisation cycle per process; two runs with different num_candidates are differenced,
    so that start-up, PCA normals and lattice generation cancel.  One thread, as upstream.  Informational: its radius
    search is the stand-in's grid index, not FLANN's kd-tree."""
    import tempfile
    exe = ROOT / "oracle" / "_ref" / "virtual_lidar_ref"
    if not exe.exists():
        return {"unavailable": "oracle/_ref/virtual_lidar_ref not built (needs /root/reference at build time)"}

    def records(xyz):
        r = np.zeros((xyz.shape[0], 8), np.float32)
        r[:, :3] = xyz[:, :3]
        return r

    def run(tmp, k):
        out = tmp / f"k{k}"
        out.mkdir()
        lines = [f"param num_candidates int {k}", f"param vertical_layers int {layers}", f"param grid_resolution double {res!r}"]
        for topic, cloud in (("/excavation_area", scene.area), ("/excavated_terrain", scene.terrain)):
            f = out / (topic.strip("/") + ".bin")
            records(cloud).tofile(f)
            lines.append(f"cloud {topic} {f} {cloud.shape[0]} 32 map 1 0 x:0 y:4 z:8 rgb:16")
        lines += ["tf map zx120/base_link 0 0 0 0 0 0 1", "timers"]
        (out / "script.txt").write_text("\n".join(lines) + "\n")
        env = dict(os.environ, MOCK_ROS_SCRIPT=str(out / "script.txt"), MOCK_ROS_OUT=str(out))
        t0 = time.perf_counter()
        subprocess.run([str(exe)], env=env, check=True, capture_output=True, timeout=600)
        dt = time.perf_counter() - t0
        n = {}
        for j in sorted(out.glob("[0-9][0-9][0-9].json")):
            m = json.loads(j.read_text())
            n[m["topic"]] = m.get("markers", 0)
        # candidate array: delete-all + fixed sensor + candidates + optimum; grid array: delete-all + cells
        return dt, n["/mobile_lidar_candidate_positions"] - 3, n["/excavation_grid_visualization"] - 1

    try:
        with tempfile.TemporaryDirectory() as tmp:
            t1, c1, m1 = run(Path(tmp), k_small)
            t2, c2, m2 = run(Path(tmp), k_large)
    except Exception as e:  # noqa: BLE001 — informational leg: report, do not fail the arm
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    if c2 <= c1 or t2 <= t1 or m1 != m2:
        return {"unavailable": f"differencing failed: {c1} poses in {t1:.2f} s, {c2} in {t2:.2f} s"}
    return {"value": (c2 - c1) * m2 / (t2 - t1), "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": f"({c2} - {c1}) poses x {m2} rays of the node's own lattices in ({t2:.1f} - {t1:.1f}) s, one optimisation cycle per run",
            "note": "virtual_lidar.cpp unmodified; KdTreeFLANN / NormalEstimation are the stand-ins of oracle/ref_shim (grid index, not FLANN)"}


def feeders_cpu_reference(n_scan: int = 1_000_000) -> dict:
    """SURVEY D5's CPU baseline of the filter and merger stages on BASELINE config 4's scans (two synthetic 1 M-point
    32-ring scans, crop 15/10/10, leaf 0.2, the launch files' transforms): the oracle port on one thread, and the
    reference's own pointcloud_filter.cpp (oracle/_ref, stand-in VoxelGrid) per message — the same scan delivered once and
    four times to one process, differenced.  The GPU figures beside them are `feeders.filter_*` / `feeders.merge` of the
    other arm's line."""
    import tempfile
    try:
        from oracle import binding as O
        from pointcloud_processor_b200 import _abi as A, scenes
        zt, zq = scenes.sensor_in_map(dict(xyz=(0.0, 0.0, 0.0), yaw=0.0), scenes.ZX120_VELODYNE)
        rt, rq = scenes.sensor_in_map(scenes.ROBOT_BASE, scenes.ROBOT_VELODYNE)
        scans = {"robot": scenes.lidar_scan(rt, rq, n_scan, 4, "rect"), "zx120": scenes.lidar_scan(zt, zq, n_scan, 5, "rect")}
        tfs = {"robot": (rt, rq), "zx120": (zt, zq)}
        box = A.CropBox(0.0, 15.0, -10.0, 10.0, -1.5, 10.0)
        out = {"config": f"cfg4: two {n_scan}-point 32-ring scans -> crop 15/10/10 -> voxel 0.2 -> merge", "cores": 1}
        cent = {}
        for who, scan in scans.items():
            best = float("inf")
            for _ in range(3):
                t0 = time.perf_counter()
                cropped = O.crop(scan, box)[0]
                cent[who] = O.voxel_grid(cropped, 0.2)[0]
                best = min(best, time.perf_counter() - t0)
            out[f"filter_{who}_port_ms"] = 1e3 * best
            out[f"filter_{who}_points"] = [int(scan.shape[0]), int(cropped.shape[0]), int(cent[who].shape[0])]
        t0 = time.perf_counter()
        for _ in range(20):
            merged = np.concatenate([O.transform(cent[w], *tfs[w]) for w in ("robot", "zx120")])
        out["merge_port_ms"] = 1e3 * (time.perf_counter() - t0) / 20
        out["merge_points"] = int(merged.shape[0])
        exe = ROOT / "oracle" / "_ref" / "pointcloud_filter_ref"
        if exe.exists():
            with tempfile.TemporaryDirectory() as tmp:
                tmp = Path(tmp)
                scans["robot"].tofile(tmp / "scan.bin")
                line = f"cloud /four_wheel_robot/velodyne_points {tmp / 'scan.bin'} {n_scan} 16 four_wheel_robot/velodyne 1 0 x:0 y:4 z:8"
                wall = {}
                for k in (1, 4):
                    d = tmp / f"k{k}"
                    d.mkdir()
                    (d / "script.txt").write_text("\n".join([line] * k) + "\n")
                    env = dict(os.environ, MOCK_ROS_SCRIPT=str(d / "script.txt"), MOCK_ROS_OUT=str(d))
                    t0 = time.perf_counter()
                    subprocess.run([str(exe)], env=env, check=True, capture_output=True, timeout=300)
                    wall[k] = time.perf_counter() - t0
                    n_out = json.loads(sorted(d.glob("[0-9][0-9][0-9].json"))[-1].read_text())["width"]
            out["filter_robot_reference_source_ms"] = 1e3 * (wall[4] - wall[1]) / 3
            out["filter_robot_reference_source_voxels"] = int(n_out)
            out["note"] = ("reference_source: pointcloud_filter.cpp unmodified per message (fromROSMsg + crop + VoxelGrid + toROSMsg; the mock's "
                           "read of the scan file included), pcl::VoxelGrid = the stand-in of oracle/ref_shim")
        return out
    except Exception as e:  # noqa: BLE001 — informational leg
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


class Rig:
    """One rank's context, stream and collectives."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        from pointcloud_processor_b200 import Context, exchange
        self.torch, self.dist, self.exchange = torch, dist, exchange
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the plc_* path")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.host_group = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
            # a CPU-side barrier for the stretch where rank 0 alone drives every GPU (group_block): an NCCL barrier
            # would leave a kernel spinning on the very devices rank 0 is about to use from another process
            self.host_group = dist.new_group(backend="gloo")
        # a non-default torch stream (handle != 0) made current: the library launches on it, NCCL orders
        # against it, and torch.cuda.Event records see exactly the kernels that are timed
        self.stream = torch.cuda.Stream(self.dev)
        torch.cuda.set_stream(self.stream)
        self.ctx = Context(self.local_rank, stream=self.stream.cuda_stream)
        assert self.stream.cuda_stream != 0
        self.pair = None
        self.gathered = torch.empty((self.world, 2), dtype=torch.int64, device=self.dev)
        self.flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.int32, device=self.dev)   # 256 MiB > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def host_barrier(self):
        self.torch.cuda.synchronize(self.dev)
        if self.world > 1:
            self.dist.barrier(group=self.host_group)

    def max_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_values(self, v: float) -> list:
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        out = self.torch.empty(self.world, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_gather_into_tensor(out, t)
        else:
            out.copy_(t)
        return [float(x) for x in out.cpu().numpy()]

    def load(self, scene):
        """Resident inputs: terrain index, cells, this rank's round-robin share of the pose list (rank r evaluates
        poses r, r + N, ...: neighbouring poses cost about the same, contiguous blocks of the lattice do not)."""
        self.ctx.set_terrain(scene.terrain)
        self.ctx.set_cells(scene.cells)
        self.ctx.upload_candidates_strided(scene.params, scene.fixed, scene.candidates, self.rank, self.world)
        self.pair = self.exchange.result_pair_tensor(self.ctx)

    def step(self, events=None):
        """One sweep of the resident workload + the 16-byte exchange.  `events` = (mid,) records between the two."""
        self.ctx.run_async(False, False)
        if events is not None:
            events[0].record(self.stream)
        if self.world > 1:
            self.exchange.all_gather_best(self.pair, self.gathered)
        else:
            self.gathered.copy_(self.pair.view(1, 2))

    def timed(self, steps: int, warmup: int, clocks: bool = False) -> dict:
        """`steps` device-timed sweeps of the resident workload, L2 flushed before each (outside the event pairs)."""
        E = lambda: self.torch.cuda.Event(enable_timing=True)
        for _ in range(warmup):
            self.step()
        self.barrier()
        ev = [(E(), E(), E()) for _ in range(steps)]
        launches0 = self.ctx.launch_count
        k_ms, fixed_ms, reduce_ms = [], [], []
        sampler = ClockSampler(self.local_rank) if clocks else None
        if sampler:
            sampler.__enter__()
        self.barrier()
        for a, m, b in ev:
            self.flush.fill_(1)
            a.record(self.stream)
            self.step((m,))
            b.record(self.stream)
            if steps <= 64:                       # per-launch duration of the dominant kernel and of what surrounds it
                t = self.ctx.device_times()
                k_ms.append(t["sweep_kernel_ms"])
                fixed_ms.append(t["sweep_fixed_ms"])
                reduce_ms.append(t["sweep_reduce_ms"])
        self.barrier()
        if sampler:
            sampler.__exit__()
        launches = self.ctx.launch_count - launches0
        step_ms = [a.elapsed_time(b) for a, _, b in ev]
        gather_ms = float(np.mean([m.elapsed_time(b) for _, m, b in ev]))
        ms_per_step = self.max_over_ranks(sum(step_ms)) / steps
        k_local = float(np.mean(k_ms)) if k_ms else float(np.mean(step_ms))
        k_all = self.all_values(k_local)
        g_all = self.all_values(gather_ms)
        return {"ms_per_step": ms_per_step, "launches": int(launches), "k_sweep_ms_local": k_local,
                "clocks": sampler.summary() if sampler else None,
                "breakdown": {"k_sweep_ms": max(k_all), "fixed_ms": self.max_over_ranks(float(np.mean(fixed_ms)) if fixed_ms else 0.0),
                              "finalize_argmax_ms": self.max_over_ranks(float(np.mean(reduce_ms)) if reduce_ms else 0.0),
                              # the rank that arrives last does not wait in the collective: its time is the exchange itself
                              "allgather_ms": min(g_all),
                              "rank_imbalance_ms": max(k_all) - float(np.mean(k_all)),
                              "k_sweep_ms_per_rank": [round(x, 4) for x in k_all]},
                "best": self.exchange.pick_best(self.gathered)}


def measure_e2e(rig: Rig, scene, steps: int) -> dict:
    """The same sweep through the C ABI with HOST buffers every step: plc_vl_set_terrain (H2D + index build) +
    plc_vl_set_cells + plc_vl_sweep_strided (H2D poses, kernels, D2H results), wall clock, max over ranks.
      pinned           page-locked inputs, scores out (round 1's figure)
      pageable         plain malloc'ed inputs — what a ROS2 subscriber's PointCloud2::data is
      node_equivalent  pageable inputs + the outputs plc_vlnode_run_optimization asks for (per-cell flags and fixed
                       scores: k_flags runs, node_core.cpp) — the headline `e2e.value`"""
    torch, ctx = rig.torch, rig.ctx
    C, M = int(scene.candidates.size), int(scene.cells.size)
    n_local = len(range(rig.rank, C, rig.world))
    keep = []

    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
        keep.append(t)
        return t.numpy().view(a.dtype).reshape(a.shape)

    def run(bufs, node: bool) -> float:
        terr, cells, cands = bufs
        rig.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            ctx.set_terrain(terr)
            ctx.set_cells(cells)
            r = ctx.sweep_strided(scene.params, scene.fixed, cands, rig.rank, rig.world, scores=True, flags=node,
                                  fixed_scores=node)
            if rig.world > 1:
                rig.exchange.all_gather_best(rig.exchange.pack_pair(r.best_score, r.best_index, rig.dev), rig.gathered)
                torch.cuda.synchronize(rig.dev)
        return C * M / (rig.max_over_ranks(time.perf_counter() - t0) / steps)

    pin = (pinned(scene.terrain), pinned(scene.cells), pinned(scene.candidates))
    page = (np.array(scene.terrain, copy=True), np.array(scene.cells, copy=True), np.array(scene.candidates, copy=True))
    run(page, True)                                    # warm-up of the staging buffers
    v_pinned, v_page, v_node = run(pin, False), run(page, False), run(page, True)
    h2d = scene.terrain.nbytes + scene.cells.nbytes + scene.candidates.nbytes + 40 + 96
    d2h = n_local * 12 + 112 + M * 9                  # scores + covered, result block, flags + fixed scores
    # sweep latency as SURVEY D1 defines it: wall time of the sweep call with terrain and cells resident,
    # host pose list in, NCCL step and D2H of the results included
    rig.barrier()
    walls = []
    for _ in range(max(3, steps)):
        rig.barrier()
        t0 = time.perf_counter()
        r = ctx.sweep_strided(scene.params, scene.fixed, page[2], rig.rank, rig.world, scores=True)
        if rig.world > 1:
            rig.exchange.all_gather_best(rig.exchange.pack_pair(r.best_score, r.best_index, rig.dev), rig.gathered)
            torch.cuda.synchronize(rig.dev)
        walls.append(time.perf_counter() - t0)
    wall = rig.max_over_ranks(float(np.median(walls))) * 1e3
    return {"value": v_node, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": steps,
            "path": "plc_vl_set_terrain + plc_vl_set_cells + plc_vl_sweep_strided with pageable host buffers, per-cell flags "
                    "and fixed scores out (what the node core asks for)",
            "variants": {"pinned": v_pinned, "pageable": v_page, "node_equivalent": v_node},
            "sweep_wall_ms_resident": wall}


def extra_block(rig: Rig, scene, workload: str, steps: int, weak: bool = False, scaling: str = "strong") -> dict:
    rig.load(scene)
    t = rig.timed(steps, 2)
    C, M = int(scene.candidates.size), int(scene.cells.size)
    return {"workload": workload_text(workload, rig.world, weak), "poses": C, "rays_per_pose": M,
            "terrain_points": int(scene.terrain.shape[0]), "scaling": scaling, "n_gpus": rig.world, "steps": steps,
            "sweep_latency_ms": t["ms_per_step"], "rays_per_s": C * M / (t["ms_per_step"] * 1e-3),
            "k_sweep_ms": t["breakdown"]["k_sweep_ms"], "best_pose_index": t["best"][1], "best_score": t["best"][0]}


def parity_block(rig: Rig, scene, n_mask_poses: int = 64) -> dict:
    """N > 1: the rank-sharded sweep against an unsharded sweep of the same scene on rank 0 — every score bit for
    bit, the optimum, and the visibility masks of this rank's first poses."""
    torch, dist, ctx = rig.torch, rig.dist, rig.ctx
    C, M = int(scene.candidates.size), int(scene.cells.size)
    rig.load(scene)
    rig.step()
    rig.barrier()
    local = ctx.fetch(scores=True, mask=False)
    sharded_best = rig.exchange.pick_best(rig.gathered)
    parts = [None] * rig.world
    dist.all_gather_object(parts, (local.scores, local.covered))
    out = None
    if rank_zero(rig):
        scores = np.empty(C, np.float64)
        covered = np.empty(C, np.int32)
        for r, (s, c) in enumerate(parts):
            scores[r::rig.world] = s
            covered[r::rig.world] = c
        t0 = time.perf_counter()
        full = ctx.sweep(scene.params, scene.fixed, scene.candidates, scores=True)
        unsharded_ms = (time.perf_counter() - t0) * 1e3
        ctx.sweep(scene.params, scene.fixed, scene.candidates, scores=False)      # warm (index at its final refinement)
        dt = ctx.device_times()      # the same workload on ONE GPU in this run: the base of the strong-scaling figure
        unsharded_dev_ms = dt["sweep_kernel_ms"] + dt["sweep_fixed_ms"] + dt["sweep_reduce_ms"]
        k = min(n_mask_poses, len(range(0, C, rig.world)))
        mine = ctx.sweep_strided(scene.params, scene.fixed, scene.candidates[: k * rig.world], 0, rig.world, mask=True)
        whole = ctx.sweep(scene.params, scene.fixed, scene.candidates[: k * rig.world], mask=True)
        mism = int(np.count_nonzero(np.unpackbits((mine.mask ^ whole.mask[:: rig.world]).view(np.uint8))))
        out = {"against": "unsharded sweep of the same scene on rank 0", "scores_compared": C,
               "scores_bit_equal": bool(np.array_equal(scores, full.scores)),
               "covered_equal": bool(np.array_equal(covered, full.covered)),
               "argmax_equal": bool(sharded_best[1] == full.best_index and sharded_best[0] == full.best_score),
               "pairs": int(k * M), "mask_mismatches": mism, "unsharded_wall_ms": unsharded_ms,
               "one_gpu_sweep_ms_same_workload": unsharded_dev_ms}
        assert out["scores_bit_equal"] and out["argmax_equal"] and mism == 0, out
    rig.barrier()
    return out


def group_block(rig: Rig, scene, reps: int = 3) -> dict | None:
    """plc_group_* — the single-process form the drop-in node uses (one process, N devices, library-side NCCL):
    rank 0 alone drives all N devices while the other ranks wait; results must equal the rank-sharded ones."""
    out = None
    rig.load(scene)
    rig.step()
    rig.barrier()
    sharded_best = rig.exchange.pick_best(rig.gathered)
    rig.barrier()
    rig.host_barrier()              # every GPU idle, every other rank waiting on the host from here on
    if rank_zero(rig):
        from pointcloud_processor_b200.context import Group
        try:
            with Group(list(range(rig.world))) as g:
                g.set_terrain(scene.terrain)
                g.set_cells(scene.cells)
                r = g.sweep(scene.params, scene.fixed, scene.candidates)
                walls = []
                for _ in range(reps):
                    t0 = time.perf_counter()
                    r = g.sweep(scene.params, scene.fixed, scene.candidates)
                    walls.append((time.perf_counter() - t0) * 1e3)
            out = {"group_sweep_wall_ms": float(np.median(walls)), "devices": rig.world,
                   "group_equals_ranks": bool(r.best_index == sharded_best[1] and r.best_score == sharded_best[0]),
                   "best_pose_index": int(r.best_index)}
        except Exception as e:          # reported, not fatal: the rank-sharded numbers above stand on their own
            out = {"error": repr(e)[:200]}
    rig.host_barrier()
    rig.barrier()
    return out


def rank_zero(rig: Rig) -> bool:
    return rig.rank == 0


def run_ours(args) -> dict | None:
    from pointcloud_processor_b200 import scenes

    rig = Rig()
    world, ctx = rig.world, rig.ctx
    workload = args.workload or default_workload(world)
    weak = workload == "cfg2" and world > 1
    scene = build_scene(workload, world, weak=weak)
    C, M = int(scene.candidates.size), int(scene.cells.size)
    n_local = len(range(rig.rank, C, world))

    rig.load(scene)
    # One sweep in counting mode, untimed: the number of sample positions the reference evaluates on this workload
    # (SURVEY D4's algorithmic bytes).  The timed steps run the library's default any-hit traversal, whose outputs
    # are the same bit for bit (tests/conftest.py checks every test sweep both ways).
    ctx.count_samples(True)
    rig.step()
    rig.barrier()
    counted = ctx.fetch(scores=True)
    counting_k_ms = ctx.last_sweep_kernel_ms()
    ctx.count_samples(False)
    main = rig.timed(args.steps, max(args.warmup, 3), clocks=True)
    ms_per_step = main["ms_per_step"]
    best_score, best_index = main["best"]
    res = ctx.fetch(scores=True)
    grid = ctx.grid_info()[0]
    if not (np.array_equal(res.scores, counted.scores) and res.best_index == counted.best_index and
            res.counters["rays_visible"] == counted.counters["rays_visible"]):
        raise SystemExit("any-hit and counting traversals disagree")
    res.counters["samples"] = counted.counters["samples"]

    e2e = measure_e2e(rig, scene, max(5, min(args.steps, 20)))

    # ---- roofline of the dominant kernel (k_sweep) on this rank
    peak, peak_src = measured_peaks()
    k_avg = main["k_sweep_ms_local"]
    alg = algorithmic_bytes(res.counters["rays"], res.counters["samples"], n_local, grid["grid_bytes"])
    achieved = alg / (k_avg * 1e-3) / 1e9
    traffic = None
    tfile = ROOT / "profiles" / "sweep_dram_traffic.json"
    if tfile.exists():
        try:
            traffic = json.loads(tfile.read_text()).get(workload if world == 1 else f"{workload}_n{world}")
        except ValueError:
            traffic = None

    extras = {}
    if world == 1:
        if workload == "cfg2" and not args.no_north_star:
            extras["north_star_cfg3"] = extra_block(rig, scenes.build("cfg3"), "cfg3", 3)
            extras["north_star_cfg3"]["target_ms_on_8_gpus"] = 50.0
    else:
        if not args.no_parity:
            extras["parity"] = parity_block(rig, scene)
            extras["group"] = group_block(rig, scene)
        if workload == "cfg3" and not args.no_north_star:
            extras["weak_cfg2_shard"] = extra_block(rig, build_scene("cfg2", world, weak=True), "cfg2", 5, weak=True,
                                                    scaling="weak")
            if world >= 8 or args.cfg5:
                extras["cfg5"] = extra_block(rig, scenes.build("cfg5"), "cfg5", 3)

    rig.barrier()
    if rig.rank != 0:
        if world > 1:
            rig.dist.destroy_process_group()
        return None

    strong = world > 1 and not weak
    line = {
        "metric": METRIC, "value": C * M / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
        "dtype": "f64 geometry / f32 distance test", "data": "synthetic",
        "config": {"workload": workload_text(workload, world, weak), "poses": C, "rays_per_pose": M,
                   "terrain_points": int(scene.terrain.shape[0]), "rays_per_step": C * M,
                   "samples_per_ray": res.counters["samples"] / max(res.counters["rays"], 1),
                   "parallelism": f"candidate poses dealt round-robin over {world} GPU(s), grid replicated",
                   "cache": "L2 flushed (256 MiB fill) before every timed step",
                   "traversal": "any-hit (library default; scores, visible rays and optimum checked equal to the "
                                "counting traversal before timing)",
                   "k_sweep_ms_counting_mode": counting_k_ms,
                   "best_pose_index": best_index, "best_score": best_score},
        "sweep_latency_ms": ms_per_step,
        "breakdown": main["breakdown"],
        "e2e": e2e,
        "gpu_launches": main["launches"],
        "clocks": main["clocks"],
        "roofline": {"bound": "hbm", "kernel": "k_sweep", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": int(alg), "kernel_ms": k_avg,
                     "kernel_share_of_step": k_avg / ms_per_step,
                     "note": "algorithmic bytes = 32 B/ray + 4 B/evaluated sample + 64 B/pose + grid (SURVEY D4); "
                             "the grid is L2-resident, so the kernel is issue/latency bound, not DRAM bound"},
    }
    if strong:
        line["config"]["note"] = ("N > 1 lines strong-scale configs[2]; the N = 1 line is configs[1] (cfg2), so the same-"
                                  "workload one-GPU time is parity.unsharded_wall_ms / north_star_cfg3 of the N = 1 line")
    line.update({k: v for k, v in extras.items() if v is not None})
    if world == 1 and not args.no_feeders:
        line["feeders"] = measure_feeders(ctx, scene, peak)
        line["feeders_large"] = measure_feeders_large(ctx, peak)
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, n, dt = cpu_reference_sample(scene, args.cpu_budget, threads)
        v1, n1, dt1 = cpu_reference_sample(scene, min(args.cpu_budget / 2, 8.0), 1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{n} seeded poses x {M} rays ({dt:.1f} s), KD-tree restatement with the "
                                          "reference loop structure (oracle/, faithful mode)",
                                "value_1_thread": v1, "sample_1_thread": f"{n1} poses ({dt1:.1f} s)",
                                "full_sweep_extrapolated_s": C * M / v}
    if world > 1:
        rig.dist.destroy_process_group()
    return line


def measure_feeders(ctx, scene, peak, reps: int = 5) -> dict:
    """The secondary kernels of the path on config 4's inputs: crop + voxel downsample of two
    1 M-point scans (K6+K7), the merge into the map frame (K8), and the occupancy build the sweep starts from
    (K1).  Kernel time is the library's own CUDA-event time per call (copies excluded); wall time is the
    C-ABI call with host buffers.  Algorithmic bytes per SURVEY D4."""
    from pointcloud_processor_b200 import scenes
    from pointcloud_processor_b200.nodes import filter_scan, merge_clouds
    zt, zq = scenes.sensor_in_map(dict(xyz=(0.0, 0.0, 0.0), yaw=0.0), scenes.ZX120_VELODYNE)
    rt, rq = scenes.sensor_in_map(scenes.ROBOT_BASE, scenes.ROBOT_VELODYNE)
    scans = [scenes.lidar_scan(rt, rq, 1_000_000, 4, "rect"), scenes.lidar_scan(zt, zq, 1_000_000, 5, "rect")]

    def best(fn, key):
        ks, ws, out = [], [], None
        for _ in range(reps):
            t0 = time.perf_counter()
            out = fn()
            ws.append((time.perf_counter() - t0) * 1e3)
            ks.append(ctx.device_times()[key])
        return out, float(np.median(ks)), float(np.median(ws))

    def row(kernel_ms, wall_ms, alg_bytes, units, unit_name):
        gbs = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
        return {"kernel_ms": kernel_ms, "wall_ms": wall_ms, f"{unit_name}_per_s": units / (kernel_ms * 1e-3) if kernel_ms > 0 else 0.0,
                "algorithmic_bytes": int(alg_bytes), "achieved_GBps": gbs, "frac_of_hbm_peak": gbs / peak}

    out = {"config": "cfg4: two 1M-point 32-ring scans -> crop 15/10/10 -> voxel 0.2 -> merge; K1 on this workload's terrain"}
    filtered = []
    for name, scan in zip(("robot", "zx120"), scans):
        f, k_ms, w_ms = best(lambda: filter_scan(ctx, scan), "filter_ms")
        p = scan[:, :3].astype(np.float64)
        m_crop = int(np.count_nonzero((p[:, 0] > 0) & (p[:, 0] < 15) & (np.abs(p[:, 1]) < 10) & (p[:, 2] > -1.5) & (p[:, 2] < 10)))
        v = int(f.shape[0])
        alg = scan.shape[0] * 12 + m_crop * 12 + m_crop * 12 + v * 12      # K6 + K7
        r = row(k_ms, w_ms, alg, scan.shape[0], "points")
        r.update(points_in=int(scan.shape[0]), cropped=m_crop, voxels=v)
        out[f"filter_{name}"] = r
        filtered.append(f)
    (merged, _, _), k_ms, w_ms = best(lambda: merge_clouds(ctx, dict(cloud=filtered[0], t=rt, q=rq),
                                                          dict(cloud=filtered[1], t=zt, q=zq)), "merge_ms")
    v = int(merged.shape[0])
    out["merge"] = row(k_ms, w_ms, v * 12 + v * 16 * 2, v, "points")
    out["merge"]["points_in"] = v
    _, k_ms, w_ms = best(lambda: (ctx.set_terrain(scene.terrain), ctx.sync()), "set_terrain_ms")
    n = int(scene.terrain.shape[0])
    out["voxelise_terrain"] = row(k_ms, w_ms, n * (12 + 16), n, "points")
    out["voxelise_terrain"]["points_in"] = n
    out["setup"] = measure_setup(ctx, scene, peak, best)
    # the nodes either side of the path on this workload's merged cloud / robot scan
    base_t, base_q = np.zeros(3), np.array([0.0, 0.0, 0.0, 1.0])
    (terrain, area, info), k_ms, w_ms = best(lambda: ctx.excavate(merged.reshape(-1), base_t, base_q, None, point_step=32, off_rgb=16),
                                             "excavate_ms")
    n_in = int(merged.shape[0])
    r = row(k_ms, w_ms, n_in * 32 + int(terrain.shape[0]) * 32 + int(area.shape[0]) * 32, n_in, "points")
    r.update(points_in=n_in, kept=int(info.n_kept), surface=int(info.n_surface), area=int(info.n_area))
    if not os.environ.get("PLC_BENCH_NO_CPU_FEEDERS"):
        from oracle import binding as O
        from pointcloud_processor_b200 import _abi as A
        prm = A.excavation_params()
        ns, na = A.excavation_counts(prm)
        mx = np.ascontiguousarray(merged[:, :12]).view(np.float32).reshape(-1, 3)
        t0 = time.perf_counter()
        O.excavate(mx, None, base_t, base_q, prm, ns, na)
        r["cpu_port_ms"] = (time.perf_counter() - t0) * 1e3      # one shared index; upstream rebuilds a KD-tree per query
    out["excavate"] = r
    # the same node on a dense cloud (1 M ground points around the pit), where the bucketed heights matter
    rng = np.random.default_rng(17)
    dense = np.zeros((1_000_000, 8), np.float32)
    dense[:, :2] = rng.uniform((-11.0, -14.0), (19.0, 16.0), (dense.shape[0], 2))
    dense[:, 2] = 0.15 * np.sin(0.4 * dense[:, 0]) * np.cos(0.3 * dense[:, 1]) + rng.uniform(-0.01, 0.01, dense.shape[0])
    dense[:, 3] = 1.0
    (terrain, area, info), k_ms, w_ms = best(lambda: ctx.excavate(dense, base_t, base_q, None), "excavate_ms")
    r = row(k_ms, w_ms, dense.shape[0] * 32 + int(terrain.shape[0]) * 32 + int(area.shape[0]) * 32, dense.shape[0], "points")
    r.update(points_in=int(dense.shape[0]), kept=int(info.n_kept), surface=int(info.n_surface), area=int(info.n_area))
    out["excavate_dense_1M"] = r
    rxy = scenes.ROBOT_BASE["xyz"][:2]
    (grid, ginfo), k_ms, w_ms = best(lambda: ctx.drivable_area(filtered[0], rt, rq, rxy, rxy), "drivable_ms")
    v = int(filtered[0].shape[0])
    r = row(k_ms, w_ms, v * 16 + int(ginfo.n_cells) * 13, v, "points")
    r.update(points_in=v, cells=int(ginfo.n_cells))
    if not os.environ.get("PLC_BENCH_NO_CPU_FEEDERS"):
        from oracle import binding as O
        t0 = time.perf_counter()
        O.drivable_area(filtered[0], rt, rq, rxy, rxy)
        r["cpu_port_ms"] = (time.perf_counter() - t0) * 1e3
    out["drivable_area"] = r
    out["launch_graph"] = measure_launch_graph(ctx)
    return out


def measure_feeders_large(ctx, peak, reps: int = 4) -> dict:
    """The same secondary kernels at a size where launch latency amortises: an 8 M-point scan through
    crop + voxel downsample at leaf 0.05, two 4 M-point clouds through the merger's transform, a 4 M-point terrain
    through the voxeliser.  Algorithmic bytes per SURVEY D4 (sort traffic excluded => a lower bound for K7)."""
    from pointcloud_processor_b200 import scenes
    from pointcloud_processor_b200.nodes import filter_scan, merge_clouds
    rng = np.random.default_rng(41)
    n = 8_000_000
    scan = np.ones((n, 4), np.float32)
    scan[:, 0] = rng.uniform(-6.0, 24.0, n)
    scan[:, 1] = rng.uniform(-14.0, 14.0, n)
    scan[:, 2] = 0.15 * np.sin(0.4 * scan[:, 0]) * np.cos(0.3 * scan[:, 1]) + rng.normal(0.0, 0.02, n).astype(np.float32) - 1.2

    def best(fn, key):
        ks, ws, out = [], [], None
        for _ in range(reps):
            t0 = time.perf_counter()
            out = fn()
            ws.append((time.perf_counter() - t0) * 1e3)
            ks.append(ctx.device_times()[key])
        return out, float(np.median(ks[1:])), float(np.median(ws[1:]))

    def row(kernel_ms, wall_ms, alg_bytes, units):
        gbs = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
        return {"kernel_ms": kernel_ms, "wall_ms": wall_ms, "points_per_s": units / (kernel_ms * 1e-3) if kernel_ms > 0 else 0.0,
                "algorithmic_bytes": int(alg_bytes), "achieved_GBps": gbs, "frac_of_hbm_peak": gbs / peak}

    out = {"config": "8M-point scan -> crop 15/10/10 -> voxel 0.05; 2 x 4M-point clouds -> merge; 4M-point terrain -> voxelise"}
    p = scan[:, :3].astype(np.float64)
    m_crop = int(np.count_nonzero((p[:, 0] > 0) & (p[:, 0] < 15) & (np.abs(p[:, 1]) < 10) & (p[:, 2] > -1.5) & (p[:, 2] < 10)))
    del p
    f, k_ms, w_ms = best(lambda: filter_scan(ctx, scan, leaf=0.0), "filter_ms")
    out["crop_only"] = row(k_ms, w_ms, n * 12 + m_crop * 12, n)
    out["crop_only"].update(points_in=n, cropped=m_crop)
    f, k_ms, w_ms = best(lambda: filter_scan(ctx, scan, leaf=0.05), "filter_ms")
    v = int(f.shape[0])
    out["filter"] = row(k_ms, w_ms, n * 12 + m_crop * 12 + m_crop * 12 + v * 12, n)
    out["filter"].update(points_in=n, cropped=m_crop, voxels=v,
                         note="K7 sorts (key, index) pairs in 8-bit passes: each pass moves 24 B per cropped point on top of "
                              "the algorithmic bytes")
    half = scan[: n // 2]
    ident_t, ident_q = np.array([1.0, 2.0, 0.5]), np.array([0.0, 0.0, 0.38268343, 0.92387953])
    (merged, _, _), k_ms, w_ms = best(lambda: merge_clouds(ctx, dict(cloud=half, t=ident_t, q=ident_q),
                                                          dict(cloud=scan[n // 2:], t=ident_t, q=ident_q)), "merge_ms")
    mv = int(merged.shape[0])
    out["merge"] = row(k_ms, w_ms, mv * 12 + mv * 16 * 2, mv)
    out["merge"]["points_in"] = mv
    terr = scenes.terrain("L", 4_000_000, 6, 40.0)
    _, k_ms, w_ms = best(lambda: (ctx.set_terrain(terr), ctx.sync()), "set_terrain_ms")
    out["voxelise_terrain"] = row(k_ms, w_ms, terr.shape[0] * (12 + 16), terr.shape[0])
    out["voxelise_terrain"]["points_in"] = int(terr.shape[0])
    return out


def measure_launch_graph(ctx, reps: int = 5) -> dict:
    """The whole dataflow of launch/start.launch.xml for one scan pair (two 1 M-point scans): wall time through the
    C ABI with host buffers between the stages, and with every intermediate cloud left in device buffers that the
    next stage reads in place.  leaf 0.2 is the node's default (a few thousand points after the
    voxel filter); leaf 0.03 keeps ~100 k points per scan and shows what the hand-over costs when clouds are dense."""
    from pointcloud_processor_b200 import scenes
    from pointcloud_processor_b200.nodes import run_launch_graph, run_launch_graph_resident
    inp = scenes.pipeline_inputs(n_scan=1_000_000, sweep="tiny")
    out = {"config": "2 x 1M-point scans -> filter -> merger -> generator -> normals/cells/poses (32x32) -> sweep, drivable grid"}
    for leaf in (0.2, 0.03):
        kw = dict(leaf=leaf, num_candidates=32 * 32, grid_resolution=0.05, vertical_layers=20)
        walls = {"host_buffers": [], "device_resident": []}
        counts = None
        for _ in range(reps + 1):          # the first round warms allocations up
            t0 = time.perf_counter()
            run_launch_graph(ctx, inp, **kw)
            walls["host_buffers"].append((time.perf_counter() - t0) * 1e3)
            r = run_launch_graph_resident(ctx, inp, **kw)
            walls["device_resident"].append(r["wall_ms"])
            counts = r["counts"]
        out[f"leaf_{leaf}"] = {"wall_ms_host_buffers": float(np.median(walls["host_buffers"][1:])),
                               "wall_ms_device_resident": float(np.median(walls["device_resident"][1:])),
                               "points": {k: int(v) for k, v in counts.items()}}
    return out


def measure_setup(ctx, scene, peak, best) -> dict:
    """The steps that produce the sweep's inputs on this workload: PCA normals of the
    /excavation_area cloud, the evaluation-cell lattice with cell normals, the pose lattice with ground heights.
    kernel_ms = the library's CUDA-event time of the call's kernels, wall_ms = the C-ABI call with host buffers,
    host_ms = the same step by the multi-threaded host implementation (plc_host_*) on this box's cores.
    Algorithmic bytes: 16 B per point a query has to look at (the points inside its ball)."""
    from pointcloud_processor_b200 import context as ctxmod
    m = scene.meta
    area = scene.area
    res, px = float(m["res"]), int(m["px"])
    layers = int(m.get("layers_used") or m.get("layers") or 10)

    def host(fn, reps=2):
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e3)
        return float(min(ts))

    out = {"config": f"{m['shape']} area cloud of {area.shape[0]} points, lattice res {res} x {layers} layers, "
                     f"{px}x{px} pose lattice over {scene.terrain.shape[0]} terrain points"}
    normals, k_ms, w_ms = best(lambda: ctx.compute_normals(area, 1.5), "normals_ms")
    out["pca_normals"] = {"kernel_ms": k_ms, "wall_ms": w_ms, "points": int(area.shape[0]),
                          "host_ms": host(lambda: ctxmod.host_compute_normals(area, 1.5))}
    (cells, _), k_ms, w_ms = best(lambda: ctx.generate_cells(area, normals, res, layers, 1.5), "cells_ms")
    out["cell_lattice"] = {"kernel_ms": k_ms, "wall_ms": w_ms, "cells": int(cells.size),
                           "host_ms": host(lambda: ctxmod.host_generate_cells(area, normals, res, layers, 1.5))}
    ctx.set_terrain(scene.terrain)
    cands, k_ms, w_ms = best(lambda: ctx.generate_candidates(None, scene.bounds, scene.fixed, px * px, 3.0, 1.1), "candidates_ms")
    out["pose_lattice"] = {"kernel_ms": k_ms, "wall_ms": w_ms, "poses": int(cands.size),
                           "host_ms": host(lambda: ctxmod.host_generate_candidates(scene.terrain, scene.bounds, scene.fixed,
                                                                                   px * px, 3.0, 1.1))}
    return out


def run_reference(args) -> dict | None:
    """The reference's own CPU implementation of the path, on the box's host cores.  The reference's
    build cannot run here (ROS2/PCL/FLANN absent), so the line's value times the oracle's
    restatement with the reference's loop structure (KD-tree radius searches, per-candidate fixed-sensor
    re-evaluation), all host threads, on a bounded seeded sample of the same workload — the harder
    baseline.  Beside it, `reference_source` times the reference's own virtual_lidar.cpp compiled against
    the stand-in headers (oracle/_ref): one thread, as upstream runs, on the stand-in's grid index."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return None                      # rank 0 alone runs and prints; the others exit 0
    workload = args.workload or default_workload(world)
    weak = workload == "cfg2" and world > 1
    # the scene comes from oracle/vl_setup.cpp (identical arrays to the plc_host_* route, checked in
    # tests/test_oracle.py): this arm never loads the product library
    scene = build_scene(workload, world, weak=weak, setup=oracle_setup())
    C, M = int(scene.candidates.size), int(scene.cells.size)
    threads = os.cpu_count() or 1
    steps = max(1, args.steps)
    per_step_budget = max(2.0, min(20.0, 120.0 / (steps + max(args.warmup, 0))))
    for _ in range(min(args.warmup, 1)):
        cpu_reference_sample(scene, per_step_budget / 2, threads)
    vals, ns, total = [], [], 0.0
    for k in range(steps):
        v, n, dt = cpu_reference_sample(scene, per_step_budget, threads, seed=1234 + k)
        vals.append(v)
        ns.append(n)
        total += dt
    value = float(np.mean(vals))
    refsrc = None
    if world == 1 and not args.no_reference_source:
        refsrc = reference_source_sample(scene, float(scene.meta["res"]), int(scene.meta["layers_found"]))
    sample = f"{int(np.mean(ns))} seeded poses x {M} rays per step (~{total / steps:.1f} s/step) of {C} poses"
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * C * M / value, "higher_is_better": True,
        "scaling": "weak" if (weak or world == 1) else "strong", "vs_baseline": None,
        "dtype": "f64 geometry / f32 distance test", "data": "synthetic",
        "config": {"workload": workload_text(workload, world, weak), "poses": C, "rays_per_pose": M,
                   "terrain_points": int(scene.terrain.shape[0]), "rays_per_step": C * M,
                   "note": "ms_per_step is the full sweep extrapolated from the timed sample"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        **({"reference_source": refsrc} if refsrc is not None else {}),
        **({"feeders_cpu": feeders_cpu_reference()} if world == 1 and not args.no_reference_source else {}),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-reference-source", action="store_true",
                    help="reference arm: skip the single-threaded run of the reference's own virtual_lidar.cpp (oracle/_ref) and the feeders' CPU baseline")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=None,
                    help="default: cfg2 (BASELINE.json configs[1]) on one GPU, cfg3 (configs[2]) strong-scaled on N > 1")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-north-star", action="store_true", help="skip the extra configs[2] measurement")
    ap.add_argument("--no-feeders", action="store_true", help="skip the crop/voxel/merge/voxelise measurements (config 4)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the sharded-vs-unsharded and plc_group checks")
    ap.add_argument("--cfg5", action="store_true", help="N > 1: add the configs[4] block below 8 GPUs too")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything libraries write to fd 1 meanwhile (NCCL prints its version
    # banner there) is sent to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        line = run_reference(args) if args.impl == "reference" else run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
