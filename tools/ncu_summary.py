"""Reads an .ncu-rep (no GPU needed) and prints headline metrics plus the hottest SASS regions."""
import csv, io, re, subprocess, sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_blocks", "launch__waves_per_multiprocessor",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]
d = dict(zip(hdr, zip(units, vals)))
for k in want:
    if k in d:
        print(f"{k:75s} {d[k][1]:>16s} {d[k][0]}")
st = sorted(((float(v[1]), k) for k, v in d.items() if "issue_stalled" in k and k.endswith("per_issue_active.ratio")), reverse=True)
print("stalls/issue:", ", ".join(f"{k.split('issue_stalled_')[1].split('_per_')[0]}={v:.2f}" for v, k in st[:8]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]; ix = {k: i for i, k in enumerate(h)}
data = [r for r in rows[2:] if len(r) == len(h)]
f = lambda r, k: float(r[ix[k]]) if r[ix[k]] not in ("", "-") else 0.0
tot = sum(f(r, "Instructions Executed") for r in data); tots = sum(f(r, "# Samples") for r in data)
# regions = maximal runs with the same executed count
regions = []; cur = None
for i, r in enumerate(data):
    ie = f(r, "Instructions Executed")
    if cur and abs(ie - cur["ie"]) <= 0.02 * max(ie, 1) :
        cur["n"] += 1; cur["inst"] += ie; cur["smp"] += f(r, "# Samples"); cur["end"] = i
    else:
        cur = dict(start=i, end=i, n=1, ie=ie, inst=ie, smp=f(r, "# Samples"), thr=f(r, "Avg. Threads Executed"), first=r[1].strip())
        regions.append(cur)
regions.sort(key=lambda c: -c["inst"])
print(f"total warp-inst {tot:.3e}, sass {len(data)}")
for c in regions[:top]:
    print(f"[{c['start']:4d}-{c['end']:4d}] n={c['n']:3d} inst {c['inst']/tot*100:5.1f}% smp {c['smp']/tots*100:5.1f}% thr {c['thr']:4.1f} | {c['first'][:60]}")
