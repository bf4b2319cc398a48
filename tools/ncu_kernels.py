"""Per-kernel summary of an .ncu-rep that holds many launches (feeder kernels): for each kernel name the launch with
the longest duration, with its DRAM bytes and throughput, occupancy, issue rate and top stall reasons.
    python tools/ncu_kernels.py rep.ncu-rep [peak_GBps]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
peak = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ix = {k: i for i, k in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
def val(r, k):
    if k not in ix or r[ix[k]] in ("", "n/a"): return 0.0
    return float(r[ix[k]].replace(",", "")) * scale.get(units[ix[k]], 1.0)
best = {}
for r in rows[2:]:
    if len(r) != len(hdr): continue
    name = re.sub(r"\(.*$", "", r[ix["Kernel Name"]])
    name = re.sub(r"^(void )?(plc::)?(\(anonymous namespace\)::)?(<unnamed>::)?", "", name)
    t = val(r, "gpu__time_duration.sum")
    e = best.setdefault(name, [0, None])
    e[0] += 1
    if e[1] is None or t > val(e[1], "gpu__time_duration.sum"): e[1] = r
stall_keys = [k for k in hdr if "issue_stalled" in k and k.endswith("per_issue_active.ratio")]
print("kernel | launches | longest us | grid x block | regs | dram rd MB | dram wr MB | dram GB/s" + (" | frac of peak" if peak else "") +
      " | dram %peak | warps active % | issue active % | top stalls per issue")
for name, (n, r) in sorted(best.items(), key=lambda kv: -val(kv[1][1], "gpu__time_duration.sum")):
    t = val(r, "gpu__time_duration.sum")
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    gbs = (rd + wr) / max(t, 1e-9) / 1e3
    st = sorted(((val(r, k), k.split("issue_stalled_")[1].split("_per_")[0]) for k in stall_keys), reverse=True)[:3]
    print(f"{name} | {n} | {t:.1f} | {r[ix['launch__grid_size']]} x {r[ix['launch__block_size']]} | {r[ix['launch__registers_per_thread']]} | "
          f"{rd / 1e6:.1f} | {wr / 1e6:.1f} | {gbs:.0f}" + (f" | {gbs / peak:.3f}" if peak else "") +
          f" | {val(r, 'dram__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | {val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} | " + ", ".join(f"{k}={v:.2f}" for v, k in st))
