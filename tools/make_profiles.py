"""Turns the raw ncu outputs in gpurun_out/ (scratch) into the tracked summaries under profiles/.

    python tools/make_profiles.py <tag> <launches.csv> <k_sweep.ncu-rep | -> [command text]

  profiles/<tag>_launches.csv         the ncu launch list as captured (gpu__time_duration per launch)
  profiles/<tag>_launch_summary.csv   per kernel: launches, total time, share of the profiled command
  profiles/<tag>_k_sweep_ncu.csv      every metric of the `--set full` capture of one k_sweep launch
  profiles/<tag>_k_sweep_summary.txt  headline metrics, stall reasons, hottest SASS regions
  profiles/sweep_dram_traffic.json    dram read + write bytes of that launch (bench.py's roofline.traffic)
"""
import csv
import io
import json
import re
import shutil
import subprocess
import sys
from collections import OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PROF = ROOT / "profiles"


def launch_summary(tag, path, command):
    rows = [r for r in csv.reader(open(path)) if len(r) >= 15 and r[0].isdigit()]
    per = OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*$", "", r[4])
        name = re.sub(r"^(void )?(plc::)?(<?unnamed>::)?", "", name)      # ncu demangles namespaces differently between runs
        t = float(r[14].replace(",", "")) / 1e3          # ns -> us
        n, tot = per.get(name, (0, 0.0))
        per[name] = (n + 1, tot + t)
    total = sum(t for _, t in per.values())
    shutil.copyfile(path, PROF / f"{tag}_launches.csv")
    with open(PROF / f"{tag}_launch_summary.csv", "w") as f:
        f.write(f"# {tag}: ncu launch list of `{command}` (1x B200)\n")
        f.write("# cold-cache serialised times under the profiler: compare SHARES, not absolutes\n")
        f.write("kernel,launches,total_us,share\n")
        for name, (n, t) in sorted(per.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name},{n},{t:.1f},{t / total:.4f}\n")
    return per, total


def full_capture(tag, rep, workload):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    with open(PROF / f"{tag}_k_sweep_ncu.csv", "w") as f:
        f.write(f"# {tag} k_sweep, {workload}, ncu --set full --clock-control none, 1 launch\n")
        f.write("metric,unit,value\n")
        for h, u, v in sorted(zip(hdr, units, vals)):
            if "__" in h or h.startswith("launch"):
                f.write(f"{h},{u},{v}\n")
    d = dict(zip(hdr, zip(units, vals)))

    def to_bytes(key):
        u, v = d[key]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        return int(round(float(v.replace(",", "")) * scale))

    traffic = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
    summary = subprocess.run([sys.executable, str(ROOT / "tools" / "ncu_summary.py"), rep, "12"],
                             capture_output=True, text=True).stdout
    (PROF / f"{tag}_k_sweep_summary.txt").write_text(
        f"# {tag} k_sweep, {workload}: tools/ncu_summary.py on the --set full capture\n" + summary)
    return traffic


def main():
    tag, launches, rep = sys.argv[1:4]
    command = sys.argv[4] if len(sys.argv) > 4 else "python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-north-star"
    PROF.mkdir(exist_ok=True)
    per, total = launch_summary(tag, launches, command)
    if rep == "-":              # launch list only (the k_sweep capture is unchanged)
        print("kernels", len(per), "total_us", round(total, 1))
        return
    traffic = full_capture(tag, rep, "cfg2 (1930 poses x 57600 rays, 200k terrain points)")
    tfile = PROF / "sweep_dram_traffic.json"
    cur = json.loads(tfile.read_text()) if tfile.exists() else {}
    cur["cfg2"] = traffic
    cur["source"] = f"profiles/{tag}_k_sweep_ncu.csv: dram__bytes_read.sum + dram__bytes_write.sum, one launch"
    tfile.write_text(json.dumps(cur) + "\n")
    print("kernels", len(per), "total_us", round(total, 1), "traffic", traffic)


if __name__ == "__main__":
    main()
