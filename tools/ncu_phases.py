"""Groups the per-line profile of a k_sweep capture (.ncu-rep with --import-source on) into the kernel's phases:
share of warp instructions, of stall samples (= time), and the dominant stall reasons per phase.
    python tools/ncu_phases.py rep.ncu-rep [rays]"""
import csv, io, subprocess, sys, re
rep = sys.argv[1]
rays = float(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
h = rows[hi]
iI, iS = h.index("Instructions Executed"), h.index("# Samples")
stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
# Phase boundaries = the lines of these statements in the CURRENT sources (the capture must come from the same
# revision: --import-source embeds line numbers, not the text), or "name:line,..." as argv[3].
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def find_lines(path, needles):
    lines = open(os.path.join(ROOT, "pointcloud_processor_b200", "csrc", path)).read().split("\n")
    out = {}
    for name, needle in needles:
        out[name] = next(i + 1 for i, l in enumerate(lines) if needle in l and name not in out)
    return out
if len(sys.argv) > 3:
    marks = [(int(b), a) for a, b in (kv.split(":") for kv in sys.argv[3].split(","))]
else:
    m = find_lines("plc_sweep.cu", [("prologue", "const unsigned wid = blockIdx.x"), ("drain", "auto drain = [&]()"),
                                    ("setup", "for (int r = 0; r < kRaysPerThread; ++r) {"),
                                    ("march", "a lane is marching while k < k_hi"), ("final", "---- score (:689-700)")])
    marks = [(l, n) for n, l in m.items()]
marks.sort()
vq = find_lines("plc_voxel_query.cuh", [("dist_lo", "float dist2_f32("), ("row_lo", "unsigned ldg_row("), ("row_hi", "unsigned row_shift(int fx)"),
                                        ("ldrow", "unsigned ldg_brick_row(")])
def phase_of(f, ln):
    if f == "plc_voxel_query.cuh":
        if vq["row_lo"] <= ln <= vq["row_hi"] + 1 or vq["ldrow"] <= ln <= vq["ldrow"] + 5: return "march"
        return "exact" if vq["dist_lo"] <= ln <= vq["dist_lo"] + 6 else "vq_other"
    if f != "plc_sweep.cu":
        return "intrinsics"
    p = "pre"
    for l, name in marks:
        if ln >= l: p = name
    return p
fname = ""
ph = {}
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]; continue
    if len(r) != len(h) or r[0] in ("", "Line No"): continue
    try: inst, smp = float(r[iI]), float(r[iS])
    except ValueError: continue
    e = ph.setdefault(phase_of(fname, int(r[0])), [0.0, 0.0, {}])
    e[0] += inst; e[1] += smp
    for i, c in stall_cols:
        try: e[2][c] = e[2].get(c, 0.0) + float(r[i])
        except ValueError: pass
ti = sum(e[0] for e in ph.values()); ts = sum(e[1] for e in ph.values())
for name, e in sorted(ph.items(), key=lambda kv: -kv[1][1]):
    top = sorted(e[2].items(), key=lambda kv: -kv[1])[:5]
    per = f" {e[0] / (rays / 32):6.1f}/batch" if rays else ""
    print(f"{name:12s} inst {e[0]/ti*100:5.1f}%{per}  time {e[1]/ts*100:5.1f}%  " + " ".join(f"{c[6:]}={v/max(e[1],1)*100:.0f}%" for c, v in top))
