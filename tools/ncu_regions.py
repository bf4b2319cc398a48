"""Instruction / stall-sample shares of named source regions of k_sweep (line ranges found from landmarks)."""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
src_path = "pointcloud_processor_b200/csrc/plc_sweep.cu"
text = open(src_path).read().split("\n")
def line_of(pat, after=0):
    for i, l in enumerate(text):
        if i + 1 > after and pat in l: return i + 1
    raise KeyError(pat)
k0 = line_of("k_sweep(SweepDev P")
marks = [("prologue", k0), ("drain (exact recompute)", line_of("auto drain")), ("load + range/n", line_of("for (int r = 0; r < kRaysPerThread", k0)),
         ("FOV", line_of("---- FOV", k0)), ("clip + init", line_of("checkVisibilityWithRaycasting (:754-800): clear", k0)),
         ("march loop", line_of("const unsigned tag_base", k0)),
         ("finalise + score", line_of("while (q_n > 0) drain", k0)), ("end", line_of("// per candidate: tiles summed"))]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
h = rows[hi]; iI, iS, iT = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
fname = ""; lines = {}
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"): fname = r[1].split("/")[-1]; continue
    if len(r) != len(h) or r[0] in ("", "Line No"): continue
    try: v = (float(r[iI]), float(r[iS]), float(r[iT]))
    except ValueError: continue
    e = lines.setdefault((fname, int(r[0])), [0, 0, 0]); e[0] += v[0]; e[1] += v[1]; e[2] += v[2]
tot = sum(e[0] for e in lines.values()); tots = sum(e[1] for e in lines.values())
print(f"total warp-inst {tot:.3e}")
def show(name, sel):
    i = sum(e[0] for k, e in lines.items() if sel(k)); s = sum(e[1] for k, e in lines.items() if sel(k)); t = sum(e[2] for k, e in lines.items() if sel(k))
    print(f"{name:22s} inst {i/tot*100:5.1f}%  smp {s/tots*100:5.1f}%  thr {t/max(i,1):4.1f}")
for (name, a), (_, b) in zip(marks, marks[1:]):
    show(name, lambda k, a=a, b=b: k[0] == "plc_sweep.cu" and a <= k[1] < b)
show("score/length helpers", lambda k: k[0] == "plc_sweep.cu" and k[1] < k0)
vq = open("pointcloud_processor_b200/csrc/plc_voxel_query.cuh").read().split("\n")
vs = next(i + 1 for i, l in enumerate(vq) if "voxel_state(const VoxelGeom" in l); cm = next(i + 1 for i, l in enumerate(vq) if "mid_voxel_of(const VoxelGeom" in l)
show("voxel_state (march)", lambda k: k[0] == "plc_voxel_query.cuh" and vs <= k[1] < cm)
show("exact test (lists)", lambda k: k[0] == "plc_voxel_query.cuh" and (k[1] < vs or k[1] >= cm))
show("other files", lambda k: k[0] not in ("plc_sweep.cu", "plc_voxel_query.cuh"))
