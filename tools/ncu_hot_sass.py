"""Prints every SASS instruction of an .ncu-rep (captured with --import-source on) that executes at least
`min_per_batch` times per 32-ray batch, with executions per batch and active threads.
    python tools/ncu_hot_sass.py rep.ncu-rep <rays> [min_per_batch]"""
import csv, io, subprocess, sys
rep, rays = sys.argv[1], float(sys.argv[2])
thr_min = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
h = rows[hi]
iI, iSrc, iT, iS = h.index("Instructions Executed"), h.index("Source"), h.index("Thread Instructions Executed"), h.index("# Samples")
data = [r for r in rows[hi + 1:] if len(r) == len(h)]
nb = rays / 32
tot = 0.0
for idx, r in enumerate(data):
    inst, thr = float(r[iI]), float(r[iT])
    tot += inst
    if inst / nb >= thr_min:
        print(f"{idx:5d} {inst / nb:6.2f} {thr / max(inst, 1):5.1f} {int(float(r[iS])):6d}  {r[iSrc][:90]}")
print(f"total per batch {tot / nb:.1f}")
