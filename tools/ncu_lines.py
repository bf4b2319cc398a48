"""Per-source-line totals of an .ncu-rep (kernel compiled with -lineinfo, captured with --import-source on)."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
h = rows[hi]
iI, iS, iT = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
fname = ""
lines = {}
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]
        continue
    if len(r) != len(h) or r[0] in ("", "Line No"):
        continue
    try:
        inst, smp, thr = float(r[iI]), float(r[iS]), float(r[iT])
    except ValueError:
        continue
    key = (fname, int(r[0]))
    e = lines.setdefault(key, [0.0, 0.0, 0.0, r[1].strip()])
    e[0] += inst; e[1] += smp; e[2] += thr
tot = sum(e[0] for e in lines.values()) or 1; tots = sum(e[1] for e in lines.values()) or 1
print(f"total warp-inst {tot:.3e}")
for (f, ln), e in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{e[0]/tot*100:5.1f}% inst {e[1]/tots*100:5.1f}% smp thr {e[2]/max(e[0],1):4.1f} {f}:{ln:<4d} {e[3][:90]}")
