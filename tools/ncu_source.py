"""Per-CUDA-source-line stall samples from an .ncu-rep (needs -lineinfo + --import-source on)."""
import csv
import io
import os
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 15
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
lines = []
h, fname = None, ""
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = os.path.basename(r[1])
        continue
    if r[0] == "Line No":
        h = r
        samp = h.index("# Samples")
        stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
        continue
    if h is None or not r[0].strip().isdigit() or len(r) <= samp or not r[samp].isdigit():
        continue
    n = int(r[samp])
    st = sorted(((int(r[i]), x[6:]) for i, x in stall_cols if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:3]
    lines.append((n, fname, int(r[0]), r[1].strip(), st))
tot = sum(l[0] for l in lines) or 1
for n, f, ln, text, st in sorted(lines, reverse=True)[:top]:
    reasons = ", ".join(f"{name} {100*v//max(n,1)}%" for v, name in st)
    print(f"{100*n/tot:5.1f}%  {f}:{ln:<4d} {text[:84]:84s} | {reasons}")
