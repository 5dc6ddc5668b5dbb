"""Per-CUDA-source-line instruction counts and stall samples of an .ncu-rep (needs -lineinfo + --import-source on).
Usage: ncu_lines.py rep [top] [first_line last_line]"""
import csv
import io
import os
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
lo = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi = int(sys.argv[4]) if len(sys.argv) > 4 else 1 << 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
agg, h, fname, cur = {}, None, "", None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Path", "File Name"):
        fname = os.path.basename(r[1])
        continue
    if r[0] == "Line No":
        h = r
        continue
    if h is None:
        continue
    if r[0].strip().isdigit():
        cur = (fname, int(r[0]), r[1].strip())
        agg.setdefault(cur, [0, 0, 0, 0, {}])
        continue
    if cur is None or len(r) < len(h) or not r[2].startswith("0x"):
        continue
    d = dict(zip(h[4:], r[4:]))
    a = agg[cur]
    num = lambda k: int(d[k]) if d.get(k, "").isdigit() else 0
    a[0] += num("Instructions Executed")
    a[1] += num("# Samples")
    a[2] += num("L1 Wavefronts Shared")
    a[3] += num("L1 Wavefronts Shared Ideal")
    for k, v in d.items():
        if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v) > 0:
            a[4][k[6:]] = a[4].get(k[6:], 0) + int(v)
rows = [(a[0], a[1], f, ln, text, a[2], a[3], sorted(((v, k) for k, v in a[4].items()), reverse=True)[:3])
        for (f, ln, text), a in agg.items()]
ti = sum(x[0] for x in rows) or 1
ts = sum(x[1] for x in rows) or 1
print(f"total warp instructions {ti}, samples {ts}")
sel = [x for x in rows if x[2].endswith(".cu") and lo <= x[3] <= hi] if lo else rows
for ie, sm, f, ln, text, wf, wi, st in sorted(sel, reverse=True)[:top]:
    reasons = ", ".join(f"{name} {100*v//max(sm,1)}%" for v, name in st)
    print(f"{100*ie/ti:5.1f}%i {100*sm/ts:5.1f}%s  {f}:{ln:<4d} {text[:70]:70s} | smem wf {wf}/{wi} | {reasons}")
