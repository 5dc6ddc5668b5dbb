"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel name and per level."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    rows.append((name, us, r.get("Grid Size", "")))
tot = sum(u for _, u, _ in rows)
agg = defaultdict(lambda: [0.0, 0])
for name, us, _ in rows:
    agg[name][0] += us
    agg[name][1] += 1
print(f"{len(rows)} launches, {tot/1e3:.3f} ms total (serialised, cold cache: compare shares)")
for name, (us, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"  {name:40s} {us/1e3:9.3f} ms  {100*us/tot:5.1f}%  x{cnt}")
if len(sys.argv) > 2:
    for name, us, grid in rows:
        print(f"{name:32s} {us:10.1f} us  grid {grid}")
