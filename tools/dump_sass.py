"""Writes profiles/sass/<kernel>.sass (cuobjdump -sass of libkdtree_b200.so, one file per kernel; for templated
kernels only the 3-D instantiation) and profiles/sass/SUMMARY.md with each kernel's opcode histogram."""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = ROOT / "parallel_kd_tree_b200" / "_C" / "libkdtree_b200.so"
out = ROOT / "profiles" / "sass"
out.mkdir(parents=True, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
keep = {}
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    short = re.sub(r"\(.*", "", dem).replace("kd::", "").replace("void ", "")
    m = re.match(r"(\w+)<(.*)>", short)
    if m:
        args = [a.strip() for a in m.group(2).split(",")]
        first = args[0].replace("(int)", "")
        if first.isdigit() and first not in ("3", "512"):
            continue
    keep[short] = f
rows = []
for short, body in sorted(keep.items()):
    lines = [l for l in body.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
    ops = collections.Counter()
    for l in lines:
        t = re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", l).split(";")[0].split()
        if not t:
            continue
        op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
        ops[op.split(".")[0]] += 1
    fn = re.sub(r"[^\w]+", "_", short).strip("_") + ".sass"
    clean = re.sub(r"\s*/\* 0x[0-9a-f]+ \*/", "", body)          # drop the encodings, keep address + instruction
    clean = "\n".join(l for l in clean.splitlines() if l.strip())
    (out / fn).write_text("Function : " + clean + "\n")
    top = ", ".join(f"{k} {v}" for k, v in ops.most_common(10))
    flags = [k for k in ("UBLKCP", "UTMALDG", "SYNCS", "MATCH", "SHFL", "ATOMS", "ATOMG", "RED", "VOTE", "BAR", "LDS", "STS",
                         "HMMA", "UTCHMMA", "LDGSTS", "MEMBAR", "ERRBAR") if ops.get(k)]
    rows.append(f"| `{short}` | {len(lines)} | {' '.join(flags)} | {top} |")
(out / "SUMMARY.md").write_text(
    "# SASS listings (sm_100a, cuobjdump -sass)\n\nOne file per kernel in this directory; templated kernels are listed "
    "for their 3-D instantiation.\nNo tensor-core mnemonics (HMMA / UTC*MMA) are expected: the workload has no GEMM "
    "(DESIGN.md).\n\n| kernel | SASS instructions | notable mnemonics | top opcodes |\n|---|---|---|---|\n" + "\n".join(rows) + "\n")
print(f"{len(keep)} kernels -> {out}")
