"""BASELINE.json config 1: kdtree_sequential, 2^20 uniform 2-D points, CPU build, tree diffed against the
unmodified reference (oracle B).  Prints a short report (stored in profiles/)."""
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import parallel_kd_tree_b200 as kd

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 2
n, seed = 1 << log2n, 1
ref = ROOT / "baseline" / "_ref"
with tempfile.TemporaryDirectory() as tmp:
    out = Path(tmp) / "inorder.bin"
    t0 = time.time()
    subprocess.run([str(ref / "oracle_seq"), "inorder", f"seed:{seed}", str(dim), str(n), "0", str(out)], check=True)
    t_ref = time.time() - t0
    ref_ids = np.fromfile(out, dtype=np.int32)
pts, qs = kd.ops.generate_problem(seed, dim, n, 10)
t0 = time.time()
tree = kd.KDTree(pts)
t_ours = time.time() - t0
ours = tree.ids_inorder().numpy()
diff = int((ours != ref_ids).sum())
print(f"config 1: N = 2^{log2n}, dim = {dim}, seed = {seed}")
print(f"reference (oracle_seq, incl. generation and dump): {t_ref:.2f} s; this library, CPU build: {t_ours:.2f} s")
print(f"in-order slots that differ from the unmodified reference: {diff} of {n}")
print(f"this library's tree, strict rule check (canonical tie-break): {tree.check()} violations")
ref_perm = torch.from_numpy(ref_ids - 1)
print(f"reference's tree, strict rule check: {kd.ops.check_tree(pts, ref_perm)} violations "
      f"(> 0 means std::sort ordered some equal coordinates differently)")
print(f"reference's tree, tie-tolerant rule check: {kd.ops.check_tree(pts, ref_perm, allow_ties=True)} violations")
print(f"this library's tree, tie-tolerant rule check: {kd.ops.check_tree(pts, tree.perm, allow_ties=True)} violations")
dup = n - len(np.unique(pts[:, 0].numpy()))
print(f"duplicate axis-0 coordinates in the input: {dup}")
res = tree.nearest(qs)
lines = "".join(kd.utils.result_line(n + q, float(res.dist[q])) + "\n" for q in range(10))
ref_lines = subprocess.run([str(ref / "oracle_seq"), "lines", f"seed:{seed}", str(dim), str(n), "10"],
                           capture_output=True, text=True, check=True).stdout
print(f"10 query lines identical to the reference: {lines == ref_lines}")
