"""BASELINE.json config 2 (CPU part): kdtree_mpi-equivalent output at world sizes 2/4/8 — the 10 printed lines of
the forest mode against the emulated reference MPI run (unmodified kdtree_mpi.cpp functions, oracle_mpi)."""
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import parallel_kd_tree_b200 as kd

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n, dim, seed = 1 << log2n, 3, 1
ref = ROOT / "baseline" / "_ref"
seq = subprocess.run([str(ref / "oracle_seq"), "lines", f"seed:{seed}", str(dim), str(n), "10"], capture_output=True,
                     text=True, check=True).stdout
print(f"config 2 (host part): N = 2^{log2n}, dim = {dim}, seed = {seed}")
for world in (2, 4, 8):
    t0 = time.time()
    want = subprocess.run([str(ref / "oracle_mpi"), "lines", str(seed), str(dim), str(n), "10", str(world)],
                          capture_output=True, text=True, check=True).stdout
    t_ref = time.time() - t0
    t0 = time.time()
    forest, qs = kd.models.Forest.from_seed(seed, dim, n, world)
    dist, _ = forest.nearest(qs)
    got = "".join(kd.utils.result_line(n + q, float(dist[q])) + "\n" for q in range(10))
    t_ours = time.time() - t0
    differ_seq = sum(a != b for a, b in zip(want.splitlines(), seq.splitlines()))
    print(f"W = {world}: forest lines identical to the emulated reference MPI run: {got == want}; "
          f"reference {t_ref:.1f} s (ranks run one after the other), this library on the CPU {t_ours:.1f} s; "
          f"{differ_seq} of 10 lines differ from the sequential program's (expected in low dimension, SURVEY.md §0)")
