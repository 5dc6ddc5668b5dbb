"""Batched query throughput (BASELINE.json config 5: 2^24 8-D points build + batched k-NN traversal).
    python tools/bench_knn.py [log2n] [dim] [log2nq]
Queries are drawn from the same distribution as the points.  Device time, CUDA events, best of 3."""
import json
import sys

sys.path.insert(0, ".")
import torch

import parallel_kd_tree_b200 as kd

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 8
log2nq = int(sys.argv[3]) if len(sys.argv) > 3 else 20
n, nq = 1 << log2n, 1 << log2nq
pts = kd.ops.generate_uniform(1, dim, n).cuda()
qs = kd.ops.generate_uniform(2, dim, nq).cuda()


def timed(fn, reps=3):
    best, out = 1e30, None
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best, out


tree = kd.KDTree(pts)
build_ms, tree = timed(lambda: kd.KDTree(pts))
rows = {"n": n, "dim": dim, "nq": nq, "build_ms": round(build_ms, 3)}
ms, res = timed(lambda: tree.nearest(qs, mode="dfs"))
rows["nearest"] = {"ms": round(ms, 3), "queries_per_s": round(nq / (ms * 1e-3)), "mean_nodes_visited": float(res.visited.float().mean())}
for k in (1, 8, 32):
    ms, _ = timed(lambda: kd.ops.knn(tree.tree_points, tree.perm, qs, k))
    rows[f"knn_k{k}"] = {"ms": round(ms, 3), "queries_per_s": round(nq / (ms * 1e-3))}
# spot check against the CPU definition on a few queries
sub = qs[:256].cpu()
cpu_tree = kd.KDTree(pts.cpu()) if n <= (1 << 22) else None
if cpu_tree is not None:
    rows["k1_equals_cpu_on_256_queries"] = bool(torch.equal(tree.nearest(qs[:256]).index.cpu(), cpu_tree.nearest(sub).index))
print(json.dumps(rows))
