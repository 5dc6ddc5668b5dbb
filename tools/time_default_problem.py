"""Device time of the reference's default problem (500 000 x 128, 10 queries, Utility.cpp:98-99) on one B200:
build (gather mode) and the 10 queries through the dense path and through the thread-per-query DFS."""
import json
import sys
import time

sys.path.insert(0, ".")
import torch

import parallel_kd_tree_b200 as kd

t0 = time.time()
pts, qs = kd.ops.generate_problem(1, 128, 500000, 10)
gen_s = time.time() - t0
dpts, dqs = pts.cuda(), qs.cuda()
for _ in range(2):
    tree = kd.KDTree(dpts)
torch.cuda.synchronize()


def timed(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best, out


build_ms, tree = timed(lambda: kd.KDTree(dpts))
dense_ms, r1 = timed(lambda: tree.nearest(dqs, mode="dense"))
dfs_ms, r2 = timed(lambda: tree.nearest(dqs, mode="dfs"), reps=1)
t0 = time.time()
tree2 = kd.KDTree(pts.cuda()); r3 = tree2.nearest(qs.cuda()); lines = [kd.utils.format_float(float(d)) for d in r3.dist.cpu()]
e2e_s = time.time() - t0
print(json.dumps({"problem": "500000 x 128, 10 queries, seed 1", "host_generate_s": round(gen_s, 3),
                  "build_ms": round(build_ms, 3), "query_dense_ms": round(dense_ms, 3), "query_dfs_ms": round(dfs_ms, 3),
                  "same_answers": bool(torch.equal(r1.index, r2.index) and torch.equal(r1.dist, r2.dist)),
                  "upload_build_query_download_wall_s": round(e2e_s, 4), "lines": lines,
                  "reference": "unmodified ./sequential: 9.15 s wall; harness phases build 2.3-2.9 s, queries 4.6 s (BASELINE.md rows 1-2)"}))
