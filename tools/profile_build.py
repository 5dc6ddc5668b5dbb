"""One warm-up build, then one build between cudaProfilerStart/Stop (for `ncu --profile-from-start off`)."""
import argparse
import sys
import time

sys.path.insert(0, ".")
import torch

import parallel_kd_tree_b200 as kd

ap = argparse.ArgumentParser()
ap.add_argument("--log2n", type=int, default=24)
ap.add_argument("--dim", type=int, default=3)
ap.add_argument("--dist", default="uniform")
args = ap.parse_args()
n = 1 << args.log2n
gen = kd.ops.generate_clustered if args.dist == "clustered" else kd.ops.generate_uniform
pts = gen(1, args.dim, n).cuda()
perm = torch.empty(n, dtype=torch.int32, device="cuda")
tp = torch.empty_like(pts)
ws = torch.empty(kd.ops.build_workspace_bytes(n, args.dim), dtype=torch.uint8, device="cuda")
kd.ops.build_into(pts, perm, tp, ws, verify=True)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
kd.ops.build_into(pts, perm, tp, ws)
b.record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f"build n=2^{args.log2n} dim={args.dim}: {a.elapsed_time(b):.3f} ms, launches {kd.ops.last_launch_count()}")
