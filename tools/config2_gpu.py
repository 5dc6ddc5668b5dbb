"""BASELINE.json config 2 on GPUs: the kdtree_mpi-equivalent (forest of W local trees, kdtree_mpi.cpp:204-253) with
one rank per B200.  `python -m torch.distributed.run --nproc-per-node W tools/config2_gpu.py [log2n] [dim] [seed]`

Every rank regenerates its slice, builds its local tree on its GPU, answers the 10 queries, and the distances are
min-reduced into rank 0 over peer memory (k_forest_min).  Rank 0 compares the 10 printed lines with the emulated
reference MPI run (`oracle_mpi lines`, the unmodified kdtree_mpi.cpp functions) and reports the build time
(CUDA events, max over ranks) next to the reference's own build time for the same W on this box's CPUs."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import torch.distributed as dist

import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel.forest import PeerMinReducer

rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 3
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 1
n, nq = 1 << log2n, 10
first, local = kd.models.slice_bounds(n, world, rank)
pts = kd.ops.generate_uniform(seed, dim, local, first).to(dev)
qs = kd.ops.generate_uniform(seed, dim, nq, n).to(dev)
perm = torch.empty(local, dtype=torch.int32, device=dev)
tp = torch.empty_like(pts)
ws = torch.empty(kd.ops.build_workspace_bytes(local, dim), dtype=torch.uint8, device=dev)
for _ in range(3):
    kd.ops.build_into(pts, perm, tp, ws)
torch.cuda.synchronize(); dist.barrier()
a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
a.record()
kd.ops.build_into(pts, perm, tp, ws, verify=False)
b.record()
idx, d2, d, vis = kd.ops.nearest(tp, perm, qs)
c.record()
torch.cuda.synchronize()
t = torch.tensor([a.elapsed_time(b), b.elapsed_time(c)], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
red = PeerMinReducer(nq)
best = red.reduce(d)
red.close()
if rank == 0:
    got = "".join(kd.utils.result_line(n + q, float(best[q])) + "\n" for q in range(nq))
    ref = ROOT / "baseline" / "_ref"
    want = subprocess.run([str(ref / "oracle_mpi"), "lines", str(seed), str(dim), str(n), str(nq), str(world)],
                          capture_output=True, text=True, check=True).stdout
    tref = json.loads(subprocess.run([str(ref / "oracle_mpi"), "time", str(seed), str(dim), str(n), str(nq), str(world)],
                                     capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1])
    print(json.dumps({"config": "kdtree_mpi-equivalent (forest)", "n": n, "dim": dim, "seed": seed, "world": world,
                      "lines_identical_to_reference_mpi_run": got == want, "lines": got.splitlines(),
                      "forest_build_ms_max_over_ranks": round(float(t[0]), 3), "queries_ms": round(float(t[1]), 3),
                      "points_per_s": round(n / (float(t[0]) * 1e-3)),
                      "reference_same_box_build_s": tref["build_s"], "reference_points_per_s": round(n / tref["build_s"]),
                      "speedup_vs_reference_build": round(tref["build_s"] / (float(t[0]) * 1e-3), 1)}), flush=True)
dist.barrier()
dist.destroy_process_group()
