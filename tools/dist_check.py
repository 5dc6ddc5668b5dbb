"""torchrun -m check of the multi-process distributed build: every rank compares the replicated tree with a
single-GPU build of the same points.  `torchrun --nproc-per-node W tools/dist_check.py [log2n ...]`"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel.dist_build import IpcDistBuild

rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
ok = True
for spec in (sys.argv[1:] or ["100003x3", "1048576x2", "4194304x3"]):
    clustered = spec.startswith("clustered:")      # Gaussian blobs (BASELINE.json config 4): uneven exchange volumes
    n, dim = (int(x) for x in spec.split(":")[-1].split("x"))
    first, local = kd.models.slice_bounds(n, world, rank)
    pts = kd.ops.generate_clustered(5, dim, n) if clustered else kd.ops.generate_uniform(5, dim, n)
    db = IpcDistBuild(n, dim)
    if rank == 0:
        print(f"heap backend: symm={db._symm is not None} multicast={db.multicast}", flush=True)
    for trial in range(2):
        perm, tp = db.build(pts[first:first + local])
        db.status()
        ref_perm, ref_tp = kd.ops.build(pts.to(dev))
        torch.cuda.synchronize()
        same = torch.equal(perm, ref_perm) and torch.equal(tp, ref_tp)
        ok &= same
        print(f"rank {rank}/{world} n={n} dim={dim} trial {trial}: {'OK' if same else 'MISMATCH'}", flush=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    a.record(); db.load(pts[first:first + local].to(dev)); db.run(); b.record(); torch.cuda.synchronize()
    print(f"rank {rank} n={n}: {a.elapsed_time(b):.3f} ms", flush=True)
    db.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
