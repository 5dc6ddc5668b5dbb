"""torchrun check of the peer-memory bounding-box all-reduce (K12 over ranks): every rank compares with the bounds of
the whole point set.  `torchrun --nproc-per-node W tools/bbox_check.py`"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel.bbox import PeerMinMax, distributed_bbox

rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
ok = True
for n, dim in [(100003, 3), (50000, 128), (1 << 20, 8)]:
    pts = kd.ops.generate_clustered(3, dim, n)
    first, local = kd.models.slice_bounds(n, world, rank)
    red = PeerMinMax(dim)
    for trial in range(3):                       # the buffers are reused: epochs grow
        mn, mx = distributed_bbox(pts[first:first + local].to(dev), reducer=red)
        same = torch.equal(mn.cpu(), pts.min(0).values) and torch.equal(mx.cpu(), pts.max(0).values)
        ok &= same
        print(f"rank {rank}/{world} n={n} dim={dim} trial {trial}: {'OK' if same else 'MISMATCH'}", flush=True)
    red.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
