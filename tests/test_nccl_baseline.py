"""The NCCL + library-sort baseline (parallel/nccl_baseline.py) must produce the same tree as the native
builder; its distributed path is exercised here with gloo on CPU (2 and 4 ranks)."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["KD_ROOT"])
import torch, torch.distributed as dist
import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200.parallel import nccl_baseline as nb
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ok = True
for n, dim, seed in [(5003, 3, 1), (20000, 2, 2), (4099, 1, 3)]:
    pts = kd.ops.generate_uniform(seed, dim, n)
    first, local = kd.models.slice_bounds(n, world, rank)
    perm = nb.nccl_build(pts[first:first + local].clone(), n)
    ref = kd.ops.build(pts)[0]
    same = torch.equal(perm, ref)
    ok &= same
    print(f"rank {rank} n={n} dim={dim}: {'OK' if same else 'MISMATCH'}", flush=True)
from parallel_kd_tree_b200.parallel.bbox import distributed_bbox
allp = kd.ops.generate_uniform(9, 5, 4001)
first, local = kd.models.slice_bounds(4001, world, rank)
mn, mx = distributed_bbox(allp[first:first + local])
ok &= torch.equal(mn, allp.min(0).values) and torch.equal(mx, allp.max(0).values)
g = torch.Generator().manual_seed(7)
pts = torch.randint(-3, 4, (9001, 3), generator=g).float()
first, local = kd.models.slice_bounds(9001, world, rank)
ok &= torch.equal(nb.nccl_build(pts[first:first + local].clone(), 9001), kd.ops.build(pts)[0])
dist.destroy_process_group()
sys.exit(0 if ok else 1)
'''


def test_library_sort_build_single_device(kd):
    from parallel_kd_tree_b200.parallel.nccl_baseline import sort_build
    for n, dim, seed in [(1, 2, 1), (2, 2, 1), (3, 3, 1), (7, 2, 3), (1000, 3, 4), (4097, 2, 5), (30000, 3, 6)]:
        pts = kd.ops.generate_uniform(seed, dim, n)
        assert torch.equal(sort_build(pts), kd.ops.build(pts)[0]), (n, dim)
    g = torch.Generator().manual_seed(0)
    pts = torch.randint(-2, 3, (5000, 3), generator=g).float()
    pts[::13] = -0.0
    assert torch.equal(sort_build(pts), kd.ops.build(pts)[0])


@pytest.mark.slow
@pytest.mark.parametrize("world", [2, 4])
def test_collective_build_with_gloo(tmp_path, world):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, KD_ROOT=str(ROOT))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "MISMATCH" not in r.stdout
