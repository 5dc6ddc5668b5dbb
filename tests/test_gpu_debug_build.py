"""Sanitizer substitute: the -DKD_DEBUG build of the library (a bounds assertion on every computed store offset of the
build kernels, local and peer) runs the small-case suite in a process of its own; no assertion may fire and the trees
must still equal the CPU builder's."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

CODE = r"""
import torch
import parallel_kd_tree_b200 as kd
from parallel_kd_tree_b200._lib import check
from parallel_kd_tree_b200.parallel.dist_build import EmulatedDistBuild
L = kd.lib()
assert L.kd_debug_enabled() == 1, "not the KD_DEBUG library"
g = torch.Generator().manual_seed(0)
cases = [(1, 1), (2, 2), (3, 3), (9, 2), (100, 3), (4095, 1), (4096, 3), (4097, 3), (8193, 2), (70000, 8), (300000, 3),
         (20000, 16), ((1 << 20) + 7, 3)]
for n, dim in cases:
    pts = kd.ops.generate_uniform(n % 97 + 1, dim, n)
    assert torch.equal(kd.ops.build(pts.cuda())[0].cpu(), kd.ops.build(pts)[0]), (n, dim)
    check(L.kd_debug_bounds_status(), f"n={n} dim={dim}")
ties = torch.randint(-3, 4, (200000, 3), generator=g).float()
assert torch.equal(kd.ops.build(ties.cuda())[0].cpu(), kd.ops.build(ties)[0])
check(L.kd_debug_bounds_status(), "ties")
for n, dim, world in [(100003, 3, 2), (250000, 4, 4), (300000, 3, 8), (150000, 8, 8), ((1 << 23) + 3, 3, 2)]:
    pts = kd.ops.generate_uniform(5, dim, n)
    ref = kd.ops.build(pts)[0]
    outs = EmulatedDistBuild(n, dim, world).build(pts)
    torch.cuda.synchronize()
    assert all(torch.equal(p.cpu(), ref) for p, _ in outs), (n, dim, world)
    check(L.kd_debug_bounds_status(), f"dist n={n} dim={dim} world={world}")
print("debug build ok")
"""


def test_debug_build_runs_the_small_suite_without_a_bounds_assertion():
    env = dict(os.environ, KD_B200_DEBUG="1")
    r = subprocess.run([sys.executable, "-c", CODE], cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "debug build ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
