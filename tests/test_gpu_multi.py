"""The real multi-process paths, one rank per GPU (skipped on boxes with a single GPU):

* IpcDistBuild — symmetric heaps (torch symmetric memory + NVLS multicast, and cudaMalloc + CUDA IPC), true
  cross-GPU flags, peer stores: the replicated tree must equal the single-GPU build bit for bit, on uniform 3-D
  and 8-D points and on clustered points, twice per heap (growing epochs);
* forest mode (`python -m parallel_kd_tree_b200 mpi`, K8 k_forest_min over peer memory): the printed lines must be
  the emulated reference MPI run's (unmodified kdtree_mpi.cpp functions, kdtree_mpi.cpp:252-260).
"""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs at least two GPUs")
WORLDS = [w for w in (2, 4, 8) if w <= NGPU]


def _torchrun(world, args, env=None, port=29600, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port)] + args
    return subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, **(env or {})), capture_output=True, text=True,
                          timeout=timeout)


@needs2
@pytest.mark.parametrize("backend", ["symm", "ipc"])
@pytest.mark.parametrize("world", WORLDS)
def test_multi_process_global_build_equals_single_gpu_build(world, backend):
    specs = ["100003x3", "1048576x2", "300000x8", "clustered:2097152x3", "60000x128"]
    r = _torchrun(world, ["tools/dist_check.py"] + specs, env={"KD_DIST_BACKEND": backend},
                  port=29600 + world + (10 if backend == "ipc" else 0))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout
    assert r.stdout.count(": OK") == 2 * world * len(specs)


@needs2
@pytest.mark.parametrize("world", WORLDS)
def test_forest_mode_prints_the_reference_mpi_lines(world, oracle):
    n, dim, seed = 300000, 3, 4
    want = oracle._run("oracle_mpi", "lines", seed, dim, n, 10, world)
    r = _torchrun(world, ["-m", "parallel_kd_tree_b200", "mpi", str(seed), str(dim), str(n)], port=29650 + world)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("ID:")]
    assert "\n".join(lines) + "\n" == want
    assert "READY" in r.stdout and r.stdout.rstrip().endswith("DONE")


@needs2
@pytest.mark.parametrize("world", WORLDS[:1] + WORLDS[-1:] if len(WORLDS) > 1 else WORLDS)
def test_peer_memory_bounding_box_all_reduce(world):
    # K12 over ranks: system-scope atomicMin / atomicMax into every rank's symmetric buffer, no collective call
    r = _torchrun(world, ["tools/bbox_check.py"], port=29680 + world)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout and r.stdout.count(": OK") == 9 * world
