import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 with `pytest -m gpu`)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Oracle:
    """The unmodified reference functions behind baseline/_ref/oracle_* (oracle B)."""

    def __init__(self, ref_dir: Path):
        self.dir = ref_dir

    def _run(self, exe, *args, **kw):
        r = subprocess.run([str(self.dir / exe), *map(str, args)], capture_output=True, text=True, check=True, **kw)
        return r.stdout

    def points(self, seed, dim, n, nq, tmp):
        out = Path(tmp) / f"pts_{seed}_{dim}_{n}_{nq}.bin"
        self._run("oracle_seq", "points", f"seed:{seed}", dim, n, nq, out)
        return np.fromfile(out, dtype=np.float32).reshape(n + nq, dim)

    def inorder_ids(self, src, dim, n, tmp):
        out = Path(tmp) / "inorder.bin"
        self._run("oracle_seq", "inorder", src, dim, n, 0, out)
        return np.fromfile(out, dtype=np.int32)

    def tree_text(self, src, dim, n):
        return self._run("oracle_seq", "tree", src, dim, n, 0)

    def head_text(self, src, dim, n):
        return self._run("oracle_seq", "head", src, dim, n, 0)

    def lines(self, src, dim, n, nq):
        return self._run("oracle_seq", "lines", src, dim, n, nq)

    def queryx(self, src, dim, n, nq):
        rows = []
        for ln in self._run("oracle_seq", "queryx", src, dim, n, nq).splitlines():
            q, nn, b2, b1 = ln.split()
            rows.append((int(q), int(nn), int(b2, 16), int(b1, 16)))
        return rows

    def forest_lines(self, seed, dim, n, nq, world):
        return self._run("oracle_mpi", "lines", seed, dim, n, nq, world)

    def forest_queryx(self, seed, dim, n, nq, world):
        rows = []
        for ln in self._run("oracle_mpi", "queryx", seed, dim, n, nq, world).splitlines():
            q, r, nn, b = ln.split()
            rows.append((int(q), int(r), int(nn), int(b, 16)))
        return rows


@pytest.fixture(scope="session")
def oracle():
    ref = ROOT / "baseline" / "_ref"
    if not (ref / "oracle_seq").exists() or not (ref / "oracle_mpi").exists():
        try:
            subprocess.run(["bash", str(ROOT / "baseline" / "build_ref.sh")], check=True, capture_output=True)
        except Exception as e:  # reference sources not on this machine
            pytest.skip(f"reference oracle unavailable: {e}")
    return Oracle(ref)


@pytest.fixture(scope="session")
def kd():
    import parallel_kd_tree_b200 as kd
    kd.lib()
    return kd


def write_points(path, arr):
    np.ascontiguousarray(arr, dtype=np.float32).tofile(path)
    return f"file:{path}"
