"""In-tree build of libkdtree_b200.so (host core + sm_100a kernels) with nvcc.

The reference builds with a 50-line Makefile (Makefile:1-50); here one nvcc
invocation per source (-gencode arch=compute_100a,code=sm_100a -lineinfo,
-fmad=false so device float math is contraction-free like the reference's
`-mavx`-only host build) and one link step.  The .so stays in the package
directory so that it travels with a repository snapshot.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OUT_DIR = PKG / "_C"
LIB = OUT_DIR / "libkdtree_b200.so"
LIB_DBG = OUT_DIR / "dbg" / "libkdtree_b200_dbg.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall,-Wno-unknown-pragmas",
    "-Xptxas", "-v",
    "-I", str(PKG.parent / "include"), "-I", str(CSRC),
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libkdtree_b200.so")
    return exe


def sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))


def _stamp() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*")) + [PKG.parent / "include" / "kdtree_b200.h", Path(__file__)]):
        if p.is_file():
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()


def needs_build(debug: bool = False) -> bool:
    lib, out = (LIB_DBG, OUT_DIR / "dbg") if debug else (LIB, OUT_DIR)
    stamp = out / "build.stamp"
    return not (lib.exists() and stamp.exists() and stamp.read_text() == _stamp())


def build_library(force: bool = False, verbose: bool = False, debug: bool = False) -> Path:
    """Compile every source under csrc/ for sm_100a and link the shared library.

    ``debug``: the -DKD_DEBUG variant (``libkdtree_b200_dbg.so``, objects under ``_C/dbg``): same kernels with a
    bounds assertion on every computed store offset (kd_common.h); ``KD_B200_DEBUG=1`` makes the package load it."""
    lib, out_dir = (LIB_DBG, OUT_DIR / "dbg") if debug else (LIB, OUT_DIR)
    if not force and not needs_build(debug):
        return lib
    out_dir.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    srcs = sources()
    log_dir = out_dir / "log"
    log_dir.mkdir(exist_ok=True)
    flags = NVCC_FLAGS + (["-DKD_DEBUG"] if debug else [])

    def compile_one(src: Path) -> Path:
        obj = out_dir / (src.stem + ".o")
        cmd = [nvcc, *flags, "-c", str(src), "-o", str(obj)]
        if src.suffix == ".cpp":
            cmd = [nvcc, *flags, "-x", "cu", "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        (log_dir / (src.stem + ".ptxas.txt")).write_text(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    tmp = out_dir / (lib.name + ".tmp")
    r = subprocess.run(
        [nvcc, "-shared", "-o", str(tmp), *map(str, objs), "-cudart", "static", "-lpthread"],
        capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, lib)
    (out_dir / "build.stamp").write_text(_stamp())
    return lib


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True, debug="--debug" in sys.argv))
