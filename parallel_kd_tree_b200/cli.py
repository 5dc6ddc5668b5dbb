"""Command-line drivers with the reference's I/O contract (kdtree_sequential.cpp:140-208,
kdtree_mpi.cpp:170-291).

    python -m parallel_kd_tree_b200 sequential              # READY, seed on stdin, 500000 x 128
    python -m parallel_kd_tree_b200 sequential SEED DIM N   # the reference's DEBUG argv mode (+ elapsed time)
    python -m parallel_kd_tree_b200 mpi [--world W] [SEED DIM N]
    torchrun --nproc-per-node W -m parallel_kd_tree_b200 mpi ...   # one rank per GPU, peer-memory min-reduce

stdout carries READY / the result lines / DONE, stderr the prompts, exactly like the reference.
Extra switches (ours): --device, --print-tree, --print-head, --trace.
"""
from __future__ import annotations

import argparse
import os
import sys

import torch

from . import ops
from .models import Forest, KDTree
from .utils import protocol, trace
from .utils.timing import PhaseTimer


def _device(arg: str | None) -> torch.device:
    if arg:
        return torch.device(arg)
    return torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")


def run_sequential(args, positional) -> int:
    timer = PhaseTimer()
    debug = len(positional) > 0
    if debug:
        seed, dim, n = protocol.specify_problem_argv(positional, prog="sequential")
    else:
        seed, dim, n = protocol.specify_problem_stdin()
    dev = _device(args.device)
    nq = protocol.NUM_QUERIES
    with trace.span("generate"):
        pts, qs = ops.generate_problem(seed, dim, n, nq)  # last rows are the queries (kdtree_sequential.cpp:156-157)
    tree = KDTree(pts.to(dev))
    res = tree.nearest(qs.to(dev))
    dist = res.dist.cpu()
    for q in range(nq):
        # query IDs are num_points + q (kdtree_sequential.cpp:170)
        print(protocol.result_line(n + q, float(dist[q])))
    if args.print_tree:
        sys.stdout.write(tree.format_tree())
    if args.print_head:
        sys.stdout.write(tree.format_head_and_leaves())
    if debug:
        print(timer.elapsed_line())
    print("DONE", flush=True)
    return 0


def run_mpi(args, positional) -> int:
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    if world_env > 1:
        from .parallel.forest import run_forest_distributed
        return run_forest_distributed(args, positional)
    world = args.world
    timer = PhaseTimer()
    debug = len(positional) > 0
    if debug:
        seed, dim, n = protocol.specify_problem_argv(positional, prog="mpi")
    else:
        seed, dim, n = protocol.specify_problem_stdin()
    nq = protocol.NUM_QUERIES
    if torch.cuda.is_available() and not args.device:
        devices = [torch.device("cuda", i) for i in range(min(world, torch.cuda.device_count()))]
    else:
        devices = [_device(args.device)]
    forest, qs = Forest.from_seed(seed, dim, n, world, devices, nq)
    dist, _ = forest.nearest(qs)
    for q in range(nq):
        print(protocol.result_line(n + q, float(dist[q])))   # kdtree_mpi.cpp:255-260
    if debug:
        print(timer.elapsed_line())
    print("DONE", flush=True)
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="parallel_kd_tree_b200", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("program", choices=["sequential", "mpi"])
    ap.add_argument("--device", default=None)
    ap.add_argument("--world", type=int, default=16, help="ranks of the forest (reference: mpirun -np 16, Makefile:36)")
    ap.add_argument("--print-tree", action="store_true")
    ap.add_argument("--print-head", action="store_true")
    ap.add_argument("--trace", action="store_true", help="per-phase timings on stderr (ours)")
    args, positional = ap.parse_known_args(argv)
    if args.trace:
        trace.enable()
        trace.reset()
    try:
        if args.program == "sequential":
            return run_sequential(args, positional)
        return run_mpi(args, positional)
    finally:
        if args.trace:
            trace.report()


if __name__ == "__main__":
    raise SystemExit(main())
