"""Multi-GPU arm of bench.py: the distributed global build (one process per GPU, peer memory)."""
from __future__ import annotations

import os
import time

import torch

from .dist_build import IpcDistBuild


def run(args, kd, dist, rank, world, dev, sampler_cls, baseline_pps):
    n, dim = 1 << args.log2n, args.dim
    first, local = kd.models.slice_bounds(n, world, rank)
    t0 = time.time()
    # every rank regenerates only its slice of the stream, like kdtree_mpi.cpp:19-41
    if args.dist == "clustered":
        host_local = kd.ops.generate_clustered(args.seed, dim, n, pin_memory=False)[first:first + local].contiguous().pin_memory()
    else:
        host_local = kd.ops.generate_uniform(args.seed, dim, local, first, pin_memory=True)
    gen_s = time.time() - t0
    local_pts = host_local.to(dev, non_blocking=True)
    db = IpcDistBuild(n, dim, with_tree_points=True)
    host_perm = torch.empty(local, dtype=torch.int32, pin_memory=True)
    torch.cuda.synchronize()

    def step():
        db.load(local_pts)      # SoA transpose of the local slice into the symmetric heap (k_init's counterpart)
        db.run()

    for _ in range(args.warmup):
        step()
    db.status()                 # surfaces device-side failures / peer time-outs (untimed)
    launches_per_step = int(kd.lib().kd_last_launch_count())
    dist.barrier()
    torch.cuda.synchronize()

    sampler = sampler_cls(int(os.environ.get("LOCAL_RANK", "0")))
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    dist.barrier()
    torch.cuda.synchronize()
    for a, b in evs:
        a.record()
        step()
        b.record()
    torch.cuda.synchronize()
    dist.barrier()
    total_ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], device=dev)
    dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)     # max over ranks of the device time
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = float(total_ms.item()) / args.steps
    db.status()

    # End to end: every step uploads this rank's slice from pinned host memory, runs the distributed build and
    # downloads this rank's share of perm.  Steps are pipelined like a stream of requests: double-buffered
    # input slices, the result snapshotted device-to-device, copies on their own streams.
    pts2 = [local_pts, torch.empty_like(local_pts)]
    stage2 = [torch.empty(local, dtype=torch.int32, device=dev) for _ in range(2)]
    host_perm2 = [host_perm, torch.empty(local, dtype=torch.int32, pin_memory=True)]
    s_h2d, s_d2h = torch.cuda.Stream(), torch.cuda.Stream()
    main = torch.cuda.current_stream()

    def e2e_run(k):
        ev_build, ev_d2h = [None] * k, [None] * k
        for i in range(k):
            b = i & 1
            with torch.cuda.stream(s_h2d):
                if i >= 2:
                    s_h2d.wait_event(ev_build[i - 2])
                pts2[b].copy_(host_local, non_blocking=True)
                ev_up = torch.cuda.Event()
                ev_up.record(s_h2d)
            main.wait_event(ev_up)
            if i >= 2:
                main.wait_event(ev_d2h[i - 2])
            db.load(pts2[b])
            db.run()
            # 0.1 ms device-to-device snapshot of this rank's share: the next build overwrites the tree arrays while
            # the 134 MB download (7 ms when eight ranks share the host's PCIe, profiles/h2d_probe_8gpu.json) is in flight
            stage2[b].copy_(db.perm[first:first + local])
            ev_build[i] = torch.cuda.Event()
            ev_build[i].record(main)
            with torch.cuda.stream(s_d2h):
                s_d2h.wait_event(ev_build[i])
                host_perm2[b].copy_(stage2[b], non_blocking=True)
                ev_d2h[i] = torch.cuda.Event()
                ev_d2h[i].record(s_d2h)
        torch.cuda.synchronize()

    e2e_run(2)
    dist.barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    e2e = torch.tensor([(time.perf_counter() - t0) / args.steps], device=dev)
    dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e.item())

    # cheap sanity check of the replicated result: every rank must hold the same permutation
    chk = torch.stack([db.perm.long().sum(), (db.perm.long() * torch.arange(n, device=dev)).sum() % 1000003])
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    consistent = bool(torch.equal(lo, hi)) and int(chk[0].item()) == n * (n - 1) // 2
    # The RIGHT tree, not only a consistent one: rank 0 builds the same points on its own GPU alone and compares the
    # two permutations element by element (the other ranks' replicas are tied to rank 0's by the checksums above).
    matches = None
    if os.environ.get("KD_BENCH_SKIP_TREE_CHECK") != "1":
        db.status()
        if rank == 0:
            full = (kd.ops.generate_clustered(args.seed, dim, n) if args.dist == "clustered"
                    else kd.ops.generate_uniform(args.seed, dim, n)).to(dev)
            ref_perm = torch.empty(n, dtype=torch.int32, device=dev)
            ws = torch.empty(kd.ops.build_workspace_bytes(n, dim), dtype=torch.uint8, device=dev)
            kd.ops.build_into(full, ref_perm, None, ws, verify=True)
            matches = bool(torch.equal(ref_perm, db.perm))
            del full, ref_perm, ws
        dist.barrier()
    value = n / (ms_per_step * 1e-3)
    out = {
        "metric": "kd-tree build points/sec", "value": value, "unit": "points/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": value / baseline_pps, "dtype": "fp32", "data": "synthetic",
        "config": {"model": "kd-tree build (reference rule, ONE global tree, perm + tree-ordered coords replicated "
                            "on every GPU)",
                   "n_points": n, "dim": dim, "distribution": args.dist, "seed": args.seed, "global_batch": n,
                   "parallelism": f"distributed-global-tree x{world} (peer-memory select / partition / all-gather, "
                                  "no NCCL on the data path)",
                   "l2": "inputs and working set exceed the 126 MB L2; no flush needed",
                   "baseline": "BASELINE.md row 15: reference forest W=8 build, 4.2e5 points/s",
                   "replicas_consistent": consistent, "tree_matches_single_gpu": matches,
                   "heap": "torch symmetric memory + NVLS multicast stores for the tree arrays" if db.multicast
                   else ("torch symmetric memory, unicast P2P stores" if db._symm is not None else "cudaMalloc + CUDA IPC, unicast P2P stores"), "host_generate_s": round(gen_s, 2)},
        "clocks": clocks,
        "e2e": {"value": n / e2e_s, "unit": "points/s", "ms_per_step": e2e_s * 1e3,
                "h2d_bytes_per_step": host_local.numel() * 4 * world, "d2h_bytes_per_step": n * 4,
                "pipelined": "steps overlap (double buffers, copy streams); wall clock over all steps / steps"},
        "gpu_launches": int(launches_per_step * args.steps),
    }
    db.close()
    return out
