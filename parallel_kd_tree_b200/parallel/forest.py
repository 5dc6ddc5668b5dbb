"""Forest mode with one process per GPU — what `mpirun -np W ./mpi` does (kdtree_mpi.cpp:170-291).

Rank r regenerates only its slice of the random stream (kdtree_mpi.cpp:19-41), builds its local tree,
answers the 10 queries on it, and the per-query distances are min-reduced into rank 0 — here by
system-scope atomicMin into rank 0's memory over NVLink (K8), not by a collective call.
``torch.distributed`` only bootstraps (configuration broadcast like the reference's 12-byte MPI_Bcast,
kdtree_mpi.cpp:199, and the IPC handle exchange).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import ops
from .._lib import check, lib
from ..models.forest import slice_bounds
from ..models.kdtree import KDTree
from ..utils import protocol
from ..utils.timing import PhaseTimer


class PeerMinReducer:
    """Owns one small symmetric buffer per rank and runs kd_forest_min_reduce."""

    def __init__(self, nq: int, group=None):
        import torch.distributed as dist
        self.group, self.nq = group, nq
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        L = lib()
        nbytes = int(L.kd_forest_buf_bytes(nq))
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        check(L.kd_symm_alloc(nbytes, C.byref(ptr), handle), "kd_symm_alloc")
        self._own = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        if self.rank == 0:
            self._root, self._opened = self._own, None
        else:
            p = C.c_void_p()
            check(L.kd_symm_open(handles[0], C.byref(p)), "kd_symm_open")
            self._root = self._opened = p.value
        self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.epoch = 0
        dist.barrier(group=group)

    def reduce(self, dist_local: torch.Tensor) -> torch.Tensor | None:
        self.epoch += 1
        out = torch.empty(self.nq, dtype=torch.float32, device=self.device) if self.rank == 0 else None
        st = torch.cuda.current_stream(self.device).cuda_stream
        d = dist_local.to(self.device, torch.float32).contiguous()
        check(lib().kd_forest_min_reduce(self._root, self._own, self.rank, self.world, d.data_ptr(), self.nq,
                                         out.data_ptr() if out is not None else None, self.epoch,
                                         self.err.data_ptr(), st), "kd_forest_min_reduce")
        torch.cuda.synchronize(self.device)
        if int(self.err.item()):
            raise RuntimeError(f"forest min-reduce timed out waiting for a peer (code {int(self.err.item())})")
        return out

    def close(self):
        import torch.distributed as dist
        dist.barrier(group=self.group)
        if self._opened:
            lib().kd_symm_close(self._opened)
        lib().kd_symm_free(self._own)


def run_forest_distributed(args, positional) -> int:
    """The `mpi` program under torchrun: rank 0 talks to stdin/stdout, every rank builds and queries."""
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and args.device != "cpu"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if use_cuda:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group("gloo")
    timer = PhaseTimer()
    debug = len(positional) > 0
    cfg = [None]
    if rank == 0:      # only rank 0 reads the problem (kdtree_mpi.cpp:185-197)
        try:
            cfg[0] = (protocol.specify_problem_argv(positional, prog="mpi") if debug
                      else protocol.specify_problem_stdin())
        except SystemExit as e:
            cfg[0] = ("exit", e.code)
    dist.broadcast_object_list(cfg, src=0)          # the reference's 12-byte MPI_Bcast (kdtree_mpi.cpp:199)
    if cfg[0][0] == "exit":
        dist.destroy_process_group()
        return int(cfg[0][1] or 1)
    seed, dim, n = cfg[0]
    nq = protocol.NUM_QUERIES
    first, local = slice_bounds(n, world, rank)
    pts = ops.generate_uniform(seed, dim, local, first)
    qs = ops.generate_uniform(seed, dim, nq, n)
    dev = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    tree = KDTree(pts.to(dev), id_offset=first)
    d_local = tree.nearest(qs.to(dev)).dist
    if use_cuda:
        red = PeerMinReducer(nq)
        best = red.reduce(d_local)
        red.close()
    else:   # CPU ranks (tests): the collective the reference itself uses
        best = d_local.clone()
        dist.reduce(best, dst=0, op=dist.ReduceOp.MIN)
    if rank == 0:
        best = best.cpu()
        for q in range(nq):
            print(protocol.result_line(n + q, float(best[q])))   # kdtree_mpi.cpp:255-260
        if debug:
            print(timer.elapsed_line())
        print("DONE", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0
