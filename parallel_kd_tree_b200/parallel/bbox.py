"""Bounding box of a point set spread over ranks (distributed form).

Local per-axis min/max with the library's kernel (``kd::bbox``: per-thread running bounds, one shared-memory atomic
pair per thread, one global atomic pair per CTA and axis), then an all-reduce over the ranks:

* CUDA ranks: ``PeerMinMax`` — every rank folds its bounds into EVERY rank's small symmetric buffer with system-scope
  ``atomicMin`` / ``atomicMax`` over NVLink peer memory (``k_peer_minmax``, one kernel per rank, no collective call);
  ``torch.distributed`` only exchanges the CUDA IPC handles of the buffers once;
* CPU ranks (gloo, tests): ``all_reduce(MIN)`` / ``all_reduce(MAX)``.

No counterpart in the reference (it never needs a bounding box).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from .. import ops
from .._lib import check, lib


class PeerMinMax:
    """Symmetric buffers (one per rank, mapped into every peer) + the peer-memory min/max all-reduce."""

    def __init__(self, dim: int, group=None):
        self.group, self.dim = group, int(dim)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        L = lib()
        nbytes = int(L.kd_peer_minmax_buf_bytes(self.dim))
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        check(L.kd_symm_alloc(nbytes, C.byref(ptr), handle), "kd_symm_alloc")      # zero-initialised
        self._own = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)              # bootstrap only
        self._opened, ptrs = [], []
        for r in range(self.world):
            if r == self.rank:
                ptrs.append(self._own)
                continue
            p = C.c_void_p()
            check(L.kd_symm_open(handles[r], C.byref(p)), "kd_symm_open")
            self._opened.append(p.value)
            ptrs.append(p.value)
        self._ptrs = (C.c_void_p * self.world)(*ptrs)
        self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.epoch = 0
        dist.barrier(group=group)

    def reduce(self, local_min: torch.Tensor, local_max: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        self.epoch += 1
        mn = local_min.to(self.device, torch.float32).contiguous()
        mx = local_max.to(self.device, torch.float32).contiguous()
        out_min, out_max = torch.empty_like(mn), torch.empty_like(mx)
        st = torch.cuda.current_stream(self.device).cuda_stream
        check(lib().kd_peer_minmax_reduce(self._ptrs, self.rank, self.world, mn.data_ptr(), mx.data_ptr(), self.dim,
                                          out_min.data_ptr(), out_max.data_ptr(), self.epoch, self.err.data_ptr(), st),
              "kd_peer_minmax_reduce")
        torch.cuda.synchronize(self.device)
        if int(self.err.item()):
            raise RuntimeError(f"peer min/max reduce timed out waiting for a peer (code {int(self.err.item())})")
        return out_min, out_max

    def close(self):
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        for p in self._opened:
            lib().kd_symm_close(p)
        lib().kd_symm_free(self._own)


def distributed_bbox(local_points: torch.Tensor, group=None, reducer: PeerMinMax | None = None):
    """(min, max) per axis over the union of every rank's ``local_points``; identical on all ranks.

    CUDA points: peer-memory all-reduce (pass a ``PeerMinMax`` to reuse its buffers across calls); CPU points: gloo."""
    mn, mx = ops.bbox(local_points)
    if not local_points.is_cuda:
        mn, mx = mn.clone(), mx.clone()
        dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
        return mn, mx
    own = reducer is None
    red = reducer or PeerMinMax(local_points.shape[1], group)
    try:
        return red.reduce(mn, mx)
    finally:
        if own:
            red.close()
