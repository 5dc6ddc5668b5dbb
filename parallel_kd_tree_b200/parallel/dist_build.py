"""Distributed build of ONE global tree on W GPUs over peer memory (csrc/dist_kernels.cu).

Two drivers share the kernels:

* ``IpcDistBuild`` — one process per GPU (torchrun).  ``torch.distributed`` is used only to bootstrap:
  exchanging the CUDA IPC handles of the per-rank symmetric heaps.  The build itself posts and waits on
  flags in peer memory; no collective is called on the data path.
* ``EmulatedDistBuild`` — W *virtual* ranks of one process on one GPU: the same kernels, every phase
  launched for all ranks (posts before waits).  Used to test the distributed logic on a single GPU.

The result is the tree the reference's sequential program builds over all N points
(kdtree_sequential.cpp:30-66); every rank ends up with the complete ``perm`` (and ``tree_points``).
"""
from __future__ import annotations

import ctypes as C

import torch

from .._lib import DistCtx, KdError, check, lib
from ..models.forest import slice_bounds

PH_STATE, PH_HIST, PH_SCAN, PH_CNT_POST, PH_CNT_WAIT, PH_PART, PH_PART_POST, PH_PART_WAIT, PH_MAX_POST, \
    PH_MAX_WAIT, PH_FIX_POST, PH_FIX_WAIT, PH_SAMPLE_POST, PH_PIVOT_WAIT, PH_BRACKET_POST, PH_BRACKET_WAIT = range(16)


def level_schedule() -> list[tuple[int, int]]:
    """(phase, pass) pairs of one distributed level in launch order, as the native library defines them."""
    ph, ps = (C.c_int * 64)(), (C.c_int * 64)()
    n = lib().kd_dist_schedule(ph, ps, 64)
    if n < 0:
        raise KdError("kd_dist_schedule failed")
    return [(int(ph[i]), int(ps[i])) for i in range(n)]


def _levels(world: int) -> int:
    return max(0, world.bit_length() - 1)


class _RawCuda:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can view it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def _offsets(n_total, dim, world, with_tp):
    po, to = C.c_size_t(), C.c_size_t()
    check(lib().kd_dist_offsets(n_total, dim, world, int(with_tp), C.byref(po), C.byref(to)), "kd_dist_offsets")
    return po.value, to.value


class _Base:
    def __init__(self, n_total: int, dim: int, world: int, with_tree_points: bool):
        if world & (world - 1) or not 1 <= world <= 8:
            raise ValueError("world must be 1, 2, 4 or 8")
        self.n_total, self.dim, self.world, self.with_tp = int(n_total), int(dim), int(world), bool(with_tree_points)
        self.heap_bytes = int(lib().kd_dist_heap_bytes(self.n_total, self.dim, self.world, int(self.with_tp)))
        if self.heap_bytes == 0:
            raise KdError("kd_dist_heap_bytes rejected the problem")
        self.perm_off, self.tp_off = _offsets(self.n_total, self.dim, self.world, self.with_tp)
        self.serial = 0

    def _ctx(self, rank: int, heap_ptrs, mc_ptr=None) -> DistCtx:
        c = DistCtx()
        c.mc_heap = mc_ptr
        c.world, c.rank, c.dim, c.with_tree_points, c.n_total = self.world, rank, self.dim, int(self.with_tp), self.n_total
        for r in range(8):
            c.heap[r] = heap_ptrs[r] if r < self.world else None
        return c

    def _views(self, heap_u8: torch.Tensor):
        perm = heap_u8[self.perm_off:self.perm_off + 4 * self.n_total].view(torch.int32)
        tp = None
        if self.with_tp:
            tp = heap_u8[self.tp_off:self.tp_off + 4 * self.n_total * self.dim].view(torch.float32).view(self.n_total,
                                                                                                      self.dim)
        return perm, tp


class EmulatedDistBuild(_Base):
    """W virtual ranks on one GPU (testing aid; the kernels and their order are the real ones)."""

    def __init__(self, n_total, dim, world, device="cuda", with_tree_points=True):
        super().__init__(n_total, dim, world, with_tree_points)
        self.device = torch.device(device)
        self.heaps = [torch.zeros(self.heap_bytes, dtype=torch.uint8, device=self.device) for _ in range(world)]
        ptrs = [h.data_ptr() for h in self.heaps]
        self.ctxs = [self._ctx(r, ptrs) for r in range(world)]

    def build(self, points: torch.Tensor):
        """``points`` = the whole [N, dim] array; rank r takes the reference's slice (kdtree_mpi.cpp:208-216)."""
        L = lib()
        self.serial += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        pts = points.to(self.device, torch.float32).contiguous()
        keep = []
        with torch.cuda.device(self.device):
            for r, ctx in enumerate(self.ctxs):
                first, local = slice_bounds(self.n_total, self.world, r)
                sl = pts[first:first + local].contiguous()
                keep.append(sl)
                check(L.kd_dist_init(C.byref(ctx), sl.data_ptr(), local, first, st), "kd_dist_init")

            def phase(level, ph, p):
                for ctx in self.ctxs:
                    check(L.kd_dist_phase(C.byref(ctx), level, ph, p, self.serial, st), "kd_dist_phase")

            steps = level_schedule()
            for level in range(_levels(self.world)):
                for ph, ps in steps:          # every step for all ranks before the next one (posts before waits)
                    phase(level, ph, ps)
            for ctx in self.ctxs:
                check(L.kd_dist_local_build(C.byref(ctx), st), "kd_dist_local_build")
            for post in (1, 0):
                for ctx in self.ctxs:
                    check(L.kd_dist_done(C.byref(ctx), post, self.serial, st), "kd_dist_done")
            for ctx in self.ctxs:
                check(L.kd_dist_finish(C.byref(ctx), st), "kd_dist_finish")
            for ctx in self.ctxs:
                check(L.kd_dist_status(C.byref(ctx), st), "kd_dist_status")
        return [self._views(h) for h in self.heaps]


class IpcDistBuild(_Base):
    """One process per GPU; heaps are mapped into every peer with CUDA IPC."""

    def __init__(self, n_total, dim, group=None, with_tree_points=True, backend="auto"):
        """``backend``: "symm" = torch symmetric memory (gives an NVLS multicast mapping when the fabric has one),
        "ipc" = cudaMalloc + CUDA IPC handles, "auto" = symm if it works, else ipc."""
        import os

        import torch.distributed as dist
        backend = os.environ.get("KD_DIST_BACKEND", backend)
        self.group = group
        self.rank = dist.get_rank(group)
        world = dist.get_world_size(group)
        super().__init__(n_total, dim, world, with_tree_points)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._own, self._peers, self._symm = None, [], None
        self.multicast = False
        if backend in ("auto", "symm"):
            try:
                self._init_symm(dist, group)
            except Exception:
                if backend == "symm":
                    raise
                self._symm = None
        if self._symm is None:
            self._init_ipc(dist, group)
        self.perm, self.tree_points = self._views(self.heap)
        dist.barrier(group=group)      # every heap is mapped (and zeroed) before anybody posts a flag

    def _init_symm(self, dist, group):
        import torch.distributed._symmetric_memory as symm_mem
        g = group if group is not None else dist.group.WORLD
        heap = symm_mem.empty(self.heap_bytes, dtype=torch.uint8, device=self.device)
        heap.zero_()
        torch.cuda.synchronize(self.device)
        hdl = symm_mem.rendezvous(heap, g.group_name)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = int(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else None
        self._symm, self.heap = hdl, heap
        self.multicast = mc is not None
        self.ctx = self._ctx(self.rank, ptrs, mc)

    def _init_ipc(self, dist, group):
        L = lib()
        world = self.world
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        check(L.kd_symm_alloc(self.heap_bytes, C.byref(ptr), handle), "kd_symm_alloc")
        self._own = ptr.value
        handles = [None] * world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)   # bootstrap only
        ptrs = []
        for r in range(world):
            if r == self.rank:
                ptrs.append(self._own)
                continue
            p = C.c_void_p()
            check(L.kd_symm_open(handles[r], C.byref(p)), "kd_symm_open")
            self._peers.append(p.value)
            ptrs.append(p.value)
        self.ctx = self._ctx(self.rank, ptrs)
        self.heap = torch.as_tensor(_RawCuda(self._own, self.heap_bytes), device=self.device)

    def slice(self):
        return slice_bounds(self.n_total, self.world, self.rank)

    def load(self, local_points: torch.Tensor) -> None:
        first, local = self.slice()
        if local_points.shape != (local, self.dim):
            raise ValueError(f"rank {self.rank} expects a [{local}, {self.dim}] slice")
        self._local = local_points.to(self.device, torch.float32).contiguous()
        st = torch.cuda.current_stream(self.device).cuda_stream
        check(lib().kd_dist_init(C.byref(self.ctx), self._local.data_ptr(), local, first, st), "kd_dist_init")

    def run(self) -> None:
        """Asynchronous on the current stream; ``perm`` / ``tree_points`` are complete on every rank afterwards."""
        self.serial += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        check(lib().kd_dist_run(C.byref(self.ctx), self.serial, st), "kd_dist_run")

    def build(self, local_points: torch.Tensor):
        self.load(local_points)
        self.run()
        return self.perm, self.tree_points

    def status(self) -> None:
        st = torch.cuda.current_stream(self.device).cuda_stream
        check(lib().kd_dist_status(C.byref(self.ctx), st), "kd_dist_status")

    def close(self) -> None:
        import torch.distributed as dist
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        for p in self._peers:
            lib().kd_symm_close(p)
        self._peers = []
        if self._own:
            lib().kd_symm_free(self._own)
            self._own = None
