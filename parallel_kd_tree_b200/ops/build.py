"""kd::build — tree construction (reference kdtree_sequential.cpp:30-70)."""
from __future__ import annotations

import ctypes as C

import torch

from .._lib import KdError, check, lib
from ._util import as_points, check_out, ptr, stream_ptr


def build_workspace_bytes(n: int, dim: int) -> int:
    return int(lib().kd_build_workspace_bytes(int(n), int(dim)))


def last_launch_count() -> int:
    """Kernel launches issued by the last GPU build on this thread."""
    return int(lib().kd_last_launch_count())


def build_into(points: torch.Tensor, perm: torch.Tensor, tree_points: torch.Tensor | None,
               workspace: torch.Tensor | None = None, verify: bool = False) -> None:
    """Builds into caller-provided outputs (no allocation when ``workspace`` is given)."""
    points = as_points(points)
    n, dim = points.shape
    check_out(perm, (n,), torch.int32, points.device, "perm")
    check_out(tree_points, (n, dim), torch.float32, points.device, "tree_points")
    if points.is_cuda:
        need = build_workspace_bytes(n, dim)
        if workspace is None:
            workspace = torch.empty(need, dtype=torch.uint8, device=points.device)
        if workspace.numel() * workspace.element_size() < need or workspace.device != points.device:
            raise KdError(f"kd::build workspace needs {need} bytes on {points.device}")
        with torch.cuda.device(points.device):
            check(lib().kd_build_gpu(points.data_ptr(), n, dim, perm.data_ptr(), ptr(tree_points),
                                     workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                                     stream_ptr(points.device)), "kd::build")
            if verify:
                check(lib().kd_build_status(workspace.data_ptr(), n, dim, stream_ptr(points.device)), "kd::build")
    else:
        check(lib().kd_build_cpu(points.data_ptr(), n, dim, perm.data_ptr(), 0), "kd::build")
        if tree_points is not None:
            torch.index_select(points, 0, perm.long(), out=tree_points)


def _build_impl(points: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    points = as_points(points)
    n, dim = points.shape
    perm = torch.empty(n, dtype=torch.int32, device=points.device)
    tree_points = torch.empty((n, dim), dtype=torch.float32, device=points.device)
    build_into(points, perm, tree_points, verify=True)
    return perm, tree_points


# registered as torch.ops.kd.build; the Python entry point below calls the implementation directly (the first
# trip through the dispatcher of a custom op costs seconds of lazy imports, which a CLI run would pay every time)
_build_op = torch.library.custom_op("kd::build", mutates_args=())(_build_impl)


@_build_op.register_fake
def _(points):
    n, dim = points.shape
    return points.new_empty(n, dtype=torch.int32), points.new_empty((n, dim))


def build(points: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns ``(perm, tree_points)``.

    ``perm`` (int32[n]) is the tree in in-order layout: the node of range ``[lo, lo+n)`` is
    ``perm[lo + n//2]``, its sub-trees are ``[lo, lo+n//2)`` and ``[lo+n//2+1, lo+n)``
    (kdtree_sequential.cpp:51-56).  ``tree_points[s] = points[perm[s]]``.
    """
    return _build_impl(points)


def build_cpu_sortref(points: torch.Tensor) -> torch.Tensor:
    """The reference's own formulation (sort the first n-1, take n/2) under the (coord, index) order."""
    points = as_points(points).cpu()
    n, dim = points.shape
    perm = torch.empty(n, dtype=torch.int32)
    check(lib().kd_build_cpu_sortref(points.data_ptr(), n, dim, perm.data_ptr()), "build_cpu_sortref")
    return perm


def check_tree(points: torch.Tensor, perm: torch.Tensor, allow_ties: bool = False) -> int:
    """Number of violations of the reference's construction rule in ``perm``.

    Strict (default): 0 = the unique tree under the canonical tie-break (coordinate, index).
    ``allow_ties``: equal coordinates may fall on either side; 0 = a possible output of the reference's
    unstable ``std::sort`` (used to validate the unmodified reference's trees at scale)."""
    points = as_points(points).cpu()
    perm = perm.to("cpu", torch.int32).contiguous()
    bad = C.c_int64(0)
    fn = lib().kd_check_tree_ties_cpu if allow_ties else lib().kd_check_tree_cpu
    check(fn(points.data_ptr(), points.shape[0], points.shape[1], perm.data_ptr(), C.addressof(bad)), "check_tree")
    return int(bad.value)


class GraphedBuild:
    """The launch sequence of one GPU build captured once in a CUDA graph and replayed per call.

    A build is ~60 (small n) to ~260 (2^28 points) dependent kernel launches with fixed parameters for fixed buffers,
    so it captures cleanly; a replay costs one launch on the host (0.03-0.06 ms instead of 0.5-1.7 ms of launch calls)
    and removes part of the gaps between the small kernels on the device (measured on a B200: 8.93 -> 8.55 ms at 2^25
    points, 68.5 -> 67.8 ms at 2^28, `tools/host_launch_time.py`).  Inputs and outputs are the object's own tensors:
    copy new points into ``.points`` and read ``.perm`` / ``.tree_points`` after ``run()``.
    """

    def __init__(self, n: int, dim: int, device="cuda", with_tree_points: bool = True):
        self.device = torch.device(device)
        self.points = torch.empty((n, dim), dtype=torch.float32, device=self.device)
        self.perm = torch.empty(n, dtype=torch.int32, device=self.device)
        self.tree_points = torch.empty((n, dim), dtype=torch.float32, device=self.device) if with_tree_points else None
        self.workspace = torch.empty(build_workspace_bytes(n, dim), dtype=torch.uint8, device=self.device)
        self._graph = None

    def _capture(self) -> None:
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            build_into(self.points, self.perm, self.tree_points, self.workspace)      # warm-up outside the capture
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                build_into(self.points, self.perm, self.tree_points, self.workspace)
        torch.cuda.current_stream(self.device).wait_stream(side)
        self._graph = graph

    def run(self, verify: bool = False) -> None:
        """Replays the build on the current stream (captures it on first use)."""
        if self._graph is None:
            self._capture()
        self._graph.replay()
        if verify:
            n, dim = self.points.shape
            with torch.cuda.device(self.device):
                check(lib().kd_build_status(self.workspace.data_ptr(), n, dim, stream_ptr(self.device)), "kd::build")
