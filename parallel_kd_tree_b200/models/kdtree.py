"""KDTree — the reference's Point/Node/tree (Node.hpp:9-45) as tensors on one device."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from .. import ops
from ..utils import trace
from .._lib import check, lib, take_string


@dataclass
class NNResult:
    """Answer of a batch of 1-NN queries.

    ``index`` is the row of the neighbour in the tree's ``points``; ``ids`` is the reference's
    1-based point ID (kdtree_sequential.cpp:161, kdtree_mpi.cpp:223); ``dist`` is what the
    reference prints (sqrt of the float sum, Node.cpp:36-38); ``visited`` counts visited nodes.
    """
    index: torch.Tensor
    ids: torch.Tensor
    dist2: torch.Tensor
    dist: torch.Tensor
    visited: torch.Tensor


class KDTree:
    """k-d tree over ``points`` ([n, dim] float32, CPU or CUDA) built with the reference's rule.

    The tree is stored as ``perm`` (in-order layout, see ``ops.build``) plus ``tree_points``
    (= ``points[perm]``); ``Node{point,left,right}`` arrays are produced on demand by ``nodes()``.
    ``id_offset`` shifts the printed IDs (a rank of the forest owns IDs first_row+1 ...).
    """

    def __init__(self, points: torch.Tensor, id_offset: int = 0, *, _prebuilt=None):
        if points.dim() != 2 or points.shape[0] < 1 or points.shape[1] < 1:
            raise ValueError("points must be a non-empty [n, dim] tensor")
        self.points = points.contiguous()
        self.n, self.dim = self.points.shape
        self.id_offset = int(id_offset)
        if _prebuilt is None:
            with trace.span(f"build n={self.n} dim={self.dim}", self.points.device):
                self.perm, self.tree_points = ops.build(self.points)
        else:
            self.perm, self.tree_points = _prebuilt

    # -- construction helpers -------------------------------------------------------------
    @classmethod
    def from_seed(cls, seed: int, dim: int, num_points: int, device: str | torch.device = "cpu",
                  num_queries: int = 10):
        """The reference driver's problem (kdtree_sequential.cpp:152-165): returns (tree, queries)."""
        pts, qs = ops.generate_problem(seed, dim, num_points, num_queries)
        return cls(pts.to(device)), qs.to(device)

    @property
    def device(self) -> torch.device:
        return self.points.device

    @property
    def root_slot(self) -> int:
        return self.n // 2

    # -- queries ---------------------------------------------------------------------------
    def nearest(self, queries: torch.Tensor, mode: str = "auto") -> NNResult:
        """Reference 1-NN (kdtree_sequential.cpp:75-136) for a batch of queries (``mode``: see ``ops.nearest``)."""
        queries = queries.to(self.device, torch.float32)
        if queries.dim() == 1:
            queries = queries[None]
        with trace.span(f"nearest nq={queries.shape[0]}", self.device):
            idx, d2, d, vis = ops.nearest(self.tree_points, self.perm, queries, mode)
        return NNResult(index=idx, ids=idx + (1 + self.id_offset), dist2=d2, dist=d, visited=vis)

    def knn(self, queries: torch.Tensor, k: int):
        """k nearest nodes per query by the reference traversal with a k-best list: (ids [nq, k], dist [nq, k])."""
        queries = queries.to(self.device, torch.float32)
        if queries.dim() == 1:
            queries = queries[None]
        idx, d2 = ops.knn(self.tree_points, self.perm, queries, k)
        ids = torch.where(idx >= 0, idx + (1 + self.id_offset), idx)
        return ids, d2.sqrt()

    # -- structure -------------------------------------------------------------------------
    def nodes(self):
        """(node_point, left, right) int32 arrays indexed by slot, -1 = nullptr (Node.hpp:35-45)."""
        return ops.materialize_nodes(self.perm)

    def ids_inorder(self) -> torch.Tensor:
        return self.perm + (1 + self.id_offset)

    def check(self) -> int:
        """Violations of the construction rule (0 for a correct tree)."""
        return ops.check_tree(self.points, self.perm)

    def bbox(self):
        return ops.bbox(self.points)

    # -- text (Utility.cpp:21-63, Node.cpp:16-28) ---------------------------------------------
    def _host(self):
        return self.points.cpu().contiguous(), self.perm.cpu().contiguous()

    def format_tree(self) -> str:
        """Exactly the text ``Utility::print_tree`` writes for this tree."""
        pts, perm = self._host()
        out, ln = C.c_void_p(), C.c_size_t()
        check(lib().kd_format_tree(pts.data_ptr(), self.n, self.dim, perm.data_ptr(), self.id_offset,
                                   C.byref(out), C.byref(ln)), "format_tree")
        return take_string(out, ln)

    def format_head_and_leaves(self) -> str:
        """Exactly the text ``Utility::print_head_and_leaves`` writes for this tree."""
        pts, perm = self._host()
        out, ln = C.c_void_p(), C.c_size_t()
        check(lib().kd_format_head_and_leaves(pts.data_ptr(), self.n, self.dim, perm.data_ptr(), self.id_offset,
                                              C.byref(out), C.byref(ln)), "format_head_and_leaves")
        return take_string(out, ln)

    def state_dict(self) -> dict:
        """Serialisable form (the reference never serialises its tree; SURVEY.md §5 checkpoint row)."""
        return {"points": self.points.cpu(), "perm": self.perm.cpu(), "id_offset": self.id_offset}

    def save(self, path) -> None:
        """Writes ``state_dict()`` with ``torch.save`` (checkpoint; the reference keeps its tree in memory only)."""
        torch.save(self.state_dict(), path)

    @classmethod
    def load(cls, path, device: str | torch.device = "cpu") -> "KDTree":
        return cls.from_state_dict(torch.load(path, map_location="cpu"), device)

    @classmethod
    def from_state_dict(cls, state: dict, device: str | torch.device = "cpu") -> "KDTree":
        pts = state["points"].to(device)
        perm = state["perm"].to(device)
        return cls(pts, state.get("id_offset", 0), _prebuilt=(perm, ops.gather_rows(pts, perm)))
