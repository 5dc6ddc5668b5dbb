"""kd::nearest and friends (reference kdtree_sequential.cpp:75-136, Node.hpp:35-45)."""
from __future__ import annotations

import torch

from .._lib import check, lib
import os

from ._util import as_points, on_device, stream_ptr


# From this dimension on the GPU path first tries the node-parallel kernels (the reference's pruning does not fire
# in high dimension, so a thread per query would walk all n nodes one after the other); "dfs" / "dense" force a path.
DENSE_MIN_DIM = int(os.environ.get("KD_B200_DENSE_MIN_DIM", "16"))


# Batches of at least this many queries are grouped by the sub-tree they enter first before the DFS kernels run
ORDER_MIN_QUERIES = int(os.environ.get("KD_B200_ORDER_MIN_QUERIES", "4096"))


def query_order(tree_points: torch.Tensor, queries: torch.Tensor, levels: int | None = None) -> torch.Tensor:
    """int32[nq] permutation that groups the queries by the sub-tree their search descends into first."""
    n, dim = tree_points.shape
    nq = queries.shape[0]
    dev = tree_points.device
    if levels is None:   # about 32 queries per group, never deeper than the tree
        levels = max(1, min(16, (max(nq, 2) - 1).bit_length() - 5, max(1, n.bit_length() - 2)))
    order = torch.empty(nq, dtype=torch.int32, device=dev)
    key = torch.empty(nq, dtype=torch.int32, device=dev)
    bins = torch.empty(1 << levels, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib().kd_query_order_gpu(tree_points.data_ptr(), n, dim, queries.data_ptr(), nq, int(levels),
                                       order.data_ptr(), key.data_ptr(), bins.data_ptr(), stream_ptr(dev)),
              "kd::query_order")
    return order


def _nearest_op_impl(tree_points: torch.Tensor, perm: torch.Tensor, queries: torch.Tensor,
                     mode: str = "auto") -> tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    tree_points = as_points(tree_points, "tree_points")
    dev = tree_points.device
    queries = on_device(as_points(queries, "queries"), dev, torch.float32, "queries")
    n, dim = tree_points.shape
    if queries.shape[1] != dim:
        # the reference exits with "Dimensions do not match!" (kdtree_sequential.cpp:15-18)
        raise ValueError("Dimensions do not match!")
    if perm.numel() != n:
        raise ValueError(f"perm has {perm.numel()} entries for a tree of {n} points")
    nq = queries.shape[0]
    perm = on_device(perm, dev, torch.int32, "perm")
    idx = torch.empty(nq, dtype=torch.int32, device=dev)
    d2 = torch.empty(nq, dtype=torch.float32, device=dev)
    d = torch.empty(nq, dtype=torch.float32, device=dev)
    if tree_points.is_cuda:
        vis = torch.empty(nq, dtype=torch.int32, device=dev)
        dense = mode == "dense" or (mode == "auto" and dim >= DENSE_MIN_DIM)
        with torch.cuda.device(dev):
            if dense:
                nbytes = int(lib().kd_nearest_dense_workspace_bytes(n, nq))
                ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
                check(lib().kd_nearest_dense_gpu(tree_points.data_ptr(), perm.data_ptr(), n, dim, queries.data_ptr(),
                                                 nq, idx.data_ptr(), d2.data_ptr(), d.data_ptr(), vis.data_ptr(),
                                                 ws.data_ptr(), nbytes, stream_ptr(dev)), "kd::nearest")
            else:
                order = query_order(tree_points, queries) if nq >= ORDER_MIN_QUERIES else None
                check(lib().kd_nearest_ordered_gpu(tree_points.data_ptr(), perm.data_ptr(), n, dim,
                                                   queries.data_ptr(), nq, None if order is None else order.data_ptr(),
                                                   idx.data_ptr(), d2.data_ptr(), d.data_ptr(), vis.data_ptr(),
                                                   stream_ptr(dev)), "kd::nearest")
        vis = vis.long()
    else:
        # the CPU core indexes the original rows through perm; tree_points[s] == points[perm[s]]
        vis = torch.empty(nq, dtype=torch.int64)
        ident = torch.arange(n, dtype=torch.int32)
        slot = torch.empty(nq, dtype=torch.int32)
        check(lib().kd_nearest_cpu(tree_points.data_ptr(), n, dim, ident.data_ptr(), queries.data_ptr(), nq,
                                   slot.data_ptr(), d2.data_ptr(), d.data_ptr(), vis.data_ptr()), "kd::nearest")
        idx = perm[slot.long()]
    return idx, d2, d, vis


def _nearest_op_auto(tree_points: torch.Tensor, perm: torch.Tensor,
                     queries: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    return _nearest_op_impl(tree_points, perm, queries, "auto")


_nearest_op = torch.library.custom_op("kd::nearest", mutates_args=())(_nearest_op_auto)


@_nearest_op.register_fake
def _(tree_points, perm, queries):
    nq = queries.shape[0]
    return (queries.new_empty(nq, dtype=torch.int32), queries.new_empty(nq), queries.new_empty(nq),
            queries.new_empty(nq, dtype=torch.int64))


def nearest(tree_points: torch.Tensor, perm: torch.Tensor, queries: torch.Tensor, mode: str = "auto"):
    """1-NN of every query with the reference's answers; returns (point index, dist², dist, nodes visited).

    ``mode`` (GPU only): "dfs" = one thread per query walks the reference's traversal; "dense" = node-parallel
    distances + a per-query proof that the reference prunes nothing, DFS for the rest (bit-identical results);
    "auto" = dense from ``DENSE_MIN_DIM`` dimensions on."""
    if mode not in ("auto", "dfs", "dense"):
        raise ValueError("mode must be auto, dfs or dense")
    return _nearest_op_impl(tree_points, perm, queries, mode)


def gather_rows(points: torch.Tensor, perm: torch.Tensor) -> torch.Tensor:
    points = as_points(points)
    if not points.is_cuda:
        return points.index_select(0, perm.to(points.device).long())
    perm = on_device(perm, points.device, torch.int32, "perm")
    if perm.numel() != points.shape[0]:
        raise ValueError("perm and points disagree in length")
    out = torch.empty_like(points)
    with torch.cuda.device(points.device):
        check(lib().kd_gather_rows_gpu(points.data_ptr(), perm.data_ptr(), points.shape[0], points.shape[1],
                                       out.data_ptr(), stream_ptr(points.device)), "kd::gather_rows")
    return out


def _materialize_op_impl(perm: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    perm = perm.to(torch.int32).contiguous()
    n = perm.numel()
    node_point, left, right = (torch.empty_like(perm) for _ in range(3))
    if perm.is_cuda:
        with torch.cuda.device(perm.device):
            check(lib().kd_materialize_nodes_gpu(n, perm.data_ptr(), node_point.data_ptr(), left.data_ptr(),
                                                 right.data_ptr(), stream_ptr(perm.device)), "kd::materialize_nodes")
    else:
        check(lib().kd_materialize_nodes_cpu(n, perm.data_ptr(), node_point.data_ptr(), left.data_ptr(),
                                             right.data_ptr()), "kd::materialize_nodes")
    return node_point, left, right


_materialize_op = torch.library.custom_op("kd::materialize_nodes", mutates_args=())(_materialize_op_impl)


@_materialize_op.register_fake
def _(perm):
    return torch.empty_like(perm), torch.empty_like(perm), torch.empty_like(perm)


def materialize_nodes(perm: torch.Tensor):
    """``Node{point,left,right}`` (Node.hpp:35-45) as int32 arrays indexed by slot; -1 = nullptr; root = n//2."""
    return _materialize_op_impl(perm)


def _bbox_op_impl(points: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    points = as_points(points)
    if not points.is_cuda:
        return points.min(0).values, points.max(0).values
    dim = points.shape[1]
    mn = torch.empty(dim, dtype=torch.float32, device=points.device)
    mx = torch.empty(dim, dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        check(lib().kd_bbox_gpu(points.data_ptr(), points.shape[0], dim, mn.data_ptr(), mx.data_ptr(),
                                stream_ptr(points.device)), "kd::bbox")
    return mn, mx


_bbox_op = torch.library.custom_op("kd::bbox", mutates_args=())(_bbox_op_impl)


@_bbox_op.register_fake
def _(points):
    return points.new_empty(points.shape[1]), points.new_empty(points.shape[1])


def bbox(points: torch.Tensor):
    """Per-axis (min, max) of a point set."""
    return _bbox_op_impl(points)


def knn(tree_points: torch.Tensor, perm: torch.Tensor, queries: torch.Tensor, k: int):
    """Batched k-NN (extension: the reference has k = 1): (point index [nq, k], dist² [nq, k]), ascending.

    Same traversal as ``nearest`` with a bounded list of the k best nodes; ``k = 1`` gives the reference's answers.
    """
    tree_points = as_points(tree_points, "tree_points")
    dev = tree_points.device
    queries = on_device(as_points(queries, "queries"), dev, torch.float32, "queries")
    n, dim = tree_points.shape
    if queries.shape[1] != dim:
        raise ValueError("Dimensions do not match!")
    if perm.numel() != n:
        raise ValueError(f"perm has {perm.numel()} entries for a tree of {n} points")
    nq = queries.shape[0]
    perm = on_device(perm, dev, torch.int32, "perm")
    idx = torch.empty((nq, k), dtype=torch.int32, device=dev)
    d2 = torch.empty((nq, k), dtype=torch.float32, device=dev)
    if tree_points.is_cuda:
        order = query_order(tree_points, queries) if nq >= ORDER_MIN_QUERIES else None
        with torch.cuda.device(dev):
            check(lib().kd_knn_ordered_gpu(tree_points.data_ptr(), perm.data_ptr(), n, dim, queries.data_ptr(), nq,
                                           int(k), None if order is None else order.data_ptr(), idx.data_ptr(),
                                           d2.data_ptr(), stream_ptr(dev)), "kd::knn")
    else:
        ident = torch.arange(n, dtype=torch.int32)
        slot = torch.empty((nq, k), dtype=torch.int32)
        check(lib().kd_knn_cpu(tree_points.data_ptr(), n, dim, ident.data_ptr(), queries.data_ptr(), nq, int(k),
                               slot.data_ptr(), d2.data_ptr()), "kd::knn")
        idx = torch.where(slot >= 0, perm[slot.clamp(min=0).long()], slot)
    return idx, d2
