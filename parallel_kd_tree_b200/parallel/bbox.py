"""Bounding box of a point set spread over ranks (distributed form).

Local per-axis min/max with the library's kernel (``kd::bbox``), then one ``all_reduce(MIN)`` and one
``all_reduce(MAX)`` of ``dim`` floats.  This is set-up work (2 * dim floats), not a hot path, so it goes through
the collective library; CPU tensors work with gloo.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .. import ops


def distributed_bbox(local_points: torch.Tensor, group=None) -> tuple[torch.Tensor, torch.Tensor]:
    """(min, max) per axis over the union of every rank's ``local_points``; identical on all ranks."""
    mn, mx = ops.bbox(local_points)
    mn, mx = mn.clone(), mx.clone()
    dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    return mn, mx
