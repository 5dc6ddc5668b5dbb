"""Forest of local trees — what the reference's MPI program computes (kdtree_mpi.cpp:204-253).

Rank r owns the contiguous slice ``[r * (N // W), ...)`` (the last rank also takes ``N % W``),
builds an independent local tree over it, answers every query on that tree, and the per-query
distances are combined with a minimum.  This class is the single-process form (W shards on one
or several devices); ``parallel.forest`` is the one-process-per-GPU form with a peer-memory reduce.
"""
from __future__ import annotations

import torch

from .. import ops
from .kdtree import KDTree


def slice_bounds(num_points: int, world: int, rank: int) -> tuple[int, int]:
    """(first_row, local_num_points) of kdtree_mpi.cpp:208-216."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("need 0 <= rank < world")
    local = num_points // world
    first = local * rank
    if rank == world - 1:
        local += num_points % world
    if local < 1:
        # the reference dereferences a null root here (SURVEY.md §2.2 "empty rank"); we refuse instead
        raise ValueError(f"world size {world} leaves rank {rank} without points (need world <= num_points)")
    return first, local


class Forest:
    def __init__(self, trees: list[KDTree], num_points: int):
        self.trees = trees
        self.num_points = num_points
        self.world = len(trees)

    @classmethod
    def from_seed(cls, seed: int, dim: int, num_points: int, world: int, devices=None, num_queries: int = 10):
        """Every shard regenerates only its slice of the stream, like kdtree_mpi.cpp:19-41."""
        devices = devices or ["cpu"]
        trees = []
        for r in range(world):
            first, local = slice_bounds(num_points, world, r)
            pts = ops.generate_uniform(seed, dim, local, first)
            trees.append(KDTree(pts.to(devices[r % len(devices)]), id_offset=first))
        queries = ops.generate_uniform(seed, dim, num_queries, num_points)
        return cls(trees, num_points), queries

    @classmethod
    def from_points(cls, points: torch.Tensor, world: int, devices=None):
        n = points.shape[0]
        devices = devices or [points.device]
        trees = []
        for r in range(world):
            first, local = slice_bounds(n, world, r)
            trees.append(KDTree(points[first:first + local].to(devices[r % len(devices)]), id_offset=first))
        return cls(trees, n)

    def nearest(self, queries: torch.Tensor):
        """(dist, ids): per-query minimum over the local trees' answers (MPI_Reduce(MIN), kdtree_mpi.cpp:252-253).

        ``ids`` is the ID of a local answer that attains the minimum (lowest rank on ties); the reference
        only reduces distances.
        """
        best_d, best_id = None, None
        for t in self.trees:
            r = t.nearest(queries)
            d, ids = r.dist.cpu(), r.ids.cpu()
            if best_d is None:
                best_d, best_id = d.clone(), ids.clone()
            else:
                better = d < best_d
                best_d = torch.where(better, d, best_d)
                best_id = torch.where(better, ids, best_id)
        return best_d, best_id
