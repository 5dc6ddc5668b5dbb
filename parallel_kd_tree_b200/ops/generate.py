"""Problem generation (host side, outside every timed region — PDF p.5 task 2 exempts it)."""
from __future__ import annotations

import torch

from .._lib import check, lib


def generate_uniform(seed: int, dim: int, n_rows: int, first_row: int = 0, pin_memory: bool = False) -> torch.Tensor:
    """Rows ``[first_row, first_row + n_rows)`` of the reference's random stream.

    Bit-exact with ``Utility::generate_problem`` (Utility.cpp:6-18): ``std::mt19937(seed)`` through
    ``std::uniform_real_distribution<float>(-100, 100)``, point-major / dimension-minor draw order.
    ``first_row > 0`` is the slice mode of kdtree_mpi.cpp:19-41 (``discard`` of the lower ranks' draws).
    """
    out = torch.empty((n_rows, dim), dtype=torch.float32, pin_memory=pin_memory)
    check(lib().kd_generate_uniform_rows(int(seed), int(dim), int(first_row), int(n_rows), out.data_ptr()),
          "generate_uniform")
    return out


def generate_clustered(seed: int, dim: int, n_rows: int, n_clusters: int = 16, sigma: float = 4.0,
                       pin_memory: bool = False) -> torch.Tensor:
    """Seeded Gaussian blobs clamped to [-100, 100] (BASELINE.json config 4; no counterpart in the reference)."""
    out = torch.empty((n_rows, dim), dtype=torch.float32, pin_memory=pin_memory)
    check(lib().kd_generate_clustered(int(seed), int(dim), int(n_rows), int(n_clusters), float(sigma),
                                      out.data_ptr()), "generate_clustered")
    return out


def generate_problem(seed: int, dim: int, num_points: int, num_queries: int = 10, first_row: int = 0,
                     local_points: int | None = None):
    """(points, queries) exactly as the reference drivers see them.

    Sequential driver (kdtree_sequential.cpp:157-170): rows 0..N-1 are points, rows N..N+Q-1 the queries.
    With ``first_row``/``local_points`` it returns one rank's slice and the (shared) queries
    like kdtree_mpi.cpp:19-41.
    """
    if local_points is None:
        local_points = num_points - first_row
    pts = generate_uniform(seed, dim, local_points, first_row)
    qs = generate_uniform(seed, dim, num_queries, num_points)
    return pts, qs
