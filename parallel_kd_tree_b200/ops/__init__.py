"""Torch-level operators of the B200 kd-tree library.

Every operator is registered as a ``torch.library`` custom op in the ``kd``
namespace (``torch.ops.kd.build`` ...; the Python functions of this module call the
same implementations directly) and dispatches on the device of its inputs: CUDA tensors run the hand-written sm_100a kernels of
``libkdtree_b200.so``, CPU tensors run the native C++ host core of the same
library (a few trivial CPU helpers — bbox, row gather — use plain torch).  CUDA
inputs never fall back to PyTorch eager code: if the library is missing the op raises.

Reference parity: ``generate_uniform`` = Utility.cpp:6-18 / kdtree_mpi.cpp:19-41,
``build`` = kdtree_sequential.cpp:30-70, ``nearest`` = kdtree_sequential.cpp:75-136,
``materialize_nodes`` = Node.hpp:35-45.
"""
from .generate import generate_clustered, generate_uniform, generate_problem
from .build import build, build_workspace_bytes, build_into, last_launch_count, check_tree, build_cpu_sortref, GraphedBuild
from .query import nearest, knn, gather_rows, materialize_nodes, bbox, query_order

__all__ = [
    "generate_uniform", "generate_clustered", "generate_problem",
    "build", "build_into", "build_workspace_bytes", "GraphedBuild", "last_launch_count", "check_tree", "build_cpu_sortref",
    "nearest", "knn", "gather_rows", "materialize_nodes", "bbox", "query_order",
]
