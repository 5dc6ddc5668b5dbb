"""parallel_kd_tree_b200 — a B200-native kd-tree library with the capabilities of Dan-Yeh/Parallel-Kd-Tree.

Layers: ``csrc`` native host core + sm_100a kernels behind a C ABI
(``include/kdtree_b200.h``), ``ops`` torch custom ops, ``models`` KDTree / Forest,
``parallel`` multi-GPU runtime, ``utils`` the reference's text protocol, ``cli`` the drivers.
"""
from . import ops, utils
from ._lib import KdError, lib
from .models import KDTree, NNResult

__version__ = "0.1.0"
__all__ = ["ops", "utils", "KDTree", "NNResult", "KdError", "lib"]
