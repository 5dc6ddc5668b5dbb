"""Tree objects: the single-device KDTree and the multi-rank Forest of local trees."""
from .forest import Forest, slice_bounds
from .kdtree import KDTree, NNResult

__all__ = ["KDTree", "NNResult", "Forest", "slice_bounds"]
