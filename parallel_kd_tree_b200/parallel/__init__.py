"""Multi-GPU layer: forest of local trees, symmetric-memory runtime, distributed global build."""
