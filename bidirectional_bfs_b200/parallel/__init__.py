"""Partitioning and peer-memory plumbing for the multi-GPU variant."""
from .partition import Partition, block_size  # noqa: F401


def __getattr__(name):   # lazy: comm imports graph, which imports this package
    if name in ("DistributedSolver", "init_distributed"):
        from . import comm
        return getattr(comm, name)
    raise AttributeError(name)
