"""Partitioning and peer-memory plumbing for the multi-GPU variant."""
from .partition import Partition, block_size  # noqa: F401
