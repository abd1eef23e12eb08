"""1-D block vertex partition (owner computes).

The reference replicates the whole graph and all O(N) state on every rank and only strides the
*work* (v2/second_try.cpp:71 ``u = rank; u += size``; v4/comp.cu:27 ``(u % size) != rank``, which
is in fact disabled by the ``(N, 0, 1)`` launch, comp.cu:99).  Here the graph itself is
partitioned: rank r owns the contiguous id range ``[r*B, min(n, (r+1)*B))`` with B a multiple of
64 so that bitmap words never straddle two owners.
"""
from __future__ import annotations

from dataclasses import dataclass

from .._lib import MAX_RANKS, lib


def block_size(n: int, nranks: int) -> int:
    return int(lib().bibfs_block_size(n, nranks))


@dataclass(frozen=True)
class Partition:
    n: int
    nranks: int

    def __post_init__(self):
        if not (1 <= self.nranks <= MAX_RANKS):
            raise ValueError(f"nranks must be 1..{MAX_RANKS}")
        if self.n <= 0:
            raise ValueError("n must be positive")

    @property
    def B(self) -> int:
        return block_size(self.n, self.nranks)

    def owner(self, v: int) -> int:
        return v // self.B

    def range(self, rank: int) -> tuple[int, int]:
        lo = min(self.n, rank * self.B)
        hi = min(self.n, (rank + 1) * self.B)
        return lo, hi
