"""Public API: ``bibfs(graph, src, dst)`` and ``bibfs_batch(graph, srcs, dsts)``.

The reference has no API — four ``main()`` programs that answer one query each
(``<graph.bin> <src> <dst>``).  Here one call answers a query with any variant, and a
solver object can be kept to answer many queries on the same graph without rebuilding state.
"""
from __future__ import annotations

from .graph import Graph
from .models import BatchSolver, make_solver
from .ops.search import PathResult

_default_variant = None


def default_variant() -> str:
    global _default_variant
    if _default_variant is None:
        import torch
        _default_variant = "gpu" if torch.cuda.is_available() else "serial"
    return _default_variant


def bibfs(graph: Graph, src: int, dst: int, variant: str | None = None, want_path: bool = True, **kw) -> PathResult:
    """Shortest s-t path by bidirectional BFS.  Returns a PathResult (length None = no path)."""
    if variant is None:
        if graph.device.type == "cuda":
            variant = "gpu" if graph.nranks == 1 else "multi-gpu"
        else:
            variant = "serial"
    solver = make_solver(graph, variant, **kw)
    try:
        return solver.query(src, dst, want_path=want_path)
    finally:
        solver.close()


def bibfs_batch(graph: Graph, srcs, dsts, device=None, engine: str = "concurrent", lanes: int = 32) -> list[int | None]:
    """Hop counts for many (src, dst) pairs on the GPU: concurrent single-CTA searches (default) or
    the bit-parallel 64-queries-per-sweep engine."""
    solver = BatchSolver(graph, device=device, engine=engine, lanes=lanes)
    try:
        return solver.query_batch(srcs, dsts)
    finally:
        solver.close()
