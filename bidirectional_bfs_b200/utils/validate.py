"""Path validity (only the LENGTH is comparable across variants; a path is checked by
endpoints, every hop being an edge, and hop count == reported length)."""
from __future__ import annotations

import numpy as np

from .._lib import BibfsError, check, lib


def validate_path(graph, src: int, dst: int, path, length: int | None = None) -> None:
    """Raise ValueError if `path` is not a valid src->dst path of `length` hops in `graph` (CPU)."""
    g = graph if (graph.device.type == "cpu" and graph.nranks == 1) else None
    if g is None:
        from ..models.solvers import _as_single
        g = _as_single(graph, "cpu")
    s = g.single()
    p = np.ascontiguousarray(np.asarray(path, dtype=np.uint32))
    if length is not None and len(p) != length + 1:
        raise ValueError(f"path has {len(p) - 1} hops, expected {length}")
    try:
        check(lib().bibfs_validate_path(g.n, s.row_ptr.data_ptr(), s.col.data_ptr(), int(src), int(dst), p.ctypes.data, len(p)))
    except BibfsError as e:
        raise ValueError(str(e)) from None
