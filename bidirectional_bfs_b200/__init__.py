"""bidirectional_bfs_b200 — Blackwell-native bidirectional-BFS shortest-path framework.

Capabilities of Tomaszbrauntsch/Bidirectional-BFS (serial, multi-rank CPU, single-GPU and
multi-GPU s-t shortest path; graph generator + path oracle; benchmark table; visualiser) rebuilt
around hand-written sm_100a kernels behind a C ABI (include/bibfs_b200.h).
"""
from ._lib import BibfsError, GenSpec, build_native, lib, native_path  # noqa: F401
from .graph import Graph, Shard  # noqa: F401
from .ops.gen import er_spec, grid_spec, preset_spec, rmat_spec, generate_edges  # noqa: F401
from .ops.search import PathResult, SearchContext  # noqa: F401
from .models import (BatchSolver, CpuRanksSolver, GpuSolver, MultiGpuSolver, NaiveGpuSolver,  # noqa: F401
                     ResidentSolver, SerialSolver, make_solver)
from .api import bibfs, bibfs_batch  # noqa: F401
from .utils.validate import validate_path  # noqa: F401
from . import io  # noqa: F401

__version__ = "0.1.0"
