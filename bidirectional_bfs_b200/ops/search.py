"""SearchContext: Python handle on a device search context (csrc/ctx.cu, csrc/bfs_dev.cu).

One context hosts the per-query state ("arena") of one or more consecutive ranks on the current
device and launches ONE persistent cooperative kernel per query.  Counterpart of the reference's
v4 C API (v4/comp.h:8-36), minus the per-level host calls.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import CtxDesc, LevelTrace, Result, check, lib


@dataclass
class PathResult:
    src: int
    dst: int
    length: int | None           # hops, None when no path
    path: list[int] | None
    meet: int | None = None
    levels: int = 0
    levels_bottom_up: int = 0
    depth_s: int = 0
    depth_t: int = 0
    edges_scanned: int = 0
    vertices_visited: int = 0
    search_ms: float = 0.0       # device-timed for GPU variants, wall for CPU variants
    path_ms: float = 0.0
    variant: str = ""
    levels_grid: int = 0         # levels that needed the whole grid
    levels_paired: int = 0       # levels run by the pair tier (both directions per round)
    status: int = 0

    @property
    def found(self) -> bool:
        return self.length is not None


def result_from_c(src, dst, r: Result, path_buf, variant) -> PathResult:
    found = r.length >= 0
    path = None
    if found and r.path_len:
        path = path_buf[: r.path_len].tolist()        # (a list comprehension over numpy scalars costs 3 us for 15 entries)
    return PathResult(src, dst, int(r.length) if found else None, path,
                      int(r.meet) if found else None, int(r.levels), int(r.levels_bottom_up),
                      int(r.depth_s), int(r.depth_t), int(r.edges_scanned), int(r.vertices_visited),
                      float(r.search_ms), float(r.path_ms), variant, int(r.levels_grid), int(r.levels_paired))


_NVTX = bool(os.environ.get("BIBFS_NVTX"))


class SearchContext:
    def __init__(self, graph, rank_lo=0, vranks=None, max_path=0, solo_threshold=0, bottom_up_div=0,
                 max_ctas=0, watchdog_ms=0, pair_tier=True, max_degree=None):
        if graph.device.type != "cuda":
            raise ValueError("SearchContext needs a CUDA graph")
        self.graph = graph
        self.device = graph.device
        self.nranks = graph.nranks
        self.rank_lo = rank_lo
        self.vranks = len(graph.shards) if vranks is None else vranks
        d = CtxDesc()
        d.n, d.nranks, d.rank_lo, d.vranks = graph.n, graph.nranks, rank_lo, self.vranks
        for i in range(self.vranks):
            s = graph.shards[rank_lo + i]
            d.d_row_ptr[i] = s.row_ptr.data_ptr()
            d.d_col[i] = s.col.data_ptr() if s.col.numel() else s.row_ptr.data_ptr()
            if pair_tier:
                vr = s.vertex_records()
                d.d_vrec[i] = vr.data_ptr() if vr is not None else None
        # longest row of the whole graph: known when every shard is hosted here, else the caller passes it (0 = unknown)
        if max_degree is None:
            hosted_all = self.vranks == graph.nranks and all(graph.shards[r].vrec is not None or graph.shards[r].hi == graph.shards[r].lo for r in graph.shards)
            max_degree = max((graph.shards[r].max_degree for r in graph.shards), default=0) if (pair_tier and hosted_all) else 0
        d.max_degree = int(max_degree)
        d.max_path = max_path or min(graph.n + 1, 1 << 24)
        d.solo_threshold, d.bottom_up_div, d.max_ctas, d.watchdog_ms = solo_threshold, bottom_up_div, max_ctas, watchdog_ms
        self.max_path = d.max_path
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().bibfs_ctx_create(C.byref(self._h), C.byref(d)))
        self._path = np.empty(self.max_path, dtype=np.uint32)
        self._desc = d
        # col / vertex-record pointers of every rank as mapped on this device (hosted shards now, peers when attached)
        self.graph_ptrs = {rank_lo + i: (d.d_col[i], d.d_vrec[i]) for i in range(self.vranks) if d.d_vrec[i]}
        # per-query constants of the hot call (each of these lookups costs 0.3-1 us in CPython)
        self._path_ptr = self._path.ctypes.data
        self._fn_query_result = lib().bibfs_ctx_query_result
        self._res = Result()
        self._res_ref = C.byref(self._res)

    # ---- peers ---------------------------------------------------------------------------
    def ipc_handle(self, slot=0) -> bytes:
        buf = C.create_string_buffer(64)
        check(lib().bibfs_ctx_ipc_handle(self._h, slot, buf))
        return buf.raw

    def attach_ipc(self, rank: int, handle: bytes) -> None:
        with torch.cuda.device(self.device):
            check(lib().bibfs_ctx_attach_ipc(self._h, rank, C.create_string_buffer(handle, 64)))

    def attach_local(self, rank: int, owner: "SearchContext") -> None:
        check(lib().bibfs_ctx_attach_local(self._h, rank, owner._h))

    def graph_ipc(self, slot=0) -> tuple:
        """(handle, offset) pairs of hosted rank `slot`'s col and vertex-record tensors for the other processes of the group
        (None where there is nothing to share): with them rank 0's front kernel reads every shard through peer mappings."""
        s = self.graph.shards[self.rank_lo + slot]
        out = []
        for t in (s.col if s.col.numel() else None, s.vrec):
            if t is None:
                out.append(None)
                continue
            buf, off = C.create_string_buffer(64), C.c_uint64(0)
            with torch.cuda.device(self.device):
                check(lib().bibfs_ipc_export(t.data_ptr(), buf, C.byref(off)))
            out.append((buf.raw, int(off.value)))
        return tuple(out)

    def _import(self, shared: tuple):
        ptrs = []
        for item in shared:
            p = C.c_void_p(0)
            if item is not None:
                check(lib().bibfs_ipc_import(C.create_string_buffer(item[0], 64), item[1], C.byref(p)))
            ptrs.append(p)
        return ptrs

    def attach_graph_ipc(self, rank: int, shared: tuple) -> None:
        """Map rank `rank`'s col / vertex-record tensors (exported by graph_ipc in its process) on this device."""
        with torch.cuda.device(self.device):
            col, vrec = self._import(shared)
            if col.value and vrec.value:
                check(lib().bibfs_ctx_set_peer_graph(self._h, rank, col, vrec))
                self.graph_ptrs[rank] = (col.value, vrec.value)

    def replicate_records(self, shared_by_rank: list) -> bool:
        """Keep a LOCAL, contiguous copy of every shard's vertex records (64 bytes per vertex: the index the front kernel
        walks) instead of reading the other shards' records over NVLink: a round of the front kernel then costs what it
        costs on one GPU, and the copy is streamed into L2 at query start like on one GPU.  The adjacency itself (col)
        stays with its owner; rows longer than six entries are still read from there."""
        B, P = self.graph.partition.B, self.nranks
        with torch.cuda.device(self.device):
            full = torch.zeros(B * P * 16, dtype=torch.int32, device=self.device)
            cols = {}
            for r in range(P):
                lo, hi = self.graph.partition.range(r)
                if r in self.graph.shards:
                    s = self.graph.shards[r]
                    src, cols[r] = (s.vrec.data_ptr() if s.vrec is not None else 0), (s.col.data_ptr() if s.col.numel() else s.row_ptr.data_ptr())
                else:
                    col, vrec = self._import(shared_by_rank[r])
                    src, cols[r] = vrec.value or 0, col.value or 0
                if hi > lo and not src:
                    return False
                if hi > lo:
                    check(lib().bibfs_copy_device(full.data_ptr() + r * B * 64, src, (hi - lo) * 64, self._stream()))
            torch.cuda.synchronize(self.device)
            self._record_replica = full
            for r in range(P):
                check(lib().bibfs_ctx_set_peer_graph(self._h, r, C.c_void_p(cols[r]), C.c_void_p(full.data_ptr() + r * B * 64)))
                self.graph_ptrs[r] = (cols[r], full.data_ptr() + r * B * 64)
        return True

    def resume(self) -> None:
        check(lib().bibfs_ctx_resume(self._h, self._stream()))

    @property
    def arena_bytes(self) -> int:
        return int(lib().bibfs_ctx_arena_bytes(self._h))

    # ---- queries -------------------------------------------------------------------------
    def _stream(self):
        # raw handle of the current stream (the public current_stream() builds a Stream object: ~3 us per query)
        try:
            return torch._C._cuda_getCurrentRawStream(self.device.index or 0)
        except AttributeError:      # pragma: no cover - other torch builds
            return torch.cuda.current_stream(self.device).cuda_stream

    def launch(self, src: int, dst: int, flags: int = 0, pinned_query: torch.Tensor | None = None) -> None:
        q = pinned_query.data_ptr() if pinned_query is not None else None
        check(lib().bibfs_ctx_query(self._h, int(src), int(dst), q, flags, self._stream()))

    def fetch(self, src: int, dst: int, want_path=True, variant="gpu") -> PathResult:
        r = Result()
        check(lib().bibfs_ctx_result(self._h, C.byref(r), self._path.ctypes.data if want_path else None,
                                     self.max_path if want_path else 0, self._stream()))
        out = result_from_c(src, dst, r, self._path, variant)
        out.status = int(r.status)
        return out

    def query(self, src, dst, flags=0, want_path=True, pinned_query=None, variant="gpu") -> PathResult:
        if _NVTX:       # BIBFS_NVTX=1: NVTX range per query for nsys timelines
            with torch.cuda.nvtx.range(f"bibfs.query {src}->{dst}"):
                return self._query(src, dst, flags, want_path, pinned_query, variant)
        return self._query(src, dst, flags, want_path, pinned_query, variant)

    def _query(self, src, dst, flags, want_path, pinned_query, variant) -> PathResult:
        q = pinned_query.data_ptr() if pinned_query is not None else None
        if self._fn_query_result(self._h, int(src), int(dst), q, flags, self._res_ref, self._path_ptr if want_path else None,
                                 self.max_path if want_path else 0, self._stream()):
            check(1)
        return result_from_c(src, dst, self._res, self._path, variant)

    def timing(self) -> tuple[float, float]:
        a, b = C.c_float(0), C.c_float(0)
        check(lib().bibfs_ctx_timing(self._h, C.byref(a), C.byref(b)))
        return float(a.value), float(b.value)

    def trace(self, slot=0) -> list[dict]:
        cnt = C.c_uint32(0)
        check(lib().bibfs_ctx_trace(self._h, slot, None, 0, C.byref(cnt)))
        arr = (LevelTrace * max(1, cnt.value))()
        check(lib().bibfs_ctx_trace(self._h, slot, arr, cnt.value, C.byref(cnt)))
        return [dict(side=int(t.side), mode=int(t.mode), frontier=int(t.frontier), discovered=int(t.discovered),
                     t_ns=int(t.t_ns), edges=int(t.edges), t_expand_ns=int(t.t_expand_ns)) for t in arr[: cnt.value]]

    def reset(self) -> None:
        check(lib().bibfs_ctx_reset(self._h, self._stream()))

    def close(self) -> None:
        if self._h:
            lib().bibfs_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
