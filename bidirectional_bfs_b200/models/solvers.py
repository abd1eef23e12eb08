"""Solver classes behind ``bibfs(graph, src, dst, variant=...)``."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .._lib import F_DEVICE_CLOCK, F_NO_PATH, STATUS_HANDED_OVER, Result, check, lib
from ..graph import Graph
from ..ops.search import PathResult, SearchContext, result_from_c

VARIANTS = ("serial", "cpu-ranks", "gpu", "multi-gpu", "naive-gpu")


class _CpuSolver:
    variant = ""

    def __init__(self, graph: Graph):
        if graph.device.type != "cpu" or graph.nranks != 1:
            graph = _as_single(graph, "cpu")
        self.graph = graph
        self._path = np.empty(graph.n + 1, dtype=np.uint32)

    def close(self):
        pass


def _as_single(graph: Graph, device) -> Graph:
    """Re-assemble / move a graph into a single shard on `device` (host-side convenience)."""
    dev = torch.device(device)
    if graph.nranks == 1:
        return graph.to(dev)
    if len(graph.shards) != graph.nranks:
        raise ValueError("cannot re-assemble a graph whose shards live in other processes")
    rps, cols, base = [], [], 0
    for r in range(graph.nranks):
        s = graph.shards[r]
        rp = s.row_ptr.cpu()
        rps.append(rp[:-1] + base)
        base += int(rp[-1])
        cols.append(s.col.cpu())
    rps.append(torch.tensor([base], dtype=torch.int64))
    g = Graph(graph.n, 1, dev, {}, graph.m_records, graph.spec, graph.name)
    from ..graph import Shard, padded_col
    g.shards[0] = Shard(0, 0, graph.n, torch.cat(rps).to(dev), padded_col(torch.cat(cols), dev))
    return g


class SerialSolver(_CpuSolver):
    """v1 equivalent.  policy 0 replays v1's schedule (early break inside the level)."""
    variant = "serial"

    def __init__(self, graph: Graph, policy: int = 0):
        super().__init__(graph)
        self.policy = policy

    def query(self, src, dst, want_path=True) -> PathResult:
        s = self.graph.single()
        r = Result()
        check(lib().bibfs_serial(self.graph.n, s.row_ptr.data_ptr(), s.col.data_ptr(), int(src), int(dst), self.policy,
                                 C.byref(r), self._path.ctypes.data if want_path else None, self._path.size if want_path else 0))
        return result_from_c(src, dst, r, self._path, self.variant)


class CpuRanksSolver(_CpuSolver):
    """v2 equivalent with `ranks` worker threads standing in for MPI ranks."""
    variant = "cpu-ranks"

    def __init__(self, graph: Graph, ranks: int = 4):
        super().__init__(graph)
        self.ranks = ranks

    def query(self, src, dst, want_path=True) -> PathResult:
        s = self.graph.single()
        r = Result()
        check(lib().bibfs_cpu_ranks(self.graph.n, s.row_ptr.data_ptr(), s.col.data_ptr(), int(src), int(dst), self.ranks,
                                    C.byref(r), self._path.ctypes.data if want_path else None, self._path.size if want_path else 0))
        return result_from_c(src, dst, r, self._path, self.variant)


class GpuSolver:
    """Single-GPU persistent kernel (one launch per query)."""
    variant = "gpu"

    def __init__(self, graph: Graph, device=None, **ctx_kw):
        if graph.device.type != "cuda" or graph.nranks != 1:
            dev = device or (graph.device if graph.device.type == "cuda" else "cuda:0")
            graph = _as_single(graph, dev)
        self.graph = graph
        self.ctx = SearchContext(graph, 0, 1, **ctx_kw)

    def query(self, src, dst, want_path=True, flags=0, pinned_query=None, timing="device") -> PathResult:
        """One launch per query.  ``timing="device"`` (default): no CUDA events around the launch, the call returns
        as soon as the kernel has written [result | path] into pinned memory, ``search_ms`` / ``path_ms`` are measured
        by the device's own clock inside the kernel.  ``timing="events"``: cudaEvent pair around the launch (what
        bench.py reports as the device-timed number; includes the launch latency of the kernel)."""
        if not want_path:
            flags |= F_NO_PATH
        if timing == "device":
            flags |= F_DEVICE_CLOCK
        elif timing != "events":
            raise ValueError("timing must be 'device' or 'events'")
        return self.ctx.query(src, dst, flags, want_path, pinned_query, self.variant)

    def close(self):
        self.ctx.close()


class ResidentSolver:
    """Single GPU, NO launch per query: one front kernel stays on the GPU (``cluster`` SMs: 1, 2, 4, 8 or 16) and takes its
    queries from a mailbox in pinned host memory (csrc/front.cuh k_front_batch in resident mode, csrc/ctx.cu
    bibfs_resident_*).  ``query`` writes {src, dst} into the mailbox, the kernel runs the search and the path walk and
    writes [result | path] back, the call returns when the completion word arrives: the search plus three PCIe trips
    instead of a launch and a teardown.  The kernel leaves by itself after ``idle_ms`` without a query (the next query
    launches it again), so a device-wide synchronisation elsewhere in the process waits at most that long; ``close()``
    sends it home at once.  Searches beyond the front kernel are answered by the general kernel (the resident kernel is
    sent home first).  The reference starts a process per query (benchmark_test.sh:47-54)."""
    variant = "gpu-resident"

    def __init__(self, graph: Graph, device=None, cluster: int = 1, idle_ms: int = 50, max_path: int = 0, watchdog_ms: int = 0):
        if graph.device.type != "cuda" or graph.nranks != 1:
            dev = device or (graph.device if graph.device.type == "cuda" else "cuda:0")
            graph = _as_single(graph, dev)
        self.graph = graph
        from .._lib import CtxDesc
        s = graph.single()
        vr = s.vertex_records()
        if vr is None:
            raise ValueError("the resident solver needs vertex records (graph without vertices?)")
        d = CtxDesc()
        d.n, d.nranks, d.rank_lo, d.vranks = graph.n, 1, 0, 1
        d.d_row_ptr[0] = s.row_ptr.data_ptr()
        d.d_col[0] = s.col.data_ptr() if s.col.numel() else s.row_ptr.data_ptr()
        d.d_vrec[0] = vr.data_ptr()
        d.max_degree = s.max_degree
        d.max_path = max_path or min(graph.n + 1, 1 << 24)
        d.watchdog_ms = watchdog_ms
        self.max_path = d.max_path
        self._desc = d
        self._path = np.empty(self.max_path, dtype=np.uint32)
        self._path_ptr = self._path.ctypes.data
        self._res = Result()
        self._res_ref = C.byref(self._res)
        self._h = C.c_void_p()
        with torch.cuda.device(graph.device):
            check(lib().bibfs_resident_start(C.byref(self._h), C.byref(d), cluster, idle_ms))
        self._fn = lib().bibfs_resident_query

    def query(self, src, dst, want_path=True) -> PathResult:
        if self._fn(self._h, int(src), int(dst), 0 if want_path else F_NO_PATH, self._res_ref,
                    self._path_ptr if want_path else None, self.max_path if want_path else 0):
            check(1)
        return result_from_c(src, dst, self._res, self._path, self.variant)

    @property
    def launches(self) -> int:
        """How often the kernel had to be (re)launched so far (1 under a steady stream of queries)."""
        n = C.c_uint64(0)
        check(lib().bibfs_resident_launches(self._h, C.byref(n)))
        return int(n.value)

    def close(self):
        if self._h:
            h, self._h = self._h, C.c_void_p()
            check(lib().bibfs_resident_stop(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiGpuSolver:
    """P ranks over 1-D partitioned shards.

    Three deployments of the same fused kernel:
      * ``MultiGpuSolver(graph)`` with all P shards on ONE device: P virtual ranks in one
        cooperative launch (protocol test / fewer GPUs than ranks);
      * ``MultiGpuSolver.across_devices(graphs)``: one process, one context per device, peers mapped
        with cudaDeviceEnablePeerAccess;
      * ``parallel.comm.DistributedSolver``: one process per GPU (torchrun), peers mapped through
        CUDA IPC handles exchanged over torch.distributed.
    """
    variant = "multi-gpu"

    def __init__(self, graph: Graph, **ctx_kw):
        if graph.device.type != "cuda":
            raise ValueError("MultiGpuSolver needs CUDA shards")
        if len(graph.shards) != graph.nranks:
            raise ValueError("virtual-rank mode hosts every shard on one device")
        self.graph = graph
        self.ctxs = [SearchContext(graph, 0, graph.nranks, **ctx_kw)]

    @classmethod
    def across_devices(cls, graphs: list[Graph], **ctx_kw) -> "MultiGpuSolver":
        self = cls.__new__(cls)
        self.graph = graphs[0]
        self.ctxs = []
        degs = []
        for g in graphs:
            (sh,) = g.shards.values()
            sh.vertex_records()
            degs.append(sh.max_degree if (sh.vrec is not None or sh.hi == sh.lo) else None)
        ctx_kw.setdefault("max_degree", 0 if None in degs else max(degs))
        for g in graphs:
            (rank,) = g.shards.keys()
            with torch.cuda.device(g.device):
                self.ctxs.append(SearchContext(g, rank, 1, **ctx_kw))
        for a in self.ctxs:
            for b in self.ctxs:
                if a is not b:
                    a.attach_local(b.rank_lo, b)
        return self

    def query(self, src, dst, want_path=True, flags=0, timing=None) -> PathResult:   # timing: always CUDA events
        if not want_path:
            flags |= F_NO_PATH
        for c in self.ctxs:
            with torch.cuda.device(c.device):
                c.launch(src, dst, flags)
        res = None
        for i, c in enumerate(self.ctxs):
            with torch.cuda.device(c.device):
                r = c.fetch(src, dst, want_path and i == 0, self.variant)
            if i == 0 and r.status == STATUS_HANDED_OVER:
                # several devices driven by this process: rank 0's front kernel handed the search over, every context
                # continues it in the general kernel
                for c2 in self.ctxs:
                    with torch.cuda.device(c2.device):
                        c2.resume()
                with torch.cuda.device(c.device):
                    r = c.fetch(src, dst, want_path, self.variant)
            elif i > 0 and r.status == STATUS_HANDED_OVER:
                continue            # rank 0 answered alone: this context launched nothing
            if res is None:
                res = r
            else:   # device time = max over ranks
                res.search_ms = max(res.search_ms, r.search_ms)
                res.path_ms = max(res.path_ms, r.path_ms)
        return res

    def close(self):
        for c in self.ctxs:
            c.close()


class NaiveGpuSolver:
    """Reference-style level-synchronous GPU search (thread per vertex, host in the loop)."""
    variant = "naive-gpu"

    def __init__(self, graph: Graph, device=None):
        if graph.device.type != "cuda" or graph.nranks != 1:
            graph = _as_single(graph, device or "cuda:0")
        self.graph = graph

    def query(self, src, dst, want_path=False) -> PathResult:
        s = self.graph.single()
        length, ms = C.c_int32(-1), C.c_float(0)
        with torch.cuda.device(self.graph.device):
            check(lib().bibfs_naive_levelsync_device(self.graph.n, s.row_ptr.data_ptr(), s.col.data_ptr(), int(src), int(dst),
                                                     C.byref(length), C.byref(ms),
                                                     torch.cuda.current_stream(self.graph.device).cuda_stream))
        return PathResult(src, dst, int(length.value) if length.value >= 0 else None, None,
                          search_ms=float(ms.value), variant=self.variant)

    def close(self):
        pass


class BatchSolver:
    """Many queries at once.

    engine="concurrent" (default): ONE launch per batch whose thread-block clusters are `lanes` front kernels side by
        side, each with its own search arena, drawing queries from a shared counter (csrc/front.cuh k_front_batch,
        csrc/ctx.cu bibfs_pool_*): independent latency-bound searches overlap on the 148 SMs.
    engine="bitparallel": bit-parallel bidirectional MS-BFS, 64 queries per sweep (csrc/msbfs_dev.cu);
        pays off only when the searches are heavy and overlap, because every level sweeps all N vertices.
    """
    variant = "batch"

    def __init__(self, graph: Graph, device=None, engine: str = "concurrent", lanes: int = 64, ctas_per_lane: int = 1):
        if graph.device.type != "cuda" or graph.nranks != 1:
            graph = _as_single(graph, device or "cuda:0")
        self.graph = graph
        self.engine = engine
        self.last_ms = 0.0
        self.last_meet: list[int] = []
        self.last_levels: list[int] = []
        self.last_edges: list[int] = []
        self._pool = C.c_void_p()
        self._ws = None
        with torch.cuda.device(graph.device):
            if engine == "bitparallel":
                self._ws = torch.empty(lib().bibfs_batch_workspace_bytes(graph.n), dtype=torch.uint8, device=graph.device)
            elif engine == "concurrent":
                from .._lib import CtxDesc
                s = graph.single()
                d = CtxDesc()
                d.n, d.nranks, d.rank_lo, d.vranks = graph.n, 1, 0, 1
                d.d_row_ptr[0] = s.row_ptr.data_ptr()
                d.d_col[0] = s.col.data_ptr() if s.col.numel() else s.row_ptr.data_ptr()
                vr = s.vertex_records()
                d.d_vrec[0] = vr.data_ptr() if vr is not None else None
                d.max_degree = s.max_degree if vr is not None else 0
                free, _ = torch.cuda.mem_get_info(graph.device)
                per_lane = 20 * graph.n + (1 << 22)            # parents + queues + bitmaps, roughly
                lanes = max(1, min(lanes, int(free * 0.6 // per_lane)))
                self.lanes = lanes
                check(lib().bibfs_pool_create(C.byref(self._pool), C.byref(d), lanes, ctas_per_lane))
            else:
                raise ValueError("engine must be 'concurrent' or 'bitparallel'")

    def query_batch(self, srcs, dsts) -> list[int | None]:
        srcs = np.ascontiguousarray(np.asarray(srcs, dtype=np.uint32))
        dsts = np.ascontiguousarray(np.asarray(dsts, dtype=np.uint32))
        assert srcs.shape == dsts.shape
        nq = int(srcs.size)
        with torch.cuda.device(self.graph.device):
            if self.engine == "concurrent":
                lens = np.empty(nq, dtype=np.int32)
                meet = np.empty(nq, dtype=np.uint32)
                lev = np.empty(nq, dtype=np.uint32)
                edg = np.empty(nq, dtype=np.uint64)
                ms = C.c_float(0)
                check(lib().bibfs_pool_query(self._pool, srcs.ctypes.data, dsts.ctypes.data, nq, lens.ctypes.data, meet.ctypes.data,
                                             lev.ctypes.data, edg.ctypes.data, C.byref(ms)))
                self.last_ms = float(ms.value)
                self.last_meet = [int(x) for x in meet]
                self.last_levels = [int(x) for x in lev]
                self.last_edges = [int(x) for x in edg]
                return [int(x) if x >= 0 else None for x in lens]
            s = self.graph.single()
            out: list[int | None] = []
            self.last_ms = 0.0
            self.last_meet = []
            st = torch.cuda.current_stream(self.graph.device).cuda_stream
            for i in range(0, nq, 64):
                a, b = srcs[i:i + 64], dsts[i:i + 64]
                lens = np.empty(a.size, dtype=np.int32)
                meet = np.empty(a.size, dtype=np.uint32)
                ms = C.c_float(0)
                check(lib().bibfs_batch_query_device(self.graph.n, s.row_ptr.data_ptr(), s.col.data_ptr(), a.ctypes.data,
                                                     b.ctypes.data, a.size, lens.ctypes.data, meet.ctypes.data,
                                                     self._ws.data_ptr(), self._ws.numel(), C.byref(ms), st))
                self.last_ms += float(ms.value)
                out += [int(x) if x >= 0 else None for x in lens]
                self.last_meet += [int(x) for x in meet]
            return out

    def close(self):
        self._ws = None
        if self._pool:
            lib().bibfs_pool_destroy(self._pool)
            self._pool = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def multi_gpu_solver(n, edges, ranks: int = 4, gpus: int = 0, device="cuda:0", **ctx_kw):
    """The multi-GPU variant for the CLI / bench harness (the reference: `mpirun -np 6 ./mpi_bibfs`, v4/Makefile:63-64).
    gpus >= 2: `gpus` ranks, one per device of this process (peer access between the devices, fused exchange over
    NVLink); else `ranks` virtual ranks on `device`.  Returns (solver, number of GPUs used)."""
    if gpus and gpus >= 2:
        if torch.cuda.device_count() < gpus:
            raise ValueError(f"--gpus {gpus}: this machine has {torch.cuda.device_count()} CUDA device(s)")
        graphs = [Graph.from_edges(n, edges, device=f"cuda:{i}", nranks=gpus, ranks=[i]) for i in range(gpus)]
        return MultiGpuSolver.across_devices(graphs, **ctx_kw), gpus
    return MultiGpuSolver(Graph.from_edges(n, edges, device=device, nranks=ranks), **ctx_kw), 1


def make_solver(graph: Graph, variant: str = "gpu", **kw):
    if variant == "serial":
        return SerialSolver(graph, **kw)
    if variant == "cpu-ranks":
        return CpuRanksSolver(graph, **kw)
    if variant == "gpu":
        return GpuSolver(graph, **kw)
    if variant == "multi-gpu":
        return MultiGpuSolver(graph, **kw)
    if variant == "naive-gpu":
        return NaiveGpuSolver(graph, **kw)
    if variant == "gpu-resident":
        return ResidentSolver(graph, **kw)
    if variant == "batch":
        return BatchSolver(graph, **kw)
    raise ValueError(f"unknown variant {variant!r}; choose from {VARIANTS + ('gpu-resident', 'batch')}")
