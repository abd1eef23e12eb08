"""One-process-per-GPU deployment of the fused multi-GPU search (torchrun / torch.distributed).

torch.distributed (NCCL or gloo) is used for PLUMBING only: exchanging the 64-byte CUDA IPC
handles of the per-rank state arenas at start-up and an occasional barrier.  The per-level
frontier exchange, meet test and termination run inside the persistent kernels over the mapped
peer memory (csrc/bfs_dev.cu) — no NCCL / MPI call per level, unlike the reference's
MPI_Allreduce(BOR) + two scalar allreduces per level (v4/mpi_bas.cpp:85,107,124).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from .._lib import F_DEVICE_CLOCK, F_NO_PATH
from ..graph import Graph
from ..ops.search import PathResult, SearchContext


def init_distributed(backend: str | None = None) -> tuple[int, int, int]:
    """Initialise torch.distributed from the torchrun environment; returns (rank, world, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, local


class DistributedSolver:
    """Rank-local handle of a P-GPU search; every rank must make the same calls in the same order."""
    variant = "multi-gpu"

    def __init__(self, graph: Graph, record_replica_mb: int | None = None, **ctx_kw):
        """record_replica_mb: rank 0 keeps a local copy of the other shards' vertex records (64 bytes per vertex, the index
        its front kernel walks) when all of them fit this many MB; beyond that it reads them through the peer mappings.
        Default 1024 (BIBFS_RECORD_REPLICA_MB overrides; 0 = never).  The CSR itself is never replicated."""
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised (use init_distributed())")
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        if graph.nranks != self.world or list(graph.shards) != [self.rank]:
            raise ValueError("graph must hold exactly this rank's shard of a world-sized partition")
        self.graph = graph
        with torch.cuda.device(graph.device):
            # longest row of the whole graph (the search skips its long-row handling when there is none)
            sh = graph.shards[self.rank]
            sh.vertex_records()
            mx = torch.tensor([sh.max_degree if (sh.vrec is not None or sh.hi == sh.lo) else 1 << 30], dtype=torch.int64, device=graph.device)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            ctx_kw.setdefault("max_degree", int(mx.item()) if int(mx.item()) < (1 << 30) else 0)
            self.ctx = SearchContext(graph, self.rank, 1, **ctx_kw)
            handles = [None] * self.world
            dist.all_gather_object(handles, self.ctx.ipc_handle(0))
            for r, h in enumerate(handles):
                if r != self.rank:
                    self.ctx.attach_ipc(r, h)
            # the graph itself: col + vertex records of every shard mapped on every device, so that rank 0's front kernel
            # can answer small queries alone (csrc/front.cuh)
            shared = [None] * self.world
            dist.all_gather_object(shared, self.ctx.graph_ipc(0))
            if record_replica_mb is None:
                record_replica_mb = int(os.environ.get("BIBFS_RECORD_REPLICA_MB", "1024"))
            self.records_replicated = False
            if self.rank == 0 and graph.n * 64 <= record_replica_mb << 20:
                self.records_replicated = self.ctx.replicate_records(shared)
            if not self.records_replicated:
                for r, sh in enumerate(shared):
                    if r != self.rank:
                        self.ctx.attach_graph_ipc(r, sh)
        dist.barrier()

    def launch(self, src, dst, flags=0, pinned_query=None):
        self.ctx.launch(src, dst, flags, pinned_query)

    def fetch(self, src, dst, want_path=True) -> PathResult:
        return self.ctx.fetch(src, dst, want_path, self.variant)

    def query(self, src, dst, want_path=True, flags=0, pinned_query=None, timing="events") -> PathResult:
        """timing="events": CUDA events around this rank's launches; "device": no events, completion by polling the pinned
        result, search_ms from the device clock (the latency path, like GpuSolver.query's default)."""
        if not want_path:
            flags |= F_NO_PATH
        if timing == "device":
            flags |= F_DEVICE_CLOCK
        return self.ctx.query(src, dst, flags, want_path, pinned_query, self.variant)

    def query_many(self, srcs, dsts, want_path=False) -> list[PathResult]:
        """A batch of queries on the partitioned graph, one after the other (every rank passes the same lists)."""
        return [self.query(int(s), int(t), want_path=want_path) for s, t in zip(srcs, dsts)]

    def query_batch(self, srcs, dsts, lanes: int = 32) -> list[int | None]:
        """Lengths of a batch of queries (BASELINE config 5: 64 random queries on a graph that is partitioned over the
        GPUs).  Rank 0 answers the batch with ONE launch whose thread-block clusters are `lanes` front kernels
        side by side — each one rank 0 working ALONE on the whole partitioned graph through the peer mappings (pair tier,
        wide rounds, long rows over its cluster), so the searches overlap their latency chains instead of queueing behind
        one another; the few searches a lane cannot finish
        (frontier beyond 65536 per side) are then answered by the whole group, one at a time.  The reference answers one
        query per process (benchmark_test.sh:47-54).  Every rank calls this with the same lists; `last_batch_ms` holds
        the device time of the pooled part on rank 0."""
        import ctypes as C
        import numpy as np
        from .._lib import check, lib
        srcs = np.ascontiguousarray(np.asarray(srcs, dtype=np.uint32))
        dsts = np.ascontiguousarray(np.asarray(dsts, dtype=np.uint32))
        nq = int(srcs.size)
        lens = np.full(nq, -2, dtype=np.int32)
        self.last_batch_ms = 0.0
        if self.rank == 0 and len(self.ctx.graph_ptrs) == self.world:
            with torch.cuda.device(self.graph.device):
                if getattr(self, "_pool", None) is None:
                    free, _ = torch.cuda.mem_get_info(self.graph.device)
                    per_lane = self.ctx.arena_bytes + (1 << 22)
                    self._lanes = max(1, min(lanes, int(free * 0.5 // per_lane)))
                    self._pool = C.c_void_p()
                    check(lib().bibfs_pool_create(C.byref(self._pool), C.byref(self.ctx._desc), self._lanes, 1))
                    for r, (col, vrec) in self.ctx.graph_ptrs.items():
                        check(lib().bibfs_pool_set_peer_graph(self._pool, r, C.c_void_p(col), C.c_void_p(vrec)))
                ms = C.c_float(0)
                check(lib().bibfs_pool_query(self._pool, srcs.ctypes.data, dsts.ctypes.data, nq, lens.ctypes.data, None, None, None, C.byref(ms)))
                self.last_batch_ms = float(ms.value)
        # the lengths go to the other ranks as ONE tensor broadcast (an object broadcast pickles and sends twice: 0.4 ms)
        buf = getattr(self, "_batch_buf", None)
        if buf is None or buf.numel() < nq:
            buf = self._batch_buf = torch.empty(max(nq, 1024), dtype=torch.int32, device=self.graph.device)
        if self.rank == 0:
            buf[:nq].copy_(torch.from_numpy(lens))
        dist.broadcast(buf[:nq], src=0)
        out = []
        for i, L in enumerate(buf[:nq].tolist()):
            if L == -2:                      # not answered by the pool: the whole group takes it
                out.append(self.query(int(srcs[i]), int(dsts[i]), want_path=False).length)
            else:
                out.append(None if L < 0 else int(L))
        return out

    def close(self):
        torch.cuda.synchronize(self.graph.device)
        dist.barrier()
        if getattr(self, "_pool", None) is not None:
            from .._lib import lib
            lib().bibfs_pool_destroy(self._pool)
            self._pool = None
        self.ctx.close()
