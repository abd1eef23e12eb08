"""Graph: 1-D vertex-partitioned CSR shards held as torch tensors.

``row_ptr`` is int64 (64-bit offsets: the reference's ``int`` offsets overflow at 2M >= 2^31,
v4/comp.cu:49-55), ``col`` is int32 carrying uint32 bit patterns.  A Graph hosts one or more of
the P shards: all of them on the CPU or for single-process multi-rank emulation, exactly one per
process in the torchrun multi-GPU mode.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from . import io as _io
from ._lib import GenSpec
from .ops import csr as _csr
from .ops import gen as _gen
from .parallel.partition import Partition


@dataclass
class Shard:
    rank: int
    lo: int
    hi: int
    row_ptr: torch.Tensor
    col: torch.Tensor
    vrec: torch.Tensor | None = None      # vertex records (csrc/vrec_dev.cu), built on first use by a search context
    max_degree: int = 0                   # longest row of the shard (known once the records are built)

    @property
    def nnz(self) -> int:
        return int(self.col.numel())

    def vertex_records(self) -> torch.Tensor | None:
        """64-byte record per owned vertex ({degree, row start, first six neighbours} + neighbours 7..14) for the latency tier of the search
        kernel (csrc/pair_tier.cuh); CUDA shards with < 2^32 stored entries only, else None."""
        if self.vrec is None and self.row_ptr.device.type == "cuda" and self.nnz < (1 << 32) and self.hi > self.lo:
            import ctypes as C
            from ._lib import check, lib
            rows = self.hi - self.lo
            dev = self.row_ptr.device
            with torch.cuda.device(dev):
                out = torch.empty(rows * 16, dtype=torch.int32, device=dev)
                colp = self.col.data_ptr() if self.col.numel() else self.row_ptr.data_ptr()
                mx = torch.zeros(1, dtype=torch.int32, device=dev)
                check(lib().bibfs_vrec_build_device(self.row_ptr.data_ptr(), colp, rows, self.nnz, out.data_ptr(), mx.data_ptr(),
                                                    torch.cuda.current_stream(dev).cuda_stream))
                self.max_degree = int(mx.item())
            self.vrec = out
        return self.vrec


@dataclass
class Graph:
    n: int
    nranks: int
    device: torch.device
    shards: dict[int, Shard] = field(default_factory=dict)
    m_records: int = 0            # edge records fed to the builder (each undirected edge once)
    spec: GenSpec | None = None
    name: str = ""

    # ---- construction ------------------------------------------------------------------
    @staticmethod
    def _build(n, edges, spec, device, nranks, ranks, sort_rows, name) -> "Graph":
        dev = torch.device(device)
        part = Partition(n, nranks)
        g = Graph(n=n, nranks=nranks, device=dev, spec=spec, name=name)
        hosted = list(range(nranks)) if ranks is None else list(ranks)
        if (spec is not None or edges is None) and _sharded_build_applies(dev, nranks, hosted, from_generator=True):
            # one process per GPU: each rank walks 1/P of the generator's chunks and routes the half-edges to their owners
            from .parallel.build import build_shard
            (r,) = hosted
            lo, hi = part.range(r)
            rp, col, _ = build_shard(n, dev, spec=spec, sort_rows=sort_rows)
            g.shards[r] = Shard(r, lo, hi, rp, col)
            return g
        if edges is not None:
            if isinstance(edges, np.ndarray):
                edges = torch.from_numpy(np.ascontiguousarray(edges.reshape(-1, 2)).view(np.int32))
            edges = edges.reshape(-1, 2).to(torch.int32).to(dev).contiguous()
            g.m_records = edges.shape[0]
        for r in hosted:
            lo, hi = part.range(r)
            rp, col = _csr.build_csr(n, edges, spec if edges is None else None, lo, hi, dev, sort_rows)
            g.shards[r] = Shard(r, lo, hi, rp, col)
        return g

    @staticmethod
    def from_edges(n, edges, device="cpu", nranks=1, ranks=None, sort_rows=False, name="") -> "Graph":
        return Graph._build(int(n), edges, None, device, nranks, ranks, sort_rows, name)

    @staticmethod
    def from_bin(path, device="cpu", nranks=1, ranks=None, sort_rows=False) -> "Graph":
        dev = torch.device(device)
        hosted = list(range(nranks)) if ranks is None else list(ranks)
        if _sharded_build_applies(dev, nranks, hosted):
            # one process per GPU: this rank reads and uploads only its M/P slice of the records
            from .parallel.build import build_shard, read_bin_slice
            (r,) = hosted
            n, rec = read_bin_slice(path, r, nranks)
            part = Partition(n, nranks)
            lo, hi = part.range(r)
            rp, col, m_slice = build_shard(n, dev, edges=torch.from_numpy(rec), sort_rows=sort_rows)
            g = Graph(n=n, nranks=nranks, device=dev, name=str(path), m_records=m_slice)
            g.shards[r] = Shard(r, lo, hi, rp, col)
            return g
        n, edges = _io.read_bin(path)
        return Graph._build(n, edges, None, device, nranks, ranks, sort_rows, str(path))

    @staticmethod
    def from_json(path, device="cpu", nranks=1, ranks=None, sort_rows=False) -> "Graph":
        n, edges = _io.read_json_graph(path)
        return Graph._build(n, edges, None, device, nranks, ranks, sort_rows, str(path))

    @staticmethod
    def from_file(path, **kw) -> "Graph":
        return Graph.from_json(path, **kw) if str(path).endswith(".json") else Graph.from_bin(path, **kw)

    @staticmethod
    def generate(spec: GenSpec, device="cpu", nranks=1, ranks=None, sort_rows=False, name="") -> "Graph":
        """Build straight from a generator spec: the edge list is never materialised; every hosted
        shard regenerates the stream and keeps the rows it owns."""
        return Graph._build(int(spec.n), None, spec, device, nranks, ranks, sort_rows, name)

    # ---- views ---------------------------------------------------------------------------
    # ---- cached partitioned CSR (the nearest thing to a checkpoint - big graphs need not be rebuilt) ----
    def save_csr(self, path) -> None:
        """Write the hosted shards to ``path`` (one uncompressed .npz: n, nranks, ranks and row_ptr / col per shard).
        In the one-process-per-GPU mode every rank saves its own file (e.g. ``f"graph.{rank}.npz"``)."""
        arrs = {"n": np.int64(self.n), "nranks": np.int64(self.nranks), "m_records": np.int64(self.m_records),
                "ranks": np.array(sorted(self.shards), dtype=np.int64)}
        for r, sh in self.shards.items():
            arrs[f"row_ptr_{r}"] = sh.row_ptr.cpu().numpy()
            arrs[f"col_{r}"] = sh.col.cpu().numpy()                               # (the device-side padding is not part of the view)
        with open(path, "wb") as f:
            np.savez(f, **arrs)

    @staticmethod
    def load_csr(path, device="cpu", ranks=None) -> "Graph":
        """Read a file written by :meth:`save_csr`; ``ranks`` selects a subset of the stored shards."""
        dev = torch.device(device)
        with np.load(path) as z:
            n, nranks = int(z["n"]), int(z["nranks"])
            part = Partition(n, nranks)
            g = Graph(n=n, nranks=nranks, device=dev, m_records=int(z["m_records"]), name=str(path))
            for r in (int(x) for x in z["ranks"]):
                if ranks is not None and r not in ranks:
                    continue
                lo, hi = part.range(r)
                rp = torch.from_numpy(z[f"row_ptr_{r}"]).to(dev)
                col = torch.from_numpy(z[f"col_{r}"])
                g.shards[r] = Shard(r, lo, hi, rp, padded_col(col, dev) if dev.type == "cuda" else col.to(dev))
        if not g.shards:
            raise ValueError(f"{path}: none of the requested ranks {ranks} is stored in the file")
        return g

    @property
    def partition(self) -> Partition:
        return Partition(self.n, self.nranks)

    @property
    def nnz(self) -> int:
        return sum(s.nnz for s in self.shards.values())

    def single(self) -> Shard:
        if self.nranks != 1:
            raise ValueError("this operation needs an unpartitioned graph (nranks=1)")
        return self.shards[0]

    def degree(self, v: int) -> int:
        s = self.shards[self.partition.owner(v)]
        i = v - s.lo
        return int(s.row_ptr[i + 1] - s.row_ptr[i])

    def neighbors(self, v: int) -> np.ndarray:
        s = self.shards[self.partition.owner(v)]
        i = v - s.lo
        a, b = int(s.row_ptr[i]), int(s.row_ptr[i + 1])
        return s.col[a:b].cpu().numpy().view(np.uint32)

    def to_scipy(self):
        from scipy.sparse import csr_matrix
        s = self.single()
        rp = s.row_ptr.cpu().numpy()
        col = s.col.cpu().numpy().view(np.uint32)
        return csr_matrix((np.ones(col.shape[0], dtype=np.uint8), col, rp), shape=(self.n, self.n))

    def to(self, device) -> "Graph":
        dev = torch.device(device)
        g = Graph(self.n, self.nranks, dev, {}, self.m_records, self.spec, self.name)
        for r, s in self.shards.items():
            g.shards[r] = Shard(r, s.lo, s.hi, s.row_ptr.to(dev), padded_col(s.col, dev))
        return g


def _sharded_build_applies(dev: torch.device, nranks: int, hosted: list, from_generator: bool = False) -> bool:
    """The routed, sharded construction (parallel/build.py) is used when this process is one rank of a torch.distributed
    group of `nranks` processes and builds exactly its own shard on a GPU.  Edge lists that exist only once (a .bin file)
    are always routed.  A GENERATOR can also be replayed by every rank (each keeps the rows it owns): the Philox stream
    costs ALU time only, the routed build costs two remote atomics and a remote store per half-edge — measured on
    Kronecker scale 26 over 4 GPUs: 0.060 s replayed, 0.12 s routed — so up to 8 ranks a generator is replayed, beyond
    that (1/P of the stream per rank vs all of it) routed.  BIBFS_SHARDED_BUILD=0 / 1 forces never / always."""
    import os
    import torch.distributed as dist
    mode = os.environ.get("BIBFS_SHARDED_BUILD")
    if dev.type != "cuda" or nranks < 2 or len(hosted) != 1 or mode == "0":
        return False
    if from_generator and mode != "1" and nranks <= 8:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() == nranks and dist.get_rank() == hosted[0]


def padded_col(col: torch.Tensor, device) -> torch.Tensor:
    """Copy of `col` on `device` whose storage has 4 spare entries after the last one: the hub phase
    fetches adjacency chunks with 16-byte bulk copies that may read up to 12 bytes past a row's end."""
    out = torch.zeros(col.numel() + 4, dtype=col.dtype, device=device)
    out[: col.numel()].copy_(col)
    return out[: col.numel()]


def edge_list(spec: GenSpec, device="cpu") -> np.ndarray:
    return _gen.edges_to_numpy(_gen.generate_edges(spec, device))
