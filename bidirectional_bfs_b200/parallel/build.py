"""Sharded construction of the partitioned CSR (one process per GPU, torch.distributed initialised).

Every rank walks only ITS 1/P slice of the edge stream — a slice of the generator's chunk space, or its M/P slice of
the records of a `.bin` file — and routes each half-edge to the rank that owns the row: remote atomics on the owner's
degree histogram, then on its fill cursors, and a remote store into its column array, all through CUDA IPC peer
mappings inside two kernels (csrc/csr_dev.cu, k_degree_routed / k_fill_routed).  torch.distributed only carries the IPC
handles and three barriers.  The reference reads the file on rank 0 and MPI_Bcasts the WHOLE edge list to every rank,
each of which builds the whole CSR (v4/mpi_bas.cpp:29-57).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from .._lib import check, lib
from .partition import Partition


def _share(t: torch.Tensor):
    buf, off = C.create_string_buffer(64), C.c_uint64(0)
    check(lib().bibfs_ipc_export(t.data_ptr(), buf, C.byref(off)))
    return buf.raw, int(off.value)


def _gather_ptrs(t: torch.Tensor, rank: int, world: int) -> list[int]:
    """Device pointers of every rank's tensor `t` as mapped on this device (own pointer for this rank)."""
    shared = [None] * world
    dist.all_gather_object(shared, _share(t))
    out = []
    for r, (h, off) in enumerate(shared):
        if r == rank:
            out.append(t.data_ptr())
        else:
            p = C.c_void_p(0)
            check(lib().bibfs_ipc_import(C.create_string_buffer(h, 64), off, C.byref(p)))
            out.append(p.value)
    return out


def build_shard(n: int, device, spec=None, edges: torch.Tensor | None = None, sort_rows: bool = False):
    """-> (row_ptr int64[rows+1], col int32[nnz], records this rank walked).  `edges`: THIS rank's slice of the edge
    records (int32 [m_slice, 2] on `device`), or None with a generator `spec` (the rank takes its slice of the chunks)."""
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device(device)
    part = Partition(n, world)
    lo, hi = part.range(rank)
    rows, B = hi - lo, part.B
    L = lib()
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        if edges is not None:
            edges = edges.reshape(-1, 2).to(torch.int32).to(dev).contiguous()
            m, c_lo, c_hi, e_ptr, s_ptr = int(edges.shape[0]), 0, int(edges.shape[0]), edges.data_ptr(), None
        else:
            chunks = int(L.bibfs_gen_chunks(C.byref(spec)))
            c_lo, c_hi = chunks * rank // world, chunks * (rank + 1) // world
            m, e_ptr, s_ptr = 0, None, C.byref(spec)
        PtrArr = C.c_void_p * world
        # ---- pass 1: degrees, routed to the owners
        deg = torch.zeros(B + 1, dtype=torch.int32, device=dev)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        deg_ptrs = _gather_ptrs(deg, rank, world)
        torch.cuda.synchronize(dev)
        dist.barrier()
        check(L.bibfs_csr_route_degree_device(n, m, e_ptr, s_ptr, c_lo, c_hi, world, PtrArr(*deg_ptrs), bad.data_ptr(), st))
        torch.cuda.synchronize(dev)
        dist.barrier()
        if int(bad.item()):
            raise ValueError("edge endpoint >= n")
        # ---- own histogram -> row offsets
        row_ptr = torch.empty(rows + 1, dtype=torch.int64, device=dev)
        ws = torch.empty(max(int(L.bibfs_scan_workspace_bytes(max(rows, 1))), 256), dtype=torch.uint8, device=dev)
        check(L.bibfs_scan_u32_device(deg.data_ptr(), row_ptr.data_ptr(), rows, ws.data_ptr(), ws.numel(), st))
        nnz = int(row_ptr[rows].item()) if rows else 0
        col = torch.zeros(nnz + 4, dtype=torch.int32, device=dev)        # 4 spare entries: 16-byte bulk reads of the last row
        cursor = torch.zeros(B + 1, dtype=torch.int64, device=dev)
        cursor[:rows].copy_(row_ptr[:rows])
        cur_ptrs = _gather_ptrs(cursor, rank, world)
        col_ptrs = _gather_ptrs(col, rank, world)
        torch.cuda.synchronize(dev)
        dist.barrier()
        # ---- pass 2: fill, routed
        check(L.bibfs_csr_route_fill_device(n, m, e_ptr, s_ptr, c_lo, c_hi, world, PtrArr(*cur_ptrs), PtrArr(*col_ptrs), st))
        torch.cuda.synchronize(dev)
        dist.barrier()
        if sort_rows and rows:
            check(L.bibfs_csr_sort_rows_device(row_ptr.data_ptr(), col.data_ptr(), rows, st))
            torch.cuda.synchronize(dev)
    return row_ptr, col[:nnz], c_hi - c_lo


def read_bin_slice(path, rank: int, world: int) -> tuple[int, np.ndarray]:
    """(n, this rank's M/P slice of the edge records of a `.bin` file) — only that slice is read and uploaded."""
    hdr = np.fromfile(path, dtype="<u4", count=2)
    n, m = int(hdr[0]), int(hdr[1])
    a, b = m * rank // world, m * (rank + 1) // world
    rec = np.memmap(path, dtype="<u4", mode="r", offset=8 + 8 * a, shape=(b - a, 2)) if b > a else np.zeros((0, 2), np.uint32)
    return n, np.ascontiguousarray(rec).view(np.int32)
