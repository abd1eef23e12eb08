"""CSR construction for one partition shard, on the host or with the sm_100a builder kernels.

Reference: host loops degree-count -> prefix-sum -> fill (v3/bibfs_cuda_only.cu:89-99,
v4/mpi_bas.cpp:45-57) for the whole graph on every rank.  Here: csrc/csr_dev.cu (atomic degree
histogram, hand-written scan, atomic-cursor fill, optional row sort) for the rows a rank owns; the
edge stream can be a device edge array or a generator spec (fused, nothing materialised).
"""
from __future__ import annotations

import ctypes as C

import torch

from .._lib import GenSpec, check, lib


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def build_csr(n: int, edges: torch.Tensor | None, spec: GenSpec | None, row_lo: int, row_hi: int,
              device="cpu", sort_rows: bool = False) -> tuple[torch.Tensor, torch.Tensor]:
    """Return (row_ptr int64[rows+1], col int32[nnz]) for rows [row_lo,row_hi) of the mirrored graph.

    ``col`` holds uint32 bit patterns (global vertex ids); offsets are relative to the shard.
    """
    dev = torch.device(device)
    rows = row_hi - row_lo
    if edges is not None:
        if edges.device != dev:
            edges = edges.to(dev)
        edges = edges.contiguous()
        m, eptr = edges.shape[0], edges.data_ptr()
    else:
        m, eptr = 0, None
    sp = C.byref(spec) if spec is not None else None
    row_ptr = torch.empty(rows + 1, dtype=torch.int64, device=dev)
    if dev.type == "cpu":
        check(lib().bibfs_csr_count_host(n, m, eptr, sp, row_lo, row_hi, row_ptr.data_ptr()))
        nnz = int(row_ptr[rows])
        col = torch.empty(nnz + 4, dtype=torch.int32)
        check(lib().bibfs_csr_fill_host(n, m, eptr, sp, row_lo, row_hi, row_ptr.data_ptr(), col.data_ptr(), int(sort_rows)))
        return row_ptr, col[:nnz]
    with torch.cuda.device(dev):
        ws_bytes = lib().bibfs_csr_workspace_bytes(rows)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        nnz = C.c_uint64(0)
        st = _stream_ptr(dev)
        check(lib().bibfs_csr_count_device(n, m, eptr, sp, row_lo, row_hi, row_ptr.data_ptr(), ws.data_ptr(), ws_bytes,
                                           C.byref(nnz), st))
        col = torch.empty(int(nnz.value) + 4, dtype=torch.int32, device=dev)
        check(lib().bibfs_csr_fill_device(n, m, eptr, sp, row_lo, row_hi, row_ptr.data_ptr(), col.data_ptr(), ws.data_ptr(),
                                          ws_bytes, int(sort_rows), st))
        torch.cuda.current_stream(dev).synchronize()
    return row_ptr, col[: int(nnz.value)]


def exclusive_scan_u32(x: torch.Tensor) -> torch.Tensor:
    """Device exclusive scan, int32/uint32 input -> int64 output with count+1 entries."""
    assert x.is_cuda and x.dtype in (torch.int32, torch.uint32)
    x = x.contiguous()
    out = torch.empty(x.numel() + 1, dtype=torch.int64, device=x.device)
    with torch.cuda.device(x.device):
        wsb = lib().bibfs_scan_workspace_bytes(x.numel())
        ws = torch.empty(wsb, dtype=torch.uint8, device=x.device)
        check(lib().bibfs_scan_u32_device(x.data_ptr(), out.data_ptr(), x.numel(), ws.data_ptr(), wsb, _stream_ptr(x.device)))
    return out
