"""Graph generators: Erdos-Renyi G(n,p), R-MAT / Graph500 Kronecker, 2-D grid.

Thin wrappers over the counter-based generator in csrc/gen.h, which runs identically on the host
(``device='cpu'``) and as sm_100a kernels (csrc/gen_dev.cu).  Reference: graphs/generate_graph.py
(networkx ``fast_gnp_random_graph``, unseeded, plus an unused dense N x N matrix) and
graphs/make_graphs (N in 1k/10k/50k/100k, p = 2.2/N).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .._lib import GenSpec, check, lib

KIND_ER, KIND_RMAT, KIND_GRID = 0, 1, 2


def er_spec(n: int, p: float, seed: int = 1) -> GenSpec:
    s = GenSpec()
    s.kind, s.n, s.p, s.seed = KIND_ER, n, p, seed
    return s


def rmat_spec(scale: int, edgefactor: int = 16, n: int | None = None, seed: int = 1,
              a: float = 0.57, b: float = 0.19, c: float = 0.19) -> GenSpec:
    """Graph500 Kronecker parameters by default (A=.57, B=.19, C=.19, edgefactor 16)."""
    s = GenSpec()
    s.kind, s.scale, s.seed = KIND_RMAT, scale, seed
    s.n = n if n is not None else (1 << scale)
    s.m = edgefactor * s.n
    s.a16 = int(a * 65536)
    s.ab16 = int((a + b) * 65536)
    s.abc16 = int((a + b + c) * 65536)
    return s


def grid_spec(width: int, height: int) -> GenSpec:
    s = GenSpec()
    s.kind, s.width, s.height, s.n = KIND_GRID, width, height, width * height
    return s


def preset_spec(name: str, seed: int = 1) -> GenSpec:
    """The reference's make_graphs presets: '1k', '10k', '50k', '100k' -> G(N, 2.2000000001/N)."""
    table = {"1k": 1000, "10k": 10000, "50k": 50000, "100k": 100000, "1m": 1000000}
    n = table[name.lower()]
    return er_spec(n, 2.2000000001 / n, seed)


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def count_edges(spec: GenSpec, device="cpu") -> int:
    m = C.c_uint64(0)
    dev = torch.device(device)
    if dev.type == "cpu":
        check(lib().bibfs_gen_count_host(C.byref(spec), C.byref(m)))
    else:
        with torch.cuda.device(dev):
            ws_bytes = lib().bibfs_gen_workspace_bytes(C.byref(spec))
            ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device=dev)
            check(lib().bibfs_gen_count_device(C.byref(spec), ws.data_ptr(), ws.numel(), C.byref(m), _stream_ptr(dev)))
    return int(m.value)


def generate_edges(spec: GenSpec, device="cpu") -> torch.Tensor:
    """Materialise the edge list: int32 tensor [m, 2] holding uint32 bit patterns, each undirected
    edge once (the layout of the reference's .bin payload)."""
    dev = torch.device(device)
    m = count_edges(spec, dev)
    out = torch.empty((max(m, 1), 2), dtype=torch.int32, device=dev)
    got = C.c_uint64(0)
    if dev.type == "cpu":
        check(lib().bibfs_gen_edges_host(C.byref(spec), out.data_ptr(), m, C.byref(got)))
    else:
        with torch.cuda.device(dev):
            ws_bytes = lib().bibfs_gen_workspace_bytes(C.byref(spec))
            ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device=dev)
            check(lib().bibfs_gen_edges_device(C.byref(spec), out.data_ptr(), m, ws.data_ptr(), ws.numel(),
                                               C.byref(got), _stream_ptr(dev)))
            torch.cuda.current_stream(dev).synchronize()
    assert got.value == m
    return out[:m]


def edges_to_numpy(edges: torch.Tensor) -> np.ndarray:
    return edges.detach().cpu().numpy().view(np.uint32)
