"""File formats.

* ``.bin`` edge list  ``<u32 N><u32 M>`` + M x ``<u32 u, u32 v>`` — written by the reference
  generator (graphs/generate_graph.py:35-39) and read by all four reference variants.
* path-oracle JSON  ``{"source","target","hop_count"|null,"nodes":[...]}``
  (graphs/generate_graph.py:55-62).
* networkx adjacency / node-link JSON graphs (new input format asked for by the north star; no
  reference program reads a JSON graph).
* legacy dense ``.bin``  ``<u32 N>`` + N*N bytes (v2/read_in.cpp:13-24).
"""
from __future__ import annotations

import ctypes as C
import json
from pathlib import Path

import numpy as np

from ._lib import check, lib


def read_bin(path) -> tuple[int, np.ndarray]:
    """Return (n, edges uint32[m,2]); vertex ids are validated against n."""
    n, m = C.c_uint32(0), C.c_uint64(0)
    p = str(path).encode()
    check(lib().bibfs_bin_header(p, C.byref(n), C.byref(m)))
    edges = np.empty((int(m.value), 2), dtype=np.uint32)
    check(lib().bibfs_bin_read(p, edges.ctypes.data, m.value))
    return int(n.value), edges


def write_bin(path, n: int, edges) -> None:
    e = np.ascontiguousarray(np.asarray(edges).reshape(-1, 2).astype(np.uint32, copy=False))
    check(lib().bibfs_bin_write(str(path).encode(), int(n), e.shape[0], e.ctypes.data))


def read_dense(path) -> np.ndarray:
    n = C.c_uint32(0)
    p = str(path).encode()
    check(lib().bibfs_dense_header(p, C.byref(n)))
    a = np.empty((n.value, n.value), dtype=np.uint8)
    check(lib().bibfs_dense_read(p, a.ctypes.data, n.value))
    return a


def oracle_path_for(bin_path) -> Path:
    """The reference names the oracle ``<stem>.json`` next to the .bin (generate_graph.py:53-54)."""
    return Path(bin_path).with_suffix(".json")


def write_oracle(path, source: int, target: int, nodes) -> None:
    nodes = [int(x) for x in (nodes or [])]
    data = {"source": int(source), "target": int(target),
            "hop_count": (len(nodes) - 1) if nodes else None, "nodes": nodes}
    with open(path, "w") as f:
        json.dump(data, f, indent=2)


def read_oracle(path) -> dict:
    with open(path) as f:
        d = json.load(f)
    for k in ("source", "target", "hop_count", "nodes"):
        if k not in d:
            raise ValueError(f"{path}: not a path-oracle JSON (missing {k!r})")
    return d


def read_json_graph(path) -> tuple[int, np.ndarray]:
    """networkx ``adjacency_data`` or ``node_link_data`` JSON -> (n, edges uint32[m,2]).

    Node ids are relabelled 0..n-1 in file order when they are not already such integers.
    Each undirected edge is emitted once (u < v), self loops dropped.
    """
    with open(path) as f:
        d = json.load(f)
    if "nodes" not in d:
        raise ValueError(f"{path}: no 'nodes' key; expected networkx adjacency or node-link JSON")
    ids = [nd["id"] if isinstance(nd, dict) else nd for nd in d["nodes"]]
    ok = all(isinstance(i, int) for i in ids) and sorted(ids) == list(range(len(ids)))
    index = {i: (i if ok else k) for k, i in enumerate(ids)}
    pairs = set()
    if "adjacency" in d:
        for k, nbrs in enumerate(d["adjacency"]):
            u = index[ids[k]]
            for nb in nbrs:
                v = index[nb["id"] if isinstance(nb, dict) else nb]
                if u != v:
                    pairs.add((min(u, v), max(u, v)))
    else:
        links = d.get("links", d.get("edges"))
        if links is None:
            raise ValueError(f"{path}: neither 'adjacency' nor 'links'/'edges' present")
        for e in links:
            u, v = index[e["source"]], index[e["target"]]
            if u != v:
                pairs.add((min(u, v), max(u, v)))
    edges = np.array(sorted(pairs), dtype=np.uint32).reshape(-1, 2)
    return len(ids), edges


def write_json_graph(path, n: int, edges) -> None:
    """Write a networkx ``adjacency_data``-style JSON graph."""
    adj = [[] for _ in range(n)]
    for u, v in np.asarray(edges).reshape(-1, 2).tolist():
        adj[u].append({"id": v})
        adj[v].append({"id": u})
    d = {"directed": False, "multigraph": False, "graph": {}, "nodes": [{"id": i} for i in range(n)], "adjacency": adj}
    with open(path, "w") as f:
        json.dump(d, f)
