import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def bb():
    import bidirectional_bfs_b200 as bb
    return bb


def nx_graph(n, edges):
    import networkx as nx
    G = nx.Graph()
    G.add_nodes_from(range(n))
    G.add_edges_from((int(u), int(v)) for u, v in np.asarray(edges).reshape(-1, 2) if u != v)
    return G


def nx_length(G, s, t):
    import networkx as nx
    try:
        return nx.shortest_path_length(G, s, t)
    except nx.NetworkXNoPath:
        return None


def special_graphs():
    """(name, n, edges) edge cases from SURVEY section 4."""
    out = []
    out.append(("single", 1, np.zeros((0, 2), np.uint32)))
    out.append(("pair", 2, np.array([[0, 1]], np.uint32)))
    out.append(("pair_disconnected", 2, np.zeros((0, 2), np.uint32)))
    out.append(("path33", 33, np.array([[i, i + 1] for i in range(32)], np.uint32)))
    star = np.array([[0, i] for i in range(1, 300)], np.uint32)
    out.append(("star300", 300, star))
    out.append(("two_components", 70, np.array([[i, i + 1] for i in range(30)] + [[i, i + 1] for i in range(40, 69)], np.uint32)))
    out.append(("dups_selfloops", 10, np.array([[0, 1], [1, 0], [0, 1], [1, 1], [1, 2], [2, 2], [2, 3], [3, 9], [9, 3]], np.uint32)))
    out.append(("isolated_dst", 65, np.array([[i, i + 1] for i in range(63)], np.uint32)))
    hub = [[0, i] for i in range(1, 5000)] + [[i, 5000 + (i % 7)] for i in range(1, 5000)]
    out.append(("hub5000", 5007, np.array(hub, np.uint32)))
    return out


def query_pairs(n, k, seed=0):
    rng = np.random.default_rng(seed)
    pairs = [(0, n - 1), (0, 0), (n - 1, 0)]
    while len(pairs) < k:
        pairs.append((int(rng.integers(n)), int(rng.integers(n))))
    return pairs[:k]
