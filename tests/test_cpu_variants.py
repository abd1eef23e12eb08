import os
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import nx_graph, nx_length, query_pairs, special_graphs

REF = Path("/root/reference")


def check_result(bb, g, G, s, t, r):
    want = nx_length(G, s, t)
    assert r.length == want, (s, t, r.length, want)
    if want is not None:
        assert r.path is not None and len(r.path) == want + 1
        bb.validate_path(g, s, t, r.path, r.length)
    else:
        assert r.path is None


@pytest.mark.parametrize("variant,kw", [("serial", {}), ("serial", {"policy": 1}), ("cpu-ranks", {"ranks": 1}),
                                        ("cpu-ranks", {"ranks": 4})])
def test_random_graphs_vs_networkx(bb, variant, kw):
    for seed, (n, deg) in enumerate([(500, 1.5), (2000, 2.2), (3000, 6.0)]):
        spec = bb.er_spec(n, deg / n, seed=seed)
        edges = bb.graph.edge_list(spec)
        g = bb.Graph.from_edges(n, edges)
        G = nx_graph(n, edges)
        solver = bb.make_solver(g, variant, **kw)
        for s, t in query_pairs(n, 25, seed):
            check_result(bb, g, G, s, t, solver.query(s, t))


@pytest.mark.parametrize("variant", ["serial", "cpu-ranks"])
def test_special_graphs(bb, variant):
    for name, n, edges in special_graphs():
        g = bb.Graph.from_edges(n, edges)
        G = nx_graph(n, edges)
        solver = bb.make_solver(g, variant)
        pairs = {(0, n - 1), (n - 1, 0), (0, 0), (n // 2, n - 1), (0, min(1, n - 1))}
        for s, t in pairs:
            check_result(bb, g, G, s, t, solver.query(s, t))


def test_src_equals_dst_is_zero(bb):
    # v1 prints 2 (or "No path") for src == dst; every variant here answers 0.
    g = bb.Graph.from_edges(4, np.array([[0, 1]], np.uint32))
    for v in ("serial", "cpu-ranks"):
        r = bb.bibfs(g, 3, 3, variant=v)
        assert r.length == 0 and r.path == [3]


def test_bad_arguments(bb):
    g = bb.Graph.from_edges(4, np.array([[0, 1]], np.uint32))
    with pytest.raises(bb.BibfsError):
        bb.bibfs(g, 0, 4, variant="serial")
    with pytest.raises(bb.BibfsError):
        bb.bibfs(g, 9, 0, variant="cpu-ranks")
    with pytest.raises(ValueError):
        bb.bibfs(g, 0, 1, variant="nope")


def test_rmat_and_grid(bb):
    g = bb.Graph.generate(bb.rmat_spec(11, edgefactor=8, seed=3))
    G = nx_graph(g.n, bb.graph.edge_list(g.spec))
    s1, s2 = bb.make_solver(g, "serial"), bb.make_solver(g, "cpu-ranks", ranks=3)
    for s, t in query_pairs(g.n, 20, 1):
        want = nx_length(G, s, t)
        assert s1.query(s, t).length == want and s2.query(s, t).length == want
    grid = bb.Graph.generate(bb.grid_spec(40, 25))
    assert bb.bibfs(grid, 0, 999, variant="cpu-ranks").length == 39 + 24


@pytest.fixture(scope="module")
def ref_v1(tmp_path_factory):
    src = REF / "v1" / "main-v1.cpp"
    if not src.exists() or shutil.which("g++") is None:
        pytest.skip("reference v1 source or g++ not available")
    d = tmp_path_factory.mktemp("refv1")
    exe = d / "bibfs_serial"
    subprocess.run(["g++", "-O2", "-std=c++11", str(src), "-o", str(exe)], check=True)
    return exe


def test_bit_exact_length_vs_reference_v1(bb, ref_v1, tmp_path):
    """Oracle A of SURVEY section 4: the reference's own v1 binary, same file, same query."""
    for name in ("1k", "10k"):
        spec = bb.preset_spec(name, seed=3)
        edges = bb.graph.edge_list(spec)
        p = tmp_path / f"{name}.bin"
        bb.io.write_bin(p, spec.n, edges)
        g = bb.Graph.from_bin(p)
        for s, t in [(0, spec.n - 1), (1, spec.n // 2), (7, 11)]:
            out = subprocess.run([str(ref_v1), str(p), str(s), str(t)], capture_output=True, text=True).stdout
            ours = bb.bibfs(g, s, t, variant="serial")
            if "No path found" in out:
                assert ours.length is None
            else:
                ref_len = int(out.split("Shortest path length = ")[1].split()[0])
                ref_path = [int(x) for x in out.split("Path: ")[1].splitlines()[0].split()]
                assert ours.length == ref_len
                assert ours.path == ref_path      # same edge order + same schedule => same tie-breaks
                assert bb.bibfs(g, s, t, variant="cpu-ranks").length == ref_len


def test_cpu_ranks_hands_over_to_parallel_levels(bb, monkeypatch):
    """The solo phase (rank 0, vertex lists) hands over to the bitset levels of all ranks at any level boundary:
    BIBFS_CPU_SOLO_MAX = 0 / 3 / 40 moves the hand-over through the search, the answers stay those of the serial variant."""
    from conftest import query_pairs
    spec = bb.er_spec(20000, 6.0 / 20000, seed=21)
    g = bb.Graph.generate(spec)
    ser = bb.SerialSolver(g, policy=0)
    pairs = query_pairs(spec.n, 10, 9) + [(5, 5)]
    want = [ser.query(s, t, want_path=False).length for s, t in pairs]
    for solo_max in ("0", "3", "40", "100000"):
        monkeypatch.setenv("BIBFS_CPU_SOLO_MAX", solo_max)
        par = bb.CpuRanksSolver(g, ranks=4)
        for (s, t), w in zip(pairs, want):
            r = par.query(s, t)
            assert r.length == w, (solo_max, s, t)
            if r.length:
                bb.validate_path(g, s, t, r.path, r.length)
