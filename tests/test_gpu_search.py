"""GPU: the persistent Bi-BFS kernel against the serial oracle / networkx, in every mode."""
import numpy as np
import pytest
import torch

from conftest import nx_graph, nx_length, query_pairs, special_graphs

pytestmark = pytest.mark.gpu

F_TRACE, F_FORCE_GRID, F_FORCE_TD, F_FORCE_BU = 1, 2, 4, 8


def check(bb, g_cpu, oracle, solver, s, t, flags=0):
    want = oracle.query(s, t, want_path=False).length
    r = solver.query(s, t, flags=flags)
    assert r.length == want, (s, t, r.length, want, flags)
    if want is not None:
        assert r.path is not None and len(r.path) == want + 1
        bb.validate_path(g_cpu, s, t, r.path, want)
    return r


GRAPHS = [
    ("er_sparse", lambda bb: bb.er_spec(30000, 2.2 / 30000, seed=1)),
    ("er_dense", lambda bb: bb.er_spec(4000, 12.0 / 4000, seed=2)),
    ("rmat", lambda bb: bb.rmat_spec(14, edgefactor=16, seed=3)),
    ("grid", lambda bb: bb.grid_spec(120, 90)),
]


@pytest.mark.parametrize("name,mk", GRAPHS)
@pytest.mark.parametrize("flags", [0, F_FORCE_GRID, F_FORCE_GRID | F_FORCE_BU, F_FORCE_TD])
def test_single_gpu_vs_oracle(bb, name, mk, flags):
    spec = mk(bb)
    g_cpu = bb.Graph.generate(spec)
    g = bb.Graph.generate(spec, device="cuda:0")
    oracle = bb.SerialSolver(g_cpu, policy=1)
    solver = bb.GpuSolver(g)
    for s, t in query_pairs(spec.n, 30, 4):
        check(bb, g_cpu, oracle, solver, s, t, flags)
    solver.close()


def test_lengths_match_networkx(bb):
    spec = bb.er_spec(5000, 3.0 / 5000, seed=8)
    edges = bb.graph.edge_list(spec)
    G = nx_graph(spec.n, edges)
    solver = bb.GpuSolver(bb.Graph.from_edges(spec.n, edges, device="cuda:0"))
    for s, t in query_pairs(spec.n, 40, 5):
        assert solver.query(s, t).length == nx_length(G, s, t)
    solver.close()


@pytest.mark.parametrize("flags", [0, F_FORCE_GRID, F_FORCE_GRID | F_FORCE_BU])
def test_special_graphs(bb, flags):
    for name, n, edges in special_graphs():
        g_cpu = bb.Graph.from_edges(n, edges)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        solver = bb.GpuSolver(bb.Graph.from_edges(n, edges, device="cuda:0"))
        for s, t in {(0, n - 1), (n - 1, 0), (0, 0), (n // 2, n - 1), (0, min(1, n - 1))}:
            check(bb, g_cpu, oracle, solver, s, t, flags)
        solver.close()


@pytest.mark.parametrize("nranks", [2, 3, 4, 8])
@pytest.mark.parametrize("flags", [0, F_FORCE_GRID, F_FORCE_GRID | F_FORCE_BU])
def test_virtual_ranks_vs_oracle(bb, nranks, flags):
    """The multi-rank protocol (remote claims, per-sender segments, cross-rank barrier, slots)
    with all P ranks emulated inside one cooperative launch on one GPU."""
    for name, mk in GRAPHS:
        spec = mk(bb)
        g_cpu = bb.Graph.generate(spec)
        g = bb.Graph.generate(spec, device="cuda:0", nranks=nranks)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        solver = bb.MultiGpuSolver(g)
        for s, t in query_pairs(spec.n, 12, nranks):
            check(bb, g_cpu, oracle, solver, s, t, flags)
        solver.close()


def test_virtual_ranks_special_graphs(bb):
    for name, n, edges in special_graphs():
        g_cpu = bb.Graph.from_edges(n, edges)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        for nranks in (2, 8):
            solver = bb.MultiGpuSolver(bb.Graph.from_edges(n, edges, device="cuda:0", nranks=nranks))
            for s, t in {(0, n - 1), (n - 1, 0), (0, 0), (n // 2, n - 1)}:
                check(bb, g_cpu, oracle, solver, s, t)
            solver.close()


def test_stress_many_small_queries(bb):
    """Thousands of launches back to back: shakes out barrier / cleanup races."""
    spec = bb.er_spec(3000, 2.5 / 3000, seed=21)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    rng = np.random.default_rng(0)
    pairs = rng.integers(0, spec.n, size=(1500, 2))
    want = [oracle.query(int(s), int(t), want_path=False).length for s, t in pairs]
    for solver in (bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0")),
                   bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=4))):
        got = [solver.query(int(s), int(t), want_path=False).length for s, t in pairs]
        assert got == want
        solver.close()


def test_trace_and_counters(bb):
    spec = bb.grid_spec(64, 64)
    solver = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
    r = solver.query(0, spec.n - 1, flags=F_TRACE)
    assert r.length == 126 and r.levels == 126 and r.depth_s + r.depth_t == 126
    tr = solver.ctx.trace()
    assert len(tr) == r.levels and sum(t["edges"] for t in tr) == r.edges_scanned
    assert all(tr[i]["t_ns"] <= tr[i + 1]["t_ns"] for i in range(len(tr) - 1))
    assert r.search_ms > 0
    solver.close()


def test_pinned_query_input(bb):
    spec = bb.er_spec(2000, 3.0 / 2000, seed=2)
    g_cpu = bb.Graph.generate(spec)
    solver = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
    q = torch.empty(2, dtype=torch.int32).pin_memory()
    for s, t in query_pairs(spec.n, 10, 3):
        q[0], q[1] = s, t
        r = solver.query(999999, 999999, pinned_query=q)   # by-value arguments are ignored
        assert r.length == bb.bibfs(g_cpu, s, t, variant="serial").length
    solver.close()


def test_timing_modes_agree(bb):
    """timing="device" (no events, pinned completion word) and timing="events" return the same answer; both report
    a positive search time; long paths (beyond the pinned staging window) are fetched from the device buffer."""
    spec = bb.grid_spec(40, 40)          # corner to corner: 78 hops
    g = bb.Graph.generate(spec, device="cuda:0")
    g_cpu = g.to("cpu")
    solver = bb.GpuSolver(g)
    for s, t in [(0, spec.n - 1), (5, 5), (3, 777), (spec.n - 1, 0)]:
        a = solver.query(s, t, timing="device")
        b = solver.query(s, t, timing="events")
        c = solver.query(s, t, want_path=False, timing="device")
        d = solver.query(s, t, want_path=False, timing="events")
        assert a.length == b.length == c.length == d.length
        assert a.path == b.path and c.path is None and d.path is None
        bb.validate_path(g_cpu, s, t, a.path, a.length)
        if s != t:
            assert a.search_ms > 0 and b.search_ms > 0 and c.search_ms > 0 and d.search_ms > 0
            assert a.edges_scanned == b.edges_scanned > 0
    solver.close()
    big = bb.grid_spec(400, 2)           # 401-hop path: longer than the 240-entry pinned window
    gb = bb.Graph.generate(big, device="cuda:0")
    sb = bb.GpuSolver(gb)
    r = sb.query(0, big.n - 1)
    assert r.length is not None and len(r.path) == r.length + 1
    bb.validate_path(gb.to("cpu"), 0, big.n - 1, r.path, r.length)
    sb.close()


def test_out_of_range_query_is_an_error(bb):
    solver = bb.GpuSolver(bb.Graph.generate(bb.grid_spec(8, 8), device="cuda:0"))
    with pytest.raises(bb.BibfsError):
        solver.query(0, 64)
    assert solver.query(0, 63).length == 14     # the context survives
    solver.close()


def test_batch_vs_oracle(bb):
    for name, mk in GRAPHS[:3]:
        spec = mk(bb)
        g_cpu = bb.Graph.generate(spec)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        pairs = query_pairs(spec.n, 100, 9)
        want = [oracle.query(s, t, want_path=False).length for s, t in pairs]
        g = bb.Graph.generate(spec, device="cuda:0")
        for engine in ("concurrent", "bitparallel"):
            got = bb.bibfs_batch(g, [p[0] for p in pairs], [p[1] for p in pairs], engine=engine)
            assert got == want, (name, engine)


def test_batch_pool_many_queries_and_lanes(bb):
    spec = bb.rmat_spec(13, edgefactor=16, seed=4)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    rng = np.random.default_rng(1)
    pairs = rng.integers(0, spec.n, size=(700, 2))
    want = [oracle.query(int(s), int(t), want_path=False).length for s, t in pairs]
    g = bb.Graph.generate(spec, device="cuda:0")
    for lanes, cpl in ((1, 1), (7, 1), (32, 1), (8, 4)):
        b = bb.BatchSolver(g, lanes=lanes, ctas_per_lane=cpl)
        assert b.query_batch(pairs[:, 0], pairs[:, 1]) == want
        assert b.query_batch(pairs[:50, 0], pairs[:50, 1]) == want[:50]      # the pool is reusable
        assert b.last_ms > 0 and len(b.last_meet) == 50
        b.close()


def test_batch_pool_wide_rounds_back_to_back(bb, monkeypatch):
    """Lanes of the one-launch batch kernel that go WIDE several queries in a row (grid: frontiers outgrow one CTA;
    R-MAT: long rows dealt over the lane's cluster), for every cluster size a lane may have."""
    for spec in (bb.grid_spec(700, 500), bb.rmat_spec(15, edgefactor=16, seed=2)):
        g_cpu = bb.Graph.generate(spec)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        rng = np.random.default_rng(5)
        pairs = rng.integers(0, spec.n, size=(48, 2))
        want = [oracle.query(int(s), int(t), want_path=False).length for s, t in pairs]
        g = bb.Graph.generate(spec, device="cuda:0")
        for cl, lanes in (("2", 3), ("4", 5), ("8", 2)):
            monkeypatch.setenv("BIBFS_POOL_CLUSTER", cl)
            b = bb.BatchSolver(g, lanes=lanes)
            assert b.query_batch(pairs[:, 0], pairs[:, 1]) == want, (spec.kind, cl, lanes)
            b.close()


def test_naive_baseline_agrees(bb):
    spec = bb.er_spec(5000, 2.5 / 5000, seed=6)
    g_cpu = bb.Graph.generate(spec)
    naive = bb.NaiveGpuSolver(bb.Graph.generate(spec, device="cuda:0"))
    for s, t in query_pairs(spec.n, 8, 2):
        assert naive.query(s, t).length == bb.bibfs(g_cpu, s, t, variant="serial").length
