"""GPU: query flow with the front kernel (csrc/front.cuh) — rank 0 answers small queries alone over the whole
partitioned graph, hands larger searches over to the fused multi-rank levels — with P virtual ranks on one device
(same protocol as P GPUs: peer pointers are plain device pointers), plus the watchdog / reset path."""
import numpy as np
import pytest
import torch

from conftest import query_pairs

pytestmark = pytest.mark.gpu
F_FORCE_GRID, F_NO_FRONT = 2, 64


def _check(bb, g_cpu, oracle, solver, s, t, flags=0):
    want = oracle.query(s, t, want_path=False).length
    r = solver.query(s, t, flags=flags)
    assert r.length == want, (s, t, r.length, want, flags)
    if want is not None:
        bb.validate_path(g_cpu, s, t, r.path, want)
    return r


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_front_kernel_answers_small_queries_alone(bb, nranks):
    spec = bb.er_spec(40000, 2.2 / 40000, seed=7)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    solver = bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=nranks))
    alone = 0
    for s, t in query_pairs(spec.n, 80, nranks):
        r = _check(bb, g_cpu, oracle, solver, s, t)
        if r.length not in (None, 0) and r.levels_paired == r.levels:
            alone += 1
        r2 = _check(bb, g_cpu, oracle, solver, s, t, F_NO_FRONT)          # the fused levels from the first level on
        assert r2.levels_paired == 0
    assert alone > 20, "the front kernel should finish most ER queries without a hand-over"
    solver.close()


@pytest.mark.parametrize("nranks,cluster", [(2, "1"), (3, "1"), (8, "1"), (2, None), (8, None)])
def test_hand_over_to_the_fused_levels(bb, nranks, cluster, monkeypatch):
    """Searches that outgrow the front kernel.  With a single-CTA front kernel (BIBFS_FRONT_CLUSTER=1): a grid whose
    diagonal passes 1024 vertices, R-MAT hubs after a level or two -> hand-over to the fused multi-rank levels.  With its
    cluster (default) the wide rounds take those too; only the answers are checked then."""
    if cluster:
        monkeypatch.setenv("BIBFS_FRONT_CLUSTER", cluster)
    for spec, pairs in ((bb.grid_spec(1400, 1100), [(0, 1400 * 1100 - 1), (1399, 1400 * 1099), (5, 700 * 1400 + 9)]),
                        (bb.rmat_spec(15, edgefactor=16, seed=3), None), (bb.er_spec(6000, 9.0 / 6000, seed=4), None)):
        g_cpu = bb.Graph.generate(spec)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        solver = bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=nranks))
        handed = 0
        for s, t in (pairs or query_pairs(spec.n, 40, nranks + 11)):
            r = _check(bb, g_cpu, oracle, solver, s, t)
            handed += int(r.length is not None and r.levels > r.levels_paired)
        assert handed > 0 or cluster is None or spec.kind == 0
        solver.close()


def test_single_rank_front_then_general_kernel(bb):
    """A grid stays with the front kernel's cluster (wide rounds); a graph with rows of >= 32 entries next to a long
    sparse approach is handed to the general kernel after some paired levels."""
    spec = bb.grid_spec(1500, 1300)
    g_cpu = bb.Graph.generate(spec)
    solver = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
    r = solver.query(0, spec.n - 1)          # a context starts with the single-CTA front kernel: this search outgrows it,
    assert r.length == 1499 + 1299 and 0 < r.levels_paired < r.levels       # is handed to the general kernel, and brings the cluster in
    for timing in ("device", "events"):
        r = solver.query(0, spec.n - 1, timing=timing)
        assert r.length == 1499 + 1299 and r.levels_paired == r.levels
        bb.validate_path(g_cpu, 0, spec.n - 1, r.path, r.length)
        r = solver.query(7, 7 + 1500 * 2 + 3, timing=timing)
        assert r.length == 5 and r.levels_paired == r.levels
        tr_flags = 1
        r = solver.query(0, spec.n - 1, flags=tr_flags, timing=timing)
        tr = solver.ctx.trace()
        assert len(tr) == r.levels and any(x["mode"] & 32 for x in tr), "no wide round in the trace"
    solver.close()
    # two paths of 40 vertices into a clique of 64: paired levels first, then rows of 63 entries (long-row chunks)
    import numpy as np
    k, L = 64, 40
    edges = [(i, j) for i in range(k) for j in range(i + 1, k)]
    a = list(range(k, k + L)); b = list(range(k + L, k + 2 * L))
    edges += [(a[i], a[i + 1]) for i in range(L - 1)] + [(b[i], b[i + 1]) for i in range(L - 1)] + [(a[-1], 0), (b[-1], 1)]
    n = k + 2 * L
    e = np.array(edges, dtype=np.uint32)
    g_cpu = bb.Graph.from_edges(n, e)
    solver = bb.GpuSolver(bb.Graph.from_edges(n, e, device="cuda:0"))
    r = solver.query(a[0], b[0])
    assert r.length == 2 * L + 1 and r.levels_paired > 0        # (the cluster's wide rounds take the rows of 63 entries)
    bb.validate_path(g_cpu, a[0], b[0], r.path, r.length)
    solver.close()


def test_back_to_back_queries_on_each_others_paths(bb):
    """Query k+1 starts on the path of query k (the parents of k must not be disturbed while they are walked)."""
    spec = bb.grid_spec(300, 40)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    solver = bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=4))
    s, t = 0, spec.n - 1
    for _ in range(12):
        r = _check(bb, g_cpu, oracle, solver, s, t)
        s, t = r.path[len(r.path) // 3], r.path[-1 - len(r.path) // 5]
    solver.close()


def test_watchdog_abort_and_reset(bb):
    """A rank whose peer never launches: the device-side watchdog ends the query with an error instead of a hang; after
    bibfs_ctx_reset on both contexts the group answers again."""
    from bidirectional_bfs_b200.ops.search import SearchContext
    from bidirectional_bfs_b200._lib import BibfsError
    spec = bb.er_spec(20000, 2.2 / 20000, seed=2)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    g = bb.Graph.generate(spec, device="cuda:0", nranks=2)
    c0 = SearchContext(g, 0, 1, max_ctas=32, watchdog_ms=60)
    c1 = SearchContext(g, 1, 1, max_ctas=32, watchdog_ms=60)
    c0.attach_local(1, c1)
    c1.attach_local(0, c0)
    s, t = next((a, b) for a, b in query_pairs(spec.n, 60, 3) if (oracle.query(a, b, want_path=False).length or 0) > 3)
    c0.launch(s, t, F_NO_FRONT)                    # rank 1 is never launched
    with pytest.raises(BibfsError, match="watchdog"):
        c0.fetch(s, t)
    c0.reset(); c1.reset()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    for flags in (F_NO_FRONT, 0):
        c0.launch(s, t, flags)
        with torch.cuda.stream(side):
            c1.launch(s, t, flags)
        r = c0.fetch(s, t)
        if r.status == 100:
            c0.resume()
            with torch.cuda.stream(side):
                c1.resume()
            r = c0.fetch(s, t)
        with torch.cuda.stream(side):
            c1.fetch(s, t, want_path=False)
        assert r.length == oracle.query(s, t, want_path=False).length
        bb.validate_path(g_cpu, s, t, r.path, r.length)
    c0.close(); c1.close()


def test_query_pool_on_a_partitioned_graph(bb):
    """Lanes of front kernels on rank 0's device answering a batch on a graph partitioned over 4 ranks (the other shards
    are reached through their pointers, here on the same device); what a lane cannot finish comes back as -2."""
    import ctypes as C
    from bidirectional_bfs_b200._lib import CtxDesc, check, lib
    for spec in (bb.er_spec(60000, 2.5 / 60000, seed=9), bb.rmat_spec(15, edgefactor=16, seed=4), bb.grid_spec(300, 200)):
        g_cpu = bb.Graph.generate(spec)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        g = bb.Graph.generate(spec, device="cuda:0", nranks=4)
        d = CtxDesc()
        d.n, d.nranks, d.rank_lo, d.vranks = g.n, 4, 0, 1
        s0 = g.shards[0]
        d.d_row_ptr[0], d.d_col[0], d.d_vrec[0] = s0.row_ptr.data_ptr(), s0.col.data_ptr(), s0.vertex_records().data_ptr()
        for r in range(4):
            g.shards[r].vertex_records()
        d.max_degree = max(g.shards[r].max_degree for r in range(4))
        pool = C.c_void_p()
        check(lib().bibfs_pool_create(C.byref(pool), C.byref(d), 8, 1))
        for r in range(1, 4):
            check(lib().bibfs_pool_set_peer_graph(pool, r, C.c_void_p(g.shards[r].col.data_ptr()), C.c_void_p(g.shards[r].vrec.data_ptr())))
        pairs = query_pairs(spec.n, 64, 21)
        src = np.array([p[0] for p in pairs], dtype=np.uint32)
        dst = np.array([p[1] for p in pairs], dtype=np.uint32)
        lens = np.empty(64, dtype=np.int32)
        ms = C.c_float(0)
        for _ in range(2):
            check(lib().bibfs_pool_query(pool, src.ctypes.data, dst.ctypes.data, 64, lens.ctypes.data, None, None, None, C.byref(ms)))
        answered = 0
        for (s, t), L in zip(pairs, lens):
            want = oracle.query(s, t, want_path=False).length
            if L != -2:
                assert (None if L < 0 else int(L)) == want, (s, t, L, want)
                answered += 1
        assert answered >= 60
        check(lib().bibfs_pool_destroy(pool))


def test_native_cli_gpu_variants(bb, tmp_path):
    """bin/bibfs (the reference's argv shape): front kernel on one GPU and on four virtual ranks."""
    import re
    import subprocess
    from pathlib import Path
    exe = Path(__file__).resolve().parent.parent / "bin" / "bibfs"
    if not exe.exists():
        pytest.skip("native CLI not built")
    spec = bb.er_spec(20000, 2.4 / 20000, seed=6)
    edges = bb.graph.edge_list(spec)
    path = tmp_path / "g.bin"
    bb.io.write_bin(path, spec.n, edges)
    g_cpu = bb.Graph.from_edges(spec.n, edges)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    s, t = next((a, b) for a, b in query_pairs(spec.n, 60, 5) if (oracle.query(a, b, want_path=False).length or 0) > 3)
    want = oracle.query(s, t, want_path=False).length
    for extra in (["--variant", "gpu"], ["--variant", "multi-gpu", "--vranks", "4"]):
        out = subprocess.run([str(exe), str(path), str(s), str(t)] + extra, capture_output=True, text=True, timeout=120)
        assert out.returncode == 0, out.stdout + out.stderr
        m = re.search(r"Shortest path length = (\d+)", out.stdout)
        assert m and int(m.group(1)) == want, out.stdout
        p = re.search(r"Path: ([\d ]+)", out.stdout)
        nodes = [int(x) for x in p.group(1).split()]
        bb.validate_path(g_cpu, s, t, nodes, want)


def test_resident_solver_mailbox(bb):
    """One resident front kernel answers a stream of queries through the pinned mailbox: lengths and paths equal to the
    serial variant (short paths from the mailbox, long ones from the arena), the kernel is launched once for the whole
    stream, idles out and comes back; wide rounds and long rows run inside it on a cluster of 4."""
    import time
    spec = bb.er_spec(20000, 2.2 / 20000, seed=3)
    g_cpu = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(g_cpu, policy=1)
    g = bb.Graph.generate(spec, device="cuda:0")
    pairs = query_pairs(spec.n, 60, 4) + [(5, 5), (0, spec.n - 1)]
    for cluster in (1, 4):
        r = bb.ResidentSolver(g, cluster=cluster, idle_ms=30)
        for s, t in pairs:
            want = oracle.query(s, t, want_path=False).length
            got = r.query(s, t)
            assert got.length == want, (cluster, s, t, got.length, want)
            if want is not None:
                bb.validate_path(g_cpu, s, t, got.path, want)
            assert r.query(s, t, want_path=False).length == want
        assert r.launches == 1, r.launches
        time.sleep(0.12)                      # the kernel idles out ...
        s, t = pairs[0]
        assert r.query(s, t).length == oracle.query(s, t, want_path=False).length      # ... and is launched again
        assert r.launches == 2, r.launches
        r.close()
    # wide rounds and long rows inside the resident kernel (cluster of 4), and a grid whose frontiers outgrow one CTA
    for spec2 in (bb.rmat_spec(14, edgefactor=16, seed=2), bb.grid_spec(400, 300)):
        h = bb.Graph.generate(spec2)
        o2 = bb.SerialSolver(h, policy=1)
        r = bb.ResidentSolver(bb.Graph.generate(spec2, device="cuda:0"), cluster=4, idle_ms=30)
        for s, t in query_pairs(spec2.n, 20, 6):
            want = o2.query(s, t, want_path=False).length
            got = r.query(s, t)
            assert got.length == want, (spec2.kind, s, t, got.length, want)
            if want is not None:
                bb.validate_path(h, s, t, got.path, want)
        r.close()
