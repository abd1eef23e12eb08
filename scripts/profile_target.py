"""Short workload for Nsight Compute: every kernel family of the library, a handful of launches each.
   python scripts/profile_target.py [search|build|all|final] [scale]
final:  one or two launches each of the kernels of the final tree, picked without probing queries on the GPU (ncu -c counts
        every launch): k_front<false> single CTA (ER 100k) and as a cluster (R-MAT, grid), k_front<true> + k_bibfs<true>
        (4 virtual ranks), k_front_batch (one launch for 24 queries).
search: k_front<false/true> (pair tier, wide rounds, long rows), k_bibfs<false/true> (general kernel, fused multi-rank
        levels with 4 virtual ranks), the query pool;  build: generators, CSR builder, scans, row sort, vertex records,
        bit-parallel batch kernel, reference-style level-synchronous baseline."""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
what = sys.argv[1] if len(sys.argv) > 1 else "all"
scale = int(sys.argv[2]) if len(sys.argv) > 2 else 20
F_NO_FRONT, F_FORCE_GRID, F_FORCE_BU = 64, 2, 8
spec = bb.rmat_spec(scale, edgefactor=16, seed=1)
g_cpu = None
if what in ("build", "all"):
    g = bb.Graph.generate(spec, device="cuda:0", sort_rows=True)         # k_degree, k_scan_*, k_fill, k_sort_rows_*
    e = bb.ops.gen.generate_edges(bb.er_spec(200000, 4.0 / 200000, seed=2), "cuda:0")        # k_gen_* (edge list writers)
    gg = bb.Graph.generate(bb.grid_spec(512, 512), device="cuda:0")
    g.shards[0].vertex_records()                                          # k_vrec_build
    rng = np.random.default_rng(5)
    b = bb.BatchSolver(g, engine="bitparallel")
    b.query_batch(rng.integers(0, spec.n, 64), rng.integers(0, spec.n, 64))                   # k_msbfs
    b.close()
    bb.NaiveGpuSolver(gg).query(0, 512 * 512 - 1)                         # reference-style loop (k_naive_*)
    print("build kernels done", g.nnz)
if what in ("search", "all"):
    g = bb.Graph.generate(spec, device="cuda:0")
    rng = np.random.default_rng(3)
    pairs = []
    s = bb.GpuSolver(g)
    while len(pairs) < 3:
        a, c = int(rng.integers(spec.n)), int(rng.integers(spec.n))
        if (s.query(a, c, want_path=False).length or 0) >= 3:
            pairs.append((a, c))
    for a, c in pairs[:2]:
        r = s.query(a, c); print("front", a, c, r.length, r.levels, r.levels_paired)                 # k_front<false>
    for a, c in pairs[:2]:
        r = s.query(a, c, flags=F_NO_FRONT); print("general", a, c, r.length, r.levels)               # k_bibfs<false>
    r = s.query(*pairs[0], flags=F_FORCE_GRID | F_FORCE_BU); print("bottom-up", r.length, r.levels_bottom_up)
    s.close()
    eg = bb.Graph.generate(bb.preset_spec("100k", seed=1), device="cuda:0")
    se = bb.GpuSolver(eg); r = se.query(0, 99999); print("er100k front", r.length, r.levels_paired); se.close()
    gr = bb.Graph.generate(bb.grid_spec(1024, 1024), device="cuda:0")
    sg = bb.GpuSolver(gr); r = sg.query(0, 1024 * 1024 - 1); print("grid front (wide rounds)", r.length, r.levels_paired); sg.close()
    m = bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=4))
    for a, c in pairs[:2]:
        r = m.query(a, c); print("front, 4 virtual ranks", r.length, r.levels_paired)                 # k_front<true>
        r = m.query(a, c, flags=F_NO_FRONT); print("fused levels, 4 virtual ranks", r.length)           # k_bibfs<true>
    m.close()
    b = bb.BatchSolver(g, lanes=8)
    b.query_batch([p[0] for p in pairs] * 3, [p[1] for p in pairs] * 3); print("pool ms", b.last_ms)    # lanes of k_front
    b.close()

if what == "final":
    host = bb.Graph.generate(spec)
    oracle = bb.SerialSolver(host, policy=1)
    rng = np.random.default_rng(3)
    pairs = []
    while len(pairs) < 24:
        a, c = int(rng.integers(spec.n)), int(rng.integers(spec.n))
        if (oracle.query(a, c, want_path=False).length or 0) >= 3:
            pairs.append((a, c))
    eg = bb.Graph.generate(bb.preset_spec("100k", seed=1), device="cuda:0")
    se = bb.GpuSolver(eg)
    for q in ((0, 99999), (17, 4242)):
        r = se.query(*q); print("er100k front (single CTA + warm-up CTAs)", r.length, r.levels_paired)       # k_front<false>
    se.close()
    m = bb.MultiGpuSolver(bb.Graph.generate(spec, device="cuda:0", nranks=4))
    r = m.query(*pairs[0]); print("front, 4 virtual ranks", r.length, r.levels_paired)                        # k_front<true>
    for a, c in pairs[:2]:
        r = m.query(a, c, flags=F_NO_FRONT); print("fused levels, 4 virtual ranks", r.length, r.levels)        # k_bibfs<true>
    m.close()
    g = bb.Graph.generate(spec, device="cuda:0")
    b = bb.BatchSolver(g, lanes=8)
    b.query_batch([p[0] for p in pairs], [p[1] for p in pairs]); print("batch of 24, one launch, ms", b.last_ms)   # k_front_batch
    b.close()
    s = bb.GpuSolver(g)
    for a, c in pairs[:3]:
        r = s.query(a, c); print("front", a, c, r.length, r.levels, r.levels_paired)                          # k_front<false>, cluster from the 2nd on
    s.close()
    gr = bb.Graph.generate(bb.grid_spec(512, 512), device="cuda:0")
    sg = bb.GpuSolver(gr)
    for q in ((0, 512 * 512 - 1), (5, 512 * 300)):
        r = sg.query(*q); print("grid front (wide rounds)", r.length, r.levels_paired)
    sg.close()
