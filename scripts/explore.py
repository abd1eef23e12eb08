"""Exploration: time the search kernel on the BASELINE.json graph shapes (run under gpurun)."""
import sys, time, json
import numpy as np
import torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb

def timeit(solver, pairs, flags=0, reps=3):
    best = []
    for s, t in pairs:
        ms = []
        for _ in range(reps):
            r = solver.query(s, t, flags=flags, want_path=False)
            ms.append(r.search_ms)
        best.append((min(ms), r))
    return best

def far_pair(g_solver, n, seed=0):
    rng = np.random.default_rng(seed)
    return [(int(rng.integers(n)), int(rng.integers(n))) for _ in range(4)]

which = sys.argv[1:] or ["er100k", "er1m", "rmat20", "grid1k"]
for name in which:
    t0 = time.time()
    if name == "er100k": spec = bb.preset_spec("100k", seed=1)
    elif name == "er1m": spec = bb.preset_spec("1m", seed=1)
    elif name.startswith("rmat"): spec = bb.rmat_spec(int(name[4:]), edgefactor=16, seed=1)
    elif name.startswith("grid"): k = int(name[4:-1]) * 1024 if name.endswith("k") else int(name[4:]); spec = bb.grid_spec(k, k)
    g = bb.Graph.generate(spec, device="cuda:0")
    torch.cuda.synchronize()
    t_build = time.time() - t0
    n = spec.n
    pairs = [(0, n - 1)] + far_pair(None, n, 1)
    print(f"== {name}: n={n} nnz={g.nnz} build={t_build:.2f}s")
    solver = bb.GpuSolver(g)
    for (ms, r), (s, t) in zip(timeit(solver, pairs), pairs):
        print(f"  gpu   {s}->{t}: len={r.length} levels={r.levels} bu={r.levels_bottom_up} edges={r.edges_scanned} visited={r.vertices_visited} ms={ms:.4f} GTEPS={r.edges_scanned/ms/1e6 if ms else 0:.3f}")
    for fl, nm in ((2, "grid"), (4, "td"), (2 | 8, "bu")):
        if name.startswith("grid") and nm == "bu": continue
        res = timeit(solver, pairs[:2], flags=fl, reps=2)
        print(f"  gpu[{nm}] " + " ".join(f"{ms:.4f}" for ms, _ in res))
    r = solver.query(pairs[0][0], pairs[0][1], flags=1, want_path=True)
    tr = solver.ctx.trace()
    if len(tr) <= 40:
        t_prev = None
        for i, t in enumerate(tr):
            dt = (t["t_ns"] - t_prev) / 1e3 if t_prev else 0
            t_prev = t["t_ns"]
            print(f"    L{i} side={t['side']} mode={t['mode']} front={t['frontier']} disc={t['discovered']} edges={t['edges']} dt_us={dt:.1f}")
    else:
        dts = np.diff([t["t_ns"] for t in tr]) / 1e3
        print(f"    {len(tr)} levels, per-level us: median={np.median(dts):.2f} mean={dts.mean():.2f} max={dts.max():.1f}; path_ms={r.path_ms:.3f}")
    solver.close()
    for P in (2, 8):
        gm = bb.Graph.generate(spec, device="cuda:0", nranks=P)
        ms_ = bb.MultiGpuSolver(gm)
        res = timeit(ms_, pairs[:2], reps=2)
        print(f"  vranks{P} " + " ".join(f"{ms:.4f}(len={r.length})" for ms, r in res))
        ms_.close(); del gm
    if n <= 2_000_000:
        gc = g.to("cpu")
        ser = bb.SerialSolver(gc)
        for s, t in pairs[:2]:
            r = min((ser.query(s, t, want_path=False) for _ in range(3)), key=lambda r: r.search_ms)
            print(f"  serial {s}->{t}: len={r.length} ms={r.search_ms:.4f} edges={r.edges_scanned}")
        nv = bb.NaiveGpuSolver(g)
        r = nv.query(*pairs[0]); r = nv.query(*pairs[0])
        print(f"  naive-gpu {pairs[0]}: len={r.length} ms={r.search_ms:.3f}")
    del g
    torch.cuda.empty_cache()
