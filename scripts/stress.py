"""Randomised stress: many queries per graph family on one GPU, every length checked against the serial CPU variant,
every path validated edge by edge.  python scripts/stress.py [queries_per_graph]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
nq = int(sys.argv[1]) if len(sys.argv) > 1 else 400
rng = np.random.default_rng(7)
cases = [("er50k", bb.preset_spec("50k", seed=3)), ("er dense 20k deg 12", bb.er_spec(20000, 12.0 / 20000, seed=4)),
         ("rmat16", bb.rmat_spec(16, edgefactor=16, seed=5)), ("rmat18 ef 8", bb.rmat_spec(18, edgefactor=8, seed=6)),
         ("grid 300x300", bb.grid_spec(300, 300)), ("grid 2000x40", bb.grid_spec(2000, 40))]
bad = 0
for name, spec in cases:
    g = bb.Graph.generate(spec, device="cuda:0"); gc = g.to("cpu")
    s = bb.GpuSolver(g); ser = bb.SerialSolver(gc, policy=0)
    t0 = time.time(); found = 0; maxlv = 0
    for i in range(nq):
        a, b = int(rng.integers(spec.n)), int(rng.integers(spec.n))
        r = s.query(a, b)
        w = ser.query(a, b, want_path=False)
        if r.length != w.length:
            bad += 1; print("MISMATCH", name, a, b, r.length, w.length)
        elif r.length is not None:
            found += 1; maxlv = max(maxlv, r.levels)
            if i % 4 == 0:
                bb.validate_path(gc, a, b, r.path, r.length)
    print(f"{name}: {nq} queries, {found} connected, max levels {maxlv}, {time.time() - t0:.1f}s", flush=True)
    s.close()
print("STRESS", "FAIL" if bad else "PASS", bad)
# ---- second suite: large frontiers (grid mode, bottom-up) and virtual ranks
cases2 = [("er 300k deg 8", bb.er_spec(300000, 8.0 / 300000, seed=8)), ("rmat20 ef 16", bb.rmat_spec(20, edgefactor=16, seed=9)),
          ("grid 1500x1500", bb.grid_spec(1500, 1500))]
for name, spec in cases2:
    gc = bb.Graph.generate(spec); ser = bb.SerialSolver(gc, policy=0)
    qs = [(int(rng.integers(spec.n)), int(rng.integers(spec.n))) for _ in range(max(20, nq // 10))]
    want = [ser.query(a, b, want_path=False).length for a, b in qs]
    for P in (1, 2, 4):
        g = bb.Graph.generate(spec, device="cuda:0", nranks=P)
        s = bb.GpuSolver(g) if P == 1 else bb.MultiGpuSolver(g)
        t0 = time.time(); nbu = 0
        for (a, b), w in zip(qs, want):
            r = s.query(a, b)
            nbu += r.levels_bottom_up
            if r.length != w:
                bad += 1; print("MISMATCH", name, P, a, b, r.length, w)
            elif r.length is not None:
                bb.validate_path(gc, a, b, r.path, r.length)
        print(f"{name} P={P}: {len(qs)} queries ok, bottom-up levels {nbu}, {time.time() - t0:.1f}s", flush=True)
        s.close(); del g
print("STRESS2", "FAIL" if bad else "PASS", bad)
