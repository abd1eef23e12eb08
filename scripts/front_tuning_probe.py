"""Front kernel tuning: cluster size x wide-round threshold, er100k (event-timed incl. path, cold L2) and a grid."""
import os, sys, subprocess
if len(sys.argv) > 1:
    import numpy as np, torch
    sys.path.insert(0, ".")
    import bidirectional_bfs_b200 as bb, bench
    info = bench.prepare("er100k", 1, 25); spec = bench.make_spec(bb, "er100k", 1); qs = info["queries"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    s = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
    for a, b, L, _, _ in qs[:5]: s.query(a, b)
    ev = []
    for a, b, L, _, _ in qs[5:25]:
        flush.fill_(1); r = s.query(a, b, timing="events"); ev.append((r.search_ms + r.path_ms) * 1e3); assert r.length == L
    s.close()
    g = bb.grid_spec(2048, 2048); s = bb.GpuSolver(bb.Graph.generate(g, device="cuda:0"))
    s.query(0, g.n - 1); r = s.query(0, g.n - 1, timing="events")
    print(sys.argv[1], "er100k us %.1f | grid2048 corner ms %.3f (%.2f us/level, paired %d of %d)" % (np.mean(ev), r.search_ms, r.search_ms * 1e3 / r.levels, r.levels_paired, r.levels), flush=True)
else:
    for cl in ("1", "8", "16"):
        for wm in ("96", "160", "320", "100000"):
            if cl == "1" and wm != "160": continue
            env = dict(os.environ, BIBFS_FRONT_CLUSTER=cl, BIBFS_WIDE_MIN=wm)
            subprocess.run([sys.executable, __file__, f"cluster={cl} wide_min={wm}"], env=env)
