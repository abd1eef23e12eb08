import sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
spec = bb.grid_spec(1400, 1100)
for P in (1, 2):
    g = bb.Graph.generate(spec, device="cuda:0", nranks=P)
    s = bb.GpuSolver(g) if P == 1 else bb.MultiGpuSolver(g)
    for a, b in ((0, spec.n - 1), (1399, 1400 * 1099), (5, 700 * 1400 + 9)):
        for fl in (0, 64):
            r = s.query(a, b, flags=fl)
            print(P, a, b, "flags", fl, "len", r.length, "levels", r.levels, "paired", r.levels_paired, "grid", r.levels_grid, "status", r.status, "ms", round(r.search_ms, 4))
    s.close()
