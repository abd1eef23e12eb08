"""Short workload for ncu: build an R-MAT graph on the device, then a few queries.
   python scripts/profile_target.py [scale] [nqueries] [flags]"""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 6
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
spec = bb.rmat_spec(scale, edgefactor=16, seed=1)
g = bb.Graph.generate(spec, device="cuda:0")
solver = bb.GpuSolver(g)
rng = np.random.default_rng(3)
done = 0
while done < nq:
    s, t = int(rng.integers(spec.n)), int(rng.integers(spec.n))
    r = solver.query(s, t, flags=flags)
    if r.length is not None and r.length >= 3:
        done += 1
        print(f"{s}->{t}: len={r.length} levels={r.levels} bu={r.levels_bottom_up} edges={r.edges_scanned} visited={r.vertices_visited} search_ms={r.search_ms:.4f} path_ms={r.path_ms:.4f}")
solver.close()
