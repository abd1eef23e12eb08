"""Corner-to-corner queries on grids: the cluster tier / grid mode per-level costs."""
import sys, numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
for w in (1024, 4096):
    spec = bb.grid_spec(w, w)
    g = bb.Graph.generate(spec, device="cuda:0")
    s = bb.GpuSolver(g)
    ms = [s.query(0, spec.n - 1, want_path=False, timing="events").search_ms for _ in range(3)]
    r = s.query(0, spec.n - 1, want_path=False, timing="events")
    print(f"grid {w}x{w} corner: len={r.length} levels={r.levels} search_ms min={min(ms):.3f} us/level={min(ms)*1e3/r.levels:.2f}")
    s.close(); del g
spec = bb.preset_spec("100k", seed=1)
s = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
ms = [s.query(0, spec.n - 1, want_path=False, timing="events").search_ms for _ in range(200)]
print("er100k 0->n-1 L2-hot min %.4f med %.4f" % (min(ms), np.median(ms)))
