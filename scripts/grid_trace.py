"""Per-level time by frontier size on a grid corner query (level trace)."""
import sys, numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
w = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
spec = bb.grid_spec(w, w)
s = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
for _ in range(2): s.query(0, spec.n - 1, want_path=False)
r = s.query(0, spec.n - 1, want_path=False, flags=1, timing="events")
tr = s.ctx.trace()
f = np.array([t["frontier"] for t in tr[1:]]); dt = np.array([tr[i]["t_ns"] - tr[i - 1]["t_ns"] for i in range(1, len(tr))]) / 1e3
ex = np.array([tr[i]["t_expand_ns"] - tr[i - 1]["t_ns"] for i in range(1, len(tr))]) / 1e3
mode = np.array([t["mode"] for t in tr[1:]])
print(f"grid {w}: levels {r.levels} search_ms {r.search_ms:.3f}")
for lo, hi in [(1, 32), (33, 128), (129, 256), (257, 512), (513, 768), (769, 1024), (1025, 1536), (1537, 2048), (2049, 3072), (3073, 4096), (4097, 10**9)]:
    m = (f >= lo) & (f <= hi)
    if m.any():
        print(f"  frontier {lo:5d}-{hi:<6d}: {m.sum():5d} levels, mean {dt[m].mean():6.2f} us (expand part {ex[m].mean():6.2f}), modes {sorted(set(mode[m].tolist()))}")
