import sys, numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
from bench import make_spec, pick_queries
cfg = sys.argv[1]
spec, desc = make_spec(bb, cfg, 1)
g = bb.Graph.generate(spec, device="cuda:0")
s = bb.GpuSolver(g)
qs = pick_queries(lambda a, b: s.query(a, b, want_path=False).length, int(spec.n), 4, 1)
for a, b, L in qs:
    for _ in range(2): r = s.query(a, b, flags=1)
    tr = s.ctx.trace()
    print(f"{a}->{b} len={L} search_ms={r.search_ms:.4f} levels={r.levels}")
    prev = None
    for i, t in enumerate(tr[:24]):
        dt = (t["t_ns"] - prev) / 1e3 if prev else 0
        prev = t["t_ns"]
        print(f"   L{i} side={t['side']} mode={t['mode']} front={t['frontier']} disc={t['discovered']} edges={t['edges']} dt_us={dt:.1f}")
