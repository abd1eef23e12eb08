"""Per-level trace of a few bench queries: python scripts/trace_levels.py <config> [flags] [nq]
(mode: 0 solo, 1 grid, 2 bottom-up, +8 hub phase, +16 paired round, +32 wide round of the front kernel's cluster)"""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb, bench
cfg = sys.argv[1]; flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0; nq = int(sys.argv[3]) if len(sys.argv) > 3 else 3
info = bench.prepare(cfg, 1, 9); spec = bench.make_spec(bb, cfg, 1)
s = bb.GpuSolver(bb.Graph.generate(spec, device="cuda:0"))
for a, b, L, _, _ in info["queries"][:3]: s.query(a, b, flags=flags)
for a, b, L, e1, lv in info["queries"][3:3 + nq]:
    s.query(a, b, flags=flags)
    r = s.query(a, b, flags=flags | 1, timing="events")
    tr = s.ctx.trace()
    print(f"{cfg} flags={flags} {a}->{b} len={r.length} levels={r.levels} paired={r.levels_paired} grid={r.levels_grid} search_ms={r.search_ms:.4f} path_ms={r.path_ms:.4f} own_edges={r.edges_scanned} v1_edges={e1}")
    prev = None
    for i, t in enumerate(tr[:40]):
        dt = (t["t_ns"] - prev) / 1e3 if prev else 0.0
        prev = t["t_ns"]
        print(f"   L{i:<3d} side={t['side']} mode={t['mode']:<3d} frontier={t['frontier']:<8d} discovered={t['discovered']:<8d} edges={t['edges']:<9d} dt_us={dt:8.2f}")
