"""Per-level roofline report (run under gpurun): for each traced level, algorithmic bytes vs measured HBM copy
bandwidth (MEASURED_PEAKS.json), i.e. the fraction BASELINE.md asks for.  Bytes per level (single rank):
  frontier vertex: 4 (queue) + 16 (two 8-byte row offsets);  inspected adjacency entry: 4 (col) + 8 (visited word RMW);
  discovered vertex: 4 (queue) + 4 (parent).   python scripts/roofline_report.py <config> [flags]"""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench
cfg = sys.argv[1]; flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
try:
    hbm = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]; src = "measured"
except Exception:
    hbm = 6650.0; src = "fallback"
spec, desc = bench.make_spec(bb, cfg, 1), bench.spec_string(cfg, 1)[1]
g = bb.Graph.generate(spec, device="cuda:0")
s = bb.GpuSolver(g)
qs = [(a, b, L) for a, b, L, _, _ in bench.prepare(cfg, 1, 3)["queries"]]
print(f"# {cfg}: {desc}; HBM peak {hbm} GB/s ({src}); flags={flags}")
for a, b, L in qs:
    for _ in range(2): r = s.query(a, b, flags=1 | flags)
    tr = s.ctx.trace()
    tot_b = 0
    print(f"query {a}->{b} len={L} levels={r.levels} search_ms={r.search_ms:.4f}")
    prev = None; rows = []
    for i, t in enumerate(tr):
        dt = (t["t_ns"] - prev) * 1e-9 if prev else None
        prev = t["t_ns"]
        byt = t["frontier"] * 20 + t["edges"] * 12 + t["discovered"] * 8
        tot_b += byt
        if dt: rows.append((i, t["side"], t["mode"], t["frontier"], t["edges"], t["discovered"], dt * 1e6, byt / dt / 1e9, byt / dt / 1e9 / hbm))
    show = rows if len(rows) <= 12 else rows[:4] + sorted(rows[4:], key=lambda x: -x[8])[:6]
    for x in show:
        print("   L%-5d side=%d mode=%d frontier=%-8d edges=%-10d disc=%-8d dt=%8.2f us  %9.3f GB/s  %.5f of %s HBM peak" % (*x, src))
    print(f"   whole query: {tot_b / (r.search_ms * 1e-3) / 1e9:.3f} GB/s = {tot_b / (r.search_ms * 1e-3) / 1e9 / hbm:.5f} of peak; {r.edges_scanned / (r.search_ms * 1e-3) / 1e9:.4f} GTEPS (own edges)")
