"""Where the microseconds of a small query go (run under gpurun).  For each flow of the single-GPU solver —
front kernel (BIBFS_FRONT=1), general kernel with the pair tier inside it (BIBFS_FRONT=0), general kernel without the
pair tier — the K bench queries are timed three ways, cold L2 (256 MB write between queries) and L2-hot:
   events   cudaEvent pair around the launch (what bench.py reports)
   device   %globaltimer inside the kernel, query start -> verdict (no launch, no teardown)
   wall     host wall clock of solver.query() with the device-clock path
python scripts/latency_probe.py [config] [K]"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench

cfg = sys.argv[1] if len(sys.argv) > 1 else "er100k"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
info = bench.prepare(cfg, 1, K + 5)
spec = bench.make_spec(bb, cfg, 1)
g = bb.Graph.generate(spec, device="cuda:0")
qs = info["queries"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
print(f"# {cfg}: n={spec.n} nnz={g.nnz}; {K} queries, lengths {min(q[2] for q in qs)}..{max(q[2] for q in qs)}, v1 levels mean {np.mean([q[4] for q in qs]):.1f}")
for name, front, pair in (("front kernel", "1", True), ("general kernel + pair tier", "0", True), ("general kernel", "0", False)):
    os.environ["BIBFS_FRONT"] = front
    s = bb.GpuSolver(g, pair_tier=pair)
    for a, b, L, _, _ in qs[:5]:
        assert s.query(a, b).length == L
    row = {}
    for cold in (True, False):
        ev, dv, wl, lv, lp = [], [], [], [], []
        for a, b, L, _, _ in qs[5:5 + K]:
            if cold: flush.fill_(1)
            r = s.query(a, b, timing="events"); ev.append(r.search_ms + r.path_ms); lv.append(r.levels); lp.append(r.levels_paired)
            if cold: flush.fill_(2)
            torch.cuda.synchronize()
            t0 = time.perf_counter(); r = s.query(a, b, timing="device"); wl.append((time.perf_counter() - t0) * 1e3); dv.append(r.search_ms)
            assert r.length == L
        row[cold] = (np.mean(ev), np.mean(dv), np.mean(wl), np.mean(lv), np.mean(lp))
    c, h = row[True], row[False]
    print(f"{name:28s} | cold: events(+path) {c[0]*1e3:7.2f} us  device {c[1]*1e3:7.2f} us  wall {c[2]*1e3:7.2f} us | hot: events {h[0]*1e3:7.2f}  device {h[1]*1e3:7.2f}  wall {h[2]*1e3:7.2f} | levels {c[3]:.1f} (paired {c[4]:.1f})")
    s.close()
