import os, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench
info = bench.prepare("er100k", 1, 25)
spec = bench.make_spec(bb, "er100k", 1)
qs = info["queries"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for P in (1, 2, 8):
    g = bb.Graph.generate(spec, device="cuda:0", nranks=P)
    s = bb.GpuSolver(g) if P == 1 else bb.MultiGpuSolver(g)
    for a, b, L, _, _ in qs[:5]: s.query(a, b)
    for cold in (False, True):
        ev = []
        for a, b, L, _, _ in qs[5:25]:
            if cold: flush.fill_(1)
            r = s.query(a, b, timing="events"); ev.append((r.search_ms + r.path_ms) * 1e3)
            assert r.length == L and r.levels == r.levels_paired
        print("P", P, "cold" if cold else "hot ", "events(+path) us mean %.1f" % np.mean(ev), " ".join(f"{x:5.1f}" for x in ev[:10]), flush=True)
    s.close()
