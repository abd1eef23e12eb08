import os, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench
info = bench.prepare("er100k", 1, 25)
spec = bench.make_spec(bb, "er100k", 1)
g = bb.Graph.generate(spec, device="cuda:0")
qs = info["queries"]
for front in ("1", "0"):
    os.environ["BIBFS_FRONT"] = front
    s = bb.GpuSolver(g)
    for a, b, L, _, _ in qs[:5]: s.query(a, b)
    for rep in range(2):
        dv = []
        for a, b, L, _, _ in qs[5:15]:
            r = s.query(a, b, timing="device"); dv.append(r.search_ms * 1e3)
        print("front", front, "hot device us per query:", " ".join(f"{x:5.1f}" for x in dv), flush=True)
    ev = [s.query(a, b, timing="events").search_ms * 1e3 for a, b, L, _, _ in qs[5:15]]
    print("front", front, "hot events us per query:", " ".join(f"{x:5.1f}" for x in ev), flush=True)
    s.close()
