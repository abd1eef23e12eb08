"""End-to-end latency of one query (wall clock around the public call, query in, result + path out): GpuSolver (one launch
per query, pinned query, result polled in pinned memory) vs ResidentSolver (mailbox).  python scripts/resident_probe.py [config]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb, bench
cfg = sys.argv[1] if len(sys.argv) > 1 else "er100k"
info = bench.prepare(cfg, 1, 25)
spec = bench.make_spec(bb, cfg, 1)
g = bb.Graph.generate(spec, device="cuda:0")
qs = [(s, t, L) for s, t, L, _, _ in info["queries"]]
side = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")

def run(name, fn, cold):
    ts = []
    for rep in range(3):
        for s, t, L in qs:
            if cold:
                with torch.cuda.stream(side):
                    flush.fill_(rep)
                side.synchronize()
            t0 = time.perf_counter()
            r = fn(s, t)
            ts.append(time.perf_counter() - t0)
            assert r.length == L, (name, s, t, r.length, L)
    ts = np.array(ts[len(qs):]) * 1e6
    print(f"{name:28s} {'cold L2' if cold else 'warm   '}: mean {ts.mean():6.2f} us  median {np.median(ts):6.2f}  min {ts.min():6.2f}  max {ts.max():6.2f}", flush=True)

one = bb.GpuSolver(g)
pinned = torch.empty(2, dtype=torch.int32).pin_memory()
def q_one(s, t):
    pinned[0] = s; pinned[1] = t
    return one.query(s, t, want_path=True, pinned_query=pinned)
for cold in (False, True):
    run("GpuSolver (launch per query)", q_one, cold)
one.close()
for cl in (1, 4, 16):
    res = bb.ResidentSolver(g, cluster=cl, idle_ms=100)
    for cold in (False, True):
        run(f"ResidentSolver cluster={cl}", lambda s, t: res.query(s, t), cold)
    print("   launches", res.launches, "device search us (last)", round(res.query(*qs[0][:2]).search_ms * 1e3, 2))
    res.close()
