"""Batch of 64 random queries on one GPU: the pool of front-kernel lanes vs one query after the other.
python scripts/batch_probe.py [config]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb, bench
cfg = sys.argv[1] if len(sys.argv) > 1 else "rmat20"
spec = bench.make_spec(bb, cfg, 1)
t0 = time.time(); g = bb.Graph.generate(spec, device="cuda:0"); torch.cuda.synchronize(); print(cfg, "n", spec.n, "nnz", g.nnz, "build s %.2f" % (time.time() - t0))
rng = np.random.default_rng(3)
src = rng.integers(0, spec.n, 64); dst = rng.integers(0, spec.n, 64)
s = bb.GpuSolver(g)
for a, c in zip(src[:4], dst[:4]): s.query(int(a), int(c), want_path=False)
torch.cuda.synchronize(); t0 = time.perf_counter()
seq = [s.query(int(a), int(c), want_path=False).length for a, c in zip(src, dst)]
torch.cuda.synchronize(); t_seq = (time.perf_counter() - t0) * 1e3
import os
for cl in ("", "2", "4"):
    if cl: os.environ["BIBFS_POOL_CLUSTER"] = cl
    else: os.environ.pop("BIBFS_POOL_CLUSTER", None)
    for lanes in (24, 32, 37, 48, 64, 74):
        b = bb.BatchSolver(g, lanes=lanes)
        for _ in range(3): lens = b.query_batch(src, dst)
        assert lens == seq, "pool and sequential answers differ"
        t0 = time.perf_counter(); b.query_batch(src, dst); wall = (time.perf_counter() - t0) * 1e3
        print(f"  cluster={cl or 'auto'} lanes={b.lanes}: 64 queries in {b.last_ms:.3f} ms (device, one launch), {wall:.3f} ms wall, vs {t_seq:.3f} ms one after the other (wall); found {sum(x is not None for x in lens)}")
        b.close()
