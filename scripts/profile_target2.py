"""Workload for ncu (bandwidth-side kernels): generator-fused CSR build, bit-parallel batch search, one forced bottom-up query."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
spec = bb.rmat_spec(scale, edgefactor=16, seed=1)
t0 = time.time(); g = bb.Graph.generate(spec, device="cuda:0"); torch.cuda.synchronize(); print("csr build s", round(time.time() - t0, 3), "nnz", g.nnz)
t0 = time.time(); g = bb.Graph.generate(spec, device="cuda:0"); torch.cuda.synchronize(); print("csr build (warm) s", round(time.time() - t0, 3))
rng = np.random.default_rng(3)
src = rng.integers(0, spec.n, 64); dst = rng.integers(0, spec.n, 64)
for engine in ("bitparallel", "concurrent"):
    b = bb.BatchSolver(g, engine=engine)
    for _ in range(2):
        lens = b.query_batch(src, dst)
    print("batch 64 queries", engine, "ms", round(b.last_ms, 3), "found", sum(x is not None for x in lens))
    b.close()
s = bb.GpuSolver(g)
t = [s.query(int(a), int(c), want_path=False).search_ms for a, c in zip(src[:16], dst[:16])]
print("16 single queries ms total", round(sum(t), 3))
r = s.query(int(src[0]), int(dst[0]), flags=2 | 8, want_path=False)
print("forced bottom-up: len", r.length, "levels", r.levels, "edges", r.edges_scanned, "ms", round(r.search_ms, 3))
