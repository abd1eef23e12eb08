"""Short workload touching every kernel family once on small graphs, every answer checked against the serial CPU variant.
Meant for `compute-sanitizer --tool memcheck|racecheck python scripts/sanitize_target.py`; on the round-1 GPU pool the
sanitizer is closed by the operators ("runs under it have left GPUs needing a reset"), so it was only run plainly there."""
import sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
spec = bb.er_spec(4000, 3.0 / 4000, seed=3)
g = bb.Graph.generate(spec, device="cuda:0")                 # generator + CSR builder kernels
gc = g.to("cpu")
s = bb.GpuSolver(g)
rng = np.random.default_rng(1)
ok = 0
for _ in range(6):
    a, b = int(rng.integers(spec.n)), int(rng.integers(spec.n))
    r = s.query(a, b)
    assert r.length == bb.bibfs(gc, a, b, variant="serial").length
    ok += 1
for flags in (2, 2 | 8):                                      # forced grid mode, forced bottom-up
    r = s.query(0, spec.n - 1, flags=flags, timing="events")
    assert r.length == bb.bibfs(gc, 0, spec.n - 1, variant="serial").length
s.close()
gg = bb.Graph.generate(bb.grid_spec(48, 48), device="cuda:0")  # cluster tier (frontier > 192 impossible here: 48) + long path
sg = bb.GpuSolver(gg); assert sg.query(0, 48 * 48 - 1).length == 94; sg.close()
gr = bb.Graph.generate(bb.rmat_spec(12, edgefactor=16, seed=2), device="cuda:0")   # hubs -> TMA hub phase
sr = bb.GpuSolver(gr); grc = gr.to("cpu")
for _ in range(4):
    a, b = int(rng.integers(4096)), int(rng.integers(4096))
    assert sr.query(a, b).length == bb.bibfs(grc, a, b, variant="serial").length
sr.close()
gm = bb.Graph.generate(spec, device="cuda:0", nranks=2)        # two virtual ranks in one launch: the cross-rank protocol
m = bb.MultiGpuSolver(gm)
for _ in range(3):
    a, b = int(rng.integers(spec.n)), int(rng.integers(spec.n))
    assert m.query(a, b).length == bb.bibfs(gc, a, b, variant="serial").length
m.close()
srcs = rng.integers(spec.n, size=8); dsts = rng.integers(spec.n, size=8)
for eng in ("concurrent", "bitparallel"):
    lens = bb.bibfs_batch(g, srcs, dsts, engine=eng)
print("SANITIZE_TARGET_OK", ok)
