import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
L = bb.lib()
spec = bb.grid_spec(256, 256)
g = bb.Graph.generate(spec, device="cuda:0")
s = bb.GpuSolver(g)
for _ in range(3): s.query(0, spec.n - 1, want_path=False, flags=2)
out = np.zeros((512, 8), dtype=np.uint64); lv = C.c_uint32(0)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
r = s.query(0, spec.n - 1, want_path=False, flags=2)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
print("levels", lv.value, "search_ms", r.search_ms, "per level us", r.search_ms * 1e3 / r.levels)
names = ["fetch", "rowptr", "scan", "col", "atom", "enq", "flush"]
for i in list(range(100, 108)):
    t = out[i].astype(np.int64); t2 = out[256 + i].astype(np.int64); nx = out[i + 1].astype(np.int64)
    print(i, " ".join(f"{names[k]}={t[k+1]-t[k]}" for k in range(7)), "| expand", t[7] - t[0],
          "| CTA1: arrived-at(rel leader expand start)", t2[4] - t[0], "released", t2[5] - t[0], "state", t2[6] - t[0], "| leader publish (approx)", t2[2] - t[0],
          "| expand_end->leader_sync", t2[0] - t[7], "reduce", t2[1] - t2[0], "lane0", t2[2] - t2[1], "| leader_sync_end->next expand", nx[0] - t2[2])
