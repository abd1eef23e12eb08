import ctypes as C, sys
import numpy as np
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
L = bb.lib()
spec = bb.preset_spec("100k", seed=1)
g = bb.Graph.generate(spec, device="cuda:0")
s = bb.GpuSolver(g)
for _ in range(3): s.query(0, spec.n - 1, want_path=False)
out = np.zeros((512, 8), dtype=np.uint64); lv = C.c_uint32(0)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
s.query(0, spec.n - 1, want_path=False)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
print("levels", lv.value)
names = ["rowptr", "reduce", "col", "atom", "enqueue", "syncthr", "levelend"]
for i in range(min(lv.value, 16)):
    t = out[i].astype(np.int64)
    print(i, " ".join(f"{names[k]}={t[k+1]-t[k]}" for k in range(7)), "| total", t[7]-t[0], "| gap to next", (out[i+1][0].astype(np.int64)-t[7]) if i+1 < lv.value else 0)
