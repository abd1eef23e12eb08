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
names = ["fetch(smem)", "rowptr", "scan+setup", "col ld", "atom", "enqueue", "flush"]
for i in range(min(lv.value, 16)):
    t = out[i].astype(np.int64)
    t2 = out[256 + i].astype(np.int64)
    print("   sync: expand_end->sync_entry", t2[0]-out[i][7].astype(np.int64), "reduce", t2[1]-t2[0], "lane0 block", t2[2]-t2[1], "to loop end", t2[3]-t2[2], "-> next expand", (out[i+1][0].astype(np.int64)-t2[3]) if i+1 < lv.value else 0)
    print(i, " ".join(f"{names[k]}={t[k+1]-t[k]}" for k in range(7)), "| total", t[7]-t[0], "| gap to next", (out[i+1][0].astype(np.int64)-t[7]) if i+1 < lv.value else 0)
