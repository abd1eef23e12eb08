"""Stage times inside the paired rounds and around them (needs the -DBIBFS_FINE_TRACE build:
make EXTRA=-DBIBFS_FINE_TRACE BUILD=build_ft OUTDIR=build_ft/lib; BIBFS_B200_LIB=build_ft/lib/libbibfs_b200.so)."""
import ctypes as C, sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench
L = bb.lib()
cfg = sys.argv[1] if len(sys.argv) > 1 else "er100k"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1
info = bench.prepare(cfg, 1, 25)
spec = bench.make_spec(bb, cfg, 1)
g = bb.Graph.generate(spec, device="cuda:0", nranks=P)
s = bb.GpuSolver(g) if P == 1 else bb.MultiGpuSolver(g)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
out = np.zeros((512, 8), dtype=np.uint64); lv = C.c_uint32(0); kt = np.zeros(32, dtype=np.uint64)
for a, b, Lq, _, _ in info["queries"][:4]: s.query(a, b)
for cold in (True, False):
    for a, b, Lq, _, _ in info["queries"][5:9]:
        L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
        if cold: flush.fill_(1)
        torch.cuda.synchronize()
        r = s.query(a, b, timing="events")
        L.bibfs_debug_read(out.ctypes.data, C.byref(lv)); L.bibfs_debug_read_kt(kt.ctypes.data)
        k = kt.astype(np.int64)
        print(f"{'cold' if cold else 'hot '} {a}->{b} len={r.length} levels={r.levels} events={r.search_ms*1e3:.1f}us+path {r.path_ms*1e3:.1f} | entry->init {k[1]-k[0]} init {k[2]-k[1]} ->tier {k[3]-k[2]} tier {k[4]-k[3]} ->path {k[5]-k[4]} path {k[6]-k[5]} cleanup {k[7]-k[6]} | total {k[7]-k[0]} ns")
        names = ["rec+issue", "claims", "reserve", "append", "barrier", "derive"]
        for i in range(min(lv.value, 12)):
            t = out[i].astype(np.int64); w = int(out[i][7])
            print("   R%-2d sides=%d n0=%-4d n1=%-4d " % (i, w & 3, (w >> 8) & 0xFFFFFF, w >> 32), " ".join(f"{names[j]}={t[j+1]-t[j]}" for j in range(6)), "| round", t[6]-t[0], "| gap", (out[i+1][0].astype(np.int64) - t[6]) if i + 1 < lv.value else 0)
