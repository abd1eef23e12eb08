import ctypes as C, sys, numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
from bidirectional_bfs_b200.parallel import comm
import torch.distributed as dist
rank, world, local = comm.init_distributed()
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
L = bb.lib()
spec = bb.preset_spec("100k", seed=1)
g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
s = comm.DistributedSolver(g)
SRC, DST = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (0, spec.n - 1)
for _ in range(3): s.query(SRC, DST, want_path=False)
out = np.zeros((512, 8), dtype=np.uint64); lv = C.c_uint32(0)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
s.query(SRC, DST, want_path=False)
L.bibfs_debug_read(out.ctypes.data, C.byref(lv))
import time; time.sleep(0.5 * rank)
if True:
    print('--- rank', rank)
    names = ["rowptr", "reduce", "col", "claims", "enqueue", "syncthr", "levelend+xchg"]
    for i in range(min(lv.value, 16)):
        t = out[i].astype(np.int64)
        print("L%-2d" % i, " ".join(f"{names[k]}={t[k+1]-t[k] if abs(t[k+1]-t[k]) < 10**7 else -1}" for k in range(7)), "| total", t[7]-t[0])
s.close(); dist.destroy_process_group()
