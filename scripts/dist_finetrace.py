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
    for i in range(min(lv.value, 16)):
        t = out[i].astype(np.int64); d2 = out[256 + i].astype(np.int64)
        print("L%-2d work %5d | sync->poll %5d | poll wait %5d (first word +%d, hit %d, spins %d) | poll->allsync %5d | ->tid0 %5d | ->end %5d | total %5d" % (
            i, t[6] - t[0], d2[0] - t[6], d2[2] - d2[0], d2[1] - d2[0], d2[4], d2[3], d2[5] - d2[2], d2[6] - d2[5], t[7] - d2[6], t[7] - t[0]))
s.close(); dist.destroy_process_group()
