import sys, numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
from bidirectional_bfs_b200.parallel import comm
import torch.distributed as dist
rank, world, local = comm.init_distributed()
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
spec = bb.preset_spec("100k", seed=1)
g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
s = comm.DistributedSolver(g)
for _ in range(5): r = s.query(0, spec.n - 1, want_path=False)
ms = [s.query(0, spec.n - 1, want_path=False).search_ms for _ in range(50)]
r = s.query(0, spec.n - 1, flags=1)
tr = s.ctx.trace()
if rank == 0:
    print(f"world={world} len={r.length} levels={r.levels} search_ms min={min(ms):.4f} med={np.median(ms):.4f}")
    print("level dts us:", " ".join("%.1f" % ((tr[i]["t_ns"] - tr[i - 1]["t_ns"]) / 1e3) for i in range(1, len(tr))))
    print("expand part :", " ".join("%.1f" % ((tr[i]["t_expand_ns"] - tr[i - 1]["t_ns"]) / 1e3) for i in range(1, len(tr))))
    print("sync part   :", " ".join("%.1f" % ((tr[i]["t_ns"] - tr[i]["t_expand_ns"]) / 1e3) for i in range(1, len(tr))))
s.close(); dist.destroy_process_group()
