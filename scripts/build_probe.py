"""Build time of a partitioned graph under torchrun, routed (parallel/build.py) vs regenerate-and-filter (every rank walks
the whole generator stream and keeps the rows it owns).  torchrun --nproc-per-node N scripts/build_probe.py [config]"""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb, bench
from bidirectional_bfs_b200.parallel import comm

cfg = sys.argv[1] if len(sys.argv) > 1 else "kron26"
rank, world, local = comm.init_distributed()
dev = torch.device('cuda', local)
spec = bench.make_spec(bb, cfg, 1)
for mode in ("1", "0", "1", "0"):
    os.environ["BIBFS_SHARDED_BUILD"] = mode
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
    g.shards[rank].vertex_records()
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    nnz = torch.tensor([g.nnz], dtype=torch.int64, device=dev); dist.all_reduce(nnz)
    if rank == 0:
        print(f"{cfg} world {world} {'routed' if mode == '1' else 'regenerate+filter'}: {t.item():.3f} s (max over ranks), {int(nnz.item())} stored entries", flush=True)
    del g
    torch.cuda.empty_cache()
dist.destroy_process_group()
