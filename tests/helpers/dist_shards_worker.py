"""Worker of tests/test_dist_cpu.py: one process per rank (gloo), each builds ONLY its shard of the partitioned CSR on
the CPU, then the ranks compare notes: the shards tile the vertex range, their sizes add up to the full graph's, the
serial search over the reassembled graph gives the same answers as over the directly built one."""
import os, sys
import numpy as np, torch
import torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import bidirectional_bfs_b200 as bb
from bidirectional_bfs_b200.parallel import comm

rank, world, _ = comm.init_distributed("gloo")
spec = bb.er_spec(6000, 3.0 / 6000, seed=11)
g = bb.Graph.generate(spec, nranks=world, ranks=[rank])            # this rank's shard only
assert list(g.shards) == [rank]
sh = g.shards[rank]
info = [None] * world
dist.all_gather_object(info, (rank, sh.lo, sh.hi, sh.nnz, sh.row_ptr.numpy(), sh.col.numpy()))
if rank == 0:
    info.sort()
    full = bb.Graph.generate(spec)                                  # single-shard reference
    assert info[0][1] == 0 and info[-1][2] >= spec.n
    for a, b in zip(info, info[1:]):
        assert a[2] == b[1], "shards must tile the vertex range"
    assert sum(x[3] for x in info) == full.nnz
    fs = full.single()
    rp, col = fs.row_ptr.numpy(), fs.col.numpy()
    for r, lo, hi, nnz, srp, scol in info:
        hi = min(hi, spec.n)
        assert np.array_equal(srp[: hi - lo + 1] - srp[0], rp[lo: hi + 1] - rp[lo])
        for v in range(lo, hi, 97):                                 # rows are the same sets of neighbours
            assert sorted(scol[srp[v - lo]: srp[v - lo + 1]].tolist()) == sorted(col[rp[v]: rp[v + 1]].tolist())
    print("DIST_SHARDS_OK", world)
dist.barrier()
dist.destroy_process_group()
