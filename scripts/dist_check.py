"""torchrun correctness check of the one-process-per-GPU path against the serial oracle.
   python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/dist_check.py"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
from bidirectional_bfs_b200.parallel import comm
import torch.distributed as dist

rank, world, local = comm.init_distributed()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
ok = True
for name, spec in (("er", bb.er_spec(30000, 2.2 / 30000, seed=1)), ("rmat", bb.rmat_spec(15, edgefactor=16, seed=3)),
                   ("grid", bb.grid_spec(150, 100)), ("er_dense", bb.er_spec(5000, 10.0 / 5000, seed=2))):
    # two ways to a shard: routed (every rank walks 1/P of the generator and sends half-edges to their owners,
    # parallel/build.py) and replayed (every rank walks the whole stream and keeps its rows: the default up to 8 ranks)
    import os
    if name in ("er", "rmat"): os.environ["BIBFS_SHARDED_BUILD"] = "1"
    else: os.environ.pop("BIBFS_SHARDED_BUILD", None)
    g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
    g_host = bb.Graph.generate(spec)
    # this rank's shard against the host builder, row by row (the order inside a row is not fixed: compare sorted)
    sh = g.shards[rank]
    rp_h = g_host.single().row_ptr.numpy(); col_h = g_host.single().col.numpy().view(np.uint32)
    rp_d = sh.row_ptr.cpu().numpy(); col_d = sh.col.cpu().numpy().view(np.uint32)
    same = np.array_equal(rp_d - rp_d[0], rp_h[sh.lo:sh.hi + 1] - rp_h[sh.lo])
    if same:
        seg_h = col_h[rp_h[sh.lo]:rp_h[sh.hi]]
        key_h = np.repeat(np.arange(sh.hi - sh.lo, dtype=np.int64), np.diff(rp_h[sh.lo:sh.hi + 1])) << 32
        key_d = np.repeat(np.arange(sh.hi - sh.lo, dtype=np.int64), np.diff(rp_d)) << 32
        same = np.array_equal(np.sort(key_h | seg_h), np.sort(key_d | col_d))
    if not same:
        ok = False
        print(f"[rank {rank}] SHARD MISMATCH {name}: the device build differs from the host builder", flush=True)
    if name == "er":
        # the same graph from a .bin file: every rank reads and uploads only its M/P slice of the records
        import os, tempfile
        path = os.path.join(tempfile.gettempdir(), "bibfs_dist_check.bin")
        if rank == 0:
            bb.io.write_bin(path, spec.n, bb.graph.edge_list(spec))
        dist.barrier()
        gb = bb.Graph.from_bin(path, device=dev, nranks=world, ranks=[rank])
        rb = gb.shards[rank].row_ptr.cpu().numpy()
        if not np.array_equal(rb, rp_d) or gb.m_records > (g_host.nnz // 2) // world + 1:
            ok = False
            print(f"[rank {rank}] BIN SHARD MISMATCH (records read here: {gb.m_records})", flush=True)
    solver = comm.DistributedSolver(g)
    oracle = bb.SerialSolver(g_host, policy=1)
    rng = np.random.default_rng(5)
    pairs = [(0, spec.n - 1), (3, 3)] + [(int(rng.integers(spec.n)), int(rng.integers(spec.n))) for _ in range(25)]
    for flags in (0, 64, 2, 2 | 8):          # default flow (front kernel), fused levels from the start, forced grid, forced bottom-up
        t0 = time.time()
        for s, t in pairs:
            want = oracle.query(s, t, want_path=False).length
            r = solver.query(s, t, flags=flags)
            if r.length != want:
                ok = False
                print(f"[rank {rank}] MISMATCH {name} flags={flags} {s}->{t}: got {r.length} want {want}", flush=True)
            elif want is not None and rank == 0:
                bb.validate_path(oracle.graph, s, t, r.path, want)
        if rank == 0:
            print(f"{name} flags={flags}: {len(pairs)} queries ok={ok} {time.time()-t0:.2f}s last search_ms={r.search_ms:.4f}", flush=True)
    # back to back: query k+1 starts on the path of query k (rank 0 walks the parents inside its kernel; nobody may
    # overwrite them meanwhile)
    s, t = 0, spec.n - 1
    for _ in range(10):
        want = oracle.query(s, t, want_path=False).length
        r = solver.query(s, t)
        if r.length != want:
            ok = False
            print(f"[rank {rank}] MISMATCH {name} chained {s}->{t}: got {r.length} want {want}", flush=True)
        if want is None or want < 4:
            break
        p = [None]
        if rank == 0:
            bb.validate_path(oracle.graph, s, t, r.path, want)
            p[0] = (r.path[len(r.path) // 3], r.path[-1 - len(r.path) // 5])
        dist.broadcast_object_list(p, src=0)
        s, t = p[0]
    solver.close()
t = torch.tensor([0 if ok else 1], device=dev)
dist.all_reduce(t)
if rank == 0:
    print("DIST_CHECK", "PASS" if t.item() == 0 else "FAIL", "world", world, flush=True)
dist.destroy_process_group()
sys.exit(0 if t.item() == 0 else 1)
