#!/usr/bin/env python3
"""bench.py — headline benchmark: time-to-shortest-path / GTEPS on the reference's own benchmark.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config er100k|rmat10m|grid16m|kron26|...]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      (the UNMODIFIED reference binaries from baseline/_ref)

A step is ONE s-t query (the reference answers one query per process, benchmark_test.sh:47-54).
Default config: the reference's largest published graph — Erdos-Renyi G(100000, 2.2/n), first
query 0 -> n-1 (graphs/make_graphs:8-13, benchmark_test.sh:7-8,43) — BASELINE.json config 2.
The other BASELINE.json shapes are measured too (`other_configs` in the JSON line, --others) or
alone with --config (rmat10m, grid16m, kron26).

Both arms get the graph, the query list and the GTEPS numerator from the SAME stand-alone helper
(baseline/tools/bench_prep.cpp, no CUDA and no libbibfs_b200.so): the counter-based generator of
csrc/gen.h, seeded; queries = 0 -> n-1, then seeded random connected pairs; numerator = directed
adjacency entries the reference's v1 schedule inspects for that query, so more
work is never rewarded and the GTEPS ratio of two implementations equals their time ratio.

Timing (ours): W warm-up queries, then K queries, each device-timed with CUDA events around the
search on the launching stream, max over ranks per step; L2 is flushed (256 MB write) between
steps when the graph fits in L2.  "e2e" repeats the K queries through the public API with the
query read from pinned host memory and the result + path copied back, wall clock per step with the
same L2 flush before every step: on one GPU through the resident solver (a front kernel that stays
on the GPU and polls a pinned mailbox; the launch-per-query path is reported beside it), on several
GPUs through DistributedSolver.query.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# published by the reference (other hardware): benchmark_results.csv:14-17, v4 MPI+CUDA on Quadro M1200 laptops + 1 GbE
PUBLISHED_V4_MS = {"er1k": 35.9098, "er10k": 99.5562, "er50k": 62.6357, "er100k": 108.312}
# the same reference binaries measured on the B200 box by the driver in round 1
# (BENCH_r01 / SCALE_r01 `reference` arm, er100k, 20 queries): v4 ms per query by rank count, v1 on the host CPU
SAME_BOX_V4_MS = {"er100k": {1: 29.66, 2: 35.78, 4: 46.05, 8: 54.56}}
SAME_BOX_V1_MS = {"er100k": {"v1_O3": 0.0155, "v1_as_shipped": 0.0364}}
L2_BYTES = 126 << 20
UNIT = "GTEPS"
UNIT_NOTE = "directed adjacency entries inspected by the reference's v1 schedule for the query / search time (the region the reference's own timers cover)"


# --------------------------------------------------------------------------------------------
# graph spec, query list and numerator: shared by both arms, computed without this repo's .so
# --------------------------------------------------------------------------------------------
def spec_string(config: str, seed: int) -> tuple[str, str]:
    """-> (bench_prep spec, description)."""
    presets = {"er1k": 1000, "er10k": 10000, "er50k": 50000, "er100k": 100000, "er1m": 1000000}
    if config in presets:
        n = presets[config]
        return f"er:{n}:{2.2000000001 / n!r}:{seed}", "ER G(n, 2.2/n) (graphs/make_graphs)"
    if config == "rmat10m":
        return f"rmat:24:8:10000000:{seed}", "R-MAT A=.57 B=.19 C=.19, 10M vertices, 80M edges"
    if config.startswith("rmat"):
        sc = int(config[4:])
        return f"rmat:{sc}:16:{1 << sc}:{seed}", f"R-MAT scale {sc} edgefactor 16"
    if config == "grid16m":
        return "grid:4096:4096", "2-D grid 4096x4096"
    if config.startswith("grid"):
        k = int(config[4:])
        return f"grid:{k}:{k}", f"2-D grid {k}x{k}"
    if config == "kron26":
        return f"rmat:26:16:{1 << 26}:{seed}", "Graph500 Kronecker scale 26 edgefactor 16"
    raise SystemExit(f"unknown config {config}")


def make_spec(bb, config: str, seed: int):
    """The same graph as spec_string(), as a generator spec of the library (our arm only)."""
    kind, *f = spec_string(config, seed)[0].split(":")
    if kind == "er":
        return bb.er_spec(int(f[0]), float(f[1]), seed=int(f[2]))
    if kind == "rmat":
        return bb.rmat_spec(int(f[0]), edgefactor=int(f[1]), n=int(f[2]), seed=int(f[3]))
    return bb.grid_spec(int(f[0]), int(f[1]))


def prep_tool() -> Path:
    exe = ROOT / "bin" / "bench_prep"
    src = ROOT / "baseline" / "tools" / "bench_prep.cpp"
    if not exe.exists() or exe.stat().st_mtime < src.stat().st_mtime:
        exe.parent.mkdir(exist_ok=True)
        subprocess.run(["g++", "-O3", "-std=c++17", "-pthread", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
    return exe


def candidate_pairs(n: int, count: int, seed: int):
    rng = np.random.default_rng(seed)
    out = [(0, n - 1)]                   # the reference's query (benchmark_test.sh:8,43)
    for _ in range(count):
        out.append((int(rng.integers(n)), int(rng.integers(n))))
    return out


def prepare(config: str, seed: int, need: int, bin_path: Path | None = None, host_limit: int = 40_000_000):
    """-> dict(n, records, nnz, queries=[(s, t, length, v1_edges, v1_levels)]) from the stand-alone helper.
    Graphs beyond `host_limit` vertices are not replayed on the host (queries = None)."""
    spec, desc = spec_string(config, seed)
    f = spec.split(":")
    n = {"er": lambda: int(f[1]), "rmat": lambda: int(f[3]), "grid": lambda: int(f[1]) * int(f[2])}[f[0]]()
    if n > host_limit:
        return {"n": n, "desc": desc, "queries": None}
    # the answer is a pure function of (spec, need): keep it for the other runs on this box (N = 1, 2, 4, 8 back to back)
    shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    cache = shm / f"bibfs_prep_{spec.replace(':', '_')}_{need}.json"
    if cache.exists() and (bin_path is None or bin_path.exists()):
        try:
            got = json.loads(cache.read_text())
            got["queries"] = [tuple(q) for q in got["queries"]]
            return got
        except (OSError, ValueError, KeyError):
            pass
    cand = candidate_pairs(n, 50 * need + 100, seed)
    cmd = [str(prep_tool()), spec, "--need", str(need)] + (["--bin", str(bin_path)] if bin_path else [])
    out = subprocess.run(cmd, input="".join(f"{s} {t}\n" for s, t in cand), capture_output=True, text=True, check=True).stdout.splitlines()
    _, n_s, rec, nnz = out[0].split()
    picked = []
    for line in out[1:]:
        s, t, length, edges, levels = (int(x) for x in line.split())
        if s != t and length >= 0 and (length >= 2 or (s, t) == (0, n - 1)):
            picked.append((s, t, length, edges, levels))
            if len(picked) == need:
                break
    if not picked:
        raise SystemExit("no connected query pairs found")
    while len(picked) < need:
        picked.append(picked[len(picked) % len(picked)])
    got = {"n": int(n_s), "desc": desc, "records": int(rec), "nnz": int(nnz), "queries": picked}
    try:
        tmp = cache.with_suffix(f".{os.getpid()}.tmp")
        tmp.write_text(json.dumps(got))
        tmp.replace(cache)
    except OSError:
        pass
    return got


def emit(obj):
    print(json.dumps(obj), flush=True)


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def timed_build(bb, spec, dev, world, rank):
    """Partitioned CSR + vertex records of this rank's shard on its GPU; seconds = max over ranks, barrier on both sides."""
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if world == 1:
        g = bb.Graph.generate(spec, device=dev)
        g.single().vertex_records()
    else:
        g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
        g.shards[rank].vertex_records()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return g, dt


def measure_ours(bb, comm, dist, torch, dev, rank, world, config, seed, K, W, do_e2e=True):
    """Build `config` on the GPUs, run W + K queries, return the measurement dict (rank 0) or None."""
    info = [None]
    if rank == 0:
        info[0] = prepare(config, seed, K + W)
    if world > 1:
        dist.broadcast_object_list(info, src=0)
    info = info[0]
    spec = make_spec(bb, config, seed)
    n = int(spec.n)
    g, build_s = timed_build(bb, spec, dev, world, rank)
    t0 = time.time()
    solver = bb.GpuSolver(g) if world == 1 else comm.DistributedSolver(g)
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    nnz = g.nnz
    if world > 1:
        t = torch.tensor([nnz], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        nnz = int(t.item())
    graph_bytes = nnz * 4 + (n + 1) * 8

    numerator = "v1-schedule edges (host replay by baseline/tools/bench_prep)"
    if info["queries"] is None:
        # too large for the host replay: connected pairs found by the GPUs themselves, own inspected edges as numerator
        numerator = "edges inspected by this implementation (graph too large for the host replay of the v1 schedule)"
        queries = []
        for s, t in candidate_pairs(n, 50 * (K + W) + 100, seed):
            if s == t:
                continue
            r = solver.query(s, t, want_path=False)
            if r.length is not None and r.length >= 2:
                queries.append((s, t, r.length, None, r.levels))
                if len(queries) == K + W:
                    break
    else:
        assert info["nnz"] == nnz, f"device graph has {nnz} stored entries, the helper's {info['nnz']}"
        queries = info["queries"]

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if graph_bytes < 2 * L2_BYTES else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def align():
        """Several ranks, before a timed step: an NCCL barrier (coarse: the ranks leave it tens of microseconds apart, and
        the max over ranks of a 40 us step would mostly measure that), then a trivial query (s == t) through the framework's
        own protocol: every peer returns when rank 0's verdict reaches it, which aligns the ranks to a few microseconds.  The
        L2 flush comes after it (the same fill on every GPU)."""
        barrier()
        solver.query(0, 0, want_path=False)

    for s, t, length, _, _ in queries[:W]:
        if flush is not None:
            flush.fill_(1)
        r = solver.query(s, t, want_path=True, timing="events")
        if r.length != length:
            raise SystemExit(f"{config}: wrong length on {s}->{t}: {r.length}, the v1 schedule gives {length}")
    barrier()
    launches = 0
    wall0 = time.perf_counter()
    ms, ms_path, own_edges, lengths = np.zeros(K), np.zeros(K), np.zeros(K), []
    for i, (s, t, length, _, _) in enumerate(queries[W:W + K]):
        if world > 1:
            align()                                    # every rank starts the step together: a peer's events bracket its wait
                                                       # kernel, which would otherwise also count rank 0's host work (path copy)
        if flush is not None:
            flush.fill_(i & 0xFF)                      # evict the graph from L2 between steps
        l0 = bb.lib().bibfs_launch_count()
        r = solver.query(s, t, want_path=True, timing="events")      # CUDA events on the launching stream
        launches += bb.lib().bibfs_launch_count() - l0               # kernels of the timed queries only (not the alignment queries)
        ms[i] = r.search_ms            # same coverage as the reference's timers: the search loop only
        ms_path[i] = r.path_ms
        own_edges[i] = r.edges_scanned
        lengths.append(r.length)
        if r.length != length:
            raise SystemExit(f"{config}: wrong length on {s}->{t}: {r.length}, the v1 schedule gives {length}")
    barrier()
    wall_dev_loop = time.perf_counter() - wall0
    if world > 1:
        if os.environ.get("BIBFS_BENCH_VERBOSE"):
            print(f"[rank {rank}] {config}: own event-timed search ms mean {ms.mean():.4f} min {ms.min():.4f} max {ms.max():.4f}", file=sys.stderr, flush=True)
        tms = torch.from_numpy(ms).to(dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)     # max over ranks, per step
        ms = tms.cpu().numpy()

    e2e_s, d2h = None, 0
    e2e_one_launch_s, e2e_what = None, ""
    if do_e2e:
        pinned = torch.empty(2, dtype=torch.int32).pin_memory()
        pinned_np = pinned.numpy()            # same memory; numpy item stores cost ~0.1 us, torch ones ~10 us
        side = torch.cuda.Stream(device=dev)

        def e2e_loop(fn, count_bytes):
            """K queries through a public call; every step: L2 flush (outside the timed region, waited for on its own stream
            - never a device-wide synchronisation, a resident kernel would hold it up), then wall clock around
            [query into pinned memory -> call -> result + path back in host memory]."""
            nonlocal d2h
            total = 0.0
            for i, (s, t, length, _, _) in enumerate(queries[W:W + K]):
                if world > 1:
                    align()
                if flush is not None:
                    with torch.cuda.stream(side):
                        flush.fill_(i & 0xFF)
                    side.synchronize()
                t0 = time.perf_counter()
                r = fn(s, t)
                total += time.perf_counter() - t0
                assert r.length == length
                if count_bytes:
                    d2h += 64 + 4 * (length + 1)
            return total

        def one_launch(s, t):
            pinned_np[0], pinned_np[1] = s, t
            return solver.query(-1 & 0x7FFFFFFF, -1 & 0x7FFFFFFF, want_path=True, pinned_query=pinned, timing="device")

        barrier()
        if world == 1:
            # one GPU: the resident solver (a front kernel that stays on the GPU and is fed through a pinned mailbox);
            # the launch-per-query path is reported next to it
            e2e_one_launch_s = e2e_loop(one_launch, False)
            torch.cuda.synchronize()
            resident = bb.ResidentSolver(g, cluster=1, idle_ms=200)
            for s, t, _, _, _ in queries[:W]:
                resident.query(s, t)
            e2e_s = e2e_loop(lambda s, t: resident.query(s, t, want_path=True), True)
            resident_launches = resident.launches
            resident.close()
            e2e_what = ("ResidentSolver.query(): (src,dst) written into a pinned mailbox that a resident front kernel polls -> search + path walk -> "
                        "[result | path] written to pinned host memory and read by the caller; wall clock per step, L2 flushed before every step; "
                        f"kernel launches during the K steps: {resident_launches - 1}")
        else:
            e2e_s = e2e_loop(one_launch, True)
            e2e_what = ("solver.query(): (src,dst) in pinned host memory read by the kernel -> search + path walk in one launch -> [result | path] "
                        "written to pinned host memory and read by the caller; wall clock per step (max over ranks of the sums), L2 flushed before every step")
        barrier()
        for _ in range(3000):                 # keep the GPU in the same state long enough for nvidia-smi to sample it under load
            solver.query(queries[W][0], queries[W][1], want_path=False)      # (a fixed count: every rank must issue the same queries)
        if world > 1:
            te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            e2e_s = float(te.item())
    solver.close()
    del solver, g
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    num = own_edges if queries[0][3] is None else np.array([q[3] for q in queries[W:W + K]], dtype=np.float64)
    total_ms = float(ms.sum())
    return {
        "config": config, "desc": info["desc"], "n": n, "nnz": nnz, "numerator": numerator,
        "gteps": float(num.sum() / (total_ms * 1e-3) / 1e9) if total_ms > 0 else 0.0,
        "ms_per_step": total_ms / K, "ms": ms, "ms_path": ms_path, "lengths": lengths,
        "own_gteps": float(own_edges.sum() / (total_ms * 1e-3) / 1e9) if total_ms > 0 else 0.0,
        "num_sum": float(num.sum()), "e2e_s": e2e_s, "e2e_one_launch_s": e2e_one_launch_s, "e2e_what": e2e_what, "d2h": d2h, "launches": int(launches), "build_s": build_s, "setup_s": setup_s,
        "wall_dev_loop": wall_dev_loop, "l2": "256 MB write between steps" if flush is not None else "inputs larger than L2",
        "levels_mean": float(np.mean([q[4] for q in queries[W:W + K]])),
    }


def measure_batch(bb, comm, dist, torch, dev, rank, world, config, seed, nq=64):
    """BASELINE.json config 5: a batch of `nq` random src/dst queries on `config` (kron26), answered by the query pool
    (lanes of front kernels; on several GPUs rank 0's lanes over the partitioned graph + the group for what is left).
    Wall clock of the whole batch, barrier + synchronize on both sides, max over ranks; answers checked against the
    one-query-at-a-time path."""
    spec = make_spec(bb, config[:-8] if config.endswith("_batch64") else config, seed)
    n = int(spec.n)
    g, build_s = timed_build(bb, spec, dev, world, rank)
    t0 = time.time()
    if world == 1:
        single = bb.GpuSolver(g)
        pool = bb.BatchSolver(g, lanes=64)
        run = lambda a, b: pool.query_batch(a, b)
    else:
        single = comm.DistributedSolver(g)
        run = lambda a, b: single.query_batch(a, b, lanes=64)
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    rng = np.random.default_rng(seed + 64)
    src = rng.integers(0, n, nq); dst = rng.integers(0, n, nq)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run(src, dst)
    times = []
    for _ in range(3):
        barrier()
        t1 = time.perf_counter()
        lens = run(src, dst)
        barrier()
        times.append(time.perf_counter() - t1)
    t_batch = min(times)
    pool_ms = pool.last_ms if world == 1 else getattr(single, "last_batch_ms", 0.0)
    barrier()
    t1 = time.perf_counter()
    seq = [single.query(int(a), int(b), want_path=False).length for a, b in zip(src, dst)]
    barrier()
    t_seq = time.perf_counter() - t1
    if lens != seq:
        raise SystemExit(f"{config}: the pool and the one-at-a-time path disagree")
    if world > 1:
        tt = torch.tensor([t_batch, t_seq], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_batch, t_seq = float(tt[0].item()), float(tt[1].item())
    nnz = g.nnz
    if world > 1:
        t = torch.tensor([nnz], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        nnz = int(t.item())
    if world == 1:
        pool.close()
    single.close()
    del single, g
    torch.cuda.empty_cache()
    return {"model": f"{nq} random queries at once on {spec_string(config[:-8] if config.endswith('_batch64') else config, seed)[1]}",
            "n": n, "directed_edges": nnz, "queries": nq, "found": sum(x is not None for x in lens),
            "batch_ms_wall": t_batch * 1e3, "pool_ms_device": pool_ms, "one_at_a_time_ms_wall": t_seq * 1e3,
            "queries_per_s": nq / t_batch, "graph_build_s": build_s, "solver_setup_s": setup_s,
            "what": "wall clock of the whole batch through the public API (barrier + synchronize on both sides, max over ranks, best of 3); "
                    "pool_ms_device = CUDA events around the pooled launches on rank 0"}


def run_ours(args):
    import torch
    import bidirectional_bfs_b200 as bb
    from bidirectional_bfs_b200.parallel import comm
    from bidirectional_bfs_b200.utils.clocks import ClockSampler

    rank, world, local = comm.init_distributed()
    if args.gpus != world and rank == 0:
        print(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (use the serial variant via the CLI for CPU numbers)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import torch.distributed as dist

    K, W = args.steps, args.warmup
    sampler = ClockSampler(local, period_ms=20).start()      # clocks / throttle reasons during warm-up + timed regions
    m = measure_ours(bb, comm, dist, torch, dev, rank, world, args.config, args.seed, K, W)
    clocks = sampler.stop()

    # ---- the other BASELINE.json shapes, a few queries each (same method; not the headline value)
    others = {}
    for cfg in [c for c in args.others.split(",") if c and c != args.config]:
        t0 = time.time()
        if cfg.endswith("_batch64"):
            try:
                ob = measure_batch(bb, comm, dist, torch, dev, rank, world, cfg, args.seed)
                if rank == 0:
                    others[cfg] = ob
            except Exception as e:  # noqa: BLE001
                if rank == 0:
                    others[cfg] = {"error": str(e)[:200]}
            continue
        try:
            o = measure_ours(bb, comm, dist, torch, dev, rank, world, cfg, args.seed, args.other_steps, 3, do_e2e=False)
        except Exception as e:  # noqa: BLE001
            o = None
            if rank == 0:
                others[cfg] = {"error": str(e)[:200]}
        if o is not None:
            others[cfg] = {"model": o["desc"], "n": o["n"], "directed_edges": o["nnz"], "steps": args.other_steps, "warmup": 3,
                           "ms_per_step": o["ms_per_step"], "value": o["gteps"], "unit": UNIT, "numerator": o["numerator"],
                           "own_edges_gteps": o["own_gteps"], "levels_mean": o["levels_mean"],
                           "path_length": {"min": int(min(o["lengths"])), "max": int(max(o["lengths"]))},
                           "us_per_level": 1e3 * o["ms_per_step"] / max(1.0, o["levels_mean"]), "l2": o["l2"],
                           "graph_build_s": o["build_s"], "solver_setup_s": o["setup_s"], "wall_s": time.time() - t0}

    if rank == 0:
        ms_per_step = m["ms_per_step"]
        v4 = SAME_BOX_V4_MS.get(args.config, {}).get(world)
        v1 = SAME_BOX_V1_MS.get(args.config)
        e2e_gteps = m["num_sum"] / m["e2e_s"] / 1e9
        emit({
            "metric": "GTEPS", "value": m["gteps"], "unit": UNIT, "unit_note": UNIT_NOTE,
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong",
            "vs_baseline": (v4 / ms_per_step) if v4 else None,
            "vs_baseline_note": "BASELINE.md §4: the reference's v4 MPI+CUDA binary (sm_100 build) on the same B200 box with the same number of ranks, "
                                "ms per query / ours (same numerator on both sides: the GTEPS ratio is the time ratio). The driver's own ratio comes from --impl reference.",
            "vs_published_v4_laptop": (PUBLISHED_V4_MS[args.config] / ms_per_step) if args.config in PUBLISHED_V4_MS else None,
            "vs_reference_v1_serial_cpu": None if not v1 else {
                "v1_O3_ms": v1["v1_O3"], "v1_as_shipped_ms": v1["v1_as_shipped"], "ours_ms": ms_per_step,
                "ratio_vs_v1_O3": v1["v1_O3"] / ms_per_step, "ratio_vs_v1_as_shipped": v1["v1_as_shipped"] / ms_per_step,
                "note": "the reference's serial v1 on the host CPU (same box, BASELINE.md §4); a ratio below 1 means the CPU variant is faster on this latency-bound graph"},
            "dtype": "uint32 vertex ids / bitmaps (integer traversal, no floating point)", "data": "synthetic",
            "impl": "ours",
            "config": {"model": f"bidirectional BFS on {m['desc']}", "name": args.config, "n": m["n"], "directed_edges": m["nnz"],
                       "global_batch": 1, "seq_len": None, "queries": "0->n-1 first, then seeded random connected pairs",
                       "parallelism": f"1-D vertex partition over {world} GPU(s), one process per GPU",
                       "l2": m["l2"], "numerator": m["numerator"], "seed": args.seed,
                       **({"step_alignment": "before every timed step: NCCL barrier + synchronize, then a trivial query (s == t) whose verdict "
                                             "releases all ranks within a few microseconds, then the L2 flush"} if world > 1 else {})},
            "search_ms": {"mean": ms_per_step, "median": float(np.median(m["ms"])), "min": float(m["ms"].min()), "max": float(m["ms"].max())},
            "path_ms_mean": float(m["ms_path"].mean()),
            "path_length": {"min": int(min(m["lengths"])), "max": int(max(m["lengths"]))},
            "own_edges_gteps": m["own_gteps"],
            "clocks": {"sm_mhz": clocks["sm_mhz"], "sm_max_mhz": clocks["sm_max_mhz"], "reasons": clocks["reasons"], "samples": clocks["samples"]},
            "e2e": {"value": e2e_gteps, "unit": UNIT, "ms_per_step": m["e2e_s"] * 1e3 / K, "h2d_bytes_per_step": 8,
                    "d2h_bytes_per_step": m["d2h"] / K,
                    "what": m["e2e_what"],
                    "one_launch_per_query_ms_per_step": (m["e2e_one_launch_s"] * 1e3 / K) if m["e2e_one_launch_s"] is not None else None},
            "gpu_launches": m["launches"], "launches_per_step": m["launches"] / K,
            "graph_build_s": m["build_s"], "solver_setup_s": m["setup_s"], "wall_s_device_loop": m["wall_dev_loop"],
            "other_configs": others,
        })
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------
# reference arm: the unmodified reference binaries, one process launch per query (their API).
# Nothing of this repository's library is imported here: graph file, queries and numerator come
# from the stand-alone helper, the measured region is baseline/_ref/bin/v4 under the MPI stand-in.
# --------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return                                       # rank 0 drives `mpirun -n N` itself
    ref = ROOT / "baseline" / "_ref"
    exe = ref / "bin" / "v4"
    if not exe.exists():
        subprocess.run(["bash", str(ROOT / "baseline" / "build_ref.sh")], capture_output=True, text=True)
    if not exe.exists():
        emit({"impl": "reference", "unavailable": "baseline/_ref/bin/v4 missing and baseline/build_ref.sh could not build it (see baseline/_ref/build.log)"})
        return
    spec, desc = spec_string(args.config, args.seed)
    K, W = args.steps, args.warmup
    shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    path = shm / f"bibfs_bench_{args.config}_{args.seed}.bin"
    info = prepare(args.config, args.seed, K + W, bin_path=path, host_limit=40_000_000 if args.force else 2_500_000)
    if info["queries"] is None:
        emit({"impl": "reference", "unavailable": f"config {args.config}: the reference reloads and rebuilds the whole graph in every process (one query per process); not run by default (--force to try)"})
        return
    n, queries = info["n"], info["queries"]
    env = dict(os.environ, MPISHIM_GPU_PER_RANK="1", MPISHIM_SLOT_BYTES=str(max(64 << 20, 4 * n + 4096)))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    mpirun = str(ROOT / "baseline" / "mpi_shim" / "mpirun")
    pat = re.compile(r"\[Time\][^=]*=\s*([0-9.eE+-]+)\s*s")
    ms, wrong = [], 0

    def one(s, t, length):
        nonlocal wrong
        out = subprocess.run([mpirun, "-n", str(world), str(exe), str(path), str(s), str(t)], env=env, capture_output=True, text=True, timeout=1800)
        m = pat.search(out.stdout)
        if out.returncode != 0 or not m:
            raise RuntimeError(f"reference v4 failed (rc={out.returncode}): {out.stdout[-300:]} {out.stderr[-300:]}")
        lm = re.search(r"Shortest path length = (\d+)", out.stdout)
        if not lm or int(lm.group(1)) != length:
            wrong += 1
        return float(m.group(1)) * 1e3

    try:
        for s, t, length, _, _ in queries[:W]:
            one(s, t, length)
        t0 = time.perf_counter()
        for s, t, length, _, _ in queries[W:W + K]:
            ms.append(one(s, t, length))
        wall = time.perf_counter() - t0
    except Exception as e:  # noqa: BLE001
        emit({"impl": "reference", "unavailable": f"reference v4 did not run: {str(e)[:200]}"})
        return
    extra = {}
    for name, cmd in (("v1_as_shipped", [str(ref / "bin" / "v1_as_shipped")]), ("v1_O3", [str(ref / "bin" / "v1_O3")]),
                      ("v3_sm100", [str(ref / "bin" / "v3")])):
        if not Path(cmd[0]).exists():
            continue
        vals = []
        for s, t, _, _, _ in queries[W:W + K]:
            o = subprocess.run(cmd + [str(path), str(s), str(t)], capture_output=True, text=True, timeout=600).stdout
            m = re.search(r"took ([0-9.eE+-]+) seconds", o)
            if m:
                vals.append(float(m.group(1)) * 1e3)
            m = re.search(r"gpu bfs time = ([0-9.eE+-]+) ms", o)
            if m:
                vals.append(float(m.group(1)))
        if vals:
            extra[name + "_ms_mean"] = float(np.mean(vals))
    num = np.array([q[3] for q in queries[W:W + K]], dtype=np.float64)
    levels = float(np.mean([q[4] for q in queries[W:W + K]]))
    total_ms = float(np.sum(ms))
    # bytes the reference moves per query and rank, from its source: CSR upload once (v4/comp.cu:49-65), then per level
    # H2D frontier 4N (comp.cu:94) and D2H next + visited 8N (comp.cu:104-105)
    h2d = (n + 1) * 4 + info["nnz"] * 4 + levels * 4 * n
    d2h = levels * 8 * n
    emit({
        "metric": "GTEPS", "value": float(num.sum() / (total_ms * 1e-3) / 1e9), "unit": UNIT,
        "unit_note": UNIT_NOTE + "; here: the time the reference prints ([Time] MPI+CUDA BI-BFS, MPI_Wtime around its level loop, rank 0)",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": (PUBLISHED_V4_MS[args.config] / (total_ms / K)) if args.config in PUBLISHED_V4_MS else None,
        "dtype": "int32 flags per vertex (reference)", "data": "synthetic", "impl": "reference",
        "config": {"model": f"reference v4 MPI+CUDA (mpi_bibfs, sm_100 build) on {desc}", "name": args.config, "n": n,
                   "directed_edges": info["nnz"], "global_batch": 1, "seq_len": None,
                   "queries": "0->n-1 first, then seeded random connected pairs",
                   "parallelism": f"{world} MPI rank(s), one GPU per rank, graph replicated (baseline/mpi_shim/mpirun: shared-memory MPI stand-in)",
                   "l2": "one process per query: every query starts with a cold GPU", "numerator": "v1-schedule edges (host replay by baseline/tools/bench_prep)",
                   "seed": args.seed},
        "wrong_lengths": wrong, "wall_s": wall, "other_reference_variants": extra,
        "e2e": {"value": float(num.sum() / wall / 1e9), "unit": UNIT, "ms_per_step": wall * 1e3 / K,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "wall clock of the K process launches of the reference's public interface (graph load + build + upload + search + exit); bytes per rank computed from v4/comp.cu:49-106"},
        "gpu_launches": int(K * levels), "gpu_launches_note": "one expand_frontier_kernel per level and rank (v4/comp.cu:99), levels from the v1 schedule",
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", default="er100k")
    ap.add_argument("--others", default="rmat10m,grid16m", help="further BASELINE.json shapes reported as other_configs ('' = none)")
    ap.add_argument("--other-steps", type=int, default=6)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--force", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
