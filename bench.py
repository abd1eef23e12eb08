#!/usr/bin/env python3
"""bench.py — headline benchmark: time-to-shortest-path / GTEPS on the reference's own benchmark.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config er100k|rmat10m|grid16m|kron26|...]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      (the UNMODIFIED reference binaries from baseline/_ref)

A step is ONE s-t query (the reference answers one query per process, benchmark_test.sh:47-54).
Default config: the reference's largest published graph — Erdos-Renyi G(100000, 2.2/n), first
query 0 -> n-1 (graphs/make_graphs:8-13, benchmark_test.sh:7-8,43) — BASELINE.json config 2.
The other BASELINE.json shapes are available with --config (rmat10m, grid16m, kron26).

Timing: W warm-up queries, then K queries, each device-timed with CUDA events around the search
and path kernels of that rank, max over ranks per step; L2 is flushed between steps when the
graph fits in L2.  GTEPS uses the COMMON numerator of BASELINE.md: directed adjacency entries the
v1 schedule inspects for that query (replayed by the serial variant, policy 0), so more work is
never rewarded.  "e2e" repeats the K queries through the public API with the query read from
pinned host memory and the result + path copied back, wall clock.
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

PUBLISHED_V4_MS = {"er1k": 35.9098, "er10k": 99.5562, "er50k": 62.6357, "er100k": 108.312}   # benchmark_results.csv:14-17
L2_BYTES = 126 << 20


def make_spec(bb, config: str, seed: int):
    if config in ("er1k", "er10k", "er50k", "er100k", "er1m"):
        return bb.preset_spec(config[2:], seed=seed), "ER G(n, 2.2/n) (graphs/make_graphs)"
    if config == "rmat10m":
        return bb.rmat_spec(24, edgefactor=8, n=10_000_000, seed=seed), "R-MAT A=.57 B=.19 C=.19, 10M vertices, 80M edges"
    if config.startswith("rmat"):
        sc = int(config[4:])
        return bb.rmat_spec(sc, edgefactor=16, seed=seed), f"R-MAT scale {sc} edgefactor 16"
    if config == "grid16m":
        return bb.grid_spec(4096, 4096), "2-D grid 4096x4096"
    if config.startswith("grid"):
        k = int(config[4:])
        return bb.grid_spec(k, k), f"2-D grid {k}x{k}"
    if config == "kron26":
        return bb.rmat_spec(26, edgefactor=16, seed=seed), "Graph500 Kronecker scale 26 edgefactor 16"
    raise SystemExit(f"unknown config {config}")


def candidate_pairs(n: int, count: int, seed: int):
    rng = np.random.default_rng(seed)
    yield 0, n - 1                       # the reference's query (benchmark_test.sh:8,43)
    for _ in range(count):
        yield int(rng.integers(n)), int(rng.integers(n))


def pick_queries(query_fn, n, need, seed):
    """First `need` candidate pairs that are connected by a path of >= 2 hops (same on every rank)."""
    out = []
    for s, t in candidate_pairs(n, 50 * need + 100, seed):
        if s == t:
            continue
        length = query_fn(s, t)
        if length is not None and (length >= 2 or (s, t) == (0, n - 1)):
            out.append((s, t, length))
            if len(out) == need:
                break
    if not out:
        raise SystemExit("no connected query pairs found")
    while len(out) < need:
        out.append(out[len(out) % max(1, len(out))])
    return out


def emit(obj):
    print(json.dumps(obj), flush=True)


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import bidirectional_bfs_b200 as bb
    from bidirectional_bfs_b200.parallel import comm
    from bidirectional_bfs_b200.utils.clocks import ClockSampler

    rank, world, local = comm.init_distributed()
    if args.gpus != world:
        if rank == 0:
            print(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (use the serial variant via the CLI for CPU numbers)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import torch.distributed as dist

    spec, desc = make_spec(bb, args.config, args.seed)
    n = int(spec.n)
    t0 = time.time()
    if world == 1:
        g = bb.Graph.generate(spec, device=dev)
        solver = bb.GpuSolver(g)
    else:
        g = bb.Graph.generate(spec, device=dev, nranks=world, ranks=[rank])
        solver = comm.DistributedSolver(g)
    torch.cuda.synchronize()
    build_s = time.time() - t0
    nnz_local = g.nnz
    nnz = nnz_local
    if world > 1:
        t = torch.tensor([nnz_local], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        nnz = int(t.item())
    graph_bytes = nnz * 4 + (n + 1) * 8

    K, W = args.steps, args.warmup
    queries = pick_queries(lambda s, t: solver.query(s, t, want_path=False).length, n, K + W, args.seed)

    # common GTEPS numerator: the v1 schedule's inspected adjacency entries (serial replay on the CPU)
    numerator = "v1-schedule edges (serial replay)"
    v1_edges = None
    if rank == 0 and n <= 2_500_000:
        g_cpu = g.to("cpu") if world == 1 else bb.Graph.generate(spec)
        ser = bb.SerialSolver(g_cpu, policy=0)
        v1_edges = []
        for s, t, length in queries:
            r = ser.query(s, t, want_path=False)
            assert r.length == length, f"length mismatch vs serial oracle on {s}->{t}: {length} vs {r.length}"
            v1_edges.append(r.edges_scanned)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if graph_bytes < 2 * L2_BYTES else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed K steps -------------------------------------------------------------
    sampler = ClockSampler(local, period_ms=20).start()      # clocks / throttle reasons during warm-up + timed regions
    for s, t, _ in queries[:W]:
        if flush is not None:
            flush.fill_(1)
        solver.query(s, t, want_path=True, timing="events")
    barrier()
    launches0 = bb.lib().bibfs_launch_count()
    wall0 = time.perf_counter()
    ms = np.zeros(K)
    ms_path = np.zeros(K)
    own_edges = np.zeros(K)
    lengths = []
    for i, (s, t, length) in enumerate(queries[W:W + K]):
        if flush is not None:
            flush.fill_(i & 0xFF)                      # evict the graph from L2 between steps
        r = solver.query(s, t, want_path=True, timing="events")      # CUDA events on the launching stream
        ms[i] = r.search_ms            # same coverage as the reference's timers: the search loop only
        ms_path[i] = r.path_ms
        own_edges[i] = r.edges_scanned
        lengths.append(r.length)
        if r.length != length:
            raise SystemExit(f"non-deterministic answer on {s}->{t}: {r.length} vs {length}")
    barrier()
    wall_dev_loop = time.perf_counter() - wall0
    launches = bb.lib().bibfs_launch_count() - launches0
    if world > 1:
        tms = torch.from_numpy(ms).to(dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)     # max over ranks, per step
        ms = tms.cpu().numpy()

    # ---- end to end through the public API -----------------------------------------------
    pinned = torch.empty(2, dtype=torch.int32).pin_memory()
    pinned_np = pinned.numpy()            # same memory; numpy item stores cost ~0.1 us, torch ones ~10 us
    barrier()
    e0 = time.perf_counter()
    d2h = 0
    for i, (s, t, length) in enumerate(queries[W:W + K]):
        pinned_np[0], pinned_np[1] = s, t
        r = solver.query(-1 & 0x7FFFFFFF, -1 & 0x7FFFFFFF, want_path=True, pinned_query=pinned)
        assert r.length == length
        d2h += 64 + 4 * (length + 1)
    barrier()
    e2e_s = time.perf_counter() - e0
    if e2e_s < 0.3:                       # keep the GPU in the same state long enough for nvidia-smi to sample it
        t_end = time.perf_counter() + 0.3
        while time.perf_counter() < t_end:
            s, t, _ = queries[W]
            solver.query(s, t, want_path=False)
    clocks = sampler.stop()
    if world > 1:
        te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_s = float(te.item())

    # ---- extra (N > 1, graph fits one GPU): the same K queries dealt over N REPLICAS of the graph, one per GPU.
    # Not the headline (that is one query answered by all ranks on the partitioned graph, the reference's
    # multi-rank semantics); reported so that the capacity and the throughput use of N GPUs can be told apart.
    replicated = None
    if world > 1 and graph_bytes < (8 << 30):
        g_full = bb.Graph.generate(spec, device=dev)
        rsolver = bb.GpuSolver(g_full)
        mine = queries[W:W + K][rank::world]
        for s, t, _ in mine[:2]:
            rsolver.query(s, t, want_path=True)
        barrier()
        r0 = time.perf_counter()
        for s, t, length in mine:
            assert rsolver.query(s, t, want_path=True).length == length
        barrier()
        rep_s = time.perf_counter() - r0
        tr = torch.tensor([rep_s], dtype=torch.float64, device=dev)
        dist.all_reduce(tr, op=dist.ReduceOp.MAX)
        replicated = {"queries_per_s": K / float(tr.item()), "ms_per_query_wall": float(tr.item()) * 1e3 / K,
                      "what": f"the K timed queries dealt round-robin over {world} GPUs, each holding the whole graph; wall clock incl. path, max over ranks"}
        rsolver.close()

    if rank == 0:
        if v1_edges is None:
            numerator = "edges inspected by this implementation (graph too large for the CPU replay)"
            num = own_edges
        else:
            num = np.array(v1_edges[W:W + K], dtype=np.float64)
        total_ms = float(ms.sum())
        gteps = float(num.sum() / (total_ms * 1e-3) / 1e9) if total_ms > 0 else 0.0
        e2e_gteps = float(num.sum() / e2e_s / 1e9)
        ms_per_step = total_ms / K
        vsb = PUBLISHED_V4_MS[args.config] / ms_per_step if args.config in PUBLISHED_V4_MS else None
        emit({
            "metric": "GTEPS", "value": gteps, "unit": "GTEPS (v1-schedule directed edges / CUDA-event-timed search time, the region the reference's timers cover)",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": vsb,
            "vs_baseline_note": "published v4 MPI+CUDA time on this graph size (benchmark_results.csv:14-17, Quadro M1200 laptops) divided by ms_per_step; GTEPS was not published, same numerator on both sides makes the GTEPS ratio equal the time ratio",
            "dtype": "uint32 vertex ids / bitmaps (integer traversal, no floating point)", "data": "synthetic",
            "impl": "ours",
            "config": {"model": f"bidirectional BFS on {desc}", "name": args.config, "n": n, "directed_edges": nnz,
                       "global_batch": 1, "seq_len": None, "queries": "0->n-1 first, then seeded random connected pairs",
                       "parallelism": f"1-D vertex partition over {world} GPU(s), one process per GPU",
                       "l2": "256 MB write between steps" if flush is not None else "inputs larger than L2",
                       "numerator": numerator, "seed": args.seed},
            "search_ms": {"mean": ms_per_step, "median": float(np.median(ms)), "min": float(ms.min()), "max": float(ms.max())},
            "path_kernel_ms_mean": float(ms_path.mean()),
            "path_length": {"min": int(min(lengths)), "max": int(max(lengths))},
            "own_edges_gteps": float(own_edges.sum() / (total_ms * 1e-3) / 1e9) if total_ms > 0 else 0.0,
            "clocks": {"sm_mhz": clocks["sm_mhz"], "sm_max_mhz": clocks["sm_max_mhz"], "reasons": clocks["reasons"], "samples": clocks["samples"]},
            "e2e": {"value": e2e_gteps, "unit": "GTEPS", "ms_per_step": e2e_s * 1e3 / K, "h2d_bytes_per_step": 8,
                    "d2h_bytes_per_step": d2h / K, "what": "solver.query(): (src,dst) in pinned host memory read by the kernel -> search + path walk in one launch -> [result | path] written to pinned host memory and read by the caller, wall clock"},
            "gpu_launches": int(launches), "launches_per_step": launches / K, "replicated_graph_throughput": replicated,
            "graph_build_s": build_s, "wall_s_device_loop": wall_dev_loop,
        })
    solver.close()
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------
# reference arm: the unmodified reference binaries, one process launch per query (their API)
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
    import bidirectional_bfs_b200 as bb
    import torch
    spec, desc = make_spec(bb, args.config, args.seed)
    n = int(spec.n)
    if n > 2_500_000 and not args.force:
        emit({"impl": "reference", "unavailable": f"config {args.config}: the reference reloads and rebuilds the whole graph in every process (one query per process); not run by default (--force to try)"})
        return
    # the same graph as our arm: same generator, same seed (host generator == device generator bit for bit)
    edges = bb.graph.edge_list(spec, "cuda:0" if torch.cuda.is_available() else "cpu")
    shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    path = shm / f"bibfs_bench_{args.config}_{args.seed}.bin"
    bb.io.write_bin(path, n, edges)
    g_cpu = bb.Graph.from_edges(n, edges)
    ser = bb.SerialSolver(g_cpu, policy=0)
    K, W = args.steps, args.warmup
    queries = pick_queries(lambda s, t: ser.query(s, t, want_path=False).length, n, K + W, args.seed)
    v1_edges = [ser.query(s, t, want_path=False).edges_scanned for s, t, _ in queries]
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
        for s, t, length in queries[:W]:
            one(s, t, length)
        t0 = time.perf_counter()
        for s, t, length in queries[W:W + K]:
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
        for s, t, _ in queries[W:W + min(K, 5)]:
            o = subprocess.run(cmd + [str(path), str(s), str(t)], capture_output=True, text=True, timeout=600).stdout
            m = re.search(r"took ([0-9.eE+-]+) seconds", o)
            if m:
                vals.append(float(m.group(1)) * 1e3)
            m = re.search(r"gpu bfs time = ([0-9.eE+-]+) ms", o)
            if m:
                vals.append(float(m.group(1)))
        if vals:
            extra[name + "_ms_mean"] = float(np.mean(vals))
    num = np.array(v1_edges[W:W + K], dtype=np.float64)
    total_ms = float(np.sum(ms))
    emit({
        "metric": "GTEPS", "value": float(num.sum() / (total_ms * 1e-3) / 1e9), "unit": "GTEPS (v1-schedule directed edges / time printed by the reference)",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": (PUBLISHED_V4_MS[args.config] / (total_ms / K)) if args.config in PUBLISHED_V4_MS else None,
        "dtype": "int32 flags per vertex (reference)", "data": "synthetic", "impl": "reference",
        "config": {"model": f"reference v4 MPI+CUDA (mpi_bibfs, sm_100 build) on {desc}", "name": args.config, "n": n,
                   "ranks": world, "launcher": "baseline/mpi_shim/mpirun (shared-memory MPI stand-in, one GPU per rank)",
                   "timer": "the reference's own '[Time] MPI+CUDA BI-BFS' line (MPI_Wtime around its level loop, rank 0)",
                   "numerator": "v1-schedule edges (serial replay)", "seed": args.seed},
        "wrong_lengths": wrong, "wall_s": wall, "other_reference_variants": extra,
        "e2e": {"value": float(num.sum() / wall / 1e9), "unit": "GTEPS", "what": "wall clock of the K process launches (graph load + build + search + exit)"},
        "gpu_launches": None,
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", default="er100k")
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
