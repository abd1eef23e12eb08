"""Benchmark driver writing ``benchmark_results.csv`` and ``benchmark_table.txt``.

Counterpart of the reference's benchmark_test.sh / single_machine_bench.sh: every variant x every
graph (1k/10k/50k/100k by default, query 0 -> N-1), one CSV row per cell and a boxed ASCII table.
Differences on purpose: times are all in SECONDS in the ``time_sec`` column (the
reference mixes ms for v3), several repetitions with the median reported instead of one cold run,
and extra columns (length, levels, per-phase ms, edges, GTEPS).  The first four columns are the
reference's exact header ``version,graph,time_sec,logfile`` (benchmark_test.sh:27), so tools that
read the reference CSV keep working; the table file shows those four columns like the original.
"""
from __future__ import annotations

import argparse
import csv
import os
import statistics
import time

HEADER = ["version", "graph", "time_sec", "logfile", "n_gpus", "length", "levels", "load_ms", "build_ms",
          "search_ms", "path_ms", "edges_scanned", "gteps"]


def boxed_table(rows: list[list[str]]) -> str:
    """Same layout as print_boxed_table in benchmark_test.sh:79-118."""
    w = [max(len(str(r[i])) for r in rows) for i in range(len(rows[0]))]
    border = "+" + "+".join("-" * (x + 2) for x in w) + "+"
    out = [border, "| " + " | ".join(str(c).ljust(w[i]) for i, c in enumerate(rows[0])) + " |", border]
    for r in rows[1:]:
        out.append("| " + " | ".join(str(c).ljust(w[i]) for i, c in enumerate(r)) + " |")
    out.append(border)
    return "\n".join(out)


def run(graph_paths, variants, repeats=5, ranks=4, outdir=".", device="cuda:0", log_dir=None, gpus=0):
    import torch
    import bidirectional_bfs_b200 as bb
    rows = []
    for variant in variants:
        if variant in ("gpu", "multi-gpu", "naive-gpu") and not torch.cuda.is_available():
            continue
        for path in graph_paths:
            t0 = time.perf_counter()
            n, edges = bb.io.read_bin(path)
            load_ms = (time.perf_counter() - t0) * 1e3
            t0 = time.perf_counter()
            if variant in ("serial", "cpu-ranks"):
                g = bb.Graph.from_edges(n, edges)
                solver = bb.make_solver(g, variant, **({"ranks": ranks} if variant == "cpu-ranks" else {}))
                ngpu = 0
            elif variant == "multi-gpu":
                from .models.solvers import multi_gpu_solver
                solver, ngpu = multi_gpu_solver(n, edges, ranks, gpus, device)      # real devices with --gpus N, else virtual ranks
            else:
                g = bb.Graph.from_edges(n, edges, device=device)
                solver = bb.make_solver(g, variant)
                ngpu = 1
            build_ms = (time.perf_counter() - t0) * 1e3
            kw = {"timing": "events"} if variant in ("gpu", "multi-gpu") else {}       # cudaEvent-timed like the reference's v3
            res = [solver.query(0, n - 1, **kw) for _ in range(repeats)]
            r = res[-1]
            sm = statistics.median(x.search_ms for x in res)
            pm = statistics.median(x.path_ms for x in res)
            logfile = ""
            if log_dir:
                os.makedirs(os.path.join(log_dir, variant), exist_ok=True)
                logfile = os.path.join(log_dir, variant, os.path.basename(path).replace(".bin", ".log"))
                with open(logfile, "w") as f:
                    f.write(f"Shortest path length = {r.length}\nsearch_ms = {sm}\npath = {r.path}\n")
            rows.append([variant, os.path.basename(path), f"{sm * 1e-3:.6g}", logfile, ngpu,
                         r.length if r.found else "", r.levels, f"{load_ms:.3f}", f"{build_ms:.3f}", f"{sm:.6f}", f"{pm:.6f}",
                         r.edges_scanned, f"{(r.edges_scanned / (sm * 1e-3) / 1e9) if sm > 0 else 0:.6f}"])
            solver.close()
    csv_path = os.path.join(outdir, "benchmark_results.csv")
    with open(csv_path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(HEADER)
        w.writerows(rows)
    table = boxed_table([HEADER[:4]] + [r[:4] for r in rows])
    with open(os.path.join(outdir, "benchmark_table.txt"), "w") as f:
        f.write(table + "\n")
    return rows, table


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="bibfs bench")
    ap.add_argument("--graphs-dir", default="graphs")
    ap.add_argument("--graphs", nargs="*", default=["1k.bin", "10k.bin", "50k.bin", "100k.bin"])   # benchmark_test.sh:7
    ap.add_argument("--variants", nargs="*", default=["serial", "cpu-ranks", "gpu", "multi-gpu"])
    ap.add_argument("--repeats", type=int, default=5)
    ap.add_argument("--ranks", type=int, default=4)                                                  # NP=4, benchmark_test.sh:9
    ap.add_argument("--gpus", type=int, default=0, help="multi-gpu variant: one rank per GPU on this many devices (default: --ranks virtual ranks on one GPU)")
    ap.add_argument("--out", default=".")
    ap.add_argument("--logs", default=None)
    a = ap.parse_args(argv)
    paths = [os.path.join(a.graphs_dir, g) for g in a.graphs]
    missing = [p for p in paths if not os.path.exists(p)]
    if missing:
        from .cli import cmd_make_graphs
        cmd_make_graphs(["--dir", a.graphs_dir])
    rows, table = run(paths, a.variants, a.repeats, a.ranks, a.out, log_dir=a.logs, gpus=a.gpus)
    print("\nBenchmark complete.\nCSV   ->", os.path.join(a.out, "benchmark_results.csv"))
    print("Table ->", os.path.join(a.out, "benchmark_table.txt"), "\n")
    print(table)
    return 0
