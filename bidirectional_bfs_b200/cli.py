"""Command line: ``python -m bidirectional_bfs_b200 <graph> <src> <dst> [--variant ...]``.

Same argv shape and the same stdout lines as the reference binaries so that the
reference's log-scraping benchmark script (benchmark_test.sh:61-69) parses our output unchanged:

  serial     "Shortest path length = d" / "Path: ..." / "Serial bidirectional BFS took t seconds"   (v1/main-v1.cpp:84-101)
  cpu-ranks  "[Time] bidir-bitset BFS = t seconds"                                                  (v2/second_try.cpp:134-135)
  gpu        "DEBUG: gpu bfs time = t ms", "GPU BFS: PATH FOUND|NO PATH"  (+ length and path)       (v3/bibfs_cuda_only.cu:208,224)
  multi-gpu  "[Time] MPI+CUDA BI-BFS = t s"                                                         (v4/mpi_bas.cpp:137-138)

plus a "[phases]" line with load / build / search / path milliseconds (new).  Sub-commands:
``gen`` (graph + oracle JSON like graphs/generate_graph.py), ``make-graphs`` (the 1k/10k/50k/100k
presets of graphs/make_graphs), ``show`` (graphs/read_graph.py, v2/read_in.cpp), ``bench``
(benchmark_test.sh), ``viz`` (visual/m3.py).
"""
from __future__ import annotations

import argparse
import sys
import time

NBH = "‑"   # the reference prints a non-breaking hyphen in its [Time] lines


def _print_path(path):
    print("Path: " + " ".join(str(v) for v in path))


def cmd_query(argv) -> int:
    ap = argparse.ArgumentParser(prog="bibfs", description="bidirectional BFS shortest path")
    ap.add_argument("graph")
    ap.add_argument("src", type=int)
    ap.add_argument("dst", type=int)
    ap.add_argument("--variant", default=None, choices=["serial", "cpu-ranks", "gpu", "multi-gpu", "naive-gpu"])
    ap.add_argument("--ranks", type=int, default=4, help="cpu-ranks: worker threads; multi-gpu: ranks (virtual ranks on one GPU unless launched with torchrun)")
    ap.add_argument("--gpus", type=int, default=0, help="multi-gpu: one rank per GPU on this many devices (0/1: --ranks virtual ranks on --device)")
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--quiet-path", action="store_true")
    ap.add_argument("--trace", action="store_true", help="print one line per BFS level (gpu variants)")
    a = ap.parse_args(argv)
    import torch
    import bidirectional_bfs_b200 as bb
    variant = a.variant or ("gpu" if torch.cuda.is_available() else "serial")
    t0 = time.perf_counter()
    try:
        n, edges = bb.io.read_json_graph(a.graph) if a.graph.endswith(".json") else bb.io.read_bin(a.graph)
    except (bb.BibfsError, OSError, ValueError) as e:
        print(f"Cannot open {a.graph}: {e}", file=sys.stderr)
        return 1
    load_ms = (time.perf_counter() - t0) * 1e3
    if not (0 <= a.src < n and 0 <= a.dst < n):
        print(f"src/dst must be in [0,{n})", file=sys.stderr)
        return 1
    t0 = time.perf_counter()
    if variant in ("serial", "cpu-ranks"):
        g = bb.Graph.from_edges(n, edges)
        solver = bb.make_solver(g, variant, **({"ranks": a.ranks} if variant == "cpu-ranks" else {}))
    elif variant == "multi-gpu":
        from .models.solvers import multi_gpu_solver
        solver, _ = multi_gpu_solver(n, edges, a.ranks, a.gpus, a.device)
    else:
        g = bb.Graph.from_edges(n, edges, device=a.device)
        solver = bb.make_solver(g, variant)
    build_ms = (time.perf_counter() - t0) * 1e3
    kw = {}
    if a.trace and variant in ("gpu", "multi-gpu"):
        kw["flags"] = 1
    r = solver.query(a.src, a.dst, **kw)
    if variant == "serial":
        if r.found:
            print(f"Shortest path length = {r.length}")
            if not a.quiet_path:
                _print_path(r.path)
        else:
            print(f"No path found between {a.src} and {a.dst}")
        print(f"Serial bidirectional BFS took {r.search_ms * 1e-3:g} seconds")
    elif variant == "cpu-ranks":
        print(f"Shortest path length = {r.length}" if r.found else "No path found")
        print(f"[Time] bidir{NBH}bitset BFS = {r.search_ms * 1e-3:g} seconds")
        if r.found and not a.quiet_path:
            _print_path(r.path)
    elif variant in ("gpu", "naive-gpu"):
        print(f"DEBUG: gpu bfs time = {r.search_ms:f} ms")
        if r.found:
            print(f"Shortest path length = {r.length}")
            if r.path and not a.quiet_path:
                _print_path(r.path)
        print("GPU BFS: " + ("PATH FOUND" if r.found else "NO PATH"))
    else:
        print(f"Shortest path length = {r.length}" if r.found else "No path found")
        print(f"[Time] MPI+CUDA BI{NBH}BFS = {r.search_ms * 1e-3:g} s")
        if r.found and not a.quiet_path:
            _print_path(r.path)
    if a.trace and variant in ("gpu", "multi-gpu"):
        ctx = solver.ctx if variant == "gpu" else solver.ctxs[0]
        prev = None
        for i, t in enumerate(ctx.trace()):
            dt = (t["t_ns"] - prev) / 1e3 if prev else 0.0
            prev = t["t_ns"]
            print(f"[level {i}] side={'ST'[t['side']]} mode={t['mode']} frontier={t['frontier']} discovered={t['discovered']} edges={t['edges']} dt_us={dt:.2f}")
    print(f"[phases] load={load_ms:.3f} build={build_ms:.3f} search={r.search_ms:.6f} path={r.path_ms:.6f} ms  "
          f"levels={r.levels} edges_scanned={r.edges_scanned}")
    solver.close()
    return 0


def cmd_gen(argv) -> int:
    """Counterpart of graphs/generate_graph.py: graph -> .bin, shortest path -> <stem>.json."""
    ap = argparse.ArgumentParser(prog="bibfs gen")
    ap.add_argument("kind", choices=["er", "rmat", "grid"])
    ap.add_argument("--n", type=int, default=50000)
    ap.add_argument("--p", type=float, default=3 / 129000)         # generate_graph.py:15 defaults
    ap.add_argument("--scale", type=int, default=16)
    ap.add_argument("--edgefactor", type=int, default=16)
    ap.add_argument("--width", type=int, default=256)
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--src", type=int, default=0)
    ap.add_argument("--dst", type=int, default=None)
    ap.add_argument("--out", default="50k.bin")
    ap.add_argument("--device", default=None)
    ap.add_argument("--json-graph", action="store_true", help="also write <stem>.graph.json (networkx adjacency format)")
    a = ap.parse_args(argv)
    import torch
    import bidirectional_bfs_b200 as bb
    dev = a.device or ("cuda:0" if torch.cuda.is_available() else "cpu")
    if a.kind == "er":
        spec = bb.er_spec(a.n, a.p, a.seed)
    elif a.kind == "rmat":
        spec = bb.rmat_spec(a.scale, a.edgefactor, None if a.n == 50000 else a.n, a.seed)
    else:
        spec = bb.grid_spec(a.width, a.height)
    n = int(spec.n)
    print(f"Generating {a.kind} graph with {n} vertices (seed {a.seed}) on {dev}...")
    edges = bb.graph.edge_list(spec, dev)
    bb.io.write_bin(a.out, n, edges)
    dst = a.dst if a.dst is not None else n - 1
    print(f"Computing shortest path from {a.src} to {dst}...")
    r = bb.bibfs(bb.Graph.from_edges(n, edges), a.src, dst, variant="serial")
    print(f"  -> hop count = {r.length}" if r.found else "  -> no path found")
    oracle = bb.io.oracle_path_for(a.out)
    bb.io.write_oracle(oracle, a.src, dst, r.path if r.found else [])
    print("Wrote path info to", oracle)
    if a.json_graph:
        bb.io.write_json_graph(str(oracle).replace(".json", ".graph.json"), n, edges)
    return 0


def cmd_make_graphs(argv) -> int:
    """graphs/make_graphs: 1k/10k/50k/100k with p = 2.2000000001/N, src 0, dst N-1."""
    ap = argparse.ArgumentParser(prog="bibfs make-graphs")
    ap.add_argument("--dir", default="graphs")
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args(argv)
    import os
    os.makedirs(a.dir, exist_ok=True)
    for name, n in (("1k", 1000), ("10k", 10000), ("50k", 50000), ("100k", 100000)):
        print(f"=== Generating G({n}, p={2.2000000001 / n}) -> {name}.bin ===")
        cmd_gen(["er", "--n", str(n), "--p", repr(2.2000000001 / n), "--seed", str(a.seed), "--out", os.path.join(a.dir, f"{name}.bin")])
    print("All done!")
    return 0


def cmd_show(argv) -> int:
    """graphs/read_graph.py (CSR + 'u -> v' listing) and v2/read_in.cpp (dense dump)."""
    ap = argparse.ArgumentParser(prog="bibfs show")
    ap.add_argument("graph")
    ap.add_argument("--dense", action="store_true", help="print the adjacency matrix (N <= 64), or read a legacy dense file")
    ap.add_argument("--legacy", action="store_true", help="file is the old <u32 N> + N*N byte format")
    ap.add_argument("--limit", type=int, default=None)
    a = ap.parse_args(argv)
    import numpy as np
    import bidirectional_bfs_b200 as bb
    if a.legacy:
        m = bb.io.read_dense(a.graph)
        print(f"Adjacency matrix ({m.shape[0]}x{m.shape[0]}):")
        for row in m:
            print(" ".join(str(int(x)) for x in row))
        return 0
    g = bb.Graph.from_file(a.graph, sort_rows=True)
    A = g.to_scipy()
    print(f"Loaded CSR with shape {A.shape} and {A.nnz} nonzeros")
    if a.dense:
        if g.n > 64:
            print("--dense is for N <= 64", file=sys.stderr)
            return 1
        for row in np.asarray(A.todense()):
            print(" ".join(str(int(x)) for x in row))
        return 0
    rows, cols = A.nonzero()
    for k, (u, v) in enumerate(zip(rows, cols)):
        if a.limit is not None and k >= a.limit:
            break
        print(f"{u} -> {v}")
    return 0


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if argv and argv[0] == "gen":
        return cmd_gen(argv[1:])
    if argv and argv[0] == "make-graphs":
        return cmd_make_graphs(argv[1:])
    if argv and argv[0] == "show":
        return cmd_show(argv[1:])
    if argv and argv[0] == "bench":
        from .bench_harness import main as bench_main
        return bench_main(argv[1:])
    if argv and argv[0] == "viz":
        from .viz import main as viz_main
        return viz_main(argv[1:])
    if len(argv) < 3 or argv[0] in ("-h", "--help"):
        print("Usage: bibfs <graph_file> <src> <dst> [--variant serial|cpu-ranks|gpu|multi-gpu|naive-gpu] [--ranks R]\n"
              "       bibfs gen er|rmat|grid ... --out F | make-graphs | show F | bench ... | viz F", file=sys.stderr)
        return 1
    return cmd_query(argv)


if __name__ == "__main__":
    sys.exit(main())
