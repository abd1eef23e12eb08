"""CLI stdout contract, generator/oracle tool, bench CSV/table format, show, viz."""
import csv
import json
import os
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
ENV = dict(os.environ, PYTHONPATH=str(ROOT))


def run_py(*args, cwd=None):
    return subprocess.run([sys.executable, "-m", "bidirectional_bfs_b200", *map(str, args)], capture_output=True, text=True, env=ENV, cwd=cwd)


@pytest.fixture(scope="module")
def graphs(tmp_path_factory):
    d = tmp_path_factory.mktemp("graphs")
    p = run_py("gen", "er", "--n", 1000, "--p", 2.2 / 1000, "--seed", 5, "--out", d / "1k.bin", "--device", "cpu", "--json-graph")
    assert p.returncode == 0, p.stderr
    return d


def test_gen_writes_bin_and_oracle(bb, graphs):
    n, edges = bb.io.read_bin(graphs / "1k.bin")
    assert n == 1000 and len(edges) > 900
    d = bb.io.read_oracle(graphs / "1k.json")                   # generate_graph.py:55-62 schema
    assert d["source"] == 0 and d["target"] == 999
    import networkx as nx
    G = nx.Graph(); G.add_nodes_from(range(n)); G.add_edges_from(edges.tolist())
    want = nx.shortest_path_length(G, 0, 999) if nx.has_path(G, 0, 999) else None
    assert d["hop_count"] == want
    n2, e2 = bb.io.read_json_graph(graphs / "1k.graph.json")
    assert n2 == n and len(e2) == len(edges)


def test_python_cli_stdout_contract(graphs):
    p = run_py(graphs / "1k.bin", 0, 999, "--variant", "serial")
    assert p.returncode == 0, p.stderr
    out = p.stdout.splitlines()
    if "No path found" in out[0]:
        assert out[0] == "No path found between 0 and 999"
    else:
        assert re.fullmatch(r"Shortest path length = \d+", out[0])
        assert out[1].startswith("Path: 0 ") and out[1].endswith(" 999")
    assert any(re.fullmatch(r"Serial bidirectional BFS took [0-9.e+-]+ seconds", l) for l in out)
    assert any(l.startswith("[phases] load=") for l in out)
    p = run_py(graphs / "1k.bin", 0, 999, "--variant", "cpu-ranks", "--ranks", 2)
    assert re.search(r"\[Time\] bidir‑bitset BFS = [0-9.e+-]+ seconds", p.stdout)
    # the reference's awk (benchmark_test.sh:61-69) takes the first numeric token of the first matching line
    line = next(l for l in p.stdout.splitlines() if l.startswith("[Time]"))
    assert any(re.fullmatch(r"[0-9]+(\.[0-9]*)?([eE][+-]?[0-9]+)?", tok) for tok in line.split())


def test_cli_errors(graphs):
    assert run_py(graphs / "1k.bin", 0).returncode == 1                      # wrong argc -> usage, rc 1 (v1:14-17)
    p = run_py(graphs / "missing.bin", 0, 1, "--variant", "serial")
    assert p.returncode == 1 and "Cannot open" in p.stderr
    assert run_py(graphs / "1k.bin", 0, 5000, "--variant", "serial").returncode == 1


def test_native_cli_matches_python(graphs):
    exe = ROOT / "bin" / "bibfs"
    if not exe.exists():
        pytest.skip("native CLI not built")
    a = subprocess.run([str(exe), str(graphs / "1k.bin"), "0", "999"], capture_output=True, text=True).stdout.splitlines()
    b = run_py(graphs / "1k.bin", 0, 999, "--variant", "serial").stdout.splitlines()
    assert a[0] == b[0]                                                       # same length (or same 'No path' line)
    if a[0].startswith("Shortest"):
        assert a[1] == b[1]                                                   # and the same path
    s = subprocess.run([str(exe), "show", str(graphs / "1k.bin"), "--limit", "2"], capture_output=True, text=True).stdout
    assert s.startswith("Loaded CSR with shape (1000, 1000)")
    assert subprocess.run([str(exe)], capture_output=True).returncode == 1


def test_show(bb, graphs, tmp_path):
    p = run_py("show", graphs / "1k.bin", "--limit", 5)
    lines = p.stdout.splitlines()
    assert lines[0].startswith("Loaded CSR with shape (1000, 1000) and ") and len(lines) == 6
    assert re.fullmatch(r"\d+ -> \d+", lines[1])                               # graphs/read_graph.py:29-30
    bb.io.write_bin(tmp_path / "tiny.bin", 4, np.array([[0, 1], [1, 2]], np.uint32))
    p = run_py("show", tmp_path / "tiny.bin", "--dense")
    assert p.stdout.splitlines()[1:] == ["0 1 0 0", "1 0 1 0", "0 1 0 0", "0 0 0 0"]
    with open(tmp_path / "legacy.bin", "wb") as f:                             # v2/read_in.cpp format
        f.write(np.array([2], "<u4").tobytes()); f.write(bytes([0, 1, 1, 0]))
    p = run_py("show", tmp_path / "legacy.bin", "--legacy")
    assert p.stdout.splitlines() == ["Adjacency matrix (2x2):", "0 1", "1 0"]


def test_bench_harness_formats(bb, graphs, tmp_path):
    from bidirectional_bfs_b200.bench_harness import HEADER, boxed_table, run
    rows, table = run([str(graphs / "1k.bin")], ["serial", "cpu-ranks"], repeats=2, ranks=2, outdir=str(tmp_path))
    with open(tmp_path / "benchmark_results.csv") as f:
        got = list(csv.reader(f))
    assert got[0][:4] == ["version", "graph", "time_sec", "logfile"]          # benchmark_test.sh:27
    assert got[0] == HEADER and len(got) == 3 and got[1][0] == "serial" and got[1][1] == "1k.bin"
    float(got[1][2])                                                           # seconds, always
    lines = (tmp_path / "benchmark_table.txt").read_text().splitlines()
    assert lines[0].startswith("+-") and lines[0] == lines[2] == lines[-1]
    assert lines[1].split("|")[1].strip() == "version"
    ref = "+---+----+\n| a | bb |\n+---+----+\n| 1 | 2  |\n+---+----+"
    assert boxed_table([["a", "bb"], ["1", "2"]]) == ref


def test_viz_writes_png(graphs, tmp_path):
    out = tmp_path / "g.png"
    p = run_py("viz", graphs / "1k.bin", "--out", out)
    assert p.returncode == 0, p.stderr
    from PIL import Image
    im = Image.open(out)
    assert im.size[0] >= 400
    d = json.load(open(graphs / "1k.json"))
    if d["nodes"]:
        px = np.asarray(im.convert("RGB")).reshape(-1, 3)
        assert ((px[:, 0] > 200) & (px[:, 1] < 60) & (px[:, 2] < 60)).sum() > 50   # red path edges were drawn
