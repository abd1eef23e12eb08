"""bench.py's stand-alone helper (baseline/tools/bench_prep.cpp: graph file, query list and the v1-schedule numerator of BOTH
bench arms, built without this repository's library) against the library itself: same graph, same path lengths."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def test_bench_prep_agrees_with_the_library(bb, tmp_path):
    import bench
    for config in ("er10k", "rmat12", "grid40"):
        path = tmp_path / f"{config}.bin"
        info = bench.prepare(config, seed=1, need=12, bin_path=path)
        spec = bench.make_spec(bb, config, 1)
        g = bb.Graph.generate(spec)                                   # the library's generator + host CSR builder
        assert info["n"] == g.n and info["nnz"] == g.nnz, (config, info["n"], g.n, info["nnz"], g.nnz)
        n, edges = bb.io.read_bin(path)                               # the file the reference arm would load
        assert n == g.n and len(edges) == info["records"]
        gf = bb.Graph.from_edges(n, edges)
        assert np.array_equal(np.sort(gf.single().row_ptr.numpy()), np.sort(g.single().row_ptr.numpy()))
        oracle = bb.SerialSolver(g, policy=1)
        assert len(info["queries"]) == 12
        for s, t, length, v1_edges, v1_levels in info["queries"]:
            r = oracle.query(s, t, want_path=False)
            assert r.length == length, (config, s, t, r.length, length)
            assert v1_edges > 0 and v1_levels >= 1
        if config == "er10k":
            assert info["queries"][0][:2] == (0, g.n - 1) or info["queries"][0][2] >= 2   # the reference's own query comes first when it is connected


def test_bench_spec_strings():
    import bench
    assert bench.spec_string("er100k", 1)[0].startswith("er:100000:")
    assert bench.spec_string("rmat10m", 3)[0] == "rmat:24:8:10000000:3"
    assert bench.spec_string("grid16m", 1)[0] == "grid:4096:4096"
    assert bench.spec_string("kron26", 1)[0] == f"rmat:26:16:{1 << 26}:1"
    big = bench.prepare("kron26", 1, 5)                               # beyond the host limit: no replay, own numerator
    assert big["queries"] is None and big["n"] == 1 << 26
