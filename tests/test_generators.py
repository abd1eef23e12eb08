import numpy as np
import pytest

from conftest import nx_graph


def test_er_deterministic_and_shape(bb):
    s1 = bb.er_spec(2000, 3.0 / 2000, seed=11)
    a = bb.graph.edge_list(s1)
    b = bb.graph.edge_list(bb.er_spec(2000, 3.0 / 2000, seed=11))
    c = bb.graph.edge_list(bb.er_spec(2000, 3.0 / 2000, seed=12))
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert (a[:, 0] < a[:, 1]).all() and a.max() < 2000            # u < v like the reference files
    assert len({(int(u), int(v)) for u, v in a}) == len(a)          # simple graph
    assert bb.ops.count_edges(s1) == len(a)


def test_er_statistics(bb):
    n, p = 20000, 2.2 / 20000
    m = [len(bb.graph.edge_list(bb.er_spec(n, p, seed=s))) for s in range(5)]
    mean = n * (n - 1) / 2 * p
    assert abs(np.mean(m) - mean) < 4 * np.sqrt(mean / 5)
    e = bb.graph.edge_list(bb.er_spec(n, p, seed=1))
    deg = np.bincount(e.ravel(), minlength=n)
    assert abs(deg.mean() - 2.2) < 0.1 and abs(deg.var() - 2.2) < 0.3   # Poisson(2.2)


def test_er_extremes(bb):
    assert len(bb.graph.edge_list(bb.er_spec(50, 0.0))) == 0
    e = bb.graph.edge_list(bb.er_spec(30, 1.0))
    assert len(e) == 30 * 29 // 2 and len({tuple(x) for x in e.tolist()}) == len(e)
    dense = bb.graph.edge_list(bb.er_spec(200, 0.5, seed=4))
    assert abs(len(dense) - 0.5 * 200 * 199 / 2) < 5 * np.sqrt(200 * 199 / 8)


def test_presets_match_make_graphs(bb):
    s = bb.preset_spec("10k")
    assert s.n == 10000 and abs(s.p - 2.2000000001 / 10000) < 1e-18   # graphs/make_graphs:8,19-22


def test_grid_closed_form(bb):
    w, h = 7, 5
    e = bb.graph.edge_list(bb.grid_spec(w, h))
    assert len(e) == h * (w - 1) + (h - 1) * w
    want = {(y * w + x, y * w + x + 1) for y in range(h) for x in range(w - 1)} | \
           {(y * w + x, (y + 1) * w + x) for y in range(h - 1) for x in range(w)}
    assert {tuple(x) for x in e.tolist()} == want
    g = bb.Graph.generate(bb.grid_spec(w, h))
    assert bb.bibfs(g, 0, w * h - 1, variant="serial").length == (w - 1) + (h - 1)


def test_rmat_basic(bb):
    s = bb.rmat_spec(10, edgefactor=8, seed=5)
    e = bb.graph.edge_list(s)
    assert e.shape == (8 * 1024, 2) and e.max() < 1024
    assert np.array_equal(e, bb.graph.edge_list(bb.rmat_spec(10, edgefactor=8, seed=5)))
    deg = np.bincount(e.ravel(), minlength=1024)
    assert deg.max() > 8 * deg.mean()                               # skewed, unlike ER
    s2 = bb.rmat_spec(10, edgefactor=8, n=1000, seed=5)             # folded id space
    e2 = bb.graph.edge_list(s2)
    assert e2.max() < 1000 and len(e2) == 8000


def test_bad_specs(bb):
    with pytest.raises(bb.BibfsError):
        bb.graph.edge_list(bb.er_spec(10, 1.5))
    with pytest.raises(bb.BibfsError):
        bb.graph.edge_list(bb.rmat_spec(4, n=100))
    bad = bb.grid_spec(3, 3)
    bad.n = 10
    with pytest.raises(bb.BibfsError):
        bb.graph.edge_list(bad)
