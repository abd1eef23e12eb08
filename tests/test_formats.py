import json

import numpy as np
import pytest

from conftest import nx_graph


def test_bin_roundtrip(bb, tmp_path):
    edges = np.array([[0, 1], [1, 2], [2, 5], [3, 4]], np.uint32)
    p = tmp_path / "g.bin"
    bb.io.write_bin(p, 6, edges)
    raw = np.fromfile(p, dtype="<u4")
    assert raw.tolist() == [6, 4, 0, 1, 1, 2, 2, 5, 3, 4]          # reference layout, generate_graph.py:35-39
    n, e = bb.io.read_bin(p)
    assert n == 6 and np.array_equal(e, edges)


def test_bin_validation(bb, tmp_path):
    p = tmp_path / "bad.bin"
    np.array([3, 1, 0, 7], dtype="<u4").tofile(p)                   # id 7 >= N
    with pytest.raises(bb.BibfsError):
        bb.io.read_bin(p)
    q = tmp_path / "short.bin"
    np.array([3, 5, 0, 1], dtype="<u4").tofile(q)                   # header says 5 edges
    with pytest.raises(bb.BibfsError):
        bb.io.read_bin(q)
    with pytest.raises(bb.BibfsError):
        bb.io.read_bin(tmp_path / "missing.bin")


def test_oracle_json(bb, tmp_path):
    p = tmp_path / "1k.bin"
    assert bb.io.oracle_path_for(p).name == "1k.json"
    bb.io.write_oracle(tmp_path / "1k.json", 0, 9, [0, 4, 9])
    d = bb.io.read_oracle(tmp_path / "1k.json")
    assert d == {"source": 0, "target": 9, "hop_count": 2, "nodes": [0, 4, 9]}
    bb.io.write_oracle(tmp_path / "none.json", 0, 9, [])
    d = json.load(open(tmp_path / "none.json"))
    assert d["hop_count"] is None and d["nodes"] == []              # generate_graph.py:47-49


def test_json_graph_inputs(bb, tmp_path):
    import networkx as nx
    from networkx.readwrite import json_graph
    G = nx.gnp_random_graph(40, 0.1, seed=3)
    for name, data in (("adj", json_graph.adjacency_data(G)), ("nl", json_graph.node_link_data(G))):
        p = tmp_path / f"{name}.json"
        json.dump(data, open(p, "w"))
        n, e = bb.io.read_json_graph(p)
        assert n == 40 and e.shape[0] == G.number_of_edges()
        g = bb.Graph.from_json(p)
        for s, t in [(0, 39), (5, 17)]:
            want = nx.shortest_path_length(G, s, t) if nx.has_path(G, s, t) else None
            assert bb.bibfs(g, s, t, variant="serial").length == want


def test_json_graph_roundtrip_and_labels(bb, tmp_path):
    p = tmp_path / "g.json"
    bb.io.write_json_graph(p, 4, [[0, 1], [1, 3]])
    n, e = bb.io.read_json_graph(p)
    assert n == 4 and e.tolist() == [[0, 1], [1, 3]]
    q = tmp_path / "labels.json"
    json.dump({"nodes": [{"id": "a"}, {"id": "b"}, {"id": "c"}], "links": [{"source": "a", "target": "c"}]}, open(q, "w"))
    n, e = bb.io.read_json_graph(q)
    assert n == 3 and e.tolist() == [[0, 2]]


def test_dense_legacy(bb, tmp_path):
    a = (np.arange(16).reshape(4, 4) % 2).astype(np.uint8)
    p = tmp_path / "dense.bin"
    with open(p, "wb") as f:
        f.write(np.array([4], "<u4").tobytes())
        f.write(a.tobytes())
    assert np.array_equal(bb.io.read_dense(p), a)
    with pytest.raises(bb.BibfsError):
        bb.io.read_dense(tmp_path / "nope.bin")


def test_csr_cache_roundtrip(bb, tmp_path):
    """Graph.save_csr / load_csr: the cached partitioned CSR gives back the same shards."""
    import numpy as np
    spec = bb.er_spec(3000, 3.0 / 3000, seed=5)
    g = bb.Graph.generate(spec, nranks=4)
    p = tmp_path / "g.npz"
    g.save_csr(p)
    h = bb.Graph.load_csr(p)
    assert (h.n, h.nranks, sorted(h.shards)) == (g.n, 4, [0, 1, 2, 3])
    for r in range(4):
        assert np.array_equal(h.shards[r].row_ptr.numpy(), g.shards[r].row_ptr.numpy())
        assert np.array_equal(h.shards[r].col.numpy(), g.shards[r].col.numpy())
    one = bb.Graph.load_csr(p, ranks=[2])
    assert list(one.shards) == [2]
    s, t = 0, spec.n - 1
    assert bb.bibfs(h, s, t, variant="cpu-ranks").length == bb.bibfs(g, s, t, variant="serial").length
