import numpy as np
import pytest
import torch

from conftest import special_graphs


def scipy_csr(n, edges):
    from scipy.sparse import csr_matrix
    e = np.asarray(edges, np.int64).reshape(-1, 2)
    e = e[e[:, 0] != e[:, 1]]
    r = np.concatenate([e[:, 0], e[:, 1]])
    c = np.concatenate([e[:, 1], e[:, 0]])
    order = np.lexsort((c, r))
    r, c = r[order], c[order]
    rp = np.zeros(n + 1, np.int64)
    np.add.at(rp, r + 1, 1)
    return np.cumsum(rp), c


@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_csr_matches_scipy(bb, nranks):
    spec = bb.er_spec(3000, 4.0 / 3000, seed=2)
    edges = bb.graph.edge_list(spec)
    rp_ref, col_ref = scipy_csr(3000, edges)
    g = bb.Graph.from_edges(3000, edges, nranks=nranks, sort_rows=True)
    gg = bb.Graph.generate(spec, nranks=nranks, sort_rows=True)     # fused generator path
    for G in (g, gg):
        assert G.nnz == len(col_ref)
        for r, s in G.shards.items():
            lo, hi = s.lo, s.hi
            rp = s.row_ptr.numpy()
            assert np.array_equal(rp, rp_ref[lo:hi + 1] - rp_ref[lo])
            assert np.array_equal(s.col.numpy().view(np.uint32), col_ref[rp_ref[lo]:rp_ref[hi]].astype(np.uint32))


def test_block_partition(bb):
    from bidirectional_bfs_b200.parallel import Partition
    p = Partition(1000, 8)
    assert p.B % 64 == 0 and p.B * 8 >= 1000
    covered = []
    for r in range(8):
        lo, hi = p.range(r)
        covered += list(range(lo, hi))
        assert all(p.owner(v) == r for v in (lo, hi - 1) if lo < hi)
    assert covered == list(range(1000))


def test_special_graphs_build(bb):
    for name, n, edges in special_graphs():
        g = bb.Graph.from_edges(n, edges, sort_rows=True)
        rp_ref, col_ref = scipy_csr(n, edges)
        assert np.array_equal(g.single().row_ptr.numpy(), rp_ref), name
        assert np.array_equal(g.single().col.numpy().view(np.uint32), col_ref.astype(np.uint32)), name


def test_to_scipy_and_neighbors(bb):
    g = bb.Graph.from_edges(5, np.array([[0, 1], [1, 2], [1, 4]], np.uint32), sort_rows=True)
    A = g.to_scipy()
    assert A.shape == (5, 5) and A.nnz == 6
    assert g.neighbors(1).tolist() == [0, 2, 4] and g.degree(3) == 0


def test_owner_reciprocal_is_exact():
    """owner_of() in csrc/bfs_dev.cu replaces v / B by umulhi64(2^64 / B rounded up, v): exact for every 32-bit v."""
    import random
    rnd = random.Random(5)
    M64 = (1 << 64) - 1
    for _ in range(300):
        B = rnd.choice([64, 128, 192, 1 << 20, 64 * rnd.randrange(1, 1 << 20), 64 * rnd.randrange(1, 1 << 26)])
        div = (M64 // B + 1) & M64
        vs = [0, 1, B - 1, B, B + 1, 2 * B - 1, 7 * B, 8 * B - 1, (1 << 32) - 1] + [rnd.randrange(1 << 32) for _ in range(200)]
        for v in vs:
            v &= (1 << 32) - 1
            assert (div * v) >> 64 == v // B, (B, v)
