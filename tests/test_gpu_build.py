"""GPU: generator / scan / CSR-builder kernels against their host twins (bit exact)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_native_library_is_loaded(bb):
    maps = open("/proc/self/maps").read()
    bb.lib()
    maps = open("/proc/self/maps").read()
    assert "libbibfs_b200.so" in maps


@pytest.mark.parametrize("mk", [
    lambda bb: bb.er_spec(5000, 2.2 / 5000, seed=3),
    lambda bb: bb.er_spec(300, 0.3, seed=9),
    lambda bb: bb.er_spec(40, 1.0),
    lambda bb: bb.rmat_spec(12, edgefactor=16, seed=2),
    lambda bb: bb.rmat_spec(12, edgefactor=4, n=3000, seed=2),
    lambda bb: bb.grid_spec(33, 17),
])
def test_generators_device_equals_host(bb, mk):
    spec = mk(bb)
    host = bb.graph.edge_list(spec, "cpu")
    dev = bb.graph.edge_list(spec, "cuda:0")
    assert host.shape == dev.shape
    assert np.array_equal(host, dev)


@pytest.mark.parametrize("count", [0, 1, 31, 4096, 4097, 100003, 3_000_000])
def test_scan(bb, count):
    from bidirectional_bfs_b200.ops import exclusive_scan_u32
    rng = np.random.default_rng(count)
    x = rng.integers(0, 1000, size=count, dtype=np.int64).astype(np.int32)
    if count:
        x[0] = 2_000_000_000           # force totals beyond 32 bits on the large case
    out = exclusive_scan_u32(torch.from_numpy(x).cuda()).cpu().numpy()
    ref = np.concatenate([[0], np.cumsum(x.astype(np.int64))])
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("nranks", [1, 2, 8])
def test_csr_device_equals_host(bb, nranks):
    for spec in (bb.er_spec(20000, 3.0 / 20000, seed=5), bb.rmat_spec(13, edgefactor=16, seed=7), bb.grid_spec(50, 41)):
        edges = bb.graph.edge_list(spec)
        h = bb.Graph.from_edges(spec.n, edges, nranks=nranks, sort_rows=True)
        d = bb.Graph.from_edges(spec.n, edges, device="cuda:0", nranks=nranks, sort_rows=True)
        f = bb.Graph.generate(spec, device="cuda:0", nranks=nranks, sort_rows=True)   # fused generator -> CSR
        for r in range(nranks):
            for G in (d, f):
                assert torch.equal(G.shards[r].row_ptr.cpu(), h.shards[r].row_ptr)
                assert torch.equal(G.shards[r].col.cpu(), h.shards[r].col)


def test_csr_device_unsorted_is_a_permutation(bb):
    spec = bb.rmat_spec(12, edgefactor=16, seed=1)
    h = bb.Graph.generate(spec, sort_rows=True)
    d = bb.Graph.generate(spec, device="cuda:0", sort_rows=False)
    rp = d.single().row_ptr.cpu().numpy()
    assert np.array_equal(rp, h.single().row_ptr.numpy())
    col = d.single().col.cpu().numpy().view(np.uint32)
    ref = h.single().col.numpy().view(np.uint32)
    for v in range(0, spec.n, 97):
        assert np.array_equal(np.sort(col[rp[v]:rp[v + 1]]), ref[rp[v]:rp[v + 1]])


def test_big_row_sort(bb):
    # one hub row far above the shared-memory sort limit (4096) exercises the global-memory path
    n = 20000
    edges = np.array([[0, i] for i in range(1, n)], np.uint32)[::-1].copy()
    d = bb.Graph.from_edges(n, edges, device="cuda:0", sort_rows=True)
    col = d.single().col.cpu().numpy().view(np.uint32)
    assert np.array_equal(col[: n - 1], np.arange(1, n, dtype=np.uint32))
