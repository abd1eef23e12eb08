"""The allocation-free scan entry point, called through the C ABI.

bibfs_scan_u32_device_noalloc exists for callers that are handed inputs and outputs only and may
not allocate (a custom op inside a serving process with a fixed memory budget, where cudaMalloc
would also stall the other streams).  Its scratch is a fixed array inside the library, shared by
every call, so this test checks what such a caller relies on: the output equals the regular entry
point's bit for bit, a second call is not disturbed by what the first one left in the scratch,
and an input larger than the built-in scratch is refused with a readable error instead of
overrunning it.  This file sorts last on purpose so a failure here cannot hide other results.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scan_with_scratch(L, x):
    out = torch.empty(x.numel() + 1, dtype=torch.int64, device=x.device)
    nbytes = L.bibfs_scan_workspace_bytes(x.numel())
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    st = torch.cuda.current_stream(x.device).cuda_stream
    assert L.bibfs_scan_u32_device(x.data_ptr(), out.data_ptr(), x.numel(), ws.data_ptr(), nbytes, st) == 0, L.bibfs_last_error()
    torch.cuda.synchronize()
    return out


def _scan_no_alloc(L, x):
    out = torch.full((x.numel() + 1,), -1, dtype=torch.int64, device=x.device)     # caller-owned output
    st = torch.cuda.current_stream(x.device).cuda_stream
    assert L.bibfs_scan_u32_device_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), st) == 0, L.bibfs_last_error()
    torch.cuda.synchronize()
    return out


def test_no_alloc_scan_matches_regular_entry_point_twice(bb):
    L = bb.lib()
    rng = np.random.default_rng(7)
    # first a large input (fills many scratch slots), then a smaller, different one right after:
    # stale tile sums from call 1 must not leak into call 2
    for count in (1_000_003, 70_001):
        x = torch.from_numpy(rng.integers(0, 5000, size=count, dtype=np.int64).astype(np.int32)).cuda()
        want = _scan_with_scratch(L, x)
        got = _scan_no_alloc(L, x)
        assert torch.equal(got, want)                                  # integers: exact, like test_scan
        assert int(got[-1]) == int(x.to(torch.int64).sum())


def test_no_alloc_scan_edge_sizes(bb):
    L = bb.lib()
    for count in (0, 1, 4096, 4097):
        x = torch.arange(count, dtype=torch.int32, device="cuda")
        assert torch.equal(_scan_no_alloc(L, x), _scan_with_scratch(L, x))


def test_no_alloc_scan_rejects_oversize(bb):
    L = bb.lib()
    limit = int(L.bibfs_scan_noalloc_max_count())
    assert limit == 1 << 26
    dummy = torch.zeros(8, dtype=torch.int64, device="cuda")
    # refused before any memory is touched, so small dummy buffers are fine here
    rc = L.bibfs_scan_u32_device_noalloc(dummy.data_ptr(), dummy.data_ptr(), limit + 1, None)
    assert rc != 0
    msg = L.bibfs_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
