"""GPU tests that need SEVERAL devices (skipped on a one-GPU box): the one-process-per-GPU deployment under torchrun
(CUDA IPC peer mappings, k_wait, hand-over across NVLink) and the single-process multi-device solver."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

from conftest import query_pairs

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least two CUDA devices")
def test_torchrun_dist_check():
    n = min(_ngpu(), 8)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / "scripts" / "dist_check.py")]
    out = subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, MASTER_ADDR="127.0.0.1"), capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and f"DIST_CHECK PASS world {n}" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least two CUDA devices")
def test_single_process_across_devices(bb):
    from bidirectional_bfs_b200.models.solvers import multi_gpu_solver
    n_dev = min(_ngpu(), 4)
    for spec in (bb.er_spec(30000, 2.2 / 30000, seed=3), bb.rmat_spec(14, edgefactor=16, seed=5), bb.grid_spec(90, 70)):
        edges = bb.graph.edge_list(spec)
        g_cpu = bb.Graph.from_edges(spec.n, edges)
        oracle = bb.SerialSolver(g_cpu, policy=1)
        solver, used = multi_gpu_solver(spec.n, edges, gpus=n_dev)
        assert used == n_dev
        for flags in (0, 64, 2):
            for s, t in query_pairs(spec.n, 15, 7):
                want = oracle.query(s, t, want_path=False).length
                r = solver.query(s, t, flags=flags)
                assert r.length == want, (s, t, flags, r.length, want)
                if want is not None:
                    bb.validate_path(g_cpu, s, t, r.path, want)
        solver.close()
