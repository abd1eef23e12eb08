"""Multi-process host-side logic on the CPU (gloo, world_size 2): torchrun rendezvous through init_distributed and the
per-rank construction of the partitioned CSR.  (The device-side exchange needs CUDA IPC: tests/test_gpu_search.py
covers it with virtual ranks, scripts/dist_check.py on real GPUs.)"""
import os
import socket
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_two_rank_shard_build_gloo():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", CUDA_VISIBLE_DEVICES="")
    with socket.socket() as sk:                      # a free port for the rendezvous
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / "tests" / "helpers" / "dist_shards_worker.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "DIST_SHARDS_OK 2" in out.stdout
