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


def test_sharded_build_policy(monkeypatch):
    """Which construction a rank uses for its shard (graph._sharded_build_applies): files are always routed between the
    ranks, a generator is replayed by every rank up to 8 ranks and routed beyond, BIBFS_SHARDED_BUILD forces either."""
    import torch
    import torch.distributed as dist
    from bidirectional_bfs_b200 import graph as G

    cuda = torch.device("cuda", 0)
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_rank", lambda: 1)
    monkeypatch.delenv("BIBFS_SHARDED_BUILD", raising=False)
    for world in (2, 8, 16):
        monkeypatch.setattr(dist, "get_world_size", lambda w=world: w)
        assert G._sharded_build_applies(cuda, world, [1]) is True                             # an edge list that exists once
        assert G._sharded_build_applies(cuda, world, [1], from_generator=True) is (world > 8)
        assert G._sharded_build_applies(cuda, world, [0]) is False                            # not this process's shard
        assert G._sharded_build_applies(cuda, world, [0, 1]) is False                         # several shards in one process
        assert G._sharded_build_applies(torch.device("cpu"), world, [1]) is False
    monkeypatch.setattr(dist, "get_world_size", lambda: 4)
    monkeypatch.setenv("BIBFS_SHARDED_BUILD", "1")
    assert G._sharded_build_applies(cuda, 4, [1], from_generator=True) is True
    monkeypatch.setenv("BIBFS_SHARDED_BUILD", "0")
    assert G._sharded_build_applies(cuda, 4, [1]) is False
    monkeypatch.setattr(dist, "is_initialized", lambda: False)
    monkeypatch.delenv("BIBFS_SHARDED_BUILD", raising=False)
    assert G._sharded_build_applies(cuda, 4, [1]) is False
