"""Sharded (N > 1) path — SURVEY.md §8(e).  The CPU tests run the real
orchestration (sparrowhawk_b200/dist.py) at world_size 2 and 3 over `gloo` with a
CPU stand-in for the device stages; the GPU test runs the real C ABI engine in 2
processes that share the box's one GPU (host-level collectives only) and checks
every rank's complete result against the oracle."""
import os
import socket
import subprocess
import sys

import pytest

from sparrowhawk_b200.dist import shard_range, split_pairs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "sharded_worker.py")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _launch(world, *args, timeout=600):
    os.environ.setdefault("SHK_SHARD_PULL_MAX_WORLD", "8")       # (the product pulls up to 2 ranks: the tests cover more)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER, *[str(a) for a in args]]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-6000:]
    assert r.stdout.count(" ok: ") == world, r.stdout


def test_ownership_rules():
    for P in (16, 64, 1 << 15):
        for world in (1, 2, 3, 4, 8):
            edges = [shard_range(P, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == P
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            assert all(hi > lo for lo, hi in edges)
    for n in (0, 1, 7, 1000):
        for world in (1, 2, 3, 8):
            parts = [split_pairs(n, world, r) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n and all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))


@pytest.mark.parametrize("world,k", [(2, 31), (3, 21), (2, 63)])
def test_sharded_orchestration_gloo_cpu(world, k):
    _launch(world, "cpu", k, 2, 3000, 12.0)


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_pull_orchestration_gloo_cpu(world, tmp_path, monkeypatch):
    """dist.py's in-place count on CPU stand-ins (the published regions are files every rank sees): all ranks must
    agree on exchanging while one is uncalibrated, pull once all are ready, and ALL fall back to the exchange in the
    step in which ONE rank's count does not stand"""
    monkeypatch.setenv("SHK_TEST_SHARE", str(tmp_path))
    _launch(world, "cpu", 31, 2, 3000, 12.0, "pull")


@pytest.mark.gpu
@pytest.mark.parametrize("world,k,G", [(2, 31, 60000), (2, 63, 40000), (3, 31, 30000), (1, 31, 30000)])
def test_sharded_gpu_matches_oracle(world, k, G):
    """gloo, every rank on cuda:0: the sharded graph stage (peer memory between the processes, chains joined by
    the chain-level ranking) must be the one that ran, and every rank must hold the oracle's graph."""
    _launch(world, "gpu", k, 3, G, 30.0, "expect=sharded", "records=pulled")


@pytest.mark.gpu
def test_sharded_gpu_pull_falls_back_to_the_exchange(monkeypatch):
    """the owners count straight out of the sources' regions (no pack, no all-to-all) from the second pass on; a rank
    whose in-place count does not stand (forced here) makes every rank take the exchange path for that step; with the
    export switched off nothing is ever pulled — same graph every time"""
    monkeypatch.setenv("SHK_SHARD_PULL_FAIL", "1")
    _launch(2, "gpu", 31, 3, 50000, 30.0, "expect=sharded", "records=exchanged")
    monkeypatch.delenv("SHK_SHARD_PULL_FAIL")
    monkeypatch.setenv("SHK_SHARD_PULL", "0")
    _launch(2, "gpu", 31, 3, 50000, 30.0, "expect=sharded", "records=exchanged")


@pytest.mark.gpu
def test_sharded_gpu_replicated_stage_still_exact(monkeypatch):
    monkeypatch.setenv("SHK_SHARD_GRAPH", "0")
    _launch(2, "gpu", 31, 3, 40000, 30.0, "expect=replicated")


@pytest.mark.gpu
@pytest.mark.parametrize("world,k", [(2, 31), (3, 63)])
def test_sharded_gpu_cycles_across_ranks(world, k):
    """circular replicons whose single unitig is spread over all ranks (S14: cut at the minimum k-mer)"""
    _launch(world, "gpu", k, 2, 6000, 0.0, "circular")


@pytest.mark.gpu
def test_sharded_gpu_do_fit():
    """the do_fit branch of the sharded run: spectrum all-reduced, fitted on every rank, counted at the fit"""
    _launch(2, "gpu", 31, 5, 100000, 25.0, "fit")


def _n_devices():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("world,k,G", [(2, 31, 200000), (2, 63, 100000)])
def test_sharded_nccl_matches_oracle(world, k, G):
    """the path bench.py times at N > 1: NCCL, one device per rank, peer memory over NVLink — against the oracle on
    every rank.  Needs as many devices as ranks."""
    if _n_devices() < world:
        pytest.skip(f"needs {world} devices")
    _launch(world, "nccl", k, 3, G, 30.0, "expect=sharded", "records=pulled")


@pytest.mark.gpu
def test_sharded_nccl_cycles_and_fit():
    if _n_devices() < 2:
        pytest.skip("needs 2 devices")
    _launch(2, "nccl", 31, 2, 6000, 0.0, "circular")
    _launch(2, "nccl", 31, 5, 100000, 25.0, "fit")
