"""The N > 1 path: bucket ownership and the row exchange, world_size 2 on CPU (gloo, oracle compute) and, where
two GPUs are present, hs_build_index_dist over NCCL."""
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(mode, world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "dist_worker.py"), mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, "\n".join(l for l in (r.stdout + r.stderr).splitlines() if "Error" in l or "error" in l or "assert" in l.lower() or "HS_" in l)[-3000:]
    assert f"DIST_OK mode={mode} world={world}" in r.stdout


def test_owner_map():
    from hyperspace_1_b200 import dist as d
    for nb, n in ((200, 8), (8192, 8), (1000, 4), (200, 1), (7, 3)):
        owners = [d.owner_of(b, n, nb) for b in range(nb)]
        assert owners == sorted(owners) and set(owners) == set(range(n))     # contiguous, every rank owns some
        for r in range(n):
            lo, hi = d.bucket_range(r, n, nb)
            assert owners[lo:hi] == [r] * (hi - lo) and owners.count(r) == hi - lo
    assert d.bucket_range(3, 8, 200) == (75, 100)      # 200 / 8 = 25 buckets per GPU


def test_sharded_build_world2_gloo():
    _run("gloo", 2)


def test_sharded_build_world2_gloo_string_rows(monkeypatch):
    monkeypatch.setenv("HS_DIST_SHAPE", "strings")
    _run("gloo", 2)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", ["strings", "wide"])
def test_rows_first_build_world2_nccl(shape, monkeypatch):
    """String columns and keys wider than 64 bits cannot ride the fused scatter (their bytes stay with the sender):
    whole rows change hands first, then every GPU builds its own buckets. Checked against the oracle's index of
    all ranks' rows."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    monkeypatch.setenv("HS_DIST_SHAPE", shape)
    _run("nccl", 2)


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["fused", "owner", "copy", "spans", "nccl", "owner-int64-key", "fused-int64-key",
                                      "spans-int64-key", "spans-groups7", "spans-groups1"])
def test_sharded_build_world2_nccl(exchange, monkeypatch):
    """hs_build_index_dist on two GPUs against the oracle, once per way the rows can travel: straight into the
    owner's buckets (fused), grouped by owner with the bucket ids alongside and a bucket scatter at the receiver
    (owner), whole bucket runs pushed after a local scatter (copy), spans of rows grouped by (owner, bucket group)
    copied group after group while the owner sorts the groups that have arrived (spans; with group counts that do
    not divide the bucket count, and with a single group), and the NCCL send/receive fallback."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    if exchange.endswith("-int64-key"):
        # one non-null int64 key spanning less than 2^32: in owner mode it crosses NVLink as 4 bytes
        monkeypatch.setenv("HS_DIST_NKEYS", "1")
        exchange = exchange.split("-")[0]
    if "-groups" in exchange:
        monkeypatch.setenv("HSGPU_XCHG_GROUPS", exchange.split("-groups")[1])
        exchange = exchange.split("-")[0]
    if exchange == "nccl":
        monkeypatch.setenv("HSGPU_NO_P2P", "1")
    else:
        monkeypatch.setenv("HSGPU_P2P_MODE", exchange)
    _run("nccl", 2)
