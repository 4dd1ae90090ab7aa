"""Worker for the multi-rank tests, launched by torch.distributed.run (one process per rank).
  mode gloo : host-side sharding logic on CPU; the per-rank compute is the ORACLE (test infrastructure)
  mode nccl : hs_build_index_dist on one GPU per rank, checked against the oracle"""
import os
import sys

import numpy as np
import pyarrow as pa
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from hyperspace_1_b200 import dist as hsdist  # noqa: E402
from oracle import oracle  # noqa: E402
import parity  # noqa: E402

NB = 64


NKEYS = int(os.environ.get("HS_DIST_NKEYS", "2"))
if os.environ.get("HS_DIST_SHAPE", "") == "strings":
    NKEYS = 1   # 1: a single non-null int64 key (the lineitem shape)


SHAPE = os.environ.get("HS_DIST_SHAPE", "")   # "strings": string key and payload; "wide": a key of two full-range int64


def shard(rank, world, n=30000):
    rng = np.random.default_rng(100 + rank)
    if SHAPE == "strings":
        # a string key (NULLs on even ranks only, shared prefixes, some multi-byte), an int64, a string payload with
        # NULLs on odd ranks only, and a boolean: everything the rows-first exchange has to take apart
        words = ["", "a", "ab", "abc", "prefix-shared-" * 2, "é日本", "zz"]
        k = [words[int(rng.integers(0, len(words)))] + str(int(rng.integers(0, 5000))) for _ in range(n)]
        kmask = rng.random(n) < (0.03 if rank % 2 == 0 else 0.0)
        pay = ["p" * int(rng.integers(0, 30)) for _ in range(n)]
        pmask = rng.random(n) < (0.05 if rank % 2 == 1 else 0.0)
        return [pa.array(k, mask=kmask), pa.array(rng.integers(-2**40, 2**40, size=n)), pa.array(pay, mask=pmask),
                pa.array(rng.random(n) < 0.5)]
    if SHAPE == "wide":
        return [pa.array(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)),
                pa.array(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)), pa.array(rng.random(n))]
    if NKEYS == 1:
        # negative and positive keys around a large offset: exercises the sign flip and the subtraction of the
        # minimum when the key crosses NVLink as 4 bytes
        k = rng.integers(-70000, 2_000_000_000, size=n) - (1 << 40)
        v = rng.random(n)
        vmask = rng.random(n) < (0.05 if rank % 2 == 1 else 0.0)
        return [pa.array(k), pa.array(v, mask=vmask), pa.array(rng.integers(0, 100, size=n).astype(np.int16))]
    k1 = rng.integers(0, 1 << 12, size=n).astype(np.int32)
    k2 = rng.integers(0, 1 << 30, size=n)
    v = rng.random(n)
    # NULL keys only on even ranks, NULL payloads only on odd ranks: the move set must still agree everywhere
    kmask = rng.random(n) < (0.03 if rank % 2 == 0 else 0.0)
    vmask = rng.random(n) < (0.05 if rank % 2 == 1 else 0.0)
    return [pa.array(k1), pa.array(k2, mask=kmask), pa.array(v, mask=vmask)]


def gather_all(cols):
    world = dist.get_world_size()
    out = [None] * world
    dist.all_gather_object(out, [c.to_pylist() for c in cols])
    return [pa.array(sum((o[i] for o in out), []), type=cols[i].type) for i in range(len(cols))]


def main():
    mode = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if mode == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank))))
    else:
        dist.init_process_group("gloo")
    cols = shard(rank, world)
    allcols = gather_all(cols)
    perm, woff = oracle.build_index(allcols, NKEYS, NB)
    want = oracle.take(allcols, perm)
    lo, hi = hsdist.bucket_range(rank, world, NB)
    assert all(hsdist.owner_of(b, world, NB) == rank for b in range(lo, hi))
    if mode == "gloo":
        # the exchange the GPU path performs, spelled out on the host: rows grouped by owner, all-to-all
        b = oracle.bucket_ids(cols[:NKEYS], NB)
        own = b.astype(np.int64) * world // NB
        pieces = [[c.take(pa.array(np.nonzero(own == r)[0])).to_pylist() for c in cols] for r in range(world)]
        recv = [None] * world
        dist.all_gather_object(recv, pieces)
        mine = [pa.array(sum((recv[src][rank][i] for src in range(world)), []), type=cols[i].type) for i in range(len(cols))]
        p2, o2 = oracle.build_index(mine, NKEYS, NB)
        got = oracle.take(mine, p2)
        goff = o2
    else:
        from hyperspace_1_b200 import native
        ctx = native.Context(int(os.environ.get("LOCAL_RANK", rank)))
        hsdist.init_comm(ctx, torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank))))
        got, goff = hsdist.build_index_dist(ctx, cols, NKEYS, NB)
        sent, ms = ctx.comm_last_exchange()
        assert sent > 0 and ms > 0
        ctx.comm_destroy()
        ctx.close()
    # this rank must hold exactly the global index restricted to its bucket range
    assert (np.diff(goff)[:lo] == 0).all() and (np.diff(goff)[hi:] == 0).all()
    assert np.array_equal(np.diff(goff)[lo:hi], np.diff(woff)[lo:hi])
    a, b_ = int(woff[lo]), int(woff[hi])
    want_mine = [c.slice(a, b_ - a) for c in want]
    parity.assert_keys_identical(got, want_mine, NKEYS, f"rank {rank}")
    parity.assert_same_multiset(got, want_mine, goff - goff[lo] * 0, np.concatenate([np.zeros(lo, np.int64), woff[lo:hi + 1] - woff[lo], np.full(NB - hi, woff[hi] - woff[lo])]), f"rank {rank}")
    dist.barrier()
    if rank == 0:
        print(f"DIST_OK mode={mode} world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
