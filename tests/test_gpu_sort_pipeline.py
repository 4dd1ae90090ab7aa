"""The pipelined sort of the large buckets (k_sort_pipeline: one launch for the MSD level and the leaves, the level
written into a small re-used scratch area) against the oracle, in the shapes that steer it: many groups per slot
ring (a tiny ring), buckets that do not fit a slot (left to the level loop), children that stay large (skewed and
constant keys), and the plain level loop with the pipeline switched off — all must give the oracle's index."""
import numpy as np
import pyarrow as pa
import pytest

import parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def _rows(n, kind, seed):
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        k = rng.integers(-2**40, 2**40, size=n)
    elif kind == "dups":          # few distinct keys: children of the MSD level stay larger than a leaf
        k = rng.integers(0, 40, size=n) * 1_000_003
    elif kind == "skew":          # half of the rows share one key, the rest is spread widely
        k = np.where(rng.random(n) < 0.5, 77, rng.integers(-2**50, 2**50, size=n))
    else:                          # "narrow": a dense range (few key bits)
        k = rng.integers(0, 3000, size=n)
    return [pa.array(k.astype(np.int64)), pa.array(rng.random(n)), pa.array(rng.integers(0, 1000, size=n).astype(np.int32))]


@pytest.mark.parametrize("kind", ["uniform", "dups", "skew", "narrow"])
@pytest.mark.parametrize("ring_mb,pipeline", [("1", "1"), ("48", "1"), ("48", "0")])
def test_large_buckets_against_oracle(ctx, oracle, monkeypatch, kind, ring_mb, pipeline):
    monkeypatch.setenv("HSGPU_RING_MB", ring_mb)
    monkeypatch.setenv("HSGPU_SORT_PIPELINE", pipeline)
    n, nb = 400_000, 7          # buckets of ~57 k rows: 14 tiles each, well above a leaf
    cols = _rows(n, kind, 11)
    got, boff = ctx.build_index(cols, 1, nb)
    parity.check_index(oracle, got, boff, cols, 1, nb, f"{kind} ring={ring_mb} pipeline={pipeline}")


def test_int32_key_and_bucket_larger_than_a_slot(ctx, oracle, monkeypatch):
    monkeypatch.setenv("HSGPU_RING_MB", "1")     # slots of ~20 k rows: the 150 k-row bucket goes to the level loop
    rng = np.random.default_rng(3)
    n = 300_000
    k = rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32)
    cols = [pa.array(k), pa.array(rng.random(n))]
    for nb in (2, 40):
        got, boff = ctx.build_index(cols, 1, nb)
        parity.check_index(oracle, got, boff, cols, 1, nb, f"int32 key nb={nb}")
