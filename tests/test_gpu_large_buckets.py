"""Buckets well above a leaf against the oracle, in the key shapes that steer the MSD levels and the leaf kernels:
uniform wide keys, few distinct keys (children of a level stay larger than a leaf and are re-ranged), half of the
rows on one key, a dense narrow range, and leaves around the sizes where the leaf kernel changes its variant
(half a tile, a full tile of LEAF_MAX = 3072 rows, one row more)."""
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
def test_large_buckets_against_oracle(ctx, oracle, kind):
    n, nb = 400_000, 7          # buckets of ~57 k rows: 14 tiles each, well above a leaf
    cols = _rows(n, kind, 11)
    got, boff = ctx.build_index(cols, 1, nb)
    parity.check_index(oracle, got, boff, cols, 1, nb, kind)


def test_int32_key(ctx, oracle):
    rng = np.random.default_rng(3)
    n = 300_000
    k = rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32)
    cols = [pa.array(k), pa.array(rng.random(n))]
    for nb in (2, 40):
        got, boff = ctx.build_index(cols, 1, nb)
        parity.check_index(oracle, got, boff, cols, 1, nb, f"int32 key nb={nb}")


@pytest.mark.parametrize("n", [1535, 1536, 1537, 3071, 3072, 3073, 4096, 6145])
@pytest.mark.parametrize("bits", [10, 19, 20, 33, 63])
def test_one_bucket_around_the_leaf_sizes(ctx, oracle, n, bits):
    """One bucket of exactly n rows: the 4-rows-per-thread and 8-rows-per-thread leaf variants, a full leaf, and
    one row more (an MSD level first); key ranges below, at and above the packed kernel's 19 key bits (wider
    ranges sort on a prefix and repair ties, or fall to the 64-bit kernel)."""
    rng = np.random.default_rng(n * 64 + bits)
    k = rng.integers(0, 2**bits, size=n, dtype=np.uint64).astype(np.int64) - (2**(bits - 1) if bits < 63 else 0)
    cols = [pa.array(k), pa.array(rng.random(n)), pa.array(rng.integers(0, 9, size=n).astype(np.int16))]
    got, boff = ctx.build_index(cols, 1, 1)
    parity.check_index(oracle, got, boff, cols, 1, 1, f"n={n} bits={bits}")
