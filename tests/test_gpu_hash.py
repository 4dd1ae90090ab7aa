"""K1 parity: every row's Murmur3 hash and bucket id, bit-exact against the oracle and the golden vectors."""
import json
import os

import numpy as np
import pyarrow as pa
import pytest

from fixtures import sample_table, string_keys

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TYPES = {"int32": pa.int32(), "int64": pa.int64(), "string": pa.string()}


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def test_golden_vectors_on_gpu(ctx):
    with open(os.path.join(HERE, "golden", "murmur3_vectors.json")) as f:
        g = json.load(f)
    for vec in g["pinned"] + g["derived"]:
        cols = [pa.array(c["values"], type=TYPES[c["type"]]) for c in vec["cols"]]
        if "hash" in vec:
            assert ctx.hash_rows(cols).tolist() == vec["hash"]
        if "bucket" in vec:
            assert ctx.hash_bucket(cols, vec["num_buckets"]).tolist() == vec["bucket"]


@pytest.mark.parametrize("n", [0, 1, 31, 1000, 100003])
def test_int_keys(ctx, oracle, n):
    rng = np.random.default_rng(n)
    k64 = pa.array(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64))
    k32 = pa.array(rng.integers(-2**31, 2**31 - 1, size=n, dtype=np.int64).astype(np.int32))
    for keys in ([k64], [k32], [k32, k64]):
        assert np.array_equal(ctx.hash_rows(keys), oracle.hash_rows(keys))
        for nb in (1, 7, 200, 8192):
            assert np.array_equal(ctx.hash_bucket(keys, nb), oracle.bucket_ids(keys, nb))


def test_all_types_with_nulls(ctx, oracle):
    rng = np.random.default_rng(5)
    n = 5000
    mask = rng.random(n) < 0.2
    f = rng.standard_normal(n)
    f[::97] = np.nan
    f[1::97] = -0.0
    f[2::97] = 0.0
    cols = [
        pa.array(rng.integers(0, 2, size=n).astype(bool), mask=mask),
        pa.array(rng.integers(-128, 127, size=n).astype(np.int8)),
        pa.array(rng.integers(-2**15, 2**15 - 1, size=n).astype(np.int16), mask=~mask),
        pa.array(rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32), mask=mask),
        pa.array(rng.integers(-2**62, 2**62, size=n), mask=rng.random(n) < 0.5),
        pa.array(f.astype(np.float32)),
        pa.array(f, mask=mask),
    ]
    for k in range(len(cols)):
        assert np.array_equal(ctx.hash_rows([cols[k]]), oracle.hash_rows([cols[k]])), f"column {k}"
    assert np.array_equal(ctx.hash_rows(cols), oracle.hash_rows(cols))
    assert np.array_equal(ctx.hash_bucket(cols, 200), oracle.bucket_ids(cols, 200))


def test_string_keys_all_tail_lengths_and_alignments(ctx, oracle):
    rng = np.random.default_rng(11)
    vals = []
    for ln in range(0, 40):
        for _ in range(8):
            vals.append(bytes(rng.integers(0, 256, size=ln, dtype=np.uint8)))
    vals += [None, b"", None]
    b = pa.array(vals, type=pa.binary())
    assert np.array_equal(ctx.hash_rows([b]), oracle.hash_rows([b]))
    s, v = string_keys(3000)
    assert np.array_equal(ctx.hash_rows([s]), oracle.hash_rows([s]))
    assert np.array_equal(ctx.hash_bucket([s, v], 1000), oracle.bucket_ids([s, v], 1000))
    t = sample_table()
    assert ctx.hash_bucket([t["Query"]], 3).tolist() == oracle.bucket_ids([t["Query"]], 3).tolist()


def test_errors_are_reported(ctx, hs):
    with pytest.raises(hs.HyperspaceException):
        ctx.hash_bucket([pa.array([1, 2, 3])], 0)
