"""Arrow large_string / large_binary columns (int64 offsets, for more than 2 GiB of key bytes — cfg4 at its full
500 M rows): they hash, order, join and filter exactly like their int32-offset twins, so every result is checked
against the oracle run on the int32-offset copy of the same data and against the library's own UTF8 result."""
import numpy as np
import pyarrow as pa
import pytest

import parity
from fixtures import string_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def _small(cols):
    out = []
    for c in cols:
        if pa.types.is_large_string(c.type):
            c = c.cast(pa.string())
        elif pa.types.is_large_binary(c.type):
            c = c.cast(pa.binary())
        out.append(c)
    return out


def test_build_large_string_key_and_included(ctx, oracle):
    s, v = string_keys(20000)
    rng = np.random.default_rng(3)
    inc = pa.array([None if rng.random() < 0.1 else bytes(rng.integers(0, 256, size=rng.integers(0, 40)).astype(np.uint8))
                    for _ in range(len(s))], type=pa.large_binary())
    keys = pa.array(s.to_pylist(), mask=rng.random(len(s)) < 0.05).cast(pa.large_string())
    cols = [keys, v, inc]
    got, boff = ctx.build_index(cols, 1, 16)
    assert pa.types.is_large_string(got[0].type) and pa.types.is_large_binary(got[2].type)
    parity.check_index(oracle, _small(got), boff, _small(cols), 1, 16, "large_string key")
    small, soff = ctx.build_index(_small(cols), 1, 16)
    assert np.array_equal(boff, soff)
    assert _small(got)[0].equals(small[0]), "LARGE and int32-offset builds order the keys differently"


def test_join_and_filter_on_large_strings(ctx, oracle):
    s, v = string_keys(6000)
    idx = np.random.default_rng(4).integers(0, 6000, size=15000)
    ls = s.take(pa.array(idx)).cast(pa.large_string())
    lg, loff = ctx.build_index([ls, pa.array(np.arange(15000))], 1, 8)
    rg, roff = ctx.build_index([s.cast(pa.large_string()), v], 1, 8)
    jg, joff = ctx.join(lg, loff, rg, roff, 1, [0, 1], [0, 1])
    li, ri, woff = oracle.join(_small(lg), _small(rg), 1, loff, roff)
    want = [_small(lg)[i].take(pa.array(li)) for i in (0, 1)] + [_small(rg)[i].take(pa.array(ri)) for i in (0, 1)]
    parity.assert_same_multiset(_small(jg), want, joff, woff, "large_string join")
    probe = s[11].as_py()
    for p in ([("eq", 0, probe)], [("ge", 0, "m"), ("lt", 0, "p"), ("and",)], [("ne", 0, probe)]):
        fg = ctx.filter_scan(lg, 1, loff, p, [0, 1])
        rows = pa.array(oracle.filter_rows(_small(lg), p))
        parity.assert_same_multiset(_small(fg), [c.take(rows) for c in _small(lg)], what=f"large_string filter {p}")


def test_merge_runs_of_large_strings(ctx, oracle):
    s, v = string_keys(9000)
    big = [s.cast(pa.large_string()), v]
    a = [c.slice(0, 6000) for c in big]
    b = [c.slice(6000) for c in big]
    ga, oa = ctx.build_index(a, 1, 12)
    gb, ob = ctx.build_index(b, 1, 12)
    mg, moff = ctx.merge_runs([(ga, 1, oa), (gb, 1, ob)])
    parity.check_index(oracle, _small(mg), moff, _small(big), 1, 12, "merged large_string runs")


@pytest.mark.parametrize("mean_len", [3, 200, 3000])
def test_gather_of_short_and_long_strings(ctx, oracle, mean_len):
    """The string gather stages tiles of rows in shared memory; rows per tile adapt to the mean length and
    tiles that do not fit take a direct path. Mixed lengths (incl. empty and a few very long) hit all of them."""
    rng = np.random.default_rng(mean_len)
    n = 6000
    lens = rng.geometric(1.0 / mean_len, size=n) - 1
    lens[rng.integers(0, n, size=5)] = 40000
    vals = [bytes(rng.integers(97, 123, size=int(l)).astype(np.uint8)).decode() for l in lens]
    s = pa.array(vals, mask=rng.random(n) < 0.03)
    v = pa.array(np.arange(n))
    for typ in (pa.string(), pa.large_string()):
        cols = [v, s.cast(typ)]          # the strings ride as an included column
        got, boff = ctx.build_index(cols, 1, 7)
        parity.check_index(oracle, _small(got), boff, _small(cols), 1, 7, f"included strings, mean {mean_len}")
        cols = [s.cast(typ), v]          # and as the key
        got, boff = ctx.build_index(cols, 1, 7)
        parity.check_index(oracle, _small(got), boff, _small(cols), 1, 7, f"string keys, mean {mean_len}")


def test_long_shared_prefixes_small_and_big_tie_groups(ctx, oracle):
    """Keys that agree on their first 7, 14, 21, ... bytes: groups of up to 64 tied rows are settled by one
    kernel comparing whole key tails, longer runs go chunk by chunk through the segmented sort; both paths, plus
    exact duplicates, composite (string, int) keys and a string that is a prefix of another."""
    rng = np.random.default_rng(77)
    base = "p" * 30
    vals = []
    for i in range(700):                                  # one big group: 700 rows share 30 bytes
        vals.append(base + "".join(chr(97 + int(x)) for x in rng.integers(0, 3, size=rng.integers(0, 12))))
    for g in range(300):                                  # many small groups: 2-9 rows share 9..40 bytes
        pre = f"g{g:04d}" + "x" * int(rng.integers(3, 35))
        for _ in range(int(rng.integers(2, 10))):
            vals.append(pre + "".join(chr(97 + int(x)) for x in rng.integers(0, 4, size=rng.integers(0, 6))))
    vals += ["dup-dup-dup-dup-dup"] * 150 + ["dup-dup-dup-dup-dup-longer"] * 5 + ["", "a", None] * 20
    order = rng.permutation(len(vals))
    vals = [vals[i] for i in order]
    s = pa.array(vals)
    k2 = pa.array(rng.integers(0, 3, size=len(vals)).astype(np.int32))
    v = pa.array(np.arange(len(vals)))
    for cols, nk in (([s, v], 1), ([s, k2, v], 2), ([k2, s, v], 2)):
        for nb in (1, 5):
            got, boff = ctx.build_index(cols, nk, nb)
            parity.check_index(oracle, got, boff, cols, nk, nb, f"tie groups nk={nk} nb={nb}")
