"""K1+K2+K3 parity: bucket of every row, sorted key sequence per bucket (bit-exact) and the multiset of
included-column rows per bucket, against the oracle. Mirrors the invariants of the reference's
DataFrameWriterExtensionsTest.scala:93-158 and IndexManagerTest."""
import numpy as np
import pyarrow as pa
import pytest

import parity
from fixtures import departments_table, employees_table, lineitem, sample_table, string_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def test_quickstart_layout(ctx, oracle):
    d, e = departments_table(), employees_table()
    got, boff = ctx.build_index([d["deptId"], d["deptName"]], 1, 200)
    ids = got[0].to_pylist()
    nonempty = {b: ids[boff[b]:boff[b + 1]] for b in range(200) if boff[b + 1] > boff[b]}
    assert nonempty == {53: [20, 40], 122: [10], 181: [30]}
    assert got[1].to_pylist() == ["Research", "Operations", "Accounting", "Sales"]
    cols = [e["deptId"], e["empName"]]
    got, boff = ctx.build_index(cols, 1, 200)
    parity.check_index(oracle, got, boff, cols, 1, 200, "empIndex")


@pytest.mark.parametrize("n_orders,nb", [(1, 200), (50, 200), (3000, 200), (40000, 200), (40000, 7), (300000, 200),
                                         (300000, 3)])
def test_lineitem_shape(ctx, oracle, n_orders, nb):
    cols, _ = lineitem(n_orders)
    got, boff = ctx.build_index(cols, 1, nb)
    parity.check_index(oracle, got, boff, cols, 1, nb, f"lineitem {n_orders}/{nb}")


def test_lineitem_with_null_keys(ctx, oracle):
    cols, _ = lineitem(60000, null_frac=0.01)
    got, boff = ctx.build_index(cols, 1, 200)
    assert got[0].null_count == cols[0].null_count
    parity.check_index(oracle, got, boff, cols, 1, 200, "null keys")


def test_empty_input(ctx):
    cols = [pa.array([], pa.int64()), pa.array([], pa.float64()), pa.array([], pa.string())]
    got, boff = ctx.build_index(cols, 1, 200)
    assert [len(c) for c in got] == [0, 0, 0] and boff.tolist() == [0] * 201


def test_full_range_int64_and_int32_keys(ctx, oracle):
    rng = np.random.default_rng(3)
    n = 200000
    k = rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)
    k[:5] = [np.iinfo(np.int64).min, np.iinfo(np.int64).max, 0, -1, 1]
    cols = [pa.array(k), pa.array(np.arange(n, dtype=np.int32))]
    got, boff = ctx.build_index(cols, 1, 16)
    parity.check_index(oracle, got, boff, cols, 1, 16, "int64 full range")
    k32 = rng.integers(-2**31, 2**31 - 1, size=n, dtype=np.int64).astype(np.int32)
    cols = [pa.array(k32), pa.array(rng.random(n))]
    got, boff = ctx.build_index(cols, 1, 5)
    parity.check_index(oracle, got, boff, cols, 1, 5, "int32 full range")


def test_heavy_duplicates_and_skew(ctx, oracle):
    rng = np.random.default_rng(4)
    n = 300000
    k = np.where(rng.random(n) < 0.6, 12345, rng.integers(0, 50, size=n)).astype(np.int64)
    cols = [pa.array(k), pa.array(np.arange(n, dtype=np.int64))]
    got, boff = ctx.build_index(cols, 1, 8)
    parity.check_index(oracle, got, boff, cols, 1, 8, "hot key")
    # a single key: one bucket holds everything, larger than a leaf, nothing to sort
    cols = [pa.array(np.full(20000, 7, dtype=np.int64)), pa.array(np.arange(20000, dtype=np.int64))]
    got, boff = ctx.build_index(cols, 1, 4)
    parity.check_index(oracle, got, boff, cols, 1, 4, "constant key")
    # zipf-like keys with outliers: the MSD recursion goes several levels deep
    k = (rng.zipf(1.3, size=n) % (1 << 40)).astype(np.int64)
    k[:3] = [2**62, -2**62, 0]
    cols = [pa.array(k), pa.array(rng.random(n))]
    got, boff = ctx.build_index(cols, 1, 2)
    parity.check_index(oracle, got, boff, cols, 1, 2, "zipf")


def test_composite_int32_int64_key(ctx, oracle):
    # cfg5 shape: (k1:int32 < 2^20, k2:int64 < 2^40, v:float64)
    rng = np.random.default_rng(20240604)
    n = 250000
    cols = [pa.array(rng.integers(0, 1 << 20, size=n).astype(np.int32)),
            pa.array(rng.integers(0, 1 << 40, size=n)), pa.array(rng.random(n))]
    got, boff = ctx.build_index(cols, 2, 1024)
    parity.check_index(oracle, got, boff, cols, 2, 1024, "composite")
    # few distinct k1 so that k2 decides the order, with nulls in both
    cols = [pa.array(rng.integers(0, 4, size=n).astype(np.int32), mask=rng.random(n) < 0.05),
            pa.array(rng.integers(-1000, 1000, size=n), mask=rng.random(n) < 0.05), pa.array(np.arange(n))]
    got, boff = ctx.build_index(cols, 2, 16)
    parity.check_index(oracle, got, boff, cols, 2, 16, "composite with nulls")


def test_wide_composite_key_needs_two_chunks(ctx, oracle):
    rng = np.random.default_rng(9)
    n = 60000
    cols = [pa.array(rng.integers(-2**62, 2**62, size=n) // (1 << 50) * (1 << 50)),
            pa.array(rng.integers(-2**62, 2**62, size=n)), pa.array(np.arange(n, dtype=np.int32))]
    got, boff = ctx.build_index(cols, 2, 4)
    parity.check_index(oracle, got, boff, cols, 2, 4, "two-chunk key")


def test_float_keys_nan_and_negative_zero(ctx, oracle):
    rng = np.random.default_rng(10)
    n = 50000
    f = rng.standard_normal(n)
    f[::50] = np.nan
    f[1::50] = -0.0
    f[2::50] = 0.0
    f[3::50] = np.inf
    f[4::50] = -np.inf
    cols = [pa.array(f, mask=rng.random(n) < 0.02), pa.array(np.arange(n))]
    got, boff = ctx.build_index(cols, 1, 4)
    # -0.0 and 0.0 hash to different buckets but compare equal: check through the oracle's comparator
    perm, woff = oracle.build_index(cols, 1, 4)
    assert np.array_equal(boff, woff)
    want = oracle.take(cols, perm)
    g, w = got[0].to_numpy(zero_copy_only=False), want[0].to_numpy(zero_copy_only=False)
    assert np.array_equal(np.isnan(g), np.isnan(w))
    assert np.array_equal(np.nan_to_num(g, nan=0.0) + 0.0, np.nan_to_num(w, nan=0.0) + 0.0)
    assert sorted(got[1].to_pylist()) == list(range(n))
    f32 = pa.array(f.astype(np.float32))
    got, boff = ctx.build_index([f32, pa.array(np.arange(n))], 1, 3)
    perm, woff = oracle.build_index([f32, pa.array(np.arange(n))], 1, 3)
    assert np.array_equal(boff, woff)


def test_string_keys(ctx, oracle):
    t = sample_table()
    for keycols in (["Query"], ["imprs", "Query"], ["Query", "clicks"], ["RGUID"]):
        cols = [t[c].combine_chunks() for c in keycols] + [t["clicks"].combine_chunks(), t["Date"].combine_chunks()]
        got, boff = ctx.build_index(cols, len(keycols), 3)
        parity.check_index(oracle, got, boff, cols, len(keycols), 3, f"SampleData {keycols}")
    s, v = string_keys(20000)
    got, boff = ctx.build_index([s, v], 1, 10)
    parity.check_index(oracle, got, boff, [s, v], 1, 10, "cfg4-shaped strings")


def test_string_edge_cases(ctx, oracle):
    vals = ["", "a", "a\x00", "a\x00\x00", "ab", "abcdefg", "abcdefgh", "abcdefg\x00", "abcdefgha", None,
            "abcdefghijklmnopqrstuvwxyz", "abcdefghijklmnopqrstuvwxy", "é", "日本語", "abcdefghijklmn", "abcdefghijklmno",
            None, ""] * 50
    rng = np.random.default_rng(2)
    rng.shuffle(vals)
    cols = [pa.array(vals), pa.array(np.arange(len(vals)))]
    got, boff = ctx.build_index(cols, 1, 2)
    parity.check_index(oracle, got, boff, cols, 1, 2, "string edge cases")
    b = pa.array([bytes([x % 256, 255, 0, x % 7]) * (x % 5) for x in range(3000)], type=pa.binary())
    got, boff = ctx.build_index([b, pa.array(np.arange(3000))], 1, 4)
    parity.check_index(oracle, got, boff, [b, pa.array(np.arange(3000))], 1, 4, "binary keys")


def test_included_columns_of_every_type_with_nulls(ctx, oracle):
    rng = np.random.default_rng(12)
    n = 30000
    m = lambda p: rng.random(n) < p
    s, _ = string_keys(n, seed=5)
    cols = [pa.array(rng.integers(0, 5000, size=n)),
            pa.array(rng.integers(0, 2, size=n).astype(bool), mask=m(0.1)),
            pa.array(rng.integers(-100, 100, size=n).astype(np.int8), mask=m(0.1)),
            pa.array(rng.integers(-1000, 1000, size=n).astype(np.int16)),
            pa.array(rng.random(n).astype(np.float32), mask=m(0.3)),
            pa.array(s.to_pylist(), mask=m(0.2)),
            pa.array(rng.integers(0, 10**9, size=n).astype("datetime64[s]"))]
    got, boff = ctx.build_index(cols, 1, 20)
    parity.check_index(oracle, got, boff, cols, 1, 20, "payload types")


def test_merge_runs_equals_rebuild(ctx, oracle):
    # an incremental refresh leaves two sorted runs per bucket (RefreshIncrementalAction.scala:129-144);
    # merging them must equal an index built over all rows
    cols, _ = lineitem(30000)
    n = len(cols[0])
    a = [c.slice(0, n * 9 // 10) for c in cols]
    b = [c.slice(n * 9 // 10) for c in cols]
    ga, oa = ctx.build_index(a, 1, 50)
    gb, ob = ctx.build_index(b, 1, 50)
    merged, moff = ctx.merge_runs([(ga, 1, oa), (gb, 1, ob)])
    parity.check_index(oracle, merged, moff, cols, 1, 50, "merge runs")


@pytest.mark.parametrize("by_sort", ["0", "1"])
def test_merge_path_of_sorted_runs(ctx, oracle, monkeypatch, by_sort):
    """Fixed-width tables merge along the merge path (no re-sort): keys duplicated inside and across runs,
    NULLs in included columns, three runs, an empty run, buckets present in one run only, composite and small
    integer keys, float keys; HSGPU_MERGE_BY_SORT=1 forces the concatenate-and-sort path on the same inputs.
    Both must equal an index built over all rows."""
    monkeypatch.setenv("HSGPU_MERGE_BY_SORT", by_sort)
    rng = np.random.default_rng(8)

    def table(n, kinds):
        cols = []
        for kind in kinds:
            if kind == "i64":
                cols.append(pa.array(rng.integers(0, 400, size=n) * 7))
            elif kind == "i32":
                cols.append(pa.array(rng.integers(-50, 50, size=n).astype(np.int32)))
            elif kind == "i8":
                cols.append(pa.array(rng.integers(-5, 5, size=n).astype(np.int8)))
            elif kind == "f64":
                cols.append(pa.array(rng.integers(-8, 8, size=n) / 4.0))
        return cols + [pa.array(rng.random(n), mask=rng.random(n) < 0.2), pa.array(np.arange(n).astype(np.int16))]

    for kinds, nb in ((["i64"], 13), (["i32", "i64"], 8), (["i8"], 4), (["f64", "i8"], 5)):
        nk = len(kinds)
        parts = [table(n, kinds) for n in (9000, 2500, 0, 4100)]
        runs = []
        for p in parts:
            g, o = ctx.build_index(p, nk, nb)
            runs.append((g, nk, o))
        merged, moff = ctx.merge_runs(runs)
        allcols = [pa.concat_arrays([p[i] for p in parts]) for i in range(len(parts[0]))]
        parity.check_index(oracle, merged, moff, allcols, nk, nb, f"merge of 4 runs, key {kinds}, by_sort={by_sort}")
    # one key only in run A, another only in run B
    a, oa = ctx.build_index([pa.array([5] * 3000), pa.array(np.arange(3000))], 1, 6)
    b, ob = ctx.build_index([pa.array([6] * 10), pa.array(np.arange(10))], 1, 6)
    merged, moff = ctx.merge_runs([(a, 1, oa), (b, 1, ob)])
    parity.check_index(oracle, merged, moff, [pa.array([5] * 3000 + [6] * 10), pa.array(list(range(3000)) + list(range(10)))],
                       1, 6, "disjoint buckets")


def test_unsupported_inputs_fail_loudly(ctx, hs):
    with pytest.raises(hs.HyperspaceException):
        ctx.build_index([pa.array([1, 2, 3])], 1, 0)
    with pytest.raises(hs.HyperspaceException):
        ctx.build_index([pa.array([1, 2, 3]), pa.array([1.0, 2.0])], 1, 4)
