"""K5-K7 parity: join and filter results equal the oracle's row multiset (the reference's verifyIndexUsage
invariant, E2EHyperspaceRulesTest.scala:1004-1019: indexed result == un-indexed result)."""
import numpy as np
import pyarrow as pa
import pytest

import parity
from fixtures import EMPLOYEES, departments_table, employees_table, lineitem, orders, string_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def _oracle_join(oracle, lcols, loff, rcols, roff, nk, lout, rout):
    li, ri, woff = oracle.join(lcols, rcols, nk, loff, roff)
    want = [lcols[i].take(pa.array(li)) for i in lout] + [rcols[i].take(pa.array(ri)) for i in rout]
    return want, woff


def test_quickstart_join_and_filter(ctx):
    d, e = departments_table(), employees_table()
    dg, doff = ctx.build_index([d["deptId"], d["deptName"]], 1, 200)
    eg, eoff = ctx.build_index([e["deptId"], e["empName"]], 1, 200)
    jg, joff = ctx.join(eg, eoff, dg, doff, 1, [1], [1])
    dept = {10: "Accounting", 20: "Research", 30: "Sales"}
    assert sorted(zip(jg[0].to_pylist(), jg[1].to_pylist())) == sorted((n, dept[k]) for _, n, k in EMPLOYEES)
    assert joff[-1] == 14
    fg = ctx.filter_scan(dg, 1, doff, [("eq", 0, 20)], [1])
    assert fg[0].to_pylist() == ["Research"]


@pytest.mark.parametrize("n_orders,nb", [(2000, 200), (60000, 200), (60000, 3)])
def test_lineitem_orders_join(ctx, oracle, n_orders, nb):
    lcols, keys = lineitem(n_orders)
    ocols = orders(keys)
    lg, loff = ctx.build_index(lcols, 1, nb)
    og, ooff = ctx.build_index(ocols, 1, nb)
    jg, joff = ctx.join(lg, loff, og, ooff, 1, [0, 1, 2], [1])
    assert len(jg[0]) == len(lcols[0])  # every lineitem row has exactly one order
    want, woff = _oracle_join(oracle, lg, loff, og, ooff, 1, [0, 1, 2], [1])
    parity.assert_same_multiset(jg, want, joff, woff, "lineitem x orders")


def test_join_many_to_many_nulls_and_misses(ctx, oracle):
    rng = np.random.default_rng(21)
    nl, nr = 40000, 30000
    lk = pa.array(rng.integers(0, 5000, size=nl), mask=rng.random(nl) < 0.05)
    rk = pa.array(rng.integers(2500, 7500, size=nr), mask=rng.random(nr) < 0.05)
    lcols, rcols = [lk, pa.array(np.arange(nl))], [rk, pa.array(np.arange(nr) * 1.5)]
    lg, loff = ctx.build_index(lcols, 1, 16)
    rg, roff = ctx.build_index(rcols, 1, 16)
    jg, joff = ctx.join(lg, loff, rg, roff, 1, [0, 1], [0, 1])
    want, woff = _oracle_join(oracle, lg, loff, rg, roff, 1, [0, 1], [0, 1])
    assert len(want[0]) > nl  # really many-to-many
    parity.assert_same_multiset(jg, want, joff, woff, "many-to-many")


@pytest.mark.parametrize("dtype", [np.int64, np.int32])
def test_fast_join_many_to_many_and_wide_windows(ctx, oracle, dtype):
    # no NULL keys, fixed-width outputs -> the merge-path tile kernels; heavy right-side duplicates make some
    # windows larger than the staged chunk (global probes), sparse keys leave tiles without any match
    rng = np.random.default_rng(77)
    nl, nr = 120000, 90000
    lk = rng.integers(0, 3000, size=nl).astype(dtype)
    rk = np.concatenate([rng.integers(1000, 5000, size=nr - 30000), np.full(30000, 1234)]).astype(dtype)
    lcols = [pa.array(lk), pa.array(np.arange(nl, dtype=np.int64)), pa.array(rng.random(nl).astype(np.float32))]
    rcols = [pa.array(rk), pa.array(np.arange(nr) * 0.5)]
    for nb in (4, 64):
        lg, loff = ctx.build_index(lcols, 1, nb)
        rg, roff = ctx.build_index(rcols, 1, nb)
        jg, joff = ctx.join(lg, loff, rg, roff, 1, [0, 1, 2], [0, 1])
        want, woff = _oracle_join(oracle, lg, loff, rg, roff, 1, [0, 1, 2], [0, 1])
        assert len(want[0]) > 2 * nl
        parity.assert_same_multiset(jg, want, joff, woff, f"fast join {dtype.__name__}/{nb}")
    # one side empty
    eg, eoff = ctx.build_index([pa.array(np.zeros(0, dtype=dtype)), pa.array(np.zeros(0))], 1, 4)
    lg, loff = ctx.build_index(lcols, 1, 4)
    jg, joff = ctx.join(lg, loff, eg, eoff, 1, [0], [1])
    assert len(jg[0]) == 0 and joff.tolist() == [0] * 5


def test_join_on_string_and_composite_keys(ctx, oracle):
    s, v = string_keys(5000)
    idx = np.random.default_rng(1).integers(0, 5000, size=12000)
    ls = s.take(pa.array(idx))
    lg, loff = ctx.build_index([ls, pa.array(np.arange(12000))], 1, 8)
    rg, roff = ctx.build_index([s, v], 1, 8)
    jg, joff = ctx.join(lg, loff, rg, roff, 1, [0, 1], [1])
    want, woff = _oracle_join(oracle, lg, loff, rg, roff, 1, [0, 1], [1])
    parity.assert_same_multiset(jg, want, joff, woff, "string join")
    rng = np.random.default_rng(2)
    mk = lambda n: [pa.array(rng.integers(0, 30, size=n).astype(np.int32)), pa.array(rng.integers(0, 40, size=n)),
                    pa.array(rng.random(n))]
    l, r = mk(20000), mk(3000)
    lg, loff = ctx.build_index(l, 2, 32)
    rg, roff = ctx.build_index(r, 2, 32)
    jg, joff = ctx.join(lg, loff, rg, roff, 2, [0, 1, 2], [2])
    want, woff = _oracle_join(oracle, lg, loff, rg, roff, 2, [0, 1, 2], [2])
    parity.assert_same_multiset(jg, want, joff, woff, "composite join")


def test_join_packed_key_tuples(ctx, oracle):
    """NULL-free fixed-width key tuples are packed into one u64 per row and take the tile-merge path: small
    integer types, floats (NaN, -0.0), booleans, three-column keys; many-to-many matches included."""
    rng = np.random.default_rng(9)

    def side(n, kinds):
        cols = []
        for kind in kinds:
            if kind == "i8":
                cols.append(pa.array(rng.integers(-5, 6, size=n).astype(np.int8)))
            elif kind == "i16":
                cols.append(pa.array(rng.integers(-300, 300, size=n).astype(np.int16)))
            elif kind == "i32":
                cols.append(pa.array(rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32) // (1 << 22)))
            elif kind == "i64":
                cols.append(pa.array(rng.integers(-40, 40, size=n) * (1 << 40)))
            elif kind == "f32":
                v = rng.integers(-4, 5, size=n).astype(np.float32) / 2
                v[rng.random(n) < 0.05] = np.nan
                v[rng.random(n) < 0.05] = -0.0
                cols.append(pa.array(v))
            elif kind == "f64":
                v = rng.integers(-4, 5, size=n).astype(np.float64) / 4
                v[rng.random(n) < 0.05] = np.nan
                v[rng.random(n) < 0.05] = -0.0
                cols.append(pa.array(v))
            elif kind == "bool":
                cols.append(pa.array(rng.random(n) < 0.5))
        return cols + [pa.array(np.arange(n)), pa.array(rng.random(n).astype(np.float32))]

    for kinds in (["i16"], ["f64"], ["f32", "i8"], ["bool", "i32", "i16"], ["i64", "f64"], ["i8", "i8", "i8"]):
        nk = len(kinds)
        l, r = side(6000, kinds), side(900, kinds)
        lg, loff = ctx.build_index(l, nk, 16)
        rg, roff = ctx.build_index(r, nk, 16)
        lout, rout = list(range(nk + 2)), [nk, nk + 1]
        jg, joff = ctx.join(lg, loff, rg, roff, nk, lout, rout)
        want, woff = _oracle_join(oracle, lg, loff, rg, roff, nk, lout, rout)
        parity.assert_same_multiset(jg, want, joff, woff, f"packed-key join {kinds}")


@pytest.mark.parametrize("how", ["left_outer", "right_outer", "full_outer", "left_semi", "left_anti"])
def test_join_types(ctx, oracle, how):
    """Outer, semi and anti joins over two indexes (JoinIndexRule accepts every join type): NULL keys, keys
    missing on either side, many-to-many matches, empty buckets, string and composite keys, string payloads."""
    rng = np.random.default_rng(41)

    def check(l, r, nk, nb, what):
        lg, loff = ctx.build_index(l, nk, nb)
        rg, roff = ctx.build_index(r, nk, nb)
        lout = list(range(len(l)))
        rout = [] if how in ("left_semi", "left_anti") else list(range(len(r)))
        jg, joff = ctx.join(lg, loff, rg, roff, nk, lout, rout, how=how)
        li, ri, woff = oracle.join_typed(lg, rg, nk, loff, roff, how)
        want = [oracle.take_or_null(lg[i], li) for i in lout] + [oracle.take_or_null(rg[i], ri) for i in rout]
        parity.assert_same_multiset(jg, want, joff, woff, f"{how} join, {what}")

    mk = lambda n, hi, nullp: [pa.array(rng.integers(0, hi, size=n), mask=rng.random(n) < nullp),
                               pa.array(rng.random(n)), pa.array([f"s{int(x)}" * int(x % 4) for x in rng.integers(0, 99, size=n)])]
    check(mk(5000, 900, 0.03), mk(1200, 1100, 0.03), 1, 16, "int keys with NULLs")
    check(mk(300, 40, 0.0), mk(20, 4000, 0.0), 1, 64, "mostly empty buckets")
    check(mk(50, 10, 0.0), [c.slice(0, 0) for c in mk(5, 10, 0.0)], 1, 4, "empty right side")
    s, v = string_keys(3000)
    ls = s.take(pa.array(rng.integers(0, 3000, size=5000)))
    check([ls, pa.array(np.arange(5000))], [s.slice(0, 2000), v.slice(0, 2000)], 1, 8, "string keys")
    comp = lambda n: [pa.array(rng.integers(0, 20, size=n).astype(np.int32)),
                      pa.array(rng.integers(0, 30, size=n), mask=rng.random(n) < 0.02), pa.array(rng.random(n))]
    check(comp(4000), comp(700), 2, 32, "composite keys")


def test_join_after_incremental_refresh_two_runs_per_bucket(ctx, oracle):
    lcols, keys = lineitem(20000)
    n = len(lcols[0])
    a = [c.slice(0, n * 9 // 10) for c in lcols]
    b = [c.slice(n * 9 // 10) for c in lcols]
    ga, oa = ctx.build_index(a, 1, 40)
    gb, ob = ctx.build_index(b, 1, 40)
    lg, loff = ctx.merge_runs([(ga, 1, oa), (gb, 1, ob)])
    og, ooff = ctx.build_index(orders(keys), 1, 40)
    jg, joff = ctx.join(lg, loff, og, ooff, 1, [0, 1, 2], [1])
    full, foff = ctx.build_index(lcols, 1, 40)
    want, woff = _oracle_join(oracle, full, foff, og, ooff, 1, [0, 1, 2], [1])
    parity.assert_same_multiset(jg, want, joff, woff, "join over merged runs")


def test_join_with_different_bucket_counts(ctx, oracle):
    """JoinIndexRule pairs indexes whose bucket counts differ too (JoinIndexRule.scala:515-516; Spark then
    re-shuffles one side): the side with fewer rows is re-bucketed on the fly to the other side's count, and the
    rows are those of a join over two indexes built with that count."""
    lcols, keys = lineitem(8000, null_frac=0.01)
    lg, loff = ctx.build_index(lcols, 1, 24)
    og, ooff = ctx.build_index(orders(keys), 1, 7)
    for how in ("inner", "left_outer", "full_outer"):
        jg, joff = ctx.join(lg, loff, og, ooff, 1, [0, 1, 2], [0, 1], how=how)
        assert len(joff) == 25                       # the larger side (lineitem) keeps its 24 buckets
        og24, ooff24 = ctx.build_index(orders(keys), 1, 24)
        li, ri, woff = oracle.join_typed(lg, og24, 1, loff, ooff24, how)
        want = [oracle.take_or_null(lg[i], li) for i in (0, 1, 2)] + [oracle.take_or_null(og24[i], ri) for i in (0, 1)]
        parity.assert_same_multiset(jg, want, joff, woff, f"{how} join, 24 x 7 buckets")
    # the other way round: the left side is the small one
    jg, joff = ctx.join(og, ooff, lg, loff, 1, [1], [0, 2])
    li, ri, woff = oracle.join(og24, lg, 1, ooff24, loff)
    want = [og24[1].take(pa.array(li)), lg[0].take(pa.array(ri)), lg[2].take(pa.array(ri))]
    parity.assert_same_multiset(jg, want, joff, woff, "7 x 24 buckets")


def test_filter_predicates(ctx, oracle):
    rng = np.random.default_rng(31)
    cols, keys = lineitem(30000, null_frac=0.02)
    s, _ = string_keys(len(cols[0]), seed=8)
    cols = cols + [pa.array(s.to_pylist(), mask=rng.random(len(s)) < 0.1)]
    got, boff = ctx.build_index(cols, 1, 64)
    k = int(keys[100])
    probe = s[17].as_py()
    preds = [
        [("eq", 0, k)],
        [("eq", 0, -5)],
        [("gt", 0, k), ("lt", 1, 50000.0), ("and",)],
        [("le", 0, k), ("ge", 2, 0.05), ("or",)],
        [("is_null", 0)],
        [("is_null", 0), ("not",)],
        [("gt", 0, k), ("not",)],
        [("eq", 3, probe)],
        [("ge", 3, "m"), ("is_not_null", 0), ("and",), ("lt", 2, 0.03), ("or",)],
        [("ne", 2, 0.05), ("is_null", 3), ("and",)],
        [("in", 0, [k, int(keys[5]), -7, int(keys[200])])],               # IN list (the lineage filter's op)
        [("in", 0, [k, int(keys[5])]), ("not",)],                           # NOT IN: NULL keys stay out (unknown)
        [("in", 0, []), ("not",), ("gt", 1, 90000.0), ("and",)],
    ]
    for p in preds:
        fg = ctx.filter_scan(got, 1, boff, p, [0, 1, 3])
        idx = pa.array(oracle.filter_rows(got, p))
        want = [got[0].take(idx), got[1].take(idx), got[3].take(idx)]
        parity.assert_same_multiset(fg, want, what=f"filter {p}")


def test_filter_packed_outputs_keep_row_order(ctx, oracle):
    """Fixed-width NULL-free output columns take the two-pass bitmask path: same rows as the oracle, and in
    the order of the index (a scan of a sorted run stays sorted). Sizes around the 2048-row tile, a pruned
    bucket (scan starting in the middle of the table), predicates that select nothing and everything."""
    for n_orders in (300, 5000, 40000):
        cols, keys = lineitem(n_orders)
        got, boff = ctx.build_index(cols, 1, 16)
        k = int(keys[len(keys) // 2])
        for p in ([("ge", 0, k)], [("eq", 0, k)], [("lt", 1, 30000.0), ("ge", 2, 0.05), ("and",)],
                  [("lt", 0, -1)], [("ge", 0, -1)], [("eq", 0, k), ("not",)],
                  # one comparison / a range on the integer key: the specialised mask kernel
                  [("ne", 0, k)], [("le", 0, k)], [("gt", 0, k)], [("ge", 0, k // 2), ("lt", 0, k), ("and",)],
                  [("gt", 0, k), ("le", 0, k // 2), ("and",)],
                  [("in", 0, [int(x) for x in keys[::7]])], [("in", 0, [int(x) for x in keys[::3]]), ("not",)],
                  [("gt", 2, 0.02), ("le", 0, k), ("or",), ("ne", 1, 0.0), ("and",)]):
            fg = ctx.filter_scan(got, 1, boff, p, [2, 0, 1])
            idx = pa.array(oracle.filter_rows(got, p))
            want = [got[2].take(idx), got[0].take(idx), got[1].take(idx)]
            for g, w in zip(fg, want):
                assert g.equals(w), f"filter {p} on {len(got[0])} rows: rows or their order differ"


def test_filter_bucket_pruning_matches_full_scan(ctx, oracle):
    s, v = string_keys(8000)
    got, boff = ctx.build_index([s, v], 1, 100)
    for probe in (s[5].as_py(), s[4000].as_py(), "no such key", ""):
        fg = ctx.filter_scan(got, 1, boff, [("eq", 0, probe)], [0, 1])
        idx = pa.array(oracle.filter_rows(got, [("eq", 0, probe)]))
        parity.assert_same_multiset(fg, [got[0].take(idx), got[1].take(idx)], what=f"pruned {probe!r}")
    k32 = pa.array(np.arange(-500, 500, dtype=np.int32))
    got, boff = ctx.build_index([k32, pa.array(np.arange(1000))], 1, 200)
    for lit in (-500, 0, 499, 1000, 2**40):
        fg = ctx.filter_scan(got, 1, boff, [("eq", 0, lit)], [1])
        idx = pa.array(oracle.filter_rows(got, [("eq", 0, lit)]))
        parity.assert_same_multiset(fg, [got[1].take(idx)], what=f"pruned int {lit}")
