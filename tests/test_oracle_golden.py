"""Pins the CPU oracle against the reference's literal known answer, Spark's documented
hash answers, SURVEY.md Appendix A.6 and the quick-start fixture (CPU only)."""
import json
import os

import numpy as np
import pyarrow as pa
import pytest

from fixtures import departments_table, employees_table, sample_table

HERE = os.path.dirname(os.path.abspath(__file__))
TYPES = {"int32": pa.int32(), "int64": pa.int64(), "string": pa.string()}


def _cols(spec):
    return [pa.array(c["values"], type=TYPES[c["type"]]) for c in spec]


def _vectors():
    with open(os.path.join(HERE, "golden", "murmur3_vectors.json")) as f:
        g = json.load(f)
    return [(k, i, v) for k in ("pinned", "derived") for i, v in enumerate(g[k])]


@pytest.mark.parametrize("kind,i,vec", _vectors())
def test_murmur3_known_answers(oracle, kind, i, vec):
    cols = _cols(vec["cols"])
    if "hash" in vec:
        assert oracle.hash_rows(cols).tolist() == vec["hash"]
    if "bucket" in vec:
        assert oracle.bucket_ids(cols, vec["num_buckets"]).tolist() == vec["bucket"]


def test_pmod_is_never_negative(oracle):
    rng = np.random.default_rng(1)
    k = pa.array(rng.integers(-2**62, 2**62, size=10000))
    for nb in (1, 3, 200, 8192):
        b = oracle.bucket_ids([k], nb)
        assert b.min() >= 0 and b.max() < nb


def test_null_key_leaves_seed(oracle):
    # A.2: nulls skip, so an all-null key hashes to the seed 42
    k1 = pa.array([None, 5, None], type=pa.int32())
    k2 = pa.array([None, None, 7], type=pa.int64())
    h = oracle.hash_rows([k1, k2])
    assert h[0] == 42
    assert h[1] == oracle.hash_rows([pa.array([5], pa.int32())])[0]
    assert h[2] == oracle.hash_rows([pa.array([7], pa.int64())])[0]


def test_float_hash_canonicalises_nan_not_negzero(oracle):
    a = pa.array([float("nan"), np.float64(np.frombuffer(b"\x01\x00\x00\x00\x00\x00\xf8\xff", dtype=np.float64)[0]), 0.0, -0.0])
    h = oracle.hash_rows([a])
    assert h[0] == h[1]
    assert h[2] != h[3]


def test_quickstart_index_layout(oracle):
    # deptIndex has buckets 53 [20, 40], 122 [10], 181 [30]
    d = departments_table()
    cols = [d["deptId"], d["deptName"]]
    perm, off = oracle.build_index(cols, 1, 200)
    ids = d["deptId"].to_numpy()[perm]
    nonempty = {b: ids[off[b]:off[b + 1]].tolist() for b in range(200) if off[b + 1] > off[b]}
    assert nonempty == {53: [20, 40], 122: [10], 181: [30]}
    e = employees_table()
    perm, off = oracle.build_index([e["deptId"], e["empName"]], 1, 200)
    names = np.array(e["empName"].to_pylist())[perm]
    got = {b: sorted(names[off[b]:off[b + 1]].tolist()) for b in range(200) if off[b + 1] > off[b]}
    assert got == {53: sorted(["SMITH", "JONES", "SCOTT", "ADAMS", "FORD"]),
                   122: sorted(["CLARK", "KING", "MILLER"]),
                   181: sorted(["ALLEN", "WARD", "BLAKE", "TURNER", "JAMES", "MARTIN"])}


def test_quickstart_join_and_filter(oracle):
    d, e = departments_table(), employees_table()
    dcols, ecols = [d["deptId"], d["deptName"]], [e["deptId"], e["empName"]]
    dperm, doff = oracle.build_index(dcols, 1, 200)
    eperm, eoff = oracle.build_index(ecols, 1, 200)
    di, ei = oracle.take(dcols, dperm), oracle.take(ecols, eperm)
    li, ri, ooff = oracle.join(ei, di, 1, eoff, doff)
    rows = sorted(zip(ei[1].take(pa.array(li)).to_pylist(), di[1].take(pa.array(ri)).to_pylist()))
    dept = {10: "Accounting", 20: "Research", 30: "Sales"}
    want = sorted((n, dept[k]) for _, n, k in __import__("fixtures").EMPLOYEES)
    assert rows == want and len(rows) == 14
    # filter deptId = 20 -> deptName (App.scala:91)
    idx = oracle.filter_rows(di, [("eq", 0, 20)])
    assert di[1].take(pa.array(idx)).to_pylist() == ["Research"]


def test_sample_data_bucketing_invariant(oracle):
    # DataFrameWriterExtensionsTest.scala:93-158: bucket of every row in a "file" == recomputed
    # bucket, rows sorted inside a bucket, multiset preserved (n=3, string and int+string keys)
    t = sample_table()
    for keycols in (["Query"], ["imprs", "Query"]):
        cols = [t[c] for c in keycols] + [t["clicks"]]
        nk = len(keycols)
        perm, off = oracle.build_index(cols, nk, 3)
        assert sorted(perm.tolist()) == list(range(10))
        out = oracle.take(cols, perm)
        b = oracle.bucket_ids(out[:nk], 3)
        for k in range(3):
            assert (b[off[k]:off[k + 1]] == k).all()
            keys = list(zip(*[c.to_pylist() for c in out[:nk]]))[off[k]:off[k + 1]]
            assert keys == sorted(keys)


def test_sort_order_nulls_first_nan_last_strings_bytewise(oracle):
    a = pa.array([3.0, None, float("nan"), -0.0, 0.0, -1.5, None])
    perm = oracle.sort_segments([a], 1, [0, 7])
    vals = [a[int(i)].as_py() for i in perm]
    assert vals[0] is None and vals[1] is None
    assert vals[2] == -1.5 and vals[5] == 3.0 and vals[6] != vals[6]
    assert perm.tolist()[3:5] == [3, 4]  # -0.0 == 0.0, stable
    s = pa.array(["b", "ab", "a", "", "é", "a\x00", "abc", None])
    perm = oracle.sort_segments([s], 1, [0, 8])
    assert [s[int(i)].as_py() for i in perm] == [None, "", "a", "a\x00", "ab", "abc", "b", "é"]


def test_radix_fast_path_equals_merge_sort(oracle):
    rng = np.random.default_rng(7)
    k = pa.array(rng.integers(-1000, 1000, size=50000))
    v = pa.array(np.arange(50000))
    p1, o1 = oracle.build_index([k, v], 1, 7, use_radix=True)
    p2, o2 = oracle.build_index([k, v], 1, 7, use_radix=False)
    assert (p1 == p2).all() and (o1 == o2).all()


def test_join_null_keys_never_match_and_cross_product(oracle):
    l = [pa.array([None, 1, 1, 2, 5], type=pa.int64())]
    r = [pa.array([None, 1, 1, 1, 3, 5], type=pa.int64())]
    li, ri, off = oracle.join(l, r, 1, [0, 5], [0, 6])
    assert len(li) == 2 * 3 + 1
    assert sorted(zip(li.tolist(), ri.tolist())) == sorted([(1, 1), (1, 2), (1, 3), (2, 1), (2, 2), (2, 3), (4, 5)])


def test_filter_three_valued_logic(oracle):
    a = pa.array([1, None, 3, 4], type=pa.int32())
    b = pa.array(["x", "y", None, "x"])
    # a > 1 OR b = 'x'
    idx = oracle.filter_rows([a, b], [("gt", 0, 1), ("eq", 1, "x"), ("or",)])
    assert idx.tolist() == [0, 2, 3]
    # NOT (a > 1): row with NULL stays unknown -> dropped
    idx = oracle.filter_rows([a, b], [("gt", 0, 1), ("not",)])
    assert idx.tolist() == [0]
    idx = oracle.filter_rows([a, b], [("is_null", 0), ("is_not_null", 1), ("and",)])
    assert idx.tolist() == [1]


def test_join_types_against_pyarrow(oracle):
    """oracle.join_typed (outer/semi/anti on top of the inner pairs) against pyarrow's hash join on the same
    rows: an independent implementation with the same NULL-key rule (NULL keys match nothing)."""
    import numpy as np
    import pyarrow as pa
    rng = np.random.default_rng(5)
    mk = lambda n, hi: [pa.array(rng.integers(0, hi, size=n), mask=rng.random(n) < 0.05), pa.array(np.arange(n))]
    l, r = mk(400, 60), mk(150, 80)
    nb = 8
    lp, loff = oracle.build_index(l, 1, nb)
    rp, roff = oracle.build_index(r, 1, nb)
    lg, rg = oracle.take(l, lp), oracle.take(r, rp)
    lt = pa.table({"k": lg[0], "lv": lg[1]})
    rt = pa.table({"k": rg[0], "rv": rg[1]})
    for how, pa_how in (("inner", "inner"), ("left_outer", "left outer"), ("right_outer", "right outer"),
                        ("full_outer", "full outer"), ("left_semi", "left semi"), ("left_anti", "left anti")):
        li, ri, ooff = oracle.join_typed(lg, rg, 1, loff, roff, how)
        got_l = oracle.take_or_null(lg[1], li).to_pylist()
        got_r = oracle.take_or_null(rg[1], ri).to_pylist() if ri is not None else [None] * len(li)
        want = lt.join(rt, keys="k", join_type=pa_how)
        want_l = want["lv"].to_pylist()
        want_r = want["rv"].to_pylist() if "rv" in want.column_names else [None] * want.num_rows
        key = lambda p: (-1 if p[0] is None else p[0], -1 if p[1] is None else p[1])
        assert sorted(zip(got_l, got_r), key=key) == sorted(zip(want_l, want_r), key=key), how
        assert ooff[-1] == len(li)


# ---- a second, independent transcription of SURVEY Appendix A.1/A.2 (plain Python integers), used only here
def _py_hash_row(values, types):
    import struct
    M = 0xFFFFFFFF
    rotl = lambda x, r: ((x << r) | (x >> (32 - r))) & M
    mix_k1 = lambda k: (rotl((k * 0xCC9E2D51) & M, 15) * 0x1B873593) & M
    mix_h1 = lambda h, k: (rotl(h ^ k, 13) * 5 + 0xE6546B64) & M

    def fmix(h, n):
        h ^= n
        h ^= h >> 16
        h = (h * 0x85EBCA6B) & M
        h ^= h >> 13
        h = (h * 0xC2B2AE35) & M
        return h ^ (h >> 16)
    hash_int = lambda v, s: fmix(mix_h1(s, mix_k1(v & M)), 4)
    hash_long = lambda v, s: fmix(mix_h1(mix_h1(s, mix_k1(v & M)), mix_k1((v >> 32) & M)), 8)

    def hash_bytes(b, s):
        h, n = s, len(b)
        for i in range(0, n - n % 4, 4):
            h = mix_h1(h, mix_k1(int.from_bytes(b[i:i + 4], "little")))
        for i in range(n - n % 4, n):
            h = mix_h1(h, mix_k1((b[i] - 256 if b[i] >= 128 else b[i]) & M))
        return fmix(h, n)
    h = 42
    for v, t in zip(values, types):
        if v is None:
            continue
        if t == "bool":
            h = hash_int(1 if v else 0, h)
        elif t in ("i8", "i16", "i32"):
            h = hash_int(v, h)
        elif t == "i64":
            h = hash_long(v & 0xFFFFFFFFFFFFFFFF, h)
        elif t == "f32":
            bits = 0x7FC00000 if v != v else struct.unpack("<I", struct.pack("<f", v))[0]
            h = hash_int(bits, h)
        elif t == "f64":
            bits = 0x7FF8000000000000 if v != v else struct.unpack("<Q", struct.pack("<d", v))[0]
            h = hash_long(bits, h)
        else:
            h = hash_bytes(v.encode("utf-8") if isinstance(v, str) else bytes(v), h)
    return h - (1 << 32) if h >= (1 << 31) else h


def test_oracle_against_a_second_transcription_on_random_rows(oracle):
    """The C oracle and a from-the-spec Python transcription of Murmur3Hash(cols, 42) + pmod agree on 3000 random
    rows of every key type, with NULLs, NaN, -0.0, empty and multi-byte strings, bytes >= 0x80 in every tail
    position (the signed-byte rule) — two independent restatements of SURVEY Appendix A.1/A.2."""
    import numpy as np
    import pyarrow as pa
    rng = np.random.default_rng(2024)
    n = 3000
    null = lambda p: rng.random(n) < p
    f32 = rng.standard_normal(n).astype(np.float32)
    f64 = rng.standard_normal(n)
    f32[:5] = [np.nan, -0.0, 0.0, np.inf, -np.inf]
    f64[:5] = [np.nan, -0.0, 0.0, np.inf, -np.inf]
    alphabet = ["a", "Z", "0", "é", "日", "\\x7f", "ÿ", " "]
    strs = ["".join(rng.choice(alphabet, size=rng.integers(0, 12))) for _ in range(n)]
    bins = [bytes(rng.integers(0, 256, size=rng.integers(0, 11)).astype(np.uint8)) for _ in range(n)]
    cols = [pa.array(rng.random(n) < 0.5, mask=null(0.1)), pa.array(rng.integers(-128, 128, size=n).astype(np.int8)),
            pa.array(rng.integers(-2**15, 2**15, size=n).astype(np.int16), mask=null(0.1)),
            pa.array(rng.integers(-2**31, 2**31, size=n).astype(np.int32)),
            pa.array(rng.integers(-2**63, 2**63 - 1, size=n), mask=null(0.1)), pa.array(f32), pa.array(f64, mask=null(0.1)),
            pa.array(strs, mask=null(0.1)), pa.array(bins, type=pa.binary())]
    types = ["bool", "i8", "i16", "i32", "i64", "f32", "f64", "str", "bin"]
    py = [c.to_pylist() for c in cols]
    for subset in ([0], [1], [2], [3], [4], [5], [6], [7], [8], [3, 7], [7, 4, 6], list(range(9))):
        want = np.array([_py_hash_row([py[j][i] for j in subset], [types[j] for j in subset]) for i in range(n)], dtype=np.int64)
        got = oracle.hash_rows([cols[j] for j in subset]).astype(np.int64)
        assert np.array_equal(got, want), f"hash of columns {subset} differs"
        for nb in (1, 7, 200, 8192):
            b = oracle.bucket_ids([cols[j] for j in subset], nb)
            assert np.array_equal(b, ((want % nb) + nb) % nb), f"bucket ids of columns {subset}, {nb} buckets"


def test_oracle_build_against_python_sorted_on_random_rows(oracle):
    """hso_build_index against bucket + sort done with Python's sorted() over the same rows: bucket of every row
    from the second Murmur3 transcription, order inside a bucket ASC NULLS FIRST with NaN greatest, -0.0 == 0.0
    and strings compared as unsigned bytes."""
    import math
    import numpy as np
    import pyarrow as pa
    rng = np.random.default_rng(77)
    n, nb = 4000, 13
    f = rng.integers(-3, 4, size=n) / 2.0
    f[rng.random(n) < 0.05] = np.nan
    f[rng.random(n) < 0.05] = -0.0
    strs = ["".join(rng.choice(["a", "b", "é", "\\x7f", "ÿ"], size=rng.integers(0, 4))) for _ in range(n)]
    k0 = pa.array(rng.integers(-5, 6, size=n).astype(np.int32), mask=rng.random(n) < 0.1)
    k1 = pa.array(f, mask=rng.random(n) < 0.1)
    k2 = pa.array(strs, mask=rng.random(n) < 0.1)
    rid = pa.array(np.arange(n))
    types = ["i32", "f64", "str"]
    py = [k0.to_pylist(), k1.to_pylist(), k2.to_pylist()]

    def cell_key(v, t):
        if v is None:
            return (0,)
        if t == "f64":
            return (1, 1, 0.0) if math.isnan(v) else (1, 0, v + 0.0)     # NaN greatest; -0.0 + 0.0 == 0.0
        if t == "str":
            return (1, v.encode("utf-8"))
        return (1, v)
    perm, off = oracle.build_index([k0, k1, k2, rid], 3, nb)
    h = [_py_hash_row([py[j][i] for j in range(3)], types) for i in range(n)]
    bucket = [((x % nb) + nb) % nb for x in h]
    want_rows = sorted(range(n), key=lambda i: (bucket[i], tuple(cell_key(py[j][i], types[j]) for j in range(3))))
    assert list(np.diff(off)) == [bucket.count(b) for b in range(nb)]
    key_of = lambda i: tuple(cell_key(py[j][i], types[j]) for j in range(3))
    assert [key_of(i) for i in perm.tolist()] == [key_of(i) for i in want_rows]
    assert sorted(perm.tolist()) == list(range(n))
