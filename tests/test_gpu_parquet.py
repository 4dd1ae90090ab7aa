"""K4 on the device: hs_parquet_plain_pages encodes the page bodies (definition levels + PLAIN values) and
parquet_plain.py frames them into files. pyarrow's Parquet reader — an independent implementation of the format —
must read back exactly the rows that went in, for every supported type, with NULLs, across page boundaries."""
import datetime
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

from fixtures import lineitem, string_keys

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def _table(n, seed):
    rng = np.random.default_rng(seed)
    m = lambda p: rng.random(n) < p
    s, _ = string_keys(n, seed=seed)
    return [
        pa.array(rng.integers(-2**62, 2**62, size=n), mask=m(0.1)),
        pa.array(rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32)),
        pa.array(rng.integers(-128, 127, size=n).astype(np.int8), mask=m(0.3)),
        pa.array(rng.integers(-30000, 30000, size=n).astype(np.int16)),
        pa.array(rng.random(n), mask=m(0.5)),
        pa.array(rng.random(n).astype(np.float32)),
        pa.array(rng.random(n) < 0.4, mask=m(0.2)),
        pa.array(s.to_pylist(), mask=m(0.15)),
        pa.array([bytes(rng.integers(0, 256, size=rng.integers(0, 9)).astype(np.uint8)) for _ in range(n)], type=pa.binary()),
        pa.array([None] * n, type=pa.int64()),                                         # a column of NULLs only
        pa.array(rng.integers(0, 20000, size=n).astype(np.int32), type=pa.int32()).cast(pa.date32()),
        pa.array(rng.integers(0, 2**40, size=n), type=pa.int64()).cast(pa.timestamp("us")),
        pa.array(["" if i % 3 else "x" * (i % 50) for i in range(n)]).cast(pa.large_string()),
    ]


@pytest.mark.parametrize("n,page_rows", [(1, 1 << 20), (777, 100), (5000, 1 << 20), (20000, 4096)])
def test_files_read_back_by_pyarrow(ctx, tmp_path, monkeypatch, n, page_rows):
    from hyperspace_1_b200 import parquet_plain
    monkeypatch.setattr(parquet_plain, "PAGE_ROWS", page_rows)
    cols = _table(n, n)
    names = [f"c{i}" for i in range(len(cols))]
    cuts = sorted(set([0, n] + [int(x) for x in np.random.default_rng(1).integers(0, n + 1, size=5)]))
    boff = np.array(cuts + [n, n], dtype=np.int64)                                   # a few empty buckets too
    files = parquet_plain.write_buckets_plain(ctx, cols, names, boff, str(tmp_path / "out"))
    nonempty = [b for b in range(len(boff) - 1) if boff[b + 1] > boff[b]]
    assert len(files) == len(nonempty)
    from hyperspace_1_b200 import log_entry as le
    for f, b in zip(sorted(files, key=le.bucket_id_of), nonempty):
        assert le.bucket_id_of(f) == b and os.path.basename(f).endswith(".c000.parquet")
        md = pq.ParquetFile(f).metadata
        assert md.num_rows == boff[b + 1] - boff[b] and md.num_row_groups == 1
        got = pq.ParquetFile(f).read()
        assert got.column_names == names
        for i, c in enumerate(cols):
            want = c.slice(int(boff[b]), int(boff[b + 1] - boff[b]))
            g = got.column(i).combine_chunks()
            if pa.types.is_large_string(want.type):
                want = want.cast(pa.string())
            if pa.types.is_timestamp(want.type):     # TIMESTAMP_MICROS is an instant (what Spark's TimestampType
                assert g.type.tz == "UTC"             # is): pyarrow hands it back with tz=UTC, same values
                g = g.cast(want.type)
            assert g.type == want.type, (names[i], g.type, want.type)
            assert g.equals(want), f"column {names[i]} of bucket {b} differs"


def _low_cardinality(n, seed):
    rng = np.random.default_rng(seed)
    m = lambda p: rng.random(n) < p
    few64 = np.array([-1, 0, 1, 2**63 - 1, -2**63, 42, 7_000_000_000], dtype=np.int64)     # -1 = the table's empty marker
    return [
        pa.array(few64[rng.integers(0, len(few64), size=n)], mask=m(0.2)),
        pa.array(rng.integers(0, 300, size=n).astype(np.int32) - 1),                        # int32 -1 as well
        pa.array(np.array([0.0, -0.0, np.nan, 1.5, -np.inf])[rng.integers(0, 5, size=n)], mask=m(0.1)),
        pa.array(np.float32([0.25, 0.5])[rng.integers(0, 2, size=n)]),
        pa.array(rng.integers(-2**62, 2**62, size=n)),                                      # high cardinality: PLAIN (if n is large)
        pa.array(np.full(n, 5, dtype=np.int64)),                                            # one distinct value
        pa.array([None] * n, type=pa.int32()),                                              # no values at all: PLAIN
        pa.array(rng.integers(0, 40, size=n).astype(np.int32), type=pa.int32()).cast(pa.date32()),
        pa.array([f"s{i % 7}" for i in range(n)]),                                          # strings stay PLAIN
        pa.array(rng.random(n) < 0.5),
    ]


@pytest.mark.parametrize("codec", [0, 1])
@pytest.mark.parametrize("n,page_rows,max_dict", [(1, 1 << 20, 16), (3000, 256, 512), (50000, 4096, 256), (50000, 1 << 20, 32768)])
def test_dictionary_encoded_files_read_back(ctx, tmp_path, monkeypatch, n, page_rows, max_dict, codec):
    """hs_parquet_encode_pages: chunks with few distinct values carry a dictionary page and bit-packed indices;
    pyarrow AND the device reader (hs_parquet_decode_pages) return the rows bit for bit (-0.0, NaN payloads and
    the all-ones value included), and chunks past max_dict fall back to PLAIN."""
    from hyperspace_1_b200 import parquet_plain, parquet_read
    monkeypatch.setattr(parquet_plain, "PAGE_ROWS", page_rows)
    cols = _low_cardinality(n, n + max_dict)
    names = [f"c{i}" for i in range(len(cols))]
    cuts = sorted(set([0, n] + [int(x) for x in np.random.default_rng(2).integers(0, n + 1, size=4)]))
    boff = np.array(cuts[:1] + cuts + [n], dtype=np.int64)
    files = parquet_plain.write_buckets_plain(ctx, cols, names, boff, str(tmp_path / "out"), max_dict=max_dict, codec=codec)
    from hyperspace_1_b200 import log_entry as le
    files = sorted(files, key=le.bucket_id_of)
    assert all(f.endswith(".c000.snappy.parquet" if codec else ".c000.parquet") for f in files)
    nonempty = [b for b in range(len(boff) - 1) if boff[b + 1] > boff[b]]
    assert len(files) == len(nonempty)
    dict_chunks = plain_chunks = 0
    for f, b in zip(files, nonempty):
        pf = pq.ParquetFile(f)
        got = pf.read()
        rg = pf.metadata.row_group(0)
        for i, c in enumerate(cols):
            assert rg.column(i).compression == ("SNAPPY" if codec else "UNCOMPRESSED")
            want = c.slice(int(boff[b]), int(boff[b + 1] - boff[b]))
            g = got.column(i).combine_chunks()
            if pa.types.is_floating(want.type):      # compare the bit patterns: NaN != NaN, -0.0 == 0.0 otherwise
                assert g.is_valid().equals(want.is_valid())
                gv = np.frombuffer(g.buffers()[1], dtype=f"<u{want.type.bit_width // 8}", count=len(g), offset=g.offset * want.type.bit_width // 8)
                wv = np.frombuffer(want.buffers()[1], dtype=gv.dtype, count=len(want), offset=want.offset * want.type.bit_width // 8)
                ok = np.asarray(want.is_valid())
                assert np.array_equal(gv[ok], wv[ok]), f"column {names[i]} of bucket {b} differs"
            else:
                assert g.equals(want), f"column {names[i]} of bucket {b} differs"
            if rg.column(i).has_dictionary_page:
                dict_chunks += 1
                assert "PLAIN_DICTIONARY" in rg.column(i).encodings or "RLE_DICTIONARY" in rg.column(i).encodings
            else:
                plain_chunks += 1
    assert plain_chunks > 0 and dict_chunks > 0
    # the device reader over the same files
    arrays, rnames, rows = parquet_read.decode_files(ctx, files)
    assert rnames == names and rows == [int(boff[b + 1] - boff[b]) for b in nonempty]
    for i, c in enumerate(cols):
        g = arrays[i]
        if pa.types.is_floating(c.type):
            assert g.is_valid().equals(c.is_valid())
            ok = np.asarray(c.is_valid())
            u = f"<u{c.type.bit_width // 8}"
            assert np.array_equal(np.frombuffer(g.buffers()[1], dtype=u, count=len(c))[ok], np.frombuffer(c.buffers()[1], dtype=u, count=len(c))[ok])
        else:
            assert g.cast(c.type).equals(c), f"device reader: column {names[i]} differs"


def _snappy_inputs():
    rng = np.random.default_rng(9)
    text = (b"the quick brown fox jumps over the lazy dog; " * 400)
    return {
        "random": rng.integers(0, 256, size=40000).astype(np.uint8).tobytes(),          # literals only, 3-byte headers
        "zeros": bytes(70000),                                                          # one long overlapping copy per fragment
        "period3": b"abc" * 9000,
        "period40": bytes(range(40)) * 700,
        "text": text,
        "mixed": b"".join(text[i * 37:i * 37 + 90] + rng.integers(0, 256, size=int(rng.integers(0, 70))).astype(np.uint8).tobytes()
                          for i in range(300)),
        "tiny": b"ab",
        "one": b"x",
        "fragment_edge": b"0123456789abcdef" * 512 + b"Z",                             # 8193 bytes: a 1-byte second fragment
        "small_ints": np.arange(20000, dtype="<i8").tobytes(),
    }


def test_snappy_pages_are_valid_snappy(ctx):
    """codec = 1: every body is a raw snappy stream that an independent decompressor (pyarrow's) turns back into
    exactly the uncompressed body of the codec = 0 call; compressible inputs do get smaller."""
    inputs = _snappy_inputs()
    arr = pa.array(list(inputs.values()), type=pa.binary())
    n = len(arr)
    rows = list(range(n + 1))                                       # one page per value
    plain, _ = ctx.parquet_plain_pages(arr, True, rows)
    pages, nulls, dbody, dcount, usize = ctx.parquet_encode_pages(arr, True, rows, list(range(n)), n, 0, 1)
    assert list(dcount) == [0] * n and all(d == b"" for d in dbody) and list(usize[n:]) == [0] * n
    codec = pa.Codec("snappy")
    for name, p, q, u in zip(inputs, plain, pages, usize[:n]):
        assert u == len(p), name
        back = codec.decompress(q, decompressed_size=len(p)).to_pybytes()
        assert back == p, f"{name}: snappy stream does not decompress to the page body"
        if name in ("zeros", "period3", "period40", "text"):
            assert len(q) < len(p) / 10, (name, len(q), len(p))
        if name == "small_ints":                 # 8-byte values with six zero bytes each
            assert len(q) < len(p) * 0.6, (name, len(q), len(p))
        assert len(q) <= 32 + len(p) + len(p) // 6, name
    # all of it as ONE page: fragments of a page are concatenated behind one length preamble
    one, _ = ctx.parquet_plain_pages(arr, True, [0, n])
    page, _, _, _, us = ctx.parquet_encode_pages(arr, True, [0, n], [0], 1, 0, 1)
    assert codec.decompress(page[0], decompressed_size=len(one[0])).to_pybytes() == one[0] and us[0] == len(one[0])


def test_encode_pages_rejects_bad_arguments(ctx):
    from hyperspace_1_b200 import HyperspaceException
    a = pa.array(np.arange(10, dtype=np.int64))
    with pytest.raises(HyperspaceException):
        ctx.parquet_encode_pages(a, True, [0, 5, 10], [1, 0], 2, 16)          # chunks out of order
    with pytest.raises(HyperspaceException):
        ctx.parquet_encode_pages(a, True, [0, 5, 10], [0, 2], 2, 16)          # chunk id past nchunks
    with pytest.raises(HyperspaceException):
        ctx.parquet_encode_pages(a, True, [0, 5, 10], [0, 1], 2, 1 << 20)     # max_dict too large
    with pytest.raises(HyperspaceException):
        ctx.parquet_encode_pages(a, True, [0, 5, 10], [0, 1], 2, 4, 2)        # codec 2 (gzip) is not written
    pages, nulls, dbody, dcount = ctx.parquet_encode_pages(a, True, [0, 5, 10], [0, 1], 2, 4)
    assert list(dcount) == [0, 0] and dbody == [b"", b""] and list(nulls) == [0, 0]     # 5 distinct > 4: PLAIN
    plain, _ = ctx.parquet_plain_pages(a, True, [0, 5, 10])
    assert pages == plain


@pytest.mark.parametrize("writer", ["gpu-plain", "gpu-dictionary", "gpu-snappy"])
def test_index_written_by_the_gpu_encoder_serves_queries(tmp_path, hs, writer):
    """createIndex with hsgpu.index.fileWriter = gpu-plain / gpu-dictionary: files are named and laid out as Spark expects, and
    the quick-start style queries over them return what the plain queries return."""
    from hyperspace_1_b200 import Hyperspace, IndexConfig, Session, log_entry as le
    s = Session(system_path=str(tmp_path / "indexes"))
    s.conf["spark.hyperspace.index.numBuckets"] = "8"
    s.conf["hsgpu.index.fileWriter"] = writer
    lcols, keys = lineitem(3000, null_frac=0.02)
    os.makedirs(tmp_path / "lineitem")
    pq.write_table(pa.table(lcols, names=["l_orderkey", "l_extendedprice", "l_discount"]), tmp_path / "lineitem" / "part-0.parquet")
    li = s.read_parquet(str(tmp_path / "lineitem"))
    hx = Hyperspace(s)
    hx.createIndex(li, IndexConfig("liIndex", ["l_orderkey"], ["l_extendedprice"]))
    files = [le.from_uri(f["name"]) for f in le.content_files(hx.index("liIndex")["content"])]
    assert files and all(f.endswith(".c000.snappy.parquet" if writer == "gpu-snappy" else ".c000.parquet") for f in files)
    for f in files:
        k = pq.ParquetFile(f).read()["l_orderkey"]
        v = k.drop_null().to_numpy()
        assert (np.diff(v) >= 0).all()
    q = li.filter(f"l_orderkey >= {int(keys[100])}").select("l_orderkey", "l_extendedprice")
    s.disable_hyperspace()
    base = q.collect()
    s.enable_hyperspace()
    assert "liIndex" in q.explain_string()
    got = q.collect()
    key = lambda t: sorted(zip(t["l_orderkey"].to_pylist(), t["l_extendedprice"].to_pylist()))
    assert key(got) == key(base)
    s._ctx.close()
