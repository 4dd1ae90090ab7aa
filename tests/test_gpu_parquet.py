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


def test_index_written_by_the_gpu_encoder_serves_queries(tmp_path, hs):
    """createIndex with hsgpu.index.fileWriter = gpu-plain: files are named and laid out as Spark expects, and
    the quick-start style queries over them return what the plain queries return."""
    from hyperspace_1_b200 import Hyperspace, IndexConfig, Session, log_entry as le
    s = Session(system_path=str(tmp_path / "indexes"))
    s.conf["spark.hyperspace.index.numBuckets"] = "8"
    s.conf["hsgpu.index.fileWriter"] = "gpu-plain"
    lcols, keys = lineitem(3000, null_frac=0.02)
    os.makedirs(tmp_path / "lineitem")
    pq.write_table(pa.table(lcols, names=["l_orderkey", "l_extendedprice", "l_discount"]), tmp_path / "lineitem" / "part-0.parquet")
    li = s.read_parquet(str(tmp_path / "lineitem"))
    hx = Hyperspace(s)
    hx.createIndex(li, IndexConfig("liIndex", ["l_orderkey"], ["l_extendedprice"]))
    files = [le.from_uri(f["name"]) for f in le.content_files(hx.index("liIndex")["content"])]
    assert files and all(f.endswith(".c000.parquet") for f in files)
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
