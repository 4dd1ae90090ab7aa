"""hs_parquet_decode_pages against pyarrow's own reader: files written the ways index files are written (pyarrow
and Spark defaults: snappy + dictionary pages + data pages v1; the device writer: PLAIN, uncompressed), every
supported type, NULLs, several pages per chunk, several row groups, several files per call."""
import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(hs):
    c = hs.Context(0)
    yield c
    c.close()


def _table(n, seed, with_nulls=True):
    rng = np.random.default_rng(seed)

    def m(frac):
        return (rng.random(n) < frac) if with_nulls else None
    words = ["", "a", "bb", "research", "é日本語", "x" * 40]
    return pa.table({
        "k64": pa.array(rng.integers(-2**62, 2**62, size=n), mask=m(0.05)),
        "few": pa.array(rng.integers(0, 11, size=n) / 100.0),                       # dictionary-friendly doubles
        "k32": pa.array(rng.integers(-2**31, 2**31 - 1, size=n).astype(np.int32), mask=m(0.5)),
        "i16": pa.array(rng.integers(-2**15, 2**15 - 1, size=n).astype(np.int16)),
        "i8": pa.array(rng.integers(-128, 127, size=n).astype(np.int8), mask=m(0.1)),
        "f32": pa.array(rng.standard_normal(n).astype(np.float32)),
        "f64": pa.array(rng.standard_normal(n), mask=m(0.2)),
        "b": pa.array(rng.random(n) < 0.3, mask=m(0.1)),
        "s": pa.array([words[int(i)] + str(int(j)) for i, j in zip(rng.integers(0, len(words), n), rng.integers(0, 50, n))], type=pa.string(), mask=m(0.1)),
        "bin": pa.array([bytes(rng.integers(0, 256, size=int(k), dtype=np.uint8)) for k in rng.integers(0, 20, n)], type=pa.binary()),
        "d": pa.array(rng.integers(0, 20000, size=n).astype(np.int32), type=pa.date32()),
        "allnull": pa.array([None] * n, type=pa.int64()),
    })


@pytest.mark.parametrize("compression", ["NONE", "SNAPPY"])
@pytest.mark.parametrize("use_dictionary", [True, False])
def test_decode_matches_pyarrow(ctx, tmp_path, compression, use_dictionary):
    from hyperspace_1_b200 import parquet_read
    files, tables = [], []
    for i, n in enumerate((1, 3000, 0, 70001)):
        t = _table(n, 100 + i)
        f = str(tmp_path / f"part-{i:05d}-x_{i:05d}.c000.parquet")
        # small pages and row groups: several pages per chunk, several chunks per file
        pq.write_table(t, f, compression=compression, use_dictionary=use_dictionary, data_page_size=16 * 1024,
                       row_group_size=25000, data_page_version="1.0", write_statistics=(i % 2 == 0))
        files.append(f)
        tables.append(t)
    want = pa.concat_tables(tables)
    got, names, rows = parquet_read.decode_files(ctx, files)
    assert names == want.column_names and rows == [t.num_rows for t in tables]
    for name, g in zip(names, got):
        w = want[name].combine_chunks()
        assert len(g) == len(w), name
        assert g.null_count == w.null_count, name
        assert g.equals(w), f"{name} ({compression}, dictionary={use_dictionary}) differs from pyarrow's reader"


def test_decode_reads_the_device_writers_files(ctx, tmp_path):
    """Files of parquet_plain.write_buckets_plain (page bodies encoded on the device) decoded on the device."""
    from hyperspace_1_b200 import parquet_plain, parquet_read
    t = _table(5000, 7).select(["k64", "few", "k32", "s", "b"])
    cols = [t[c].combine_chunks() for c in t.column_names]
    boff = np.array([0, 1200, 1200, 5000], dtype=np.int64)
    files = parquet_plain.write_buckets_plain(ctx, cols, t.column_names, boff, str(tmp_path / "v__=0"))
    got, names, rows = parquet_read.decode_files(ctx, sorted(files))
    assert rows == [1200, 3800]
    for name, g in zip(names, got):
        assert g.equals(t[name].combine_chunks()), name


def test_unsupported_codec_is_reported(ctx, tmp_path, hs):
    from hyperspace_1_b200 import parquet_read
    f = str(tmp_path / "z.parquet")
    pq.write_table(_table(100, 1), f, compression="GZIP")
    with pytest.raises(hs.HyperspaceException):
        parquet_read.decode_files(ctx, [f])
