"""The host half of the device Parquet reader (parquet_read.py): the Thrift compact page-header parser against
files written by pyarrow — every chunk's pages are found, their value counts add up, their bodies tile the chunk."""
import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest


@pytest.mark.parametrize("compression,use_dictionary", [("NONE", False), ("SNAPPY", True)])
def test_page_headers_tile_every_chunk(tmp_path, compression, use_dictionary):
    from hyperspace_1_b200 import parquet_read
    rng = np.random.default_rng(1)
    n = 50000
    t = pa.table({"k": pa.array(rng.integers(0, 10**9, size=n), mask=rng.random(n) < 0.1),
                  "few": pa.array(rng.integers(0, 11, size=n) / 100.0),
                  "s": pa.array([f"k{int(x)}" for x in rng.integers(0, 1000, n)])})
    f = str(tmp_path / "x.parquet")
    pq.write_table(t, f, compression=compression, use_dictionary=use_dictionary, data_page_size=8192, row_group_size=20000,
                   data_page_version="1.0")
    buf = np.fromfile(f, dtype=np.uint8)
    md = pq.ParquetFile(f).metadata
    multi = False
    for rg in range(md.num_row_groups):
        for ci in range(md.num_columns):
            col = md.row_group(rg).column(ci)
            start = min(col.dictionary_page_offset or col.data_page_offset, col.data_page_offset)
            pages = parquet_read.chunk_pages(buf, start, col.total_compressed_size, col.num_values)
            data = [p for p in pages if p[0] == 0]
            assert sum(p[4] for p in data) == col.num_values
            assert len(data) >= 1
            multi = multi or len(data) > 1
            assert (len(pages) - len(data) == 1) == (use_dictionary and col.has_dictionary_page)
            # bodies lie inside the chunk, in order, without overlap
            last = start
            for kind, body, csize, usize, nvals, enc in pages:
                assert body > last - 1 and body + csize <= start + col.total_compressed_size
                last = body + csize
            assert last == start + col.total_compressed_size
            if use_dictionary:
                assert all(p[5] in (2, 8) for p in data) or any(p[5] == 0 for p in data)
            else:
                assert all(p[5] == 0 for p in data)
    assert multi   # small pages: some chunk has several
