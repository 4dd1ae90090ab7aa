"""GpuIndexWriter — the replacement of CreateActionBase.write (HS/actions/CreateActionBase.scala:122-140):
select(indexed ++ included) -> hs_build_index (hash-partition + per-bucket sort on the GPU) -> one Parquet file
per non-empty bucket, named as Spark's bucketed writer names them, so that
BucketingUtils.getBucketId recovers the bucket from the file name. Parquet encode/decode is host work
(pyarrow) and is timed separately from the kernels."""
import os
import time
import uuid

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

from . import log_entry as le

LINEAGE_COL = "_data_file_id"      # IndexConstants.scala:73 (DATA_FILE_NAME_ID)


def has_lineage(entry):
    return str(entry["derivedDataset"]["properties"]["properties"].get("lineage", "false")).lower() == "true"


def index_column_names(entry):
    """indexed ++ included (++ _data_file_id when the index carries lineage): the columns of its files."""
    c = entry["derivedDataset"]["properties"]["columns"]
    return list(c["indexed"]) + list(c["included"]) + ([LINEAGE_COL] if has_lineage(entry) else [])


def write_buckets(columns, names, bucket_offsets, out_dir, task_of_bucket=None, append=False):
    """Write rows [off[b], off[b+1]) of every non-empty bucket b to
    part-<task%05d>-<uuid>_<bucket%05d>.c000.snappy.parquet. Returns (files, seconds spent encoding)."""
    os.makedirs(out_dir, exist_ok=True)
    job = uuid.uuid4()
    table = pa.table(columns, names=names)
    files = []
    t0 = time.perf_counter()
    off = np.asarray(bucket_offsets, dtype=np.int64)
    for b in range(len(off) - 1):
        n = int(off[b + 1] - off[b])
        if n == 0:
            continue      # empty buckets produce no file
        task = b if task_of_bucket is None else task_of_bucket(b)
        path = os.path.join(out_dir, f"part-{task:05d}-{job}_{b:05d}.c000.snappy.parquet")
        pq.write_table(table.slice(int(off[b]), n), path, compression="snappy", version="1.0",
                       data_page_version="1.0", use_dictionary=True)
        files.append(path)
    return files, time.perf_counter() - t0


def write_index_files(session, arrays, names, boff, out_dir):
    """Parquet files of one index version. hsgpu.index.fileWriter = "pyarrow-snappy" (default: the host library
    encodes, files are named and compressed as Spark's default writer would), "gpu-plain" (page bodies encoded
    on the device by hs_parquet_plain_pages, PLAIN + uncompressed; parquet_plain.py) or "gpu-dictionary" (the same
    through hs_parquet_encode_pages: chunks with at most hsgpu.index.dictionaryEntries distinct values — default
    4096 — are dictionary-encoded, the rest PLAIN) or "gpu-snappy" (dictionary as before and every page
    snappy-compressed on the device: Spark's default file format end to end)."""
    t0 = time.perf_counter()
    writer = session.conf.get("hsgpu.index.fileWriter", "pyarrow-snappy")
    if writer in ("gpu-plain", "gpu-dictionary", "gpu-snappy"):
        from .parquet_plain import write_buckets_plain
        max_dict = int(session.conf.get("hsgpu.index.dictionaryEntries", "4096")) if writer != "gpu-plain" else 0
        files = write_buckets_plain(session.ctx, arrays, names, boff, out_dir, max_dict=max_dict,
                                    codec=1 if writer == "gpu-snappy" else 0)
        return files, time.perf_counter() - t0
    return write_buckets(arrays, names, boff, out_dir)


def build_and_write(session, table, indexed, included, num_buckets, out_dir):
    """table: pyarrow.Table holding at least indexed ++ included (resolved spellings)."""
    names = list(indexed) + list(included)
    cols = [table[c].combine_chunks() if isinstance(table[c], pa.ChunkedArray) else table[c] for c in names]
    arrays, boff = session.ctx.build_index(cols, len(indexed), num_buckets)
    files, enc_s = write_index_files(session, arrays, names, boff, out_dir)
    return files, {"rows": len(cols[0]) if cols else 0, "parquet_encode_s": enc_s,
                   "gpu_stages": session.ctx.last_timings()}


def read_index(entry, columns=None, files=None, session=None):
    """Read the files of an index entry grouped by bucket. Returns (runs, names) where every run is
    (arrays, bucket_offsets): run k holds the k-th file of every bucket, so an index with one file per bucket
    is one run and an incrementally refreshed index is several.
    With a session, the pages are decoded on the device (hsgpu.index.fileReader = "gpu", the default there:
    parquet_read.decode_files — the host only parses footers and page headers); files the device reader does not
    cover (other codecs, data pages v2) and hsgpu.index.fileReader = "pyarrow" go through pyarrow."""
    props = entry["derivedDataset"]["properties"]
    nb = props["numBuckets"]
    names = index_column_names(entry)
    want = names if columns is None else [n for n in names if n in columns]
    infos = le.content_files(entry["content"]) if files is None else files
    per_bucket = {}
    for fi in sorted(infos, key=lambda f: f["name"]):
        b = le.bucket_id_of(fi["name"])
        per_bucket.setdefault(b, []).append(le.from_uri(fi["name"]))
    nruns = max((len(v) for v in per_bucket.values()), default=0)
    use_gpu = session is not None and session.conf.get("hsgpu.index.fileReader", "gpu") == "gpu"
    runs = []
    t0 = time.perf_counter()
    for k in range(max(nruns, 1)):
        flist, fbucket = [], []
        for b in range(nb):
            fl = per_bucket.get(b, [])
            if k < len(fl):
                flist.append(fl[k])
                fbucket.append(b)
        if not flist:
            continue
        arrays = None
        if use_gpu:
            from . import parquet_read
            from .native import HyperspaceException
            try:
                got, gnames, rows = parquet_read.decode_files(session.ctx, flist, want)
                arrays = [got[gnames.index(c)] for c in want]
            except HyperspaceException:
                arrays = None      # not covered by the device reader: the host library reads the files
        if arrays is None:
            tables = [pq.ParquetFile(f).read(columns=want) for f in flist]
            rows = [t.num_rows for t in tables]
            t = pa.concat_tables(tables)
            arrays = [t[c].combine_chunks() for c in want]
        off = np.zeros(nb + 1, dtype=np.int64)
        for b, r in zip(fbucket, rows):
            off[b + 1] = r
        runs.append((arrays, np.cumsum(off)))
    return runs, want, time.perf_counter() - t0
