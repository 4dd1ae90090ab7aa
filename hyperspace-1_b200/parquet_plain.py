"""Parquet files whose page bodies come from the GPU (hs_parquet_plain_pages): the K4 step behind saveWithBuckets
(HS/index/DataFrameWriterExtensions.scala:58-65, Spark's DynamicPartitionDataWriter + ParquetFileFormat).

Format written: Parquet 1.0, one row group per file, data pages v1, UNCOMPRESSED, every column OPTIONAL under a
`spark_schema` root — what parquet-mr 1.10 (Spark 2.4.2) and pyarrow read. Values are PLAIN, or — max_dict > 0,
parquet-mr's default behaviour — PLAIN_DICTIONARY for the column chunks whose distinct values fit a dictionary
(hs_parquet_encode_pages decides per chunk on the device). This module only frames
the bytes: Thrift compact-protocol page headers and the footer (a few hundred bytes per file); every value is
encoded on the device, one call per column for ALL bucket files of an index version. The Spark schema JSON goes
into the footer under the key Spark looks for, so types come back as Spark wrote them."""
import os
import struct
import uuid

import numpy as np
import pyarrow as pa

from . import log_entry as le

MAGIC = b"PAR1"
# parquet.thrift enums
T_BOOLEAN, T_INT32, T_INT64, T_FLOAT, T_DOUBLE, T_BYTE_ARRAY = 0, 1, 2, 4, 5, 6
CT_UTF8, CT_DATE, CT_TS_MILLIS, CT_TS_MICROS, CT_INT_8, CT_INT_16 = 0, 6, 9, 10, 15, 16
ENC_PLAIN, ENC_PLAIN_DICTIONARY, ENC_RLE, ENC_BIT_PACKED = 0, 2, 3, 4
OPTIONAL = 1
PAGE_ROWS = 1 << 20          # rows per data page at most


class _Thrift:
    """Thrift compact protocol, the subset the Parquet footer and page headers need."""
    I32, I64, BINARY, LIST, STRUCT = 5, 6, 8, 9, 12

    def __init__(self):
        self.b = bytearray()
        self.last = [0]

    def _varint(self, v):
        while v >= 0x80:
            self.b.append((v & 0x7F) | 0x80)
            v >>= 7
        self.b.append(v)

    def _zigzag(self, v):
        self._varint((v << 1) ^ (v >> 63) if v >= 0 else ((-v) << 1) - 1)

    def _field(self, fid, ftype):
        delta = fid - self.last[-1]
        if 0 < delta <= 15:
            self.b.append((delta << 4) | ftype)
        else:
            self.b.append(ftype)
            self._zigzag(fid)
        self.last[-1] = fid

    def i32(self, fid, v):
        self._field(fid, self.I32)
        self._zigzag(int(v))

    def i64(self, fid, v):
        self._field(fid, self.I64)
        self._zigzag(int(v))

    def string(self, fid, s):
        self._field(fid, self.BINARY)
        raw = s.encode("utf-8") if isinstance(s, str) else bytes(s)
        self._varint(len(raw))
        self.b += raw

    def list_begin(self, fid, etype, n):
        self._field(fid, self.LIST)
        if n < 15:
            self.b.append((n << 4) | etype)
        else:
            self.b.append(0xF0 | etype)
            self._varint(n)

    def struct_begin(self, fid=None):
        if fid is not None:
            self._field(fid, self.STRUCT)
        self.last.append(0)

    def struct_end(self):
        self.b.append(0)
        self.last.pop()

    def elem_i32(self, v):
        self._zigzag(int(v))

    def elem_string(self, s):
        raw = s.encode("utf-8")
        self._varint(len(raw))
        self.b += raw


def physical_type(t):
    """Arrow type -> (Parquet physical type, converted type or None, Arrow type the encoder is fed)."""
    if pa.types.is_boolean(t):
        return T_BOOLEAN, None, t
    if pa.types.is_int8(t):
        return T_INT32, CT_INT_8, pa.int32()
    if pa.types.is_int16(t):
        return T_INT32, CT_INT_16, pa.int32()
    if pa.types.is_int32(t):
        return T_INT32, None, t
    if pa.types.is_date32(t):
        return T_INT32, CT_DATE, t
    if pa.types.is_int64(t):
        return T_INT64, None, t
    if pa.types.is_timestamp(t):
        if t.unit == "ms":
            return T_INT64, CT_TS_MILLIS, t
        if t.unit == "us":
            return T_INT64, CT_TS_MICROS, t
        raise TypeError(f"timestamp unit {t.unit} has no Parquet 1.0 logical type")
    if pa.types.is_float32(t):
        return T_FLOAT, None, t
    if pa.types.is_float64(t):
        return T_DOUBLE, None, t
    if pa.types.is_string(t) or pa.types.is_large_string(t):
        return T_BYTE_ARRAY, CT_UTF8, t
    if pa.types.is_binary(t) or pa.types.is_large_binary(t):
        return T_BYTE_ARRAY, None, t
    raise TypeError(f"no PLAIN Parquet mapping for {t}")


def dictionary_page_header(num_entries, body_len, uncompressed_len=None):
    t = _Thrift()
    t.i32(1, 2)                       # type = DICTIONARY_PAGE
    t.i32(2, body_len if uncompressed_len is None else uncompressed_len)
    t.i32(3, body_len)
    t.struct_begin(7)                 # dictionary_page_header
    t.i32(1, num_entries)
    t.i32(2, ENC_PLAIN_DICTIONARY)    # what parquet-mr 1.10 writes for a v1 dictionary page
    t.struct_end()
    t.b.append(0)
    return bytes(t.b)


def page_header(num_values, body_len, encoding=ENC_PLAIN, uncompressed_len=None):
    t = _Thrift()
    t.i32(1, 0)                       # type = DATA_PAGE
    t.i32(2, body_len if uncompressed_len is None else uncompressed_len)   # uncompressed_page_size
    t.i32(3, body_len)                # compressed_page_size (= the body as stored)
    t.struct_begin(5)                 # data_page_header
    t.i32(1, num_values)
    t.i32(2, encoding)
    t.i32(3, ENC_RLE)                 # definition levels
    t.i32(4, ENC_BIT_PACKED)          # repetition levels (none: flat schema)
    t.struct_end()
    t.b.append(0)
    return bytes(t.b)


def footer(names, types, num_rows, chunks, spark_schema_json, codec=0):
    """chunks: per column (chunk start, total bytes, num_values, first data page offset if the chunk starts with a
    dictionary page else None[, total uncompressed bytes])."""
    t = _Thrift()
    t.i32(1, 1)                                           # version
    t.list_begin(2, _Thrift.STRUCT, len(names) + 1)       # schema
    t.struct_begin()
    t.string(4, "spark_schema")
    t.i32(5, len(names))
    t.struct_end()
    for name, ty in zip(names, types):
        phys, conv, _ = physical_type(ty)
        t.struct_begin()
        t.i32(1, phys)
        t.i32(3, OPTIONAL)
        t.string(4, name)
        if conv is not None:
            t.i32(6, conv)
        t.struct_end()
    t.i64(3, num_rows)
    t.list_begin(4, _Thrift.STRUCT, 1)                    # row_groups
    t.struct_begin()
    t.list_begin(1, _Thrift.STRUCT, len(names))           # columns
    total = 0
    for name, ty, chunk in zip(names, types, chunks):
        off, nbytes, nvals, data_off = chunk[:4]
        ubytes = chunk[4] if len(chunk) > 4 else nbytes
        phys, _, _ = physical_type(ty)
        t.struct_begin()
        t.i64(2, off)                                     # file_offset
        t.struct_begin(3)                                 # meta_data
        t.i32(1, phys)
        encs = (ENC_PLAIN, ENC_RLE, ENC_BIT_PACKED) if data_off is None else (ENC_PLAIN_DICTIONARY, ENC_RLE, ENC_BIT_PACKED)
        t.list_begin(2, _Thrift.I32, 3)
        for e in encs:
            t.elem_i32(e)
        t.list_begin(3, _Thrift.BINARY, 1)
        t.elem_string(name)
        t.i32(4, codec)                                   # 0 UNCOMPRESSED, 1 SNAPPY
        t.i64(5, nvals)
        t.i64(6, ubytes)                                  # total_uncompressed_size (page headers included)
        t.i64(7, nbytes)                                  # total_compressed_size
        t.i64(9, off if data_off is None else data_off)   # data_page_offset
        if data_off is not None:
            t.i64(11, off)                                # dictionary_page_offset
        t.struct_end()
        t.struct_end()
        total += ubytes
    t.i64(2, total)
    t.i64(3, num_rows)
    t.struct_end()
    t.list_begin(5, _Thrift.STRUCT, 1)                    # key_value_metadata
    t.struct_begin()
    t.string(1, "org.apache.spark.sql.parquet.row.metadata")
    t.string(2, spark_schema_json)
    t.struct_end()
    t.string(6, "hsgpu (parquet_plain.py; page bodies by hs_parquet_plain_pages)")
    t.b.append(0)
    return bytes(t.b)


def write_buckets_plain(ctx, columns, names, bucket_offsets, out_dir, task_of_bucket=None, max_dict=0, codec=0):
    """The GPU counterpart of index_writer.write_buckets: rows [off[b], off[b+1]) of every non-empty bucket b go
    to part-<task%05d>-<uuid>_<bucket%05d>.c000.parquet (Spark's name for an uncompressed bucketed file; the
    bucket id is parsed from it by BucketingUtils.getBucketId either way). max_dict > 0: column chunks with at most
    that many distinct values are dictionary-encoded. codec = 1: pages snappy-compressed on the device and the
    files named .c000.snappy.parquet, as Spark's default codec names them. Returns the files."""
    os.makedirs(out_dir, exist_ok=True)
    job = uuid.uuid4()
    off = np.asarray(bucket_offsets, dtype=np.int64)
    types = [c.type for c in columns]
    # pages never cross a bucket: one call per column encodes the pages of all files
    page_rows, page_bucket, page_chunk, chunk_of = [0], [], [], {}
    for b in range(len(off) - 1):
        a, e = int(off[b]), int(off[b + 1])
        while a < e:
            a = min(e, a + PAGE_ROWS)
            page_rows.append(a)
            page_bucket.append(b)
            page_chunk.append(chunk_of.setdefault(b, len(chunk_of)))
    if not page_bucket:
        return []
    bodies, nvals, dicts, dict_entries, usizes = [], [], [], [], []
    npages = len(page_bucket)
    for c in columns:
        _, _, feed = physical_type(c.type)
        arr = c if feed == c.type else c.cast(feed)
        usize = None
        if codec:
            pages, _, dbody, dcount, usize = ctx.parquet_encode_pages(arr, True, page_rows, page_chunk, len(chunk_of),
                                                                       max_dict, codec)
        elif max_dict > 0:
            pages, _, dbody, dcount = ctx.parquet_encode_pages(arr, True, page_rows, page_chunk, len(chunk_of), max_dict)
        else:
            pages, _ = ctx.parquet_plain_pages(arr, True, page_rows)
            dbody, dcount = [b""] * len(chunk_of), np.zeros(len(chunk_of), dtype=np.int64)
        bodies.append(pages)
        dicts.append(dbody)
        dict_entries.append(dcount)
        usizes.append(usize)
        nvals.append(np.diff(np.asarray(page_rows, dtype=np.int64)))
    schema_json = le.schema_json(pa.schema([pa.field(n, t) for n, t in zip(names, types)]))
    files = []
    pages_of = {}
    for p, b in enumerate(page_bucket):
        pages_of.setdefault(b, []).append(p)
    for b, plist in pages_of.items():
        task = b if task_of_bucket is None else task_of_bucket(b)
        path = os.path.join(out_dir, f"part-{task:05d}-{job}_{b:05d}.c000{'.snappy' if codec else ''}.parquet")
        with open(path, "wb") as f:
            f.write(MAGIC)
            pos = 4
            chunks = []
            for ci in range(len(columns)):
                start, nv, data_off, upos = pos, 0, None, 0
                us = usizes[ci]
                entries = int(dict_entries[ci][chunk_of[b]])
                if entries:
                    dbody = dicts[ci][chunk_of[b]]
                    ulen = len(dbody) if us is None else int(us[npages + chunk_of[b]])
                    hdr = dictionary_page_header(entries, len(dbody), ulen)
                    f.write(hdr)
                    f.write(dbody)
                    pos += len(hdr) + len(dbody)
                    upos += len(hdr) + ulen
                    data_off = pos
                for p in plist:
                    body = bodies[ci][p]
                    ulen = len(body) if us is None else int(us[p])
                    hdr = page_header(int(nvals[ci][p]), len(body), ENC_PLAIN_DICTIONARY if entries else ENC_PLAIN, ulen)
                    f.write(hdr)
                    f.write(body)
                    pos += len(hdr) + len(body)
                    upos += len(hdr) + ulen
                    nv += int(nvals[ci][p])
                chunks.append((start, pos - start, nv, data_off, upos))
            foot = footer(names, types, int(off[b + 1] - off[b]), chunks, schema_json, codec)
            f.write(foot)
            f.write(struct.pack("<I", len(foot)))
            f.write(MAGIC)
        files.append(path)
    return files
