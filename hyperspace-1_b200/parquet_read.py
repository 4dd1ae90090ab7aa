"""Index files read back through the GPU: the host side of hs_parquet_decode_pages.

Where the reference hands the index files to Spark's ParquetFileFormat reader
(HS/index/rules/RuleUtils.scala:280-286), this module parses only the framing — the footer (through pyarrow's
metadata reader: offsets, codec, physical types) and the Thrift compact page headers inside every column chunk (a
few dozen bytes per page, parsed here) — and passes the raw page bytes of all bucket files of one column to the
device in ONE call; definition levels, dictionaries, snappy and values are decoded there (csrc/hs_parquet_read.cu).
Files written by pyarrow / Spark 2.4 (snappy, dictionary pages, data pages v1) and by parquet_plain.py
(PLAIN, uncompressed) are both covered; anything else raises HyperspaceException(HS_EUNSUPPORTED) and the caller
falls back to the host reader."""
import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

from . import abi

PAGE_DATA, PAGE_INDEX, PAGE_DICT, PAGE_DATA_V2 = 0, 1, 2, 3
CODECS = {"UNCOMPRESSED": 0, "SNAPPY": 1}


class _ThriftReader:
    """Thrift compact protocol, generic: a struct becomes {field id: value}; enough for PageHeader."""

    def __init__(self, buf, pos):
        self.b, self.p = memoryview(buf), pos   # a memoryview yields Python ints (numpy uint8 arithmetic would wrap)

    def _varint(self):
        v, sh = 0, 0
        while True:
            x = self.b[self.p]
            self.p += 1
            v |= (x & 0x7F) << sh
            sh += 7
            if not x & 0x80:
                return v

    def _zigzag(self):
        v = self._varint()
        return (v >> 1) ^ -(v & 1)

    def _value(self, t):
        if t == 1:
            return True
        if t == 2:
            return False
        if t == 3:
            v = self.b[self.p]
            self.p += 1
            return v
        if t in (4, 5, 6):
            return self._zigzag()
        if t == 7:
            v = np.frombuffer(bytes(self.b[self.p:self.p + 8]), dtype="<f8")[0]
            self.p += 8
            return float(v)
        if t == 8:
            n = self._varint()
            v = bytes(self.b[self.p:self.p + n])
            self.p += n
            return v
        if t in (9, 10):
            h = self.b[self.p]
            self.p += 1
            n, et = h >> 4, h & 0x0F
            if n == 15:
                n = self._varint()
            return [self._elem(et) for _ in range(n)]
        if t == 12:
            return self.struct()
        raise ValueError(f"thrift compact type {t} not handled")

    def _elem(self, et):
        if et in (1, 2):     # booleans inside a list are one byte each
            v = self.b[self.p]
            self.p += 1
            return v == 1
        return self._value(et)

    def struct(self):
        out, last = {}, 0
        while True:
            h = self.b[self.p]
            self.p += 1
            if h == 0:
                return out
            delta, t = h >> 4, h & 0x0F
            fid = last + delta if delta else self._zigzag()
            last = fid
            out[fid] = self._value(t)


def chunk_pages(buf, start, total_compressed, num_values):
    """Pages of one column chunk: [(kind, body offset, compressed size, uncompressed size, num_values, encoding)]."""
    pages, pos, seen = [], start, 0
    end = start + total_compressed
    while pos < end and seen < num_values:
        r = _ThriftReader(buf, pos)
        h = r.struct()
        kind, usize, csize = h[1], h[2], h[3]
        body = r.p
        if kind == PAGE_DATA:
            dh = h[5]
            pages.append((0, body, csize, usize, dh[1], dh[2]))
            seen += dh[1]
        elif kind == PAGE_DICT:
            dh = h[7]
            pages.append((2, body, csize, usize, dh[1], 0))
        elif kind == PAGE_DATA_V2:
            raise abi_unsupported("data pages v2")
        pos = body + csize
    return pages


def abi_unsupported(what):
    from .native import HyperspaceException
    return HyperspaceException(f"HS_EUNSUPPORTED: {what} are not decoded on the device")


def _hs_type(t):
    return abi.arrow_type_code(pa.string() if pa.types.is_large_string(t) else pa.binary() if pa.types.is_large_binary(t) else t)


def decode_files(ctx, files, columns=None, device=False):
    """All rows of `files` (in that order) decoded on the device, one call per column.
    Returns (arrays, names, rows_per_file): host Arrow arrays, or — device=True — device-resident Tables of one
    column each (the caller frees them) that can be handed to the join / filter without leaving the GPU."""
    metas = [pq.ParquetFile(f) for f in files]
    if not metas:
        return [], [], []
    schema = metas[0].schema_arrow
    names = [n for n in schema.names if columns is None or n in columns]
    rows_per_file = [m.metadata.num_rows for m in metas]
    handles = [open(f, "rb", buffering=0) for f in files]
    arrays = []
    try:
        for name in names:
            field = schema.field(name)
            ci = schema.get_field_index(name)
            # the column's chunks of all files, read straight into one buffer (page headers stay in between the
            # bodies; the page descriptors point past them)
            chunks, total, codec = [], 0, None
            for fi, m in enumerate(metas):
                md = m.metadata
                for rg in range(md.num_row_groups):
                    col = md.row_group(rg).column(ci)
                    c = CODECS.get(col.compression)
                    if c is None:
                        raise abi_unsupported(f"{col.compression} pages")
                    if codec is not None and c != codec:
                        raise abi_unsupported("files with different codecs in one call")
                    codec = c
                    start = col.data_page_offset
                    if col.has_dictionary_page and col.dictionary_page_offset:
                        start = min(start, col.dictionary_page_offset)
                    chunks.append((fi, start, col.total_compressed_size, col.num_values))
                    total += col.total_compressed_size
            blob = np.empty(total, dtype=np.uint8)
            view = memoryview(blob)
            pages, at = [], 0
            for chunk, (fi, start, size, nvals) in enumerate(chunks):
                h = handles[fi]
                h.seek(start)
                got = h.readinto(view[at:at + size])
                if got != size:
                    raise IOError(f"{files[fi]}: short read of a column chunk")
                for kind, body, csize, usize, nv, enc in chunk_pages(blob, at, size, nvals):
                    pages.append((body, csize, usize, nv, kind, enc, chunk))
                at += size
            optional = metas[0].schema.column(ci).max_definition_level > 0
            if device:
                ctx.set_output_location(abi.HS_MEM_DEVICE)
                try:
                    arrays.append(ctx.parquet_decode_table(blob, pages, _hs_type(field.type), optional, codec or 0))
                finally:
                    ctx.set_output_location(abi.HS_MEM_HOST)
            else:
                arrays.append(ctx.parquet_decode_pages(blob, pages, _hs_type(field.type), optional, codec or 0, field.type))
    except Exception:
        if device:
            for t in arrays:
                t.free()
        raise
    finally:
        for h in handles:
            h.close()
    return arrays, names, rows_per_file
