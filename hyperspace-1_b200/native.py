"""ctypes binding of libhsgpu.so (include/hsgpu.h). The library is the product; this module only
marshals Arrow buffers in and out. There is no CPU fallback: if the shared library is missing or no
CUDA device is present, calls raise."""
import ctypes as C
import os

import numpy as np
import pyarrow as pa

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhsgpu.so")
_lib = None

EXPORTS = [
    "hs_ctx_create", "hs_ctx_destroy", "hs_last_error", "hs_ctx_set_output_location", "hs_ctx_get_stream",
    "hs_ctx_enable_timing", "hs_ctx_last_timings", "hs_ctx_launch_count", "hs_table_free", "hs_version",
    "hs_ctx_set_deferred_results", "hs_table_wait", "hs_ctx_trim",
    "hs_parquet_plain_pages", "hs_parquet_encode_pages", "hs_parquet_decode_pages",
    "hs_hash_bucket", "hs_hash_rows", "hs_build_index", "hs_merge_runs", "hs_join_bucketed", "hs_filter_scan",
    "hs_bucket_histogram_noalloc", "hs_comm_unique_id", "hs_comm_init", "hs_comm_destroy", "hs_build_index_dist", "hs_comm_last_exchange",
]


class HyperspaceException(Exception):
    """Mirror of com.microsoft.hyperspace.HyperspaceException (HS/HyperspaceException.scala:19):
    what the JNI shim raises for a non-zero hs_status."""


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HyperspaceException(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        getattr(lib, name).restype = C.c_int32
    lib.hs_last_error.restype = C.c_char_p
    lib.hs_version.restype = C.c_char_p
    _lib = lib
    return lib


class Table:
    """A library-owned hs_table. Convert with to_arrow() (host tables) and release with free()."""

    def __init__(self, ctx, ptr):
        self.ctx, self.ptr = ctx, ptr

    @property
    def t(self):
        return self.ptr.contents

    @property
    def nrows(self):
        return self.t.nrows

    def bucket_offsets(self):
        t = self.t
        if not t.bucket_offsets:
            return None
        return np.ctypeslib.as_array(t.bucket_offsets, shape=(t.num_buckets + 1,)).copy()

    def column(self, i):
        return self.t.cols[i]

    def to_arrow(self, types=None):
        t = self.t
        if t.location != abi.HS_MEM_HOST:
            raise HyperspaceException("to_arrow needs a host-resident table")
        return [abi.host_column_to_arrow(t.cols[i], types[i] if types else None) for i in range(t.ncols)]

    def view(self, n_indexed):
        """IndexView over this table's own buffers (borrowed: keep the table alive while the view is used)."""
        t = self.t
        spec = [{"type": t.cols[i].type, "length": t.cols[i].length, "null_count": t.cols[i].null_count,
                 "validity": t.cols[i].validity, "offsets": t.cols[i].offsets, "data": t.cols[i].data}
                for i in range(t.ncols)]
        return IndexView(abi.DeviceColumnSet(spec), n_indexed, self.bucket_offsets())

    def wait(self):
        """hs_table_wait: with deferred results on, blocks until the table's host buffers are filled."""
        self.ctx._check(self.ctx._lib.hs_table_wait(self.ctx._h, self.ptr), "hs_table_wait")
        return self

    def free(self):
        if self.ptr is not None:
            self.ctx._lib.hs_table_free(self.ctx._h, self.ptr)
            self.ptr = None


class IndexView:
    """Borrowed view of one sorted run per bucket (hs_index_view)."""

    def __init__(self, colset, n_indexed, bucket_offsets):
        self.colset = colset
        self.off = np.ascontiguousarray(bucket_offsets, dtype=np.int64)
        self.c = abi.hs_index_view(colset.ptr(), colset.n, n_indexed, len(self.off) - 1, 0,
                                   self.off.ctypes.data_as(C.POINTER(C.c_int64)))


def make_predicate(terms):
    """postfix list of tuples -> (hs_predicate, keepalive)."""
    ops = {"eq": 1, "ne": 2, "lt": 3, "le": 4, "gt": 5, "ge": 6, "is_null": 7, "is_not_null": 8, "in": 9,
           "and": 16, "or": 17, "not": 18}
    keep = []
    arr = (abi.hs_pred_term * max(len(terms), 1))()
    for i, t in enumerate(terms):
        tm = arr[i]
        tm.op = ops[t[0]]
        tm.col = t[1] if len(t) > 1 else 0
        if len(t) > 2:
            lit = t[2]
            if t[0] == "in":      # list of integers -> int64 array behind lit_bytes, lit_len = count
                vals = np.ascontiguousarray(list(lit), dtype=np.int64)
                keep.append(vals)
                tm.lit_bytes = C.cast(vals.ctypes.data, C.c_char_p)
                tm.lit_len = len(vals)
                continue
            if isinstance(lit, str):
                lit = lit.encode("utf-8")
            if isinstance(lit, (bytes, bytearray)):
                b = bytes(lit)
                keep.append(b)
                tm.lit_bytes = b
                tm.lit_len = len(b)
            elif isinstance(lit, float):
                tm.lit_f64 = lit
                tm.lit_i64 = int(lit) if lit == lit and abs(lit) < 2 ** 62 else 0
            else:
                tm.lit_i64 = int(lit)
                tm.lit_f64 = float(lit)
    keep.append(arr)
    return abi.hs_predicate(len(terms), 0, C.cast(arr, C.POINTER(abi.hs_pred_term))), keep


class Context:
    """One hs_ctx: bound to a device, not thread-safe (mirrors the thread-local HyperspaceContext,
    HS/Hyperspace.scala:168-204)."""

    def __init__(self, device=0):
        self._lib = load()
        self._h = C.c_void_p()
        st = self._lib.hs_ctx_create(C.c_int32(device), C.byref(self._h))
        if st != 0:
            raise HyperspaceException(f"hs_ctx_create(device={device}) failed: {abi.STATUS_NAMES.get(st, st)} "
                                      "(a CUDA device is required; there is no CPU fallback)")
        self.device = device

    def close(self):
        if self._h:
            self._lib.hs_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st, what):
        if st != 0:
            msg = self._lib.hs_last_error(self._h)
            raise HyperspaceException(f"{what}: {abi.STATUS_NAMES.get(st, st)}: {msg.decode() if msg else ''}")

    # ---- configuration / introspection
    def set_output_location(self, loc):
        self._check(self._lib.hs_ctx_set_output_location(self._h, C.c_int32(loc)), "set_output_location")

    def set_deferred_results(self, on=True):
        """hs_ctx_set_deferred_results: host results are downloaded after the call returns; Table.wait() first."""
        self._check(self._lib.hs_ctx_set_deferred_results(self._h, C.c_int32(1 if on else 0)), "set_deferred_results")

    def trim(self):
        """hs_ctx_trim: give every cached device block and pinned buffer back to the driver."""
        self._check(self._lib.hs_ctx_trim(self._h), "hs_ctx_trim")

    def enable_timing(self, on=True):
        self._check(self._lib.hs_ctx_enable_timing(self._h, C.c_int32(1 if on else 0)), "enable_timing")

    def stream(self):
        s = C.c_void_p()
        self._check(self._lib.hs_ctx_get_stream(self._h, C.byref(s)), "get_stream")
        return s.value

    def last_timings(self):
        arr = (abi.hs_stage_timing * 64)()
        n = C.c_int32(0)
        self._check(self._lib.hs_ctx_last_timings(self._h, arr, C.c_int32(64), C.byref(n)), "last_timings")
        return [(arr[i].name.decode(), float(arr[i].ms), int(arr[i].launches)) for i in range(n.value)]

    def launch_count(self):
        v = C.c_int64(0)
        self._check(self._lib.hs_ctx_launch_count(self._h, C.byref(v)), "launch_count")
        return v.value

    # ---- K1
    def hash_bucket(self, keys, num_buckets):
        cs = keys if isinstance(keys, (abi.ColumnSet, abi.DeviceColumnSet)) else abi.ColumnSet(keys)
        out = np.empty(cs.c[0].length, dtype=np.int32)
        self._check(self._lib.hs_hash_bucket(self._h, cs.ptr(), C.c_int32(cs.n), C.c_int32(num_buckets),
                                             out.ctypes.data_as(C.c_void_p)), "hs_hash_bucket")
        return out

    def hash_rows(self, keys):
        cs = keys if isinstance(keys, (abi.ColumnSet, abi.DeviceColumnSet)) else abi.ColumnSet(keys)
        out = np.empty(cs.c[0].length, dtype=np.int32)
        self._check(self._lib.hs_hash_rows(self._h, cs.ptr(), C.c_int32(cs.n), out.ctypes.data_as(C.c_void_p)),
                    "hs_hash_rows")
        return out

    def bucket_histogram_noalloc(self, keys, num_buckets, out_ids_ptr, out_counts_ptr):
        """hs_bucket_histogram_noalloc: `keys` is a DeviceColumnSet, the outputs are raw addresses
        (int32 ids in device memory; uint32 counts in host or device memory). Allocates nothing."""
        self._check(self._lib.hs_bucket_histogram_noalloc(self._h, keys.ptr(), C.c_int32(keys.n),
                                                          C.c_int32(num_buckets), C.c_void_p(out_ids_ptr),
                                                          C.c_void_p(out_counts_ptr)), "hs_bucket_histogram_noalloc")

    # ---- build
    def build_index_table(self, cols, n_indexed, num_buckets, dist=False):
        cs = cols if isinstance(cols, (abi.ColumnSet, abi.DeviceColumnSet)) else abi.ColumnSet(cols)
        out = C.POINTER(abi.hs_table)()
        fn = self._lib.hs_build_index_dist if dist else self._lib.hs_build_index
        self._check(fn(self._h, cs.ptr(), C.c_int32(cs.n), C.c_int32(n_indexed), C.c_int32(num_buckets),
                       C.byref(out)), "hs_build_index_dist" if dist else "hs_build_index")
        return Table(self, out)

    def build_index(self, cols, n_indexed, num_buckets, dist=False):
        """-> (list of pyarrow arrays, bucket_offsets int64[nb+1])"""
        cs = cols if isinstance(cols, abi.ColumnSet) else abi.ColumnSet(cols)
        t = self.build_index_table(cs, n_indexed, num_buckets, dist=dist)
        try:
            return t.to_arrow([a.type for a in cs.arrays]), t.bucket_offsets()
        finally:
            t.free()

    def merge_runs(self, runs):
        """runs: list of (arrow columns, n_indexed, bucket_offsets) -> (arrays, bucket_offsets)"""
        views, types = [], None
        for cols, n_indexed, off in runs:
            cs = abi.ColumnSet(cols)
            types = [a.type for a in cs.arrays]
            views.append(IndexView(cs, n_indexed, off))
        arr = (abi.hs_index_view * len(views))(*[v.c for v in views])
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_merge_runs(self._h, arr, C.c_int32(len(views)), C.byref(out)), "hs_merge_runs")
        t = Table(self, out)
        try:
            return t.to_arrow(types), t.bucket_offsets()
        finally:
            t.free()

    def merge_tables(self, views):
        """hs_merge_runs over IndexViews (host or device buffers) -> Table."""
        arr = (abi.hs_index_view * len(views))(*[v.c for v in views])
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_merge_runs(self._h, arr, C.c_int32(len(views)), C.byref(out)), "hs_merge_runs")
        return Table(self, out)

    def parquet_plain_pages(self, array, optional, page_rows):
        """hs_parquet_plain_pages for one Arrow array -> (list of page bodies as bytes, NULLs per page)."""
        cs = abi.ColumnSet([array])
        pr = np.ascontiguousarray(page_rows, dtype=np.int64)
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_parquet_plain_pages(self._h, cs.ptr(), C.c_int32(1 if optional else 0),
                                                     pr.ctypes.data_as(C.POINTER(C.c_int64)), C.c_int32(len(pr) - 1),
                                                     C.byref(out)), "hs_parquet_plain_pages")
        t = Table(self, out)
        try:
            bodies, nulls = t.to_arrow()
            return [b.as_py() for b in bodies], nulls.to_numpy()
        finally:
            t.free()

    def parquet_encode_pages(self, array, optional, page_rows, page_chunk, nchunks, max_dict, codec=0):
        """hs_parquet_encode_pages -> (page bodies, NULLs per page, dictionary page bodies, entries per chunk); with
        codec = 1 (snappy) a fifth element: uncompressed sizes of the pages ++ of the dictionary pages."""
        cs = abi.ColumnSet([array])
        pr = np.ascontiguousarray(page_rows, dtype=np.int64)
        pc = np.ascontiguousarray(page_chunk, dtype=np.int32)
        npages = len(pr) - 1
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_parquet_encode_pages(self._h, cs.ptr(), C.c_int32(1 if optional else 0),
                                                      pr.ctypes.data_as(C.POINTER(C.c_int64)), C.c_int32(npages),
                                                      pc.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int32(nchunks),
                                                      C.c_int32(max_dict), C.c_int32(codec), C.byref(out)),
                    "hs_parquet_encode_pages")
        t = Table(self, out)
        try:
            cols = t.to_arrow()
            bodies = [b.as_py() for b in cols[0]]
            counts = cols[1].to_numpy()
            res = (bodies[:npages], counts[:npages], bodies[npages:], counts[npages:])
            return res + (cols[2].to_numpy(),) if codec else res
        finally:
            t.free()

    def parquet_decode_table(self, blob, pages, hs_type, optional, codec):
        """hs_parquet_decode_pages: `blob` = numpy uint8 array with the page bodies as stored in the files, `pages`
        = [(offset, size, uncompressed size, num_values, kind, encoding, chunk)] in row order -> Table of one column."""
        arr = (abi.hs_parquet_page * max(len(pages), 1))()
        for i, (off, size, usize, nvals, kind, enc, chunk) in enumerate(pages):
            a = arr[i]
            a.offset, a.size, a.uncompressed_size, a.num_values = int(off), int(size), int(usize), int(nvals)
            a.kind, a.encoding, a.chunk = int(kind), int(enc), int(chunk)
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_parquet_decode_pages(self._h, C.c_void_p(blob.ctypes.data if blob.size else None),
                                                      C.c_int64(blob.size), arr, C.c_int32(len(pages)), C.c_int32(hs_type),
                                                      C.c_int32(1 if optional else 0), C.c_int32(codec), C.byref(out)),
                    "hs_parquet_decode_pages")
        return Table(self, out)

    def parquet_decode_pages(self, blob, pages, hs_type, optional, codec, arrow_type=None):
        """The same, as one host Arrow array."""
        t = self.parquet_decode_table(blob, pages, hs_type, optional, codec)
        try:
            return t.to_arrow([arrow_type] if arrow_type is not None else None)[0]
        finally:
            t.free()

    # ---- join / filter
    def join_table(self, left, right, left_out, right_out, how="inner"):
        lo = (C.c_int32 * max(len(left_out), 1))(*left_out)
        ro = (C.c_int32 * max(len(right_out), 1))(*right_out)
        out = C.POINTER(abi.hs_table)()
        jt = abi.JOIN_TYPES[how] if isinstance(how, str) else int(how)
        self._check(self._lib.hs_join_bucketed(self._h, C.byref(left.c), C.byref(right.c), C.c_int32(jt),
                                               lo, C.c_int32(len(left_out)), ro, C.c_int32(len(right_out)),
                                               C.byref(out)), "hs_join_bucketed")
        return Table(self, out)

    def join(self, left_cols, left_off, right_cols, right_off, n_indexed, left_out, right_out, how="inner"):
        lcs, rcs = abi.ColumnSet(left_cols), abi.ColumnSet(right_cols)
        lv, rv = IndexView(lcs, n_indexed, left_off), IndexView(rcs, n_indexed, right_off)
        t = self.join_table(lv, rv, left_out, right_out, how)
        try:
            types = [lcs.arrays[i].type for i in left_out] + [rcs.arrays[i].type for i in right_out]
            return t.to_arrow(types), t.bucket_offsets()
        finally:
            t.free()

    def filter_table(self, view, terms, out_cols):
        pred, keep = make_predicate(terms)
        oc = (C.c_int32 * max(len(out_cols), 1))(*out_cols)
        out = C.POINTER(abi.hs_table)()
        self._check(self._lib.hs_filter_scan(self._h, C.byref(view.c), C.byref(pred), oc, C.c_int32(len(out_cols)),
                                             C.byref(out)), "hs_filter_scan")
        del keep
        return Table(self, out)

    def filter_scan(self, cols, n_indexed, bucket_offsets, terms, out_cols):
        cs = abi.ColumnSet(cols)
        if bucket_offsets is None:
            bucket_offsets = [0, len(cs.arrays[0])]
        view = IndexView(cs, n_indexed, bucket_offsets)
        t = self.filter_table(view, terms, out_cols)
        try:
            return t.to_arrow([cs.arrays[i].type for i in out_cols])
        finally:
            t.free()

    # ---- multi-GPU
    def comm_unique_id(self):
        buf = (C.c_uint8 * 128)()
        self._check(self._lib.hs_comm_unique_id(self._h, buf), "hs_comm_unique_id")
        return bytes(buf)

    def comm_init(self, rank, nranks, uid):
        b = (C.c_uint8 * 128).from_buffer_copy(uid) if uid is not None else None
        self._check(self._lib.hs_comm_init(self._h, C.c_int32(rank), C.c_int32(nranks), b), "hs_comm_init")

    def comm_destroy(self):
        self._lib.hs_comm_destroy(self._h)

    def comm_last_exchange(self):
        b, ms = C.c_int64(0), C.c_float(0)
        self._lib.hs_comm_last_exchange(self._h, C.byref(b), C.byref(ms))
        return b.value, ms.value
