"""ctypes wrapper of oracle/libhs_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
Inputs are pyarrow arrays / numpy arrays; outputs are numpy arrays.

The marshalling below is the oracle's own (it does not import the product's abi module): columns are rebuilt
from their LOGICAL content — validity from is_valid(), string offsets from binary_length() — rather than by
passing Arrow's raw validity/offset buffers through, so a mistake in the product's buffer handling (bit order,
offset base, slices) is not shared by the checker.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

_HERE = os.path.dirname(os.path.abspath(__file__))

# include/hsgpu.h: hs_type, hs_status, struct hs_column (48 bytes), hs_pred_term, hs_predicate
(T_BOOL, T_INT8, T_INT16, T_INT32, T_INT64, T_FLOAT32, T_FLOAT64, T_UTF8, T_BINARY, T_LARGE_UTF8,
 T_LARGE_BINARY) = range(1, 12)
_STATUS = {0: "HS_OK", 1: "HS_EINVAL", 2: "HS_ENOMEM", 3: "HS_ECUDA", 4: "HS_ENCCL", 5: "HS_EUNSUPPORTED"}


class OColumn(C.Structure):
    _fields_ = [("type", C.c_int32), ("reserved", C.c_int32), ("length", C.c_int64), ("null_count", C.c_int64),
                ("validity", C.c_void_p), ("offsets", C.c_void_p), ("data", C.c_void_p)]


class OPredTerm(C.Structure):
    _fields_ = [("op", C.c_int32), ("col", C.c_int32), ("lit_i64", C.c_int64), ("lit_f64", C.c_double),
                ("lit_bytes", C.c_char_p), ("lit_len", C.c_int32), ("reserved", C.c_int32)]


class OPredicate(C.Structure):
    _fields_ = [("nterms", C.c_int32), ("reserved", C.c_int32), ("terms", C.POINTER(OPredTerm))]


def _as_array(a):
    if isinstance(a, np.ndarray):
        a = pa.array(a)
    if isinstance(a, pa.ChunkedArray):
        a = a.combine_chunks() if a.num_chunks != 1 else a.chunk(0)
    return a


def _type_code(t):
    if pa.types.is_boolean(t): return T_BOOL
    if pa.types.is_int8(t): return T_INT8
    if pa.types.is_int16(t): return T_INT16
    if pa.types.is_int32(t) or pa.types.is_date32(t): return T_INT32
    if pa.types.is_int64(t) or pa.types.is_timestamp(t) or pa.types.is_date64(t): return T_INT64
    if pa.types.is_float32(t): return T_FLOAT32
    if pa.types.is_float64(t): return T_FLOAT64
    if pa.types.is_string(t): return T_UTF8
    if pa.types.is_binary(t): return T_BINARY
    if pa.types.is_large_string(t): return T_LARGE_UTF8
    if pa.types.is_large_binary(t): return T_LARGE_BINARY
    raise TypeError(f"oracle: unsupported column type {t}")


_NP_OF = {T_INT8: np.int8, T_INT16: np.int16, T_INT32: np.int32, T_INT64: np.int64, T_FLOAT32: np.float32,
          T_FLOAT64: np.float64}
_STORAGE_OF = {T_INT8: pa.int8(), T_INT16: pa.int16(), T_INT32: pa.int32(), T_INT64: pa.int64(),
               T_FLOAT32: pa.float32(), T_FLOAT64: pa.float64()}


class Columns:
    """An array of hs_column structs over numpy buffers rebuilt from the arrays' logical content."""

    def __init__(self, arrays):
        self.arrays = [_as_array(a) for a in arrays]
        self.n = len(self.arrays)
        self.c = (OColumn * max(self.n, 1))()
        self._keep = []
        for col, a in zip(self.c, self.arrays):
            self._fill(col, a)

    def _fill(self, col, a):
        code = _type_code(a.type)
        n = len(a)
        col.type, col.length, col.null_count = code, n, a.null_count
        col.validity = col.offsets = col.data = None
        if a.null_count > 0:
            bits = np.packbits(a.is_valid().to_numpy(zero_copy_only=False).astype(np.uint8), bitorder="little")
            self._keep.append(bits)
            col.validity = bits.ctypes.data
        if code in (T_UTF8, T_BINARY, T_LARGE_UTF8, T_LARGE_BINARY):
            large = code in (T_LARGE_UTF8, T_LARGE_BINARY)
            lens = pc.fill_null(pc.binary_length(a), 0).to_numpy(zero_copy_only=False).astype(np.int64)
            offs = np.zeros(n + 1, dtype=np.int64)
            np.cumsum(lens, out=offs[1:])
            # the value bytes in row order, NULL rows contributing nothing
            raw = a.cast(pa.large_binary() if large else pa.binary())
            bufs = raw.buffers()
            odt = np.int64 if large else np.int32
            ro = np.frombuffer(bufs[1], dtype=odt)[raw.offset: raw.offset + n + 1] if n else np.zeros(1, dtype=odt)
            body = np.frombuffer(bufs[2], dtype=np.uint8) if (len(bufs) > 2 and bufs[2] is not None) else np.zeros(0, np.uint8)
            if a.null_count == 0 and n:
                data = np.ascontiguousarray(body[int(ro[0]): int(ro[-1])])
            elif n:
                rstart = ro[:-1].astype(np.int64)
                idx = np.repeat(rstart - offs[:-1], lens) + np.arange(int(offs[-1]), dtype=np.int64)
                data = np.ascontiguousarray(body[idx])
            else:
                data = np.zeros(0, np.uint8)
            if not large:
                assert int(offs[-1]) < 2**31, "oracle: use large_string past 2 GiB"
                offs = offs.astype(np.int32)
            if data.size == 0:
                data = np.zeros(1, np.uint8)
            self._keep += [offs, data]
            col.offsets, col.data = offs.ctypes.data, data.ctypes.data
        elif code == T_BOOL:
            vals = pc.fill_null(a, False).to_numpy(zero_copy_only=False).astype(np.uint8)
            bits = np.packbits(vals, bitorder="little") if n else np.zeros(1, np.uint8)
            self._keep.append(bits)
            col.data = bits.ctypes.data
        else:
            st = a.cast(_STORAGE_OF[code]) if a.type != _STORAGE_OF[code] else a
            if st.null_count:
                st = pc.fill_null(st, 0)
            vals = np.ascontiguousarray(st.to_numpy(zero_copy_only=False), dtype=_NP_OF[code])
            if vals.size == 0:
                vals = np.zeros(1, _NP_OF[code])
            self._keep.append(vals)
            col.data = vals.ctypes.data

    def ptr(self):
        return C.cast(self.c, C.POINTER(OColumn))


_lib = None


def build():
    subprocess.check_call(["sh", os.path.join(_HERE, "build.sh")])


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(_HERE, "libhs_oracle.so")
        src = os.path.join(_HERE, "hs_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            build()
        _lib = C.CDLL(so)
        for name in ("hso_hash_rows", "hso_bucket_ids", "hso_sort_segments", "hso_build_index",
                     "hso_join", "hso_filter", "hso_num_threads", "hso_set_threads", "hso_take_fixed"):
            getattr(_lib, name).restype = C.c_int32
    return _lib


def _check(st, what):
    if st != 0:
        raise RuntimeError(f"oracle {what}: {_STATUS.get(st, st)}")


def _i64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def hash_rows(keys):
    cs = Columns(keys)
    out = np.empty(len(cs.arrays[0]), dtype=np.int32)
    _check(lib().hso_hash_rows(cs.ptr(), C.c_int32(cs.n), out.ctypes.data_as(C.c_void_p)), "hash_rows")
    return out


def bucket_ids(keys, num_buckets):
    cs = Columns(keys)
    out = np.empty(len(cs.arrays[0]), dtype=np.int32)
    _check(lib().hso_bucket_ids(cs.ptr(), C.c_int32(cs.n), C.c_int32(num_buckets),
                                out.ctypes.data_as(C.c_void_p)), "bucket_ids")
    return out


def build_index(cols, n_indexed, num_buckets, use_radix=True):
    """-> (perm int64[n], bucket_offsets int64[nb+1]); output row i = input row perm[i]."""
    cs = Columns(cols)
    n = len(cs.arrays[0])
    perm = np.empty(n, dtype=np.int64)
    off = np.empty(num_buckets + 1, dtype=np.int64)
    _check(lib().hso_build_index(cs.ptr(), C.c_int32(cs.n), C.c_int32(n_indexed), C.c_int32(num_buckets),
                                 _i64p(perm), _i64p(off), C.c_int32(1 if use_radix else 0)), "build_index")
    return perm, off


def sort_segments(cols, n_keys, seg_offsets, perm=None, use_radix=True):
    cs = Columns(cols)
    seg = np.ascontiguousarray(seg_offsets, dtype=np.int64)
    if perm is None:
        perm = np.arange(int(seg[-1]), dtype=np.int64)
    perm = np.ascontiguousarray(perm, dtype=np.int64).copy()
    _check(lib().hso_sort_segments(cs.ptr(), C.c_int32(n_keys), _i64p(seg), C.c_int64(len(seg) - 1),
                                   _i64p(perm), C.c_int32(1 if use_radix else 0)), "sort_segments")
    return perm


def join(left_cols, right_cols, n_keys, left_off, right_off):
    """Inner bucketed join -> (li, ri, out_bucket_offsets)."""
    lcs, rcs = Columns(left_cols), Columns(right_cols)
    loff = np.ascontiguousarray(left_off, dtype=np.int64)
    roff = np.ascontiguousarray(right_off, dtype=np.int64)
    nb = len(loff) - 1
    assert len(roff) - 1 == nb
    ooff = np.empty(nb + 1, dtype=np.int64)
    null = C.POINTER(C.c_int64)()
    _check(lib().hso_join(lcs.ptr(), rcs.ptr(), C.c_int32(n_keys), C.c_int32(nb), _i64p(loff), _i64p(roff),
                          _i64p(ooff), null, null), "join(count)")
    m = int(ooff[-1])
    li = np.empty(m, dtype=np.int64)
    ri = np.empty(m, dtype=np.int64)
    _check(lib().hso_join(lcs.ptr(), rcs.ptr(), C.c_int32(n_keys), C.c_int32(nb), _i64p(loff), _i64p(roff),
                          _i64p(ooff), _i64p(li), _i64p(ri)), "join(fill)")
    return li, ri, ooff


def join_typed(left_cols, right_cols, n_keys, left_off, right_off, how):
    """Outer / semi / anti joins restated on top of the inner pairs, following what Spark's SortMergeJoinExec
    emits for the plan JoinIndexRule produces (any join type, JoinIndexRule.scala:57-60): a row whose key finds
    no partner (NULL keys never do) is padded with NULLs in the outer joins on its own side, dropped by
    left_semi, and is all that left_anti keeps. -> (li, ri, out_bucket_offsets); -1 = no row (NULLs)."""
    li, ri, ooff = join(left_cols, right_cols, n_keys, left_off, right_off)
    if how == "inner":
        return li, ri, ooff
    loff, roff = np.asarray(left_off, dtype=np.int64), np.asarray(right_off, dtype=np.int64)
    nb, nl, nr = len(loff) - 1, int(loff[-1]), int(roff[-1])
    lbucket = np.repeat(np.arange(nb), np.diff(loff))
    rbucket = np.repeat(np.arange(nb), np.diff(roff))
    lmatched = np.zeros(nl, dtype=bool)
    lmatched[li] = True
    rmatched = np.zeros(nr, dtype=bool)
    rmatched[ri] = True
    lonly, ronly = np.nonzero(~lmatched)[0], np.nonzero(~rmatched)[0]
    pair_bucket = lbucket[li]
    if how == "left_semi":
        a, b, bk = np.nonzero(lmatched)[0], None, lbucket[lmatched]
    elif how == "left_anti":
        a, b, bk = lonly, None, lbucket[lonly]
    else:
        parts_a, parts_b, parts_k = [li], [ri], [pair_bucket]
        if how in ("left_outer", "full_outer"):
            parts_a.append(lonly); parts_b.append(np.full(len(lonly), -1, dtype=np.int64)); parts_k.append(lbucket[lonly])
        if how in ("right_outer", "full_outer"):
            parts_a.append(np.full(len(ronly), -1, dtype=np.int64)); parts_b.append(ronly); parts_k.append(rbucket[ronly])
        a, b, bk = np.concatenate(parts_a), np.concatenate(parts_b), np.concatenate(parts_k)
    order = np.argsort(bk, kind="stable")
    out_off = np.zeros(nb + 1, dtype=np.int64)
    np.cumsum(np.bincount(bk, minlength=nb), out=out_off[1:])
    return a[order], (None if b is None else b[order]), out_off


def take_or_null(col, idx):
    """col.take(idx) with idx == -1 giving NULL."""
    idx = np.asarray(idx, dtype=np.int64)
    return col.take(pa.array(idx, mask=idx < 0))


class Predicate:
    """Builds an hs_predicate from a postfix list of tuples:
    ("eq"|"ne"|"lt"|"le"|"gt"|"ge", col, literal) | ("is_null"|"is_not_null", col) | ("and",)|("or",)|("not",)"""
    OPS = {"eq": 1, "ne": 2, "lt": 3, "le": 4, "gt": 5, "ge": 6, "is_null": 7, "is_not_null": 8, "in": 9,
           "and": 16, "or": 17, "not": 18}

    def __init__(self, terms):
        self.terms = terms
        self._keep = []
        self.arr = (OPredTerm * max(len(terms), 1))()
        for i, t in enumerate(terms):
            tm = self.arr[i]
            tm.op = self.OPS[t[0]]
            tm.col = t[1] if len(t) > 1 else 0
            if len(t) > 2:
                lit = t[2]
                if t[0] == "in":      # list of integers -> int64 array behind lit_bytes, lit_len = count
                    vals = np.ascontiguousarray(list(lit), dtype=np.int64)
                    self._keep.append(vals)
                    tm.lit_bytes = C.cast(vals.ctypes.data, C.c_char_p)
                    tm.lit_len = len(vals)
                    continue
                if isinstance(lit, str):
                    lit = lit.encode("utf-8")
                if isinstance(lit, (bytes, bytearray)):
                    b = bytes(lit)
                    self._keep.append(b)
                    tm.lit_bytes = b
                    tm.lit_len = len(b)
                elif isinstance(lit, float):
                    tm.lit_f64 = lit
                    tm.lit_i64 = int(lit) if lit == lit and abs(lit) < 2**62 else 0
                else:
                    tm.lit_i64 = int(lit)
                    tm.lit_f64 = float(lit)
        self.c = OPredicate(len(terms), 0, C.cast(self.arr, C.POINTER(OPredTerm)))


def filter_rows(cols, terms):
    cs = Columns(cols)
    pred = Predicate(terms)
    n = len(cs.arrays[0])
    idx = np.empty(n, dtype=np.int64)
    m = C.c_int64(0)
    _check(lib().hso_filter(cs.ptr(), C.c_int32(cs.n), C.byref(pred.c), _i64p(idx), C.byref(m)), "filter")
    return idx[: m.value].copy()


def num_threads():
    return int(lib().hso_num_threads())


def set_threads(n):
    """Use n OpenMP threads from now on (launchers such as torchrun export OMP_NUM_THREADS=1)."""
    return int(lib().hso_set_threads(C.c_int32(int(n))))


def take_fixed(a, perm, out=None):
    """numpy fixed-width array a reordered by perm (int64), with every OpenMP thread."""
    a = np.ascontiguousarray(a)
    perm = np.ascontiguousarray(perm, dtype=np.int64)
    if out is None:
        out = np.empty(len(perm), dtype=a.dtype)
    _check(lib().hso_take_fixed(C.c_void_p(a.ctypes.data), C.c_int32(a.dtype.itemsize), _i64p(perm),
                                C.c_int64(len(perm)), C.c_void_p(out.ctypes.data)), "take_fixed")
    return out


def take(cols, perm):
    """Apply a row permutation to a list of Arrow arrays (helper for tests)."""
    idx = pa.array(perm, type=pa.int64())
    return [_as_array(c).take(idx) for c in cols]
