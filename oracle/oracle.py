"""ctypes wrapper of oracle/libhs_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
Inputs are pyarrow arrays / numpy arrays; outputs are numpy arrays.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from hyperspace_1_b200 import abi  # noqa: E402

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
                     "hso_join", "hso_filter", "hso_num_threads"):
            getattr(_lib, name).restype = C.c_int32
    return _lib


def _check(st, what):
    if st != 0:
        raise RuntimeError(f"oracle {what}: {abi.STATUS_NAMES.get(st, st)}")


def _i64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def hash_rows(keys):
    cs = abi.ColumnSet(keys)
    out = np.empty(len(cs.arrays[0]), dtype=np.int32)
    _check(lib().hso_hash_rows(cs.ptr(), C.c_int32(cs.n), out.ctypes.data_as(C.c_void_p)), "hash_rows")
    return out


def bucket_ids(keys, num_buckets):
    cs = abi.ColumnSet(keys)
    out = np.empty(len(cs.arrays[0]), dtype=np.int32)
    _check(lib().hso_bucket_ids(cs.ptr(), C.c_int32(cs.n), C.c_int32(num_buckets),
                                out.ctypes.data_as(C.c_void_p)), "bucket_ids")
    return out


def build_index(cols, n_indexed, num_buckets, use_radix=True):
    """-> (perm int64[n], bucket_offsets int64[nb+1]); output row i = input row perm[i]."""
    cs = abi.ColumnSet(cols)
    n = len(cs.arrays[0])
    perm = np.empty(n, dtype=np.int64)
    off = np.empty(num_buckets + 1, dtype=np.int64)
    _check(lib().hso_build_index(cs.ptr(), C.c_int32(cs.n), C.c_int32(n_indexed), C.c_int32(num_buckets),
                                 _i64p(perm), _i64p(off), C.c_int32(1 if use_radix else 0)), "build_index")
    return perm, off


def sort_segments(cols, n_keys, seg_offsets, perm=None, use_radix=True):
    cs = abi.ColumnSet(cols)
    seg = np.ascontiguousarray(seg_offsets, dtype=np.int64)
    if perm is None:
        perm = np.arange(int(seg[-1]), dtype=np.int64)
    perm = np.ascontiguousarray(perm, dtype=np.int64).copy()
    _check(lib().hso_sort_segments(cs.ptr(), C.c_int32(n_keys), _i64p(seg), C.c_int64(len(seg) - 1),
                                   _i64p(perm), C.c_int32(1 if use_radix else 0)), "sort_segments")
    return perm


def join(left_cols, right_cols, n_keys, left_off, right_off):
    """Inner bucketed join -> (li, ri, out_bucket_offsets)."""
    lcs, rcs = abi.ColumnSet(left_cols), abi.ColumnSet(right_cols)
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
        self.arr = (abi.hs_pred_term * max(len(terms), 1))()
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
        self.c = abi.hs_predicate(len(terms), 0, C.cast(self.arr, C.POINTER(abi.hs_pred_term)))


def filter_rows(cols, terms):
    cs = abi.ColumnSet(cols)
    pred = Predicate(terms)
    n = len(cs.arrays[0])
    idx = np.empty(n, dtype=np.int64)
    m = C.c_int64(0)
    _check(lib().hso_filter(cs.ptr(), C.c_int32(cs.n), C.byref(pred.c), _i64p(idx), C.byref(m)), "filter")
    return idx[: m.value].copy()


def num_threads():
    return int(lib().hso_num_threads())


def take(cols, perm):
    """Apply a row permutation to a list of Arrow arrays (helper for tests)."""
    idx = pa.array(perm, type=pa.int64())
    return [abi.ColumnSet._normalise(c).take(idx) for c in cols]
