"""ctypes mirror of include/hsgpu.h plus Arrow/numpy -> hs_column marshalling.

The structs here are byte-for-byte the PODs of the C ABI; nothing in this module
computes on data. Arrow arrays are passed zero-copy (buffer addresses).
"""
import ctypes as C

import numpy as np
import pyarrow as pa

HS_OK, HS_EINVAL, HS_ENOMEM, HS_ECUDA, HS_ENCCL, HS_EUNSUPPORTED = range(6)
HS_BOOL, HS_INT8, HS_INT16, HS_INT32, HS_INT64, HS_FLOAT32, HS_FLOAT64, HS_UTF8, HS_BINARY = range(1, 10)
HS_LARGE_UTF8, HS_LARGE_BINARY = 10, 11   # Arrow large_string / large_binary: int64 offsets
VARLEN = (HS_UTF8, HS_BINARY, HS_LARGE_UTF8, HS_LARGE_BINARY)
HS_MEM_HOST, HS_MEM_DEVICE = 0, 1
(HS_PRED_EQ, HS_PRED_NE, HS_PRED_LT, HS_PRED_LE, HS_PRED_GT, HS_PRED_GE,
 HS_PRED_IS_NULL, HS_PRED_IS_NOT_NULL) = range(1, 9)
HS_PRED_IN = 9
HS_PRED_AND, HS_PRED_OR, HS_PRED_NOT = 16, 17, 18
HS_JOIN_INNER, HS_JOIN_LEFT_OUTER, HS_JOIN_RIGHT_OUTER, HS_JOIN_FULL_OUTER, HS_JOIN_LEFT_SEMI, HS_JOIN_LEFT_ANTI = range(6)
JOIN_TYPES = {"inner": 0, "left_outer": 1, "left": 1, "right_outer": 2, "right": 2, "full_outer": 3, "outer": 3,
              "full": 3, "left_semi": 4, "semi": 4, "left_anti": 5, "anti": 5}   # Spark's joinType spellings

STATUS_NAMES = {0: "HS_OK", 1: "HS_EINVAL", 2: "HS_ENOMEM", 3: "HS_ECUDA", 4: "HS_ENCCL",
                5: "HS_EUNSUPPORTED"}

FIXED_WIDTH = {HS_INT8: 1, HS_INT16: 2, HS_INT32: 4, HS_INT64: 8, HS_FLOAT32: 4, HS_FLOAT64: 8}
NUMPY_OF = {HS_INT8: np.int8, HS_INT16: np.int16, HS_INT32: np.int32, HS_INT64: np.int64,
            HS_FLOAT32: np.float32, HS_FLOAT64: np.float64}


class hs_column(C.Structure):
    _fields_ = [("type", C.c_int32), ("reserved", C.c_int32), ("length", C.c_int64),
                ("null_count", C.c_int64), ("validity", C.c_void_p), ("offsets", C.c_void_p),
                ("data", C.c_void_p)]


class hs_table(C.Structure):
    _fields_ = [("ncols", C.c_int32), ("location", C.c_int32), ("nrows", C.c_int64),
                ("cols", C.POINTER(hs_column)), ("num_buckets", C.c_int32), ("reserved", C.c_int32),
                ("bucket_offsets", C.POINTER(C.c_int64)), ("impl", C.c_void_p)]


class hs_index_view(C.Structure):
    _fields_ = [("cols", C.POINTER(hs_column)), ("ncols", C.c_int32), ("n_indexed", C.c_int32),
                ("num_buckets", C.c_int32), ("reserved", C.c_int32),
                ("bucket_offsets", C.POINTER(C.c_int64))]


class hs_pred_term(C.Structure):
    _fields_ = [("op", C.c_int32), ("col", C.c_int32), ("lit_i64", C.c_int64),
                ("lit_f64", C.c_double), ("lit_bytes", C.c_char_p), ("lit_len", C.c_int32),
                ("reserved", C.c_int32)]


class hs_predicate(C.Structure):
    _fields_ = [("nterms", C.c_int32), ("reserved", C.c_int32), ("terms", C.POINTER(hs_pred_term))]


class hs_parquet_page(C.Structure):
    _fields_ = [("offset", C.c_int64), ("size", C.c_int64), ("uncompressed_size", C.c_int64), ("num_values", C.c_int64),
                ("kind", C.c_int32), ("encoding", C.c_int32), ("chunk", C.c_int32), ("reserved", C.c_int32)]


class hs_stage_timing(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("ms", C.c_float), ("launches", C.c_int32)]


def arrow_type_code(t):
    """Spark SQL type (as its Arrow carrier) -> hs_type. DATE rides as INT32 and
    TIMESTAMP as INT64: that is how Spark hashes and orders them."""
    if pa.types.is_boolean(t):
        return HS_BOOL
    if pa.types.is_int8(t):
        return HS_INT8
    if pa.types.is_int16(t):
        return HS_INT16
    if pa.types.is_int32(t) or pa.types.is_date32(t):
        return HS_INT32
    if pa.types.is_int64(t) or pa.types.is_timestamp(t) or pa.types.is_date64(t):
        return HS_INT64
    if pa.types.is_float32(t):
        return HS_FLOAT32
    if pa.types.is_float64(t):
        return HS_FLOAT64
    if pa.types.is_string(t):
        return HS_UTF8
    if pa.types.is_binary(t):
        return HS_BINARY
    if pa.types.is_large_string(t):
        return HS_LARGE_UTF8
    if pa.types.is_large_binary(t):
        return HS_LARGE_BINARY
    raise TypeError(f"unsupported column type {t}")


class ColumnSet:
    """An array of hs_column built from Arrow arrays; keeps the Arrow buffers alive."""

    def __init__(self, arrays):
        self._keep = []
        self.arrays = []
        n = len(arrays)
        self.c = (hs_column * max(n, 1))()
        for i, a in enumerate(arrays):
            a = self._normalise(a)
            self.arrays.append(a)
            self._fill(self.c[i], a)
        self.n = n

    @staticmethod
    def _normalise(a):
        if isinstance(a, np.ndarray):
            a = pa.array(a)
        if isinstance(a, pa.ChunkedArray):
            a = a.combine_chunks() if a.num_chunks != 1 else a.chunk(0)
        if a.offset != 0:
            # the ABI wants bit offset 0; re-materialise the slice
            a = pa.concat_arrays([a])
            if a.offset != 0:
                a = pa.array(a.to_pylist(), type=a.type)
        return a

    def _fill(self, col, a):
        code = arrow_type_code(a.type)
        bufs = a.buffers()
        self._keep.append((a, bufs))
        col.type = code
        col.length = len(a)
        col.null_count = a.null_count
        col.validity = bufs[0].address if (bufs[0] is not None and a.null_count > 0) else None
        if code in VARLEN:
            col.offsets = bufs[1].address if bufs[1] is not None else None
            col.data = bufs[2].address if (len(bufs) > 2 and bufs[2] is not None) else None
            if col.offsets is None:  # zero-length array without buffers
                z = np.zeros(1, dtype=np.int64)
                self._keep.append(z)
                col.offsets = z.ctypes.data
        else:
            col.offsets = None
            col.data = bufs[1].address if bufs[1] is not None else None

    def ptr(self):
        return C.cast(self.c, C.POINTER(hs_column))


class DeviceColumnSet:
    """hs_column array over caller-owned DEVICE buffers (addresses as ints).

    spec: list of dicts {type, length, data, offsets=None, validity=None, null_count=0}
    """

    def __init__(self, spec):
        self.n = len(spec)
        self.c = (hs_column * max(self.n, 1))()
        for i, s in enumerate(spec):
            col = self.c[i]
            col.type = s["type"]
            col.length = s["length"]
            col.null_count = s.get("null_count", 0)
            col.validity = s.get("validity")
            col.offsets = s.get("offsets")
            col.data = s["data"]

    def ptr(self):
        return C.cast(self.c, C.POINTER(hs_column))


def _np_from_addr(addr, nbytes):
    if nbytes == 0 or not addr:
        return np.zeros(0, dtype=np.uint8)
    return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(addr))


def host_column_to_arrow(col, arrow_type=None):
    """Copy one HOST hs_column into a pyarrow array (the table can then be freed)."""
    n = col.length
    code = col.type
    validity = None
    if col.validity and col.null_count > 0:
        validity = pa.py_buffer(_np_from_addr(col.validity, (n + 7) // 8).tobytes())
    if code in VARLEN:
        large = code in (HS_LARGE_UTF8, HS_LARGE_BINARY)
        ow, odt = (8, np.int64) if large else (4, np.int32)
        offs = _np_from_addr(col.offsets, ow * (n + 1)).view(odt).copy()
        nbytes = int(offs[-1]) if n >= 0 and len(offs) else 0
        data = _np_from_addr(col.data, nbytes).tobytes()
        t = {HS_UTF8: pa.string(), HS_BINARY: pa.binary(), HS_LARGE_UTF8: pa.large_string(),
             HS_LARGE_BINARY: pa.large_binary()}[code]
        arr = pa.Array.from_buffers(t, n, [validity, pa.py_buffer(offs.tobytes()), pa.py_buffer(data)],
                                    null_count=col.null_count if validity is not None else 0)
    elif code == HS_BOOL:
        data = _np_from_addr(col.data, (n + 7) // 8).tobytes()
        arr = pa.Array.from_buffers(pa.bool_(), n, [validity, pa.py_buffer(data)],
                                    null_count=col.null_count if validity is not None else 0)
    else:
        w = FIXED_WIDTH[code]
        data = _np_from_addr(col.data, n * w).tobytes()
        t = pa.from_numpy_dtype(NUMPY_OF[code])
        arr = pa.Array.from_buffers(t, n, [validity, pa.py_buffer(data)],
                                    null_count=col.null_count if validity is not None else 0)
    if arrow_type is not None and code in VARLEN:
        # a string column that outgrew int32 offsets comes back as its LARGE twin and stays that way
        if arr.type != arrow_type and code not in (HS_LARGE_UTF8, HS_LARGE_BINARY):
            arr = arr.cast(arrow_type)
    elif arrow_type is not None and arr.type != arrow_type:
        arr = arr.view(arrow_type) if not pa.types.is_date64(arrow_type) else arr.cast(arrow_type)
    return arr
