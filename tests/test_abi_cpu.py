"""CPU-side checks of the drop-in boundary: the library exists, exports every symbol that
include/hsgpu.h declares, and fails loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "hsgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hs_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    from hyperspace_1_b200 import native
    assert sorted(native.EXPORTS) == _declared()


def test_library_exports_every_declared_symbol():
    from hyperspace_1_b200 import native
    if not os.path.exists(native.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(native.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), f"libhsgpu.so does not export {name}"
    native.load()
    assert b"sm_100a" in native._lib.hs_version()


def test_struct_layouts_match_header():
    from hyperspace_1_b200 import abi
    assert ctypes.sizeof(abi.hs_column) == 48
    assert ctypes.sizeof(abi.hs_table) == 48
    assert ctypes.sizeof(abi.hs_index_view) == 32
    assert ctypes.sizeof(abi.hs_pred_term) == 40
    assert ctypes.sizeof(abi.hs_predicate) == 16
    assert ctypes.sizeof(abi.hs_stage_timing) == 40


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from hyperspace_1_b200 import native
    with pytest.raises(native.HyperspaceException):
        native.Context(0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "hyperspace-1_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, f)).read()
                assert "oracle" not in text, f"{f} mentions the oracle"


def test_enum_values_match_header():
    """The numeric codes of hs_type, hs_pred_op, hs_join_type and hs_status in abi.py are those of hsgpu.h."""
    import re
    from hyperspace_1_b200 import abi
    text = open(os.path.join(ROOT, "include", "hsgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    decl = dict((m.group(1), int(m.group(2))) for m in re.finditer(r"\b(HS_[A-Z0-9_]+)\s*=\s*(\d+)", text))
    for name, value in decl.items():
        if name.startswith(("HS_PRED_", "HS_JOIN_", "HS_MEM_")) or name in (
                "HS_OK", "HS_EINVAL", "HS_ENOMEM", "HS_ECUDA", "HS_ENCCL", "HS_EUNSUPPORTED", "HS_BOOL", "HS_INT8", "HS_INT16",
                "HS_INT32", "HS_INT64", "HS_FLOAT32", "HS_FLOAT64", "HS_UTF8", "HS_BINARY", "HS_LARGE_UTF8", "HS_LARGE_BINARY"):
            assert getattr(abi, name) == value, name
    assert len(decl) >= 30


def test_arrow_marshalling_of_every_column_kind():
    """ColumnSet fills hs_column from Arrow buffers without copying: type code, length, NULL count, bitmap only
    when there are NULLs, int32 or int64 offsets for the string kinds, date/timestamp carried as their integers."""
    import numpy as np
    import pyarrow as pa
    from hyperspace_1_b200 import abi
    arrays = [pa.array([True, None, False]), pa.array([1, 2, 3], pa.int8()), pa.array([1, 2, 3], pa.int16()),
              pa.array([1, 2, 3], pa.int32()), pa.array([1, None, 3], pa.int64()), pa.array([1.0, 2.0, 3.0], pa.float32()),
              pa.array([1.0, 2.0, None]), pa.array(["a", "", None]), pa.array([b"x", b"yy", b""], pa.binary()),
              pa.array(["a", "bb", "ccc"], pa.large_string()), pa.array([b"a", None, b"c"], pa.large_binary()),
              pa.array([1, 2, 3], pa.int32()).cast(pa.date32()), pa.array([1, 2, 3], pa.int64()).cast(pa.timestamp("us"))]
    codes = [abi.HS_BOOL, abi.HS_INT8, abi.HS_INT16, abi.HS_INT32, abi.HS_INT64, abi.HS_FLOAT32, abi.HS_FLOAT64, abi.HS_UTF8,
             abi.HS_BINARY, abi.HS_LARGE_UTF8, abi.HS_LARGE_BINARY, abi.HS_INT32, abi.HS_INT64]
    cs = abi.ColumnSet(arrays)
    for i, (a, code) in enumerate(zip(arrays, codes)):
        c = cs.c[i]
        assert c.type == code and c.length == 3 and c.null_count == a.null_count
        assert bool(c.validity) == (a.null_count > 0)
        assert bool(c.offsets) == (code in abi.VARLEN)
    # offsets really are what the type says
    off32 = np.ctypeslib.as_array((ctypes.c_int32 * 4).from_address(cs.c[7].offsets))
    off64 = np.ctypeslib.as_array((ctypes.c_int64 * 4).from_address(cs.c[9].offsets))
    assert off32.tolist() == [0, 1, 1, 1] and off64.tolist() == [0, 1, 3, 6]
    # a sliced array is re-materialised so that bit offset 0 holds
    s = pa.array([5, None, 7, 8]).slice(1, 3)
    cs2 = abi.ColumnSet([s])
    assert cs2.c[0].length == 3 and cs2.c[0].null_count == 1 and cs2.arrays[0].offset == 0
