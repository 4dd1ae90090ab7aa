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
