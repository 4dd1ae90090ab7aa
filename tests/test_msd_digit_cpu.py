"""The digit arithmetic of the MSD levels (hs_sort_shared.cuh: msd_digit, msd_cell_lo, msd_mul_for) checked on the
CPU: tests/native/msd_digit_check.cu includes the header as it is, is built with nvcc for the host only and walks
random segments (see the invariants listed there). No GPU needed."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_msd_digit_invariants(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "msd_digit_check")
    subprocess.run([nvcc, "-O2", "-std=c++17", "-I", os.path.join(ROOT, "hyperspace-1_b200", "csrc"), "-o", exe,
                    os.path.join(ROOT, "tests", "native", "msd_digit_check.cu")], check=True, capture_output=True, timeout=600)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.startswith("ok"), out.stdout + out.stderr
