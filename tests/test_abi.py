"""The C-ABI library loads and exports every symbol include/dsmatch.h declares (no GPU needed)."""

import os

import pytest
import re

from dstack_b200 import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "dsmatch.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dsm_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = native.lib()
    names = _declared()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/dsmatch.h but not exported"
    assert set(native.EXPORTS) <= set(names)


def test_build_info_names_the_target():
    assert b"sm_100a" in native.lib().dsm_build_info()


def test_struct_sizes_match_header():
    import ctypes as C

    assert C.sizeof(native.TableDesc) == 8 * 4
    # u32 n_used + 5 arrays of 16 u32 + u32 rec_words (+4 padding) + 16 + 1 pointers + 16 u32 roles
    assert C.sizeof(native.Layout) == 4 + 5 * 64 + 4 + 17 * 8 + 64
    assert native.RAW_JOB.itemsize == 168  # dsm_raw_job


def test_piece_arithmetic_covers_every_block_once(tmp_path):
    """The row axis is cut into coarse and fine pieces per segment (``Pieces`` in kernels.cuh, used by
    k_match and, through ``slot_start``, by k_extract).  ``tests/pieces_check.cu`` is a host program
    over the same struct: for 20 000 random shapes every 32-row block must be covered exactly once,
    in row order, without a share straddling the segment boundary."""
    import shutil
    import subprocess

    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = tmp_path / "pieces_check"
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pieces_check.cu")
    subprocess.run([nvcc, "-O1", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(exe), src],
                   check=True, cwd=os.path.dirname(src), capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    assert "pieces ok" in out
