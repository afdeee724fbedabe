"""CPU tests: the C-ABI library loads, exports every symbol include/sphexa_b200.h declares, and fails loudly
(no fallback) when no CUDA device is present."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "sphexa_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sphb200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_library_agree(pkg):
    syms = declared_symbols()
    assert len(syms) >= 60
    for s in syms:
        assert hasattr(pkg.binding.lib, s), f"libsphexa_b200.so lacks {s} declared in include/sphexa_b200.h"
    for s in pkg.binding.SYMBOLS:
        assert hasattr(pkg.binding.lib, s), s


def test_struct_layouts_match_header(pkg):
    """ctypes mirrors must have the C sizes (field order is checked by the GPU parity tests)"""
    import ctypes as C
    B = pkg.binding
    assert C.sizeof(B.Box) == 6 * 8 + 3 * 8 + 3 * 8 + 3 * 4 + 4
    assert C.sizeof(B.Fields) == 38 * 8 + 8
    assert C.sizeof(B.TreeView) == 8 + 8 * 8
    assert C.sizeof(B.HydroConst) == 8 + 8 + 12 * 8 + 16 + 8


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.SphError, match="no CPU fallback"):
        pkg.Context(0)
    import ctypes as C
    h = C.c_void_p()
    rc = pkg.binding.lib.sphb200_create(0, None, C.byref(h))
    assert rc != 0
    assert b"no CUDA device" in pkg.binding.lib.sphb200_last_error()


def test_make_box_host_helper(pkg):
    b = pkg.make_box([-0.5, 0.5, 0.0, 2.0, 1.0, 1.0625], [1, 0, 2])
    assert list(b.len) == [1.0, 2.0, 0.0625]
    assert list(b.inv) == [1.0, 0.5, 16.0]
    assert list(b.bc) == [1, 0, 2]


def test_product_path_does_not_import_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may touch oracle/"""
    pk = os.path.join(ROOT, "sph-exa_b200")
    for dirpath, _, files in os.walk(pk):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in text.replace("SURVEY", ""), f"{fn} mentions the oracle"
