"""CPU-side checks of the drop-in boundary: libspatial_b200.so builds for sm_100a, loads, and exports every entry
point include/spatial_b200.h declares (no compute calls: this container has no GPU)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "spatial_b200.h")).read()
    return sorted(set(re.findall(r"SB_API[^;(]*?\b(sb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(sb):
    sb.build_native()
    lib = ctypes.CDLL(sb.capi.library_path())
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    # the ctypes binding covers the whole header and nothing else
    assert sorted(sb.capi.SIGNATURES) == names


def test_library_is_sm100a_only(sb):
    out = subprocess.run(["cuobjdump", "-lelf", sb.capi.library_path()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback(sb):
    """Without a GPU the product must fail loudly, not compute on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(sb.SpatialB200Error) as e:
        sb.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "duckdb-spatial_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "liboracle" not in txt and "libref" not in txt, f
