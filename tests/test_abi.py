"""The C-ABI library loads on a CPU-only box and exports every symbol include/gt4py_b200.h declares."""
import ctypes
import os
import re

import pytest

from gt4py_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "gt4py_b200.h")).read()
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", text)))


def test_library_builds_and_loads():
    path = build.build()
    assert os.path.exists(path)
    handle = ctypes.CDLL(path)
    assert handle.b200_abi_version() == 1


def test_every_declared_symbol_is_exported():
    syms = _header_symbols()
    assert len(syms) >= 30
    handle = ctypes.CDLL(build.build())
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in the header but not exported"


def test_python_prototypes_match_header():
    assert _header_symbols() == _lib.declared_symbols()


def test_no_silent_fallback_without_gpu():
    """Compute entry points must fail loudly (negative status + message) when there is no device."""
    if _lib.device_count() > 0:
        pytest.skip("a GPU is present")
    from gt4py_b200.stencil_object import ArrayView, make_field_desc
    import numpy as np

    v = ArrayView(0x1000, (8, 8, 4), (8, 64, 512), np.dtype("float64"))
    f = make_field_desc(v, (2, 2, 0))
    dom = (ctypes.c_int64 * 3)(4, 4, 4)
    rc = _lib.lib().b200_hdiff(ctypes.byref(f), ctypes.byref(f), ctypes.byref(f), dom, 1, 0, None)
    assert rc < 0
    assert _lib.lib().b200_last_error()


def test_argument_validation_in_c():
    from gt4py_b200.stencil_object import ArrayView, make_field_desc
    import numpy as np

    v = ArrayView(0x1000, (8, 8, 4), (8, 64, 512), np.dtype("float64"))
    f = make_field_desc(v, (1, 2, 0))  # halo of 2 required in I
    dom = (ctypes.c_int64 * 3)(4, 4, 4)
    rc = _lib.lib().b200_hdiff(ctypes.byref(f), ctypes.byref(f), ctypes.byref(f), dom, 1, 0, None)
    assert rc == _lib.EINVAL
    assert b"origin" in _lib.lib().b200_last_error()
    g = make_field_desc(v, (2, 2, 0))
    dom = (ctypes.c_int64 * 3)(5, 4, 4)  # 2 + 5 + 2 > 8
    rc = _lib.lib().b200_hdiff(ctypes.byref(g), ctypes.byref(g), ctypes.byref(g), dom, 1, 0, None)
    assert rc == _lib.EINVAL and b"too small" in _lib.lib().b200_last_error()
