"""ctypes binding of oracle/_build/liboracle.so (TEST INFRASTRUCTURE: the parity checker).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package gt4py_b200 never does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "_build", "liboracle.so")
    src = os.path.join(_HERE, "stencils.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "_build/liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
    return _LIB


def _ptr(a: np.ndarray, origin):
    """Address of the origin element of `a` (any strides), as c_void_p."""
    off = sum(int(o) * int(s) for o, s in zip(origin, a.strides))
    return ctypes.c_void_p(a.ctypes.data + off)


def _strides(a: np.ndarray):
    return [ctypes.c_ssize_t(s // a.itemsize) for s in a.strides]


def _same_strides(*arrays):
    s0 = arrays[0].strides
    for a in arrays[1:]:
        if a.strides != s0 or a.dtype != arrays[0].dtype:
            raise ValueError("oracle: fields of one call must share dtype and strides")


def hdiff(in_field, out_field, coeff, *, origin=(2, 2, 0), domain=None):
    """Reference semantics of horizontal_diffusion (stencil_definitions.py:316-328); out in place."""
    _same_strides(in_field, out_field, coeff)
    ni, nj, nk = domain
    if in_field.dtype != np.float64:
        raise TypeError("oracle hdiff: float64 only")
    fn = lib().oracle_hdiff_f64
    fn(_ptr(in_field, origin), _ptr(out_field, origin), _ptr(coeff, origin),
       ctypes.c_int(ni), ctypes.c_int(nj), ctypes.c_int(nk), *_strides(in_field))
    return out_field


def simple_hdiff(in_field, coeff, out_field, *, origin=(2, 2, 0), domain=None):
    _same_strides(in_field, out_field, coeff)
    ni, nj, nk = domain
    lib().oracle_simple_hdiff_f64(_ptr(in_field, origin), _ptr(coeff, origin),
                                  _ptr(out_field, origin), ctypes.c_int(ni), ctypes.c_int(nj),
                                  ctypes.c_int(nk), *_strides(in_field))
    return out_field


def vadv(utens_stage, u_stage, wcon, u_pos, utens, *, dtr_stage, origin=(0, 0, 0), domain=None):
    """Reference semantics of vertical_advection_dycore (:235-313); utens_stage in place."""
    _same_strides(utens_stage, u_stage, wcon, u_pos, utens)
    ni, nj, nk = domain
    if nk < 3 or nk > 1024:
        raise ValueError("oracle vadv: 3 <= nk <= 1024")
    lib().oracle_vadv_f64(_ptr(utens_stage, origin), _ptr(u_stage, origin), _ptr(wcon, origin),
                          _ptr(u_pos, origin), _ptr(utens, origin), ctypes.c_double(dtr_stage),
                          ctypes.c_int(ni), ctypes.c_int(nj), ctypes.c_int(nk),
                          *_strides(utens_stage))
    return utens_stage


def tridiag(inf, diag, sup, rhs, out, *, origin=(0, 0, 0), domain=None):
    _same_strides(inf, diag, sup, rhs, out)
    ni, nj, nk = domain
    lib().oracle_tridiag_f64(_ptr(inf, origin), _ptr(diag, origin), _ptr(sup, origin),
                             _ptr(rhs, origin), _ptr(out, origin), ctypes.c_int(ni),
                             ctypes.c_int(nj), ctypes.c_int(nk), *_strides(inf))
    return out
