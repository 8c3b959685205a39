"""b200_vadv_noalloc -- the entry point that allocates nothing (include/gt4py_b200.h): its scratch is a fixed
`static __device__` array inside the library.  No counterpart in the reference.
Named test_zz_* so that pytest collects it LAST.

Called through the C ABI (ctypes on libgt4py_b200.so) and compared with the existing entry point b200_vadv on the
same inputs: bit-exact (np.array_equal, the bar of tests/test_gpu_parity.py), on two calls in a row -- the second with
other inputs and another domain, so that contents the first call left in the shared array would show up if they were
ever observed.  Passes alone in a fresh process and after the other tests (it shares no state with them).
"""
import ctypes
from ctypes import byref

import numpy as np
import pytest

from gt4py_b200 import _lib, storage
from gt4py_b200.stencil_object import as_view, make_field_desc
from oracle import inputs as gen

NAMES = ("utens_stage", "u_stage", "wcon", "u_pos", "utens")
MAX_POINTS = 2 * 1024 * 1024  # B200_VADV_NOALLOC_MAX_POINTS


def _device_fields(host):
    arrays = {k: storage.from_array(host[k]) for k in NAMES}
    descs = {k: make_field_desc(as_view(arrays[k]), (0, 0, 0)) for k in NAMES}
    return arrays, descs


def _call(fn, descs, dtr, dom, *extra):
    d = (ctypes.c_int64 * 3)(*dom)
    return fn(*[byref(descs[k]) for k in NAMES], ctypes.c_double(dtr), d, *extra, None)


@pytest.mark.gpu
def test_noalloc_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    L = _lib.lib()
    _lib.set_device(0)
    for dom, seed in (((96, 70, 80), 21), ((130, 33, 47), 22)):  # second call: other inputs, other domain
        assert dom[0] * dom[1] * dom[2] <= MAX_POINTS
        host = gen.vadv_inputs(*dom, seed=seed)
        dtr = float(host["dtr_stage"])
        a_new, d_new = _device_fields(host)
        a_gmem, d_gmem = _device_fields(host)  # b200_vadv variant 1: the same kernel with a cudaMalloc'd LIFO
        a_auto, d_auto = _device_fields(host)  # b200_vadv variant 0: the library's own choice
        _lib.check(_call(L.b200_vadv_noalloc, d_new, dtr, dom))
        assert _lib.last_kernel() == "vadv_column_gmem"
        _lib.check(_call(L.b200_vadv, d_gmem, dtr, dom, 1))
        _lib.check(_call(L.b200_vadv, d_auto, dtr, dom, 0))
        _lib.synchronize()
        got = a_new["utens_stage"].asnumpy()
        assert np.isfinite(got).all() and not np.array_equal(got, host["utens_stage"])
        assert np.array_equal(got, a_gmem["utens_stage"].asnumpy())
        assert np.array_equal(got, a_auto["utens_stage"].asnumpy())
        for k in NAMES[1:]:  # inputs untouched
            assert np.array_equal(a_new[k].asnumpy(), host[k])


@pytest.mark.gpu
def test_noalloc_entry_point_rejects_domains_beyond_its_fixed_scratch():
    L = _lib.lib()
    _lib.set_device(0)
    dom = (8, 8, 16)
    host = gen.vadv_inputs(*dom, seed=23)
    arrays, descs = _device_fields(host)
    too_big = (1024, 1024, 3)  # 3 Mi points > 2 Mi: rejected before any field is looked at
    rc = _call(L.b200_vadv_noalloc, descs, 0.15, too_big)
    assert rc == _lib.EINVAL
    msg = L.b200_last_error().decode()
    assert "exceeds the fixed scratch" in msg and str(MAX_POINTS) in msg
    _lib.check(_call(L.b200_vadv_noalloc, descs, float(host["dtr_stage"]), dom))  # and it still works afterwards
    _lib.synchronize()
