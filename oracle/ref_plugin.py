"""pytest plugin (TEST INFRASTRUCTURE) for running the reference's OWN test modules against the ``b200`` backend:

    python -m pytest -p oracle.ref_plugin oracle/_ref/tests/cartesian_tests/... -k b200

* puts the staged reference (oracle/_ref, or /root/reference in the build container), the import shims and -- on a
  machine with a CUDA device -- the cupy stand-in on sys.path, then imports ``gt4py_b200.backend`` so that the name is
  in the registry BEFORE the reference's ``cartesian_tests/definitions.py`` derives ALL_BACKENDS / GPU_BACKENDS from
  it (definitions.py:34-55);
* ``StencilTestSuite._run_test_implementation`` (testing/suites.py:494-606) compares outputs only for fields that
  are ``np.ndarray`` instances, i.e. it silently skips the value check for every GPU backend; the wrapper below
  performs the same ``assert_allclose(rtol=1e-5, atol=1e-8)`` on the device arrays, so the run is a real check;
* B200_REF_AS_CPU_BACKEND=1 additionally offers the name to the tests that the reference parametrises over
  CPU_BACKENDS only (call interface: default arguments, halo checks, ...), which are backend-agnostic in what they
  assert.
"""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

if os.environ.get("GT4PY_B200_EMU") == "1":  # GPU-less container: the host double of the C ABI (tests/emu_abi.py)
    from tests import emu_abi  # noqa: E402

    emu_abi.activate()

from oracle import ref  # noqa: E402

if not ref.activate(gpu=True):
    raise RuntimeError("oracle.ref_plugin: the reference is not available (run oracle/make_ref.py in the build container)")

import numpy as np  # noqa: E402
import pytest  # noqa: E402

import gt4py_b200.backend  # noqa: E402,F401  (registers "b200")


def _install_value_check():
    from gt4py.cartesian.testing import suites

    original = suites.StencilTestSuite._run_test_implementation.__func__
    if getattr(original, "__b200_wrapped__", False):
        return

    def run(cls, parameters_dict, implementation):
        input_data, _ = parameters_dict
        made, expected = {}, {}
        real_from_array, real_validation = suites.gt_storage.from_array, cls.validation

        def from_array(data=None, *a, **kw):
            res = real_from_array(data, *a, **kw) if data is not None else real_from_array(*a, **kw)
            made[id(data)] = res
            return res

        def validation(*a, **kw):
            expected.update({k: v for k, v in kw.items() if isinstance(v, np.ndarray)})
            return real_validation(*a, **kw)

        suites.gt_storage.from_array = from_array
        cls.validation = staticmethod(validation)
        try:
            original(cls, parameters_dict, implementation)
        finally:
            suites.gt_storage.from_array = real_from_array
            cls.validation = staticmethod(real_validation)
        checked = 0
        for name, data in input_data.items():
            value = made.get(id(data))
            if value is None or isinstance(value, np.ndarray) or name not in expected:
                continue  # numpy arrays were compared by the reference's own loop
            want = expected[name]
            want = want.base if want.base is not None and isinstance(want.base, np.ndarray) else want
            np.testing.assert_allclose(np.asarray(value), want, rtol=suites.RTOL, atol=suites.ATOL,
                                       equal_nan=suites.EQUAL_NAN,
                                       err_msg=f"Wrong data in output field '{name}' (device array)")
            checked += 1
        assert checked or not made, "no device field was compared"

    run.__b200_wrapped__ = True
    suites.StencilTestSuite._run_test_implementation = classmethod(run)


_install_value_check()

if os.environ.get("B200_REF_AS_CPU_BACKEND") == "1":
    from cartesian_tests import definitions as _defs

    if not any(getattr(p, "values", (None,))[0] == "b200" for p in _defs.CPU_BACKENDS):
        _defs.CPU_BACKENDS.append(pytest.param("b200"))


def pytest_configure(config):
    for m in ("requires_gpu", "requires_dace", "requires_atlas", "uses_cartesian_shift"):
        config.addinivalue_line("markers", f"{m}: marker of the reference's test suite")
