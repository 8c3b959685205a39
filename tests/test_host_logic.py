"""CPU tests of the host-side call path (no GPU needed): validation, origin/domain handling, caching,
exec_info -- modelled on the reference's tests/cartesian_tests/integration_tests/feature_tests/
test_call_interface.py (origin/domain :36-165, halo checks :221-286) and test_exec_info.py."""
import numpy as np
import pytest

from gt4py_b200 import stencils, storage
from gt4py_b200.stencil_object import FrozenStencil


class FakeDeviceArray:
    """Only __cuda_array_interface__ (like the reference's tests/cartesian_tests/utils.py wrappers)."""

    def __init__(self, shape, dtype=np.float64, order="F", dims=None, origin=None):
        self.a = np.zeros(shape, dtype=dtype, order=order)
        if dims is not None:
            self.__gt_dims__ = dims
        if origin is not None:
            self.__gt_origin__ = origin

    @property
    def __cuda_array_interface__(self):
        return {"shape": self.a.shape, "typestr": self.a.dtype.str, "data": (0x7F0000000000, False),
                "version": 3, "strides": self.a.strides}


class Recorder(stencils.HorizontalDiffusion):
    calls = []
    _device_sync_ = False

    def _launch(self, domain, fields, params, stream):
        type(self).calls.append((tuple(domain), {k: (tuple(v.origin)[:3], tuple(v.stride)[:3])
                                                 for k, v in fields.items()}))


@pytest.fixture()
def st():
    Recorder.calls.clear()
    s = Recorder()
    s.clean_call_args_cache()
    return s


def arrs(shape=(12, 11, 5), **kw):
    return [FakeDeviceArray(shape, **kw) for _ in range(3)]


def test_max_domain_is_inferred(st):
    a, b, c = arrs()
    st(a, b, c, origin=(2, 2, 0))
    assert Recorder.calls[-1][0] == (8, 7, 5)  # in_field halo (2,2) limits the domain


def test_origin_dict_and_all(st):
    a, b, c = arrs()
    st(a, b, c, origin={"_all_": (2, 2, 0), "out_field": (0, 0, 0)}, domain=(4, 4, 2))
    dom, f = Recorder.calls[-1]
    assert dom == (4, 4, 2)
    assert f["in_field"][0] == (2, 2, 0) and f["out_field"][0] == (0, 0, 0)
    assert f["in_field"][1] == (1, 12, 132)  # element strides, I contiguous


def test_default_origin_too_small(st):
    a, b, c = arrs()
    with pytest.raises(ValueError, match="Origin for field in_field too small"):
        st(a, b, c, domain=(4, 4, 2))


def test_domain_too_large(st):
    a, b, c = arrs()
    with pytest.raises(ValueError, match="Compute domain too large"):
        st(a, b, c, origin=(2, 2, 0), domain=(9, 7, 5))


def test_zero_domain(st):
    a, b, c = arrs()
    with pytest.raises(ValueError, match="zero sizes"):
        st(a, b, c, origin=(2, 2, 0), domain=(0, 4, 2))


def test_wrong_dtype(st):
    a, b, c = arrs()
    a32 = FakeDeviceArray((12, 11, 5), dtype=np.float32)
    with pytest.raises(TypeError, match="dtype of field 'in_field'"):
        st(a32, b, c, origin=(2, 2, 0))


def test_wrong_ndim(st):
    a, b, c = arrs()
    with pytest.raises((ValueError, AssertionError, IndexError)):
        st(FakeDeviceArray((12, 11)), b, c, origin=(2, 2, 0))


def test_host_array_rejected(st):
    a, b, c = arrs()
    with pytest.raises(TypeError, match="host array"):
        st(np.zeros((12, 11, 5)), b, c, origin=(2, 2, 0))


def test_layout_warning(st):
    a, b, c = arrs()
    with pytest.warns(UserWarning, match="layout of the field 'coeff' is not recommended"):
        st(a, b, FakeDeviceArray((12, 11, 5), order="C"), origin=(2, 2, 0))


def test_gt_origin_and_dims_wrappers(st):
    a = FakeDeviceArray((12, 11, 5), origin=(2, 2, 0))
    b = FakeDeviceArray((12, 11, 5), origin=(2, 2, 0))
    c = FakeDeviceArray((12, 11, 5), origin=(2, 2, 0))
    st(a, b, c)
    assert Recorder.calls[-1][0] == (8, 7, 5)
    # permuted axes: array stored as (K, J, I) with __gt_dims__
    p = FakeDeviceArray((5, 11, 12), order="C", dims=("K", "J", "I"))
    st(p, b, c, origin=(2, 2, 0), domain=(4, 4, 2))
    assert Recorder.calls[-1][1]["in_field"][1] == (1, 12, 132)
    with pytest.raises(ValueError, match="has dimensions"):
        st(FakeDeviceArray((12, 11, 5), dims=("I", "J", "0")), b, c, origin=(2, 2, 0))


def test_domain_origin_cache_and_validate_flag(st):
    a, b, c = arrs()
    st(a, b, c, origin=(2, 2, 0), domain=(4, 4, 2))
    n = len(type(st)._domain_origin_cache)
    st(a, b, c, origin=(2, 2, 0), domain=(4, 4, 2))
    assert len(type(st)._domain_origin_cache) == n
    st(a, b, c, origin=(2, 2, 0), domain=(3, 4, 2))
    assert len(type(st)._domain_origin_cache) == n + 1
    st(a, b, c, domain=(4, 4, 2), validate_args=False)  # too-small origin not checked
    assert Recorder.calls[-1][0] == (4, 4, 2)


def test_exec_info_keys(st):
    a, b, c = arrs()
    info = {"__aggregate_data": True}
    st(a, b, c, origin=(2, 2, 0), exec_info=info)
    for k in ("call_start_time", "call_end_time", "call_run_start_time", "call_run_end_time",
              "run_start_time", "run_end_time", "domain", "origin"):
        assert k in info
    assert info["call_start_time"] <= info["call_run_start_time"] <= info["run_start_time"]
    assert info["run_end_time"] <= info["call_run_end_time"] <= info["call_end_time"]
    agg = info["Recorder"]
    assert agg["ncalls"] == 1 and agg["total_call_time"] >= agg["total_run_time"] >= 0
    st(a, b, c, origin=(2, 2, 0), exec_info=info)
    assert info["Recorder"]["ncalls"] == 2


def test_frozen_stencil(st):
    a, b, c = arrs()
    origin = {"in_field": (2, 2, 0), "out_field": (2, 2, 0), "coeff": (2, 2, 0)}
    fz = st.freeze(origin=origin, domain=(4, 4, 2))
    assert isinstance(fz, FrozenStencil)
    fz(in_field=a, out_field=b, coeff=c)
    assert Recorder.calls[-1][0] == (4, 4, 2)
    with pytest.raises(ValueError, match="3-dimensional integer tuple"):
        st.freeze(origin={"in_field": (2, 2)}, domain=(4, 4, 2))


def test_singleton_and_immutability(st):
    assert Recorder() is st
    with pytest.raises(AttributeError):
        st.foo = 1


def test_vadv_parameter_validation():
    class V(stencils.VerticalAdvectionDycore):
        _device_sync_ = False

        def _launch(self, *a):
            pass

    v = V()
    f = [FakeDeviceArray((9, 8, 6)) for _ in range(5)]
    v(*f, dtr_stage=np.float64(0.1), domain=(8, 8, 6))
    v(*f, dtr_stage=0.1, domain=(8, 8, 6))
    with pytest.raises(TypeError, match="type of parameter 'dtr_stage'"):
        v(*f, dtr_stage=np.float32(0.1), domain=(7, 8, 6))
    with pytest.raises(ValueError, match="Compute domain too small"):
        v(*f, dtr_stage=0.1, domain=(8, 8, 2))
    with pytest.raises(ValueError, match="Compute domain too large"):
        v(*f, dtr_stage=0.1, domain=(9, 8, 6))  # wcon needs i+1


def test_storage_layout_logic():
    assert storage.layout_map(("I", "J", "K")) == (2, 1, 0)
    assert storage.layout_map(("I", "J")) == (1, 0)
    assert storage.layout_map(("I", "J", "K", "0")) == (3, 2, 1, 0)
    with pytest.raises(ValueError):
        storage.normalize_storage_spec(None, (3, 0, 2), np.float64, None)
    with pytest.raises(TypeError):
        storage.normalize_storage_spec(None, (3, 2.5), np.float64, None)
    with pytest.raises(ValueError):
        storage.normalize_storage_spec((0, 0), (3, 2, 1), np.float64, None)
    with pytest.raises(RuntimeError, match="not registered"):
        storage.empty((2, 2, 2), backend="nope")


# ---- origin normalisation: int / iterable / dict, as Index.from_k / Index.from_value (gtc/definitions.py:68-82) ----
def test_int_origin_applies_to_every_axis(st):
    a, b, c = arrs(shape=(12, 12, 9))
    st(a, b, c, origin=2, domain=(4, 4, 2))
    _, f = Recorder.calls[-1]
    assert f["in_field"][0] == (2, 2, 2) and f["out_field"][0] == (2, 2, 2)


@pytest.mark.parametrize("origin", [2, (2, 2, 0), [2, 3, 1], {"_all_": (2, 2, 0)}, {"in_field": (2, 2, 0)}, None,
                                    (2, 2), "abc", 1.5, (1.0, 2, 3)])
def test_origin_dict_matches_the_reference(origin):
    """Same normalised dict (or the same ValueError) as the reference's StencilObject._make_origin_dict
    (stencil_object.py:278-296) -- checked against the reference itself where it is importable."""
    import os
    import sys

    from gt4py_b200.stencil_object import StencilObject

    def ours():
        try:
            return {k: tuple(v) for k, v in StencilObject._make_origin_dict(origin).items()}
        except ValueError as exc:
            return ("ValueError", str(exc))

    got = ours()
    if isinstance(origin, int):
        assert got == {"_all_": (origin,) * 3}
    if not os.path.isdir("/root/reference/src"):
        return
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in ("/root/reference/src", os.path.join(root, "oracle", "ref_shims")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from gt4py.cartesian.stencil_object import StencilObject as Ref

    try:
        want = {k: tuple(v) for k, v in Ref._make_origin_dict(origin).items()}
    except ValueError as exc:
        want = ("ValueError", str(exc))
    assert got == want
