"""Stand-in for the ``cupy`` module -- TEST INFRASTRUCTURE ONLY (never imported by gt4py_b200).

This image has no cupy, but the reference's own tests take ``cupy`` as "the array library" of every backend whose
storage device is "gpu" (/root/reference/tests/cartesian_tests/definitions.py:57-65): ``xp.arange``, ``xp.array(...,
order=)``, ``A[:, :, :] = values``, ``(B[:, :, 0] == 0).all()``, ``xp.testing.assert_array_equal`` ...  This module
provides that small surface on top of ``gt4py_b200.storage.DeviceArray``: the data lives in device memory (so the
stencils under test run on the GPU on real device pointers), while element-wise arithmetic, comparisons and
reductions -- which only the TESTS perform, never a stencil -- are evaluated by copying to the host, using numpy,
and copying back.  It is put on sys.path (by oracle/ref.py, ``activate(gpu=True)``) only when a CUDA device exists.
"""
from __future__ import annotations

import numpy as _np

from gt4py_b200 import _lib as _b200_lib
from gt4py_b200 import storage as _st

__gt4py_b200_standin__ = True
__version__ = "0.0+gt4py_b200.standin"

if _b200_lib.device_count() < 1:  # like cupy without a device: the reference then treats cupy as absent
    raise ImportError("cupy stand-in: no CUDA device")

# dtypes / constants the tests spell as xp.<name>
float64, float32, int64, int32, int16, int8, bool_ = (_np.float64, _np.float32, _np.int64, _np.int32, _np.int16,
                                                      _np.int8, _np.bool_)
uint8, uint16, uint32, uint64 = _np.uint8, _np.uint16, _np.uint32, _np.uint64
newaxis, pi, inf, nan = _np.newaxis, _np.pi, _np.inf, _np.nan
dtype = _np.dtype


class _MemPtr:
    def __init__(self, ptr):
        self.ptr = ptr


def _host(x):
    return x.asnumpy() if isinstance(x, _st.DeviceArray) else x


def _dev(host, like=None):
    host = _np.asarray(host)
    if host.ndim == 0:
        return host[()]
    return asarray(host)


def _binary(name, reflected=False):
    def op(self, other):
        a, b = _host(self), _host(other)
        return _dev(getattr(_np.ndarray, name)(_np.asarray(a), b))

    op.__name__ = name
    return op


def _inplace(name):
    def op(self, other):
        res = getattr(_np.ndarray, name.replace("__i", "__", 1))(_host(self), _host(other))
        self.set(res)
        return self

    op.__name__ = name
    return op


class ndarray(_st.DeviceArray):
    """DeviceArray + host-evaluated numpy conveniences (tests only)."""

    __slots__ = ()
    __array_priority__ = 200.0

    @property
    def data(self):
        return _MemPtr(self.ptr)

    @property
    def T(self):
        return self.transpose()

    @property
    def flags(self):
        return _Flags(self)

    def get(self):
        return self.asnumpy()

    def item(self, *args):
        return self.asnumpy().item(*args)

    def tolist(self):
        return self.asnumpy().tolist()

    def astype(self, dt, **kw):
        return asarray(self.asnumpy().astype(dt))

    def copy(self, order="K"):
        return _st.DeviceArray.copy(self)

    def reshape(self, *shape, **kw):
        return asarray(self.asnumpy().reshape(*shape, **kw))

    def ravel(self, *a, **kw):
        return asarray(self.asnumpy().ravel(*a, **kw))

    def view(self, *a, **kw):
        if not a and not kw:
            return self[...]
        raise NotImplementedError("cupy stand-in: ndarray.view(dtype/type)")

    def __getitem__(self, key):
        try:
            return _st.DeviceArray.__getitem__(self, key)
        except TypeError:  # fancy / boolean / None indexing: on the host
            key = tuple(_host(k) for k in key) if isinstance(key, tuple) else _host(key)
            return _dev(self.asnumpy()[key])

    def __setitem__(self, key, value):
        value = _host(value)
        try:
            _st.DeviceArray.__getitem__(self, key).set(value)
        except TypeError:
            h = self.asnumpy()
            h[tuple(_host(k) for k in key) if isinstance(key, tuple) else _host(key)] = value
            self.set(h)

    def __iter__(self):
        for n in range(self.shape[0]):
            yield self[n]

    def __bool__(self):
        return bool(self.asnumpy())

    def __float__(self):
        return float(self.asnumpy())

    def __int__(self):
        return int(self.asnumpy())

    def __neg__(self):
        return _dev(-self.asnumpy())

    def __abs__(self):
        return _dev(abs(self.asnumpy()))

    def __invert__(self):
        return _dev(~self.asnumpy())

    def __repr__(self):
        return "cupy_standin." + repr(self.asnumpy())

    __hash__ = None


for _n in ("add", "sub", "mul", "truediv", "floordiv", "mod", "pow", "and", "or", "xor", "lshift", "rshift", "matmul"):
    setattr(ndarray, f"__{_n}__", _binary(f"__{_n}__"))
    setattr(ndarray, f"__r{_n}__", _binary(f"__r{_n}__"))
    setattr(ndarray, f"__i{_n}__", _inplace(f"__i{_n}__"))
for _n in ("eq", "ne", "lt", "le", "gt", "ge"):
    setattr(ndarray, f"__{_n}__", _binary(f"__{_n}__"))


def _reduction(name):
    def red(self, *a, **kw):
        return _dev(getattr(self.asnumpy(), name)(*a, **kw))

    red.__name__ = name
    return red


for _n in ("all", "any", "sum", "min", "max", "mean", "prod", "std", "var", "argmin", "argmax", "cumsum", "round",
           "squeeze", "flatten", "swapaxes", "nonzero", "clip"):
    setattr(ndarray, _n, _reduction(_n))


class _Flags:
    def __init__(self, a):
        item, shape, strides = a.itemsize, a.shape, a.strides
        acc, c = item, True
        for n, s in zip(reversed(shape), reversed(strides)):
            c = c and (n == 1 or s == acc)
            acc *= n
        acc, f = item, True
        for n, s in zip(shape, strides):
            f = f and (n == 1 or s == acc)
            acc *= n
        self.c_contiguous, self.f_contiguous = c, f
        self.owndata, self.writeable = False, True

    def __getitem__(self, k):
        return getattr(self, k.lower())


_st.set_array_class(ndarray)  # gt4py.storage.*(backend="b200") now returns this class (see backend.install_storage_hooks)


# ---- creation / conversion ---------------------------------------------------------------------------------------
def asarray(a, dtype=None, order=None):
    if isinstance(a, _st.DeviceArray) and dtype is None:
        return a if isinstance(a, ndarray) else ndarray(a._buf, a.ptr, a.shape, a.strides, a.dtype)
    if hasattr(a, "__cuda_array_interface__") and not isinstance(a, _st.DeviceArray) and dtype is None:
        return _st.as_device_array(a)
    host = _np.asarray(_host(a), dtype=dtype, order=order)
    return _st.as_device_array(host)


def array(obj, dtype=None, copy=True, order="K", ndmin=0):
    return _st.as_device_array(_np.array(_host(obj), dtype=dtype, order=order, ndmin=ndmin))


def asnumpy(a, stream=None, order="C", out=None):
    h = _host(a)
    h = _np.asarray(h)
    if out is not None:
        out[...] = h
        return out
    return h


def _creator(name):
    fn = getattr(_np, name)

    def make(*a, **kw):
        return _dev(fn(*[_host(x) for x in a], **{k: _host(v) for k, v in kw.items()}))

    make.__name__ = name
    return make


for _n in ("zeros", "ones", "empty", "full", "arange", "linspace", "zeros_like", "ones_like", "empty_like", "full_like",
           "transpose", "where", "abs", "sqrt", "exp", "log", "sin", "cos", "maximum", "minimum", "sum", "prod",
           "stack", "concatenate", "reshape", "ravel", "isclose", "allclose", "array_equal", "all", "any", "max", "min",
           "meshgrid", "broadcast_to", "moveaxis", "swapaxes", "squeeze", "expand_dims", "indices", "eye", "fromfunction",
           "isnan", "isinf", "isfinite", "logical_and", "logical_or", "logical_not", "roll", "diff", "cumsum", "power",
           "floor", "ceil", "round", "sign", "mean", "copy"):
    globals()[_n] = _creator(_n)


def byte_bounds(a):
    lo, hi = a._extent()
    return a.ptr + lo, a.ptr + hi


class _RandomNS:
    def __getattr__(self, name):
        fn = getattr(_np.random, name)

        def call(*a, **kw):
            return _dev(fn(*a, **kw))

        return call


random = _RandomNS()


class _TestingNS:
    def __getattr__(self, name):
        fn = getattr(_np.testing, name)

        def call(*a, **kw):
            return fn(*[_host(x) for x in a], **{k: _host(v) for k, v in kw.items()})

        return call


testing = _TestingNS()


class _LibNS:
    class stride_tricks:
        @staticmethod
        def as_strided(x, shape=None, strides=None):
            return ndarray(x._buf, x.ptr, shape if shape is not None else x.shape,
                           strides if strides is not None else x.strides, x.dtype)

    class array_utils:
        byte_bounds = staticmethod(byte_bounds)


lib = _LibNS()

from . import cuda  # noqa: E402,F401
