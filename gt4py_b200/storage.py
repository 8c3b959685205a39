"""Device storage for the b200 backend: the cupy-free counterpart of ``gt4py.storage``.

Mirrors /root/reference/src/gt4py/storage/cartesian/interface.py:40-327 (``empty / zeros / ones /
full / from_array`` with ``backend=, aligned_index=, dimensions=``) and the padded / aligned
allocation algorithm of /root/reference/src/gt4py/storage/allocators.py:189-275:

* the layout is I-contiguous -- ``layout_map (2, 1, 0)`` like ``gt:gpu``
  (storage/cartesian/layout_registry.py:105-113), data dimensions outermost;
* only the contiguous dimension is padded, to a multiple of ``ALIGNMENT`` elements;
* the allocation start is shifted so that ``aligned_index`` (the origin of the compute domain)
  falls on an ``ALIGNMENT * itemsize`` byte boundary.

sm_100a defaults: ``ALIGNMENT = 16`` elements, i.e. one 128-byte L2 line for float64, which makes
every I-row of the compute domain start on a line boundary and keeps the 128-bit vector loads of
the kernels aligned.

Arrays are ``DeviceArray`` objects: a thin owner of ``cudaMalloc`` memory exposing
``__cuda_array_interface__`` (v3), so they interoperate with cupy / torch / numba without any of
them being imported here.
"""
from __future__ import annotations

import math
import numbers
from ctypes import byref, c_void_p
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib

ALIGNMENT = 16  # elements
BASE_LAYOUT = (2, 1, 0)
BACKEND_NAME = "b200"


# ---- layout helpers (layout.py:28-57 semantics) ---------------------------------------------------
def layout_map(dimensions: Sequence[str]) -> Tuple[int, ...]:
    mask = [d in dimensions for d in "IJK"]
    ranks = [bl for m, bl in zip(mask, BASE_LAYOUT) if m]
    ndata = len(dimensions) - sum(mask)
    ranks = [ndata + r for r in ranks] + list(range(ndata))
    res = [0] * len(ranks)
    for i, idx in enumerate(np.argsort(ranks)):
        res[idx] = i
    return tuple(res)


def is_optimal_layout(array, dimensions: Sequence[str]) -> bool:
    """True when the strides of `array` are ordered like ``layout_map(dimensions)`` (layout.py:60-80)."""
    lm = layout_map(tuple(dimensions))
    strides = _strides_of(array)
    if len(strides) != len(lm):
        return False
    order = sorted(range(len(lm)), key=lambda d: lm[d])  # slowest ... fastest
    prev = None
    for d in order:
        if prev is not None and strides[d] > prev:
            return False
        prev = strides[d]
    return True


def _strides_of(array):
    if isinstance(array, DeviceArray):
        return array.strides
    cai = getattr(array, "__cuda_array_interface__", None) or getattr(array, "__array_interface__", None)
    if cai is not None:
        st = cai.get("strides")
        if st is None:
            shape, item = cai["shape"], np.dtype(cai["typestr"]).itemsize
            st, acc = [0] * len(shape), item
            for d in range(len(shape) - 1, -1, -1):
                st[d] = acc
                acc *= shape[d]
        return tuple(st)
    return tuple(array.strides)


def normalize_storage_spec(aligned_index, shape, dtype, dimensions):
    """utils.py:86-164 of the reference storage package."""
    if shape is None or isinstance(shape, (str, bytes)) or not all(
            isinstance(s, numbers.Integral) for s in shape):
        raise TypeError("shape must be an iterable of ints.")
    shape = tuple(int(s) for s in shape)
    if dimensions is None:
        dimensions = list("IJK"[: len(shape)]) if len(shape) <= 3 else list("IJK") + [
            str(d) for d in range(len(shape) - 3)]
    dimensions = tuple(getattr(d, "__gt_axis_name__", d) for d in dimensions)
    if not all(isinstance(d, str) and (d.isdigit() or d in "IJK") for d in dimensions):
        raise ValueError(f"Invalid dimensions definition: '{dimensions}'")
    if len(shape) != len(dimensions):
        raise ValueError(f"Dimensions ({dimensions}) and shape ({shape}) have non-matching sizes.")
    if any(s <= 0 for s in shape):
        raise ValueError(f"shape ({shape}) contains non-positive value.")
    if aligned_index is None:
        aligned_index = (0,) * len(shape)
    if isinstance(aligned_index, (str, bytes)) or not all(
            isinstance(i, numbers.Integral) for i in aligned_index):
        raise TypeError("aligned_index must be an iterable of ints.")
    aligned_index = tuple(int(i) for i in aligned_index)
    if len(aligned_index) != len(shape):
        raise ValueError(f"Shape ({shape}) and aligned_index ({aligned_index}) have non-matching sizes.")
    if any(i < 0 for i in aligned_index):
        raise ValueError(f"aligned_index ({aligned_index}) contains negative value.")
    dtype = np.dtype(dtype)
    if dtype.shape:
        sub_dtype, sub_shape = dtype.subdtype
        aligned_index = (*aligned_index, *((0,) * dtype.ndim))
        shape = (*shape, *sub_shape)
        dimensions = (*dimensions, *(str(d) for d in range(dtype.ndim)))
        dtype = sub_dtype
    return aligned_index, shape, dtype, dimensions


class _Buffer:
    """Owner of one cudaMalloc allocation."""

    def __init__(self, nbytes: int):
        p = c_void_p()
        _lib.check(_lib.lib().b200_malloc(byref(p), nbytes))
        self.ptr = p.value
        self.nbytes = nbytes

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().b200_free(c_void_p(self.ptr))
                self.ptr = None
        except Exception:
            pass


class DeviceArray:
    """N-d strided view of device memory (no arithmetic: data lives on the GPU for the kernels).

    Exposes ``__cuda_array_interface__`` v3, ``shape / strides (bytes) / dtype / ndim``, basic
    slicing (views), ``asnumpy()``, ``set()`` and ``fill()``.
    """

    __slots__ = ("_buf", "ptr", "shape", "strides", "dtype", "__gt_dims__", "__gt_origin__", "__weakref__")
    __array_priority__ = 100.0

    def __init__(self, buf, ptr, shape, strides, dtype, dims=None, origin=None):
        self._buf = buf
        self.ptr = int(ptr)
        self.shape = tuple(int(s) for s in shape)
        self.strides = tuple(int(s) for s in strides)
        self.dtype = np.dtype(dtype)
        self.__gt_dims__ = dims
        self.__gt_origin__ = origin

    # -- protocol ---------------------------------------------------------------------------
    @property
    def __cuda_array_interface__(self):
        return {
            "shape": self.shape,
            "typestr": self.dtype.str,
            "data": (self.ptr, False),
            "version": 3,
            "strides": self.strides,
            "descr": [("", self.dtype.str)],
            "stream": None,
        }

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def size(self):
        return int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1

    @property
    def itemsize(self):
        return self.dtype.itemsize

    @property
    def nbytes(self):
        return self.size * self.itemsize

    @property
    def device(self):
        return "gpu"

    def __len__(self):
        return self.shape[0]

    def __repr__(self):
        return f"DeviceArray(shape={self.shape}, dtype={self.dtype}, strides={self.strides}, ptr=0x{self.ptr:x})"

    # -- byte extent of the view ---------------------------------------------------------------
    def _extent(self):
        """(lowest byte offset, one-past-highest byte offset) relative to self.ptr."""
        lo = hi = 0
        if self.size == 0:
            return 0, 0
        for n, s in zip(self.shape, self.strides):
            if s >= 0:
                hi += (n - 1) * s
            else:
                lo += (n - 1) * s
        return lo, hi + self.itemsize

    # -- host transfer ---------------------------------------------------------------------------
    def asnumpy(self, stream=None) -> np.ndarray:
        """Copy to host; the result has the same shape and stride ORDER.  A dense (or nearly dense) view is one
        copy of its byte extent; a sparse window of a large array (a few rows of every level) is gathered with one
        pitched copy per index of its slower axes instead of dragging the whole extent over PCIe."""
        lo, hi = self._extent()
        if self.size and self.ndim >= 2 and (hi - lo) > 4 * self.nbytes and min(self.strides) > 0:
            gathered = self._gather(stream)
            if gathered is not None:
                return gathered
        raw = np.empty(hi - lo, dtype=np.uint8)
        if hi > lo:
            _lib.check(_lib.lib().b200_memcpy_d2h(raw.ctypes.data, c_void_p(self.ptr + lo), hi - lo,
                                                  _lib.stream_handle(stream)))
            if stream is not None:
                _lib.check(_lib.lib().b200_stream_sync(_lib.stream_handle(stream)))
        base = raw[-lo:] if lo else raw
        return np.ndarray(self.shape, dtype=self.dtype, buffer=base.data if self.size else None,
                          strides=self.strides) if self.size else np.empty(self.shape, self.dtype)

    def _gather(self, stream=None):
        order = sorted(range(self.ndim), key=lambda d: self.strides[d])  # fastest ... slowest
        item = self.itemsize
        if self.strides[order[0]] != item:
            return None
        # dense host array with the same axis order
        dense = np.empty([self.shape[d] for d in reversed(order)], dtype=self.dtype)  # C order: slowest first
        res = dense.transpose([len(order) - 1 - order.index(d) for d in range(self.ndim)])
        d0, d1 = order[0], order[1]
        width, height = self.shape[d0] * item, self.shape[d1]
        outer = order[2:]
        lib, sh = _lib.lib(), _lib.stream_handle(stream)
        for idx in np.ndindex(*[self.shape[d] for d in outer]):
            src = self.ptr + sum(i * self.strides[d] for i, d in zip(idx, outer))
            dst = res.ctypes.data + sum(i * res.strides[d] for i, d in zip(idx, outer))
            _lib.check(lib.b200_memcpy2d_d2h(c_void_p(dst), res.strides[d1], c_void_p(src), self.strides[d1], width, height, sh))
        if stream is not None:
            _lib.check(lib.b200_stream_sync(sh))
        return res

    def __array__(self, dtype=None, copy=None):
        a = self.asnumpy()
        return a.astype(dtype) if dtype is not None else a

    def set(self, data, stream=None):
        """Copy host `data` (broadcastable to self.shape) into the view."""
        lo, hi = self._extent()
        if hi <= lo:
            return self
        dense = self._is_dense()
        raw = np.empty(hi - lo, dtype=np.uint8)
        if not dense:  # keep the bytes in the gaps (padding / other views)
            _lib.check(_lib.lib().b200_memcpy_d2h(raw.ctypes.data, c_void_p(self.ptr + lo), hi - lo, None))
        view = np.ndarray(self.shape, dtype=self.dtype, buffer=(raw[-lo:] if lo else raw).data,
                          strides=self.strides)
        view[...] = data
        _lib.check(_lib.lib().b200_memcpy_h2d(c_void_p(self.ptr + lo), raw.ctypes.data, hi - lo,
                                              _lib.stream_handle(stream)))
        if stream is not None:  # `raw` is pageable and dies with this frame
            _lib.check(_lib.lib().b200_stream_sync(_lib.stream_handle(stream)))
        return self

    def _is_dense(self):
        lo, hi = self._extent()
        return lo == 0 and hi == self.nbytes

    # -- asynchronous raw transfers to / from pinned host mirrors (see `pinned_like`) ------------
    def copy_from_host(self, host: np.ndarray, stream=None):
        """Asynchronous H2D copy of the whole byte extent from a host mirror with identical
        shape / strides / dtype (``pinned_like(self)``).  One cudaMemcpyAsync, no staging."""
        self._check_mirror(host)
        lo, hi = self._extent()
        if hi > lo:
            _lib.check(_lib.lib().b200_memcpy_h2d(c_void_p(self.ptr + lo), c_void_p(host.ctypes.data + lo),
                                                  hi - lo, _lib.stream_handle(stream)))
        return self

    def copy_to_host(self, host: np.ndarray, stream=None):
        """Asynchronous D2H copy of the whole byte extent into a host mirror (``pinned_like``)."""
        self._check_mirror(host)
        lo, hi = self._extent()
        if hi > lo:
            _lib.check(_lib.lib().b200_memcpy_d2h(c_void_p(host.ctypes.data + lo), c_void_p(self.ptr + lo),
                                                  hi - lo, _lib.stream_handle(stream)))
        return host

    def _check_mirror(self, host):
        if not isinstance(host, np.ndarray) or host.shape != self.shape or \
                host.strides != self.strides or host.dtype != self.dtype:
            raise ValueError("host mirror must be a numpy array with the shape, strides and dtype of "
                             "the device array (use gt4py_b200.storage.pinned_like)")

    def fill(self, value):
        if self._is_dense() and (value == 0 or self.dtype.itemsize == 1):
            byte = int(np.array(value, dtype=self.dtype).view(np.uint8)[0]) if value != 0 else 0
            _lib.check(_lib.lib().b200_memset(c_void_p(self.ptr), byte, self.nbytes, None))
            return self
        return self.set(np.asarray(value, dtype=self.dtype))

    def copy(self) -> "DeviceArray":
        """Device-to-device copy of the whole byte extent (same strides)."""
        lo, hi = self._extent()
        buf = _Buffer(max(hi - lo, 1))
        if hi > lo:
            _lib.check(_lib.lib().b200_memcpy_d2d(c_void_p(buf.ptr), c_void_p(self.ptr + lo), hi - lo, None))
        return type(self)(buf, buf.ptr - lo, self.shape, self.strides, self.dtype,
                          self.__gt_dims__, self.__gt_origin__)

    # -- views -----------------------------------------------------------------------------------
    def __getitem__(self, key):
        if not isinstance(key, tuple):
            key = (key,)
        if any(k is Ellipsis for k in key):
            idx = key.index(Ellipsis)
            fill = self.ndim - (len(key) - 1)
            key = key[:idx] + (slice(None),) * fill + key[idx + 1:]
        key = key + (slice(None),) * (self.ndim - len(key))
        if len(key) != self.ndim:
            raise IndexError("too many indices for DeviceArray")
        ptr, shape, strides = self.ptr, [], []
        for k, n, s in zip(key, self.shape, self.strides):
            if isinstance(k, numbers.Integral):
                k = int(k)
                if k < 0:
                    k += n
                if not 0 <= k < n:
                    raise IndexError("index out of range")
                ptr += k * s
            elif isinstance(k, slice):
                start, stop, step = k.indices(n)
                ptr += start * s
                shape.append(len(range(start, stop, step)))
                strides.append(s * step)
            else:
                raise TypeError("DeviceArray supports only int / slice / Ellipsis indexing")
        return type(self)(self._buf, ptr, shape, strides, self.dtype)

    def __setitem__(self, key, value):
        self[key].set(value.asnumpy() if isinstance(value, DeviceArray) else value)

    def transpose(self, *axes):
        if len(axes) == 1 and not isinstance(axes[0], numbers.Integral):
            axes = tuple(axes[0])
        if not axes:
            axes = tuple(reversed(range(self.ndim)))
        return type(self)(self._buf, self.ptr, [self.shape[a] for a in axes],
                          [self.strides[a] for a in axes], self.dtype)


# ---- public interface (interface.py:40-327) ------------------------------------------------------
def _check_backend(backend):
    if backend != BACKEND_NAME:
        raise RuntimeError(f"Storage preset '{backend}' is not registered in gt4py_b200.storage "
                           f"(only '{BACKEND_NAME}').")


#: class of the arrays the allocators return; a subclass of DeviceArray may be installed (test harnesses that need
#: numpy-like conveniences on top of the plain device view)
_ARRAY_CLASS = [None]


def set_array_class(cls) -> None:
    if cls is not None and not issubclass(cls, DeviceArray):
        raise TypeError("array class must derive from DeviceArray")
    _ARRAY_CLASS[0] = cls


def array_class():
    return _ARRAY_CLASS[0] or DeviceArray


def allocate(shape, lm, dtype, byte_alignment: int, aligned_index) -> "DeviceArray":
    """The reference's padded / aligned allocation (storage/allocators.py:189-275) over cudaMalloc, for any layout
    map and alignment: only the contiguous dimension is padded (to whole aligned blocks), and the start is shifted
    so that `aligned_index` along the contiguous dimension lands on a `byte_alignment` boundary."""
    dtype = np.dtype(dtype)
    item = dtype.itemsize
    ndim = len(shape)
    if len(lm) != ndim or sorted(lm) != list(range(ndim)):
        raise ValueError(f"Invalid layout_map {lm} for shape {shape}")
    if math.gcd(byte_alignment, item) not in (byte_alignment, item):
        raise ValueError(f"Incompatible 'byte_alignment' ({byte_alignment}) and 'dtype' size ({item})")
    per_block = (byte_alignment // item) or 1
    dims_layout = [list(lm).index(i) for i in range(ndim)]  # slowest ... fastest dimension
    padded = list(shape)
    if ndim:
        c = dims_layout[-1]
        padded[c] = math.ceil(shape[c] / per_block) * per_block
    total = item * int(np.prod(padded, dtype=np.int64)) + (byte_alignment - 1)
    strides = [item] * ndim
    acc = item
    for d in range(ndim - 2, -1, -1):
        acc = strides[dims_layout[d]] = acc * padded[dims_layout[d + 1]]
    buf = _Buffer(total)
    if ndim:
        c = dims_layout[-1]
        aligned_off = (per_block * math.ceil(aligned_index[c] / per_block) - aligned_index[c]) * item
    else:
        aligned_off = 0
    mismatch = (byte_alignment - buf.ptr % byte_alignment) % byte_alignment
    byte_offset = (aligned_off + mismatch) % byte_alignment
    return array_class()(buf, buf.ptr + byte_offset, shape, strides, dtype)


def empty(shape, dtype=np.float64, *, backend: str = BACKEND_NAME,
          aligned_index: Optional[Sequence[int]] = None,
          dimensions: Optional[Sequence[str]] = None) -> DeviceArray:
    _check_backend(backend)
    aligned_index, shape, dtype, dimensions = normalize_storage_spec(aligned_index, shape, dtype, dimensions)
    return allocate(shape, layout_map(dimensions), dtype, ALIGNMENT * dtype.itemsize, aligned_index)


def zeros(shape, dtype=np.float64, *, backend: str = BACKEND_NAME, aligned_index=None, dimensions=None):
    arr = empty(shape, dtype, backend=backend, aligned_index=aligned_index, dimensions=dimensions)
    _lib.check(_lib.lib().b200_memset(c_void_p(arr._buf.ptr), 0, arr._buf.nbytes, None))
    return arr


def full(shape, fill_value, dtype=np.float64, *, backend: str = BACKEND_NAME, aligned_index=None, dimensions=None):
    arr = zeros(shape, dtype, backend=backend, aligned_index=aligned_index, dimensions=dimensions)
    arr.set(np.asarray(fill_value, dtype=arr.dtype))
    return arr


def ones(shape, dtype=np.float64, *, backend: str = BACKEND_NAME, aligned_index=None, dimensions=None):
    return full(shape, 1, dtype, backend=backend, aligned_index=aligned_index, dimensions=dimensions)


def from_array(data, dtype=None, *, backend: str = BACKEND_NAME, aligned_index=None, dimensions=None):
    host = data.asnumpy() if isinstance(data, DeviceArray) else np.asarray(data)
    if dtype is None:
        dtype = host.dtype
    dtype = np.dtype(dtype)
    shape = host.shape
    if dtype.shape:
        if shape[-dtype.ndim:] != dtype.shape:
            raise ValueError(f"Incompatible data shape {shape} with dtype of shape {dtype.shape}.")
        shape = shape[: -dtype.ndim]
    arr = zeros(shape, dtype, backend=backend, aligned_index=aligned_index, dimensions=dimensions)
    arr.set(host)
    return arr


class _PinnedBuffer:
    """Owner of one cudaMallocHost allocation, exported to numpy through the buffer protocol."""

    def __init__(self, nbytes: int):
        p = c_void_p()
        _lib.check(_lib.lib().b200_host_alloc(byref(p), max(nbytes, 1)))
        self.ptr = p.value
        self.nbytes = nbytes
        import ctypes

        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * max(nbytes, 1)).from_address(self.ptr))

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().b200_host_free(c_void_p(self.ptr))
                self.ptr = None
        except Exception:
            pass


_PINNED_OWNERS = {}


def pinned_like(array: DeviceArray, data=None) -> np.ndarray:
    """A page-locked host mirror of `array`: a numpy array with the same shape, strides and dtype
    (so that H2D / D2H are single contiguous copies of the byte extent), optionally filled with
    `data`.  Pair with ``DeviceArray.copy_from_host`` / ``copy_to_host``."""
    lo, hi = array._extent()
    owner = _PinnedBuffer(hi - lo)
    raw = owner.array
    host = np.ndarray(array.shape, dtype=array.dtype, buffer=raw.data, offset=-lo, strides=array.strides)
    _PINNED_OWNERS[host.ctypes.data] = owner  # keep the allocation alive with the process
    if data is not None:
        host[...] = data
    return host


def free_pinned(host: np.ndarray) -> None:
    """Release a mirror returned by `pinned_like` (it must not be used afterwards)."""
    owner = _PINNED_OWNERS.pop(host.ctypes.data, None)
    if owner is not None:
        owner.__del__()


def as_device_array(obj) -> DeviceArray:
    """Device view / copy of `obj` -- what ``cupy.asarray`` is to the reference (storage/cartesian/utils.py:176-213):
    an object exposing ``__cuda_array_interface__`` is wrapped without a copy (the wrapper keeps it alive), a host
    array is copied to a fresh allocation with the same element order (C or Fortran)."""
    if isinstance(obj, DeviceArray):
        return obj
    cai = getattr(obj, "__cuda_array_interface__", None)
    if cai is not None:
        shape = tuple(int(x) for x in cai["shape"])
        dtype = np.dtype(cai["typestr"])
        strides = cai.get("strides")
        if strides is None:
            strides, acc = [0] * len(shape), dtype.itemsize
            for d in range(len(shape) - 1, -1, -1):
                strides[d] = acc
                acc *= shape[d]
        return array_class()(obj, int(cai["data"][0] or 0), shape, strides, dtype,
                             getattr(obj, "__gt_dims__", None), getattr(obj, "__gt_origin__", None))
    host = np.asarray(obj)
    if host.dtype == object:
        raise TypeError(f"Cannot convert {type(obj)} to a device array")
    fortran = host.ndim > 1 and host.flags.f_contiguous and not host.flags.c_contiguous
    host = np.asfortranarray(host) if fortran else np.ascontiguousarray(host)
    buf = _Buffer(max(host.nbytes, 1))
    if host.nbytes:
        _lib.check(_lib.lib().b200_memcpy_h2d(c_void_p(buf.ptr), host.ctypes.data, host.nbytes, None))
    return array_class()(buf, buf.ptr, host.shape, host.strides, host.dtype)


def fill_synthetic(array: DeviceArray, seed: int, lo: float, hi: float, *, global_offset=(0, 0), base2d=None,
                   stream=None) -> DeviceArray:
    """Fill a 3-D float64 device array (every element, halo cells included) with counter-based synthetic values in
    global coordinates: element (i, j, k) gets base2d[i, j] + lo + (hi - lo) * u(global_offset + (i, j), k, seed)
    (b200_fill_synthetic, csrc/synthetic.cu).  Bench / test plumbing."""
    from .stencil_object import as_view, make_field_desc

    f = make_field_desc(as_view(array), (0,) * array.ndim)
    b = None
    if base2d is not None:
        b = make_field_desc(as_view(base2d), (0,) * base2d.ndim)
    _lib.check(_lib.lib().b200_fill_synthetic(byref(f), byref(b) if b is not None else None, int(seed),
                                              int(global_offset[0]), int(global_offset[1]), float(lo), float(hi),
                                              _lib.stream_handle(stream)))
    return array


def asnumpy(array) -> np.ndarray:
    """Host copy of a DeviceArray (or pass-through of a host array)."""
    return array.asnumpy() if isinstance(array, DeviceArray) else np.asarray(array)
