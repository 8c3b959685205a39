"""ctypes binding of libgt4py_b200.so (the C ABI declared in include/gt4py_b200.h).

The library is the only compute path of this package: if it cannot be loaded an ImportError-like
RuntimeError is raised on first use -- there is no CPU or PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, byref, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32, c_void_p

import numpy as np

MAX_DIMS = 5

OK, EINVAL, ECUDA, EUNSUP, ENOMEM = 0, -1, -2, -3, -4

# dtype codes of include/gt4py_b200.h
DTYPE_CODES = {
    np.dtype("bool"): 1,
    np.dtype("int8"): 2,
    np.dtype("int16"): 3,
    np.dtype("int32"): 4,
    np.dtype("int64"): 5,
    np.dtype("float32"): 6,
    np.dtype("float64"): 7,
}


class B200Error(RuntimeError):
    """A C-ABI call returned a negative status."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libgt4py_b200 error {code}: {message}")
        self.code = code


class HaloBox(Structure):
    """struct b200_halo_box"""

    _fields_ = [
        ("src", c_void_p),
        ("dst", c_void_p),
        ("src_stride", c_int64 * 3),
        ("dst_stride", c_int64 * 3),
        ("extent", c_int32 * 3),
        ("elem_bytes", c_int32),
    ]


HALO_MAX_PEERS = 8


class HaloSync(Structure):
    """struct b200_halo_sync"""

    _fields_ = [
        ("post_ready", c_void_p * HALO_MAX_PEERS),
        ("post_done", c_void_p * HALO_MAX_PEERS),
        ("wait_ready", c_void_p * HALO_MAX_PEERS),
        ("wait_done", c_void_p * HALO_MAX_PEERS),
        ("npeers", c_int32),
    ]


class _ScalarValue(ctypes.Union):
    _fields_ = [("f64", c_double), ("f32", c_float), ("i64", c_int64), ("i32", c_int32), ("i16", ctypes.c_int16),
                ("i8", ctypes.c_int8), ("b", ctypes.c_uint8)]


class Scalar(Structure):
    """struct b200_scalar"""

    _fields_ = [("value", _ScalarValue), ("dtype", c_int32), ("reserved", c_int32)]


SCALAR_MEMBER = {1: "b", 2: "i8", 3: "i16", 4: "i32", 5: "i64", 6: "f32", 7: "f64"}


class Field(Structure):
    """struct b200_field"""

    _fields_ = [
        ("ptr", c_void_p),
        ("stride", c_int64 * MAX_DIMS),
        ("shape", c_int64 * MAX_DIMS),
        ("origin", c_int32 * MAX_DIMS),
        ("ndim", c_int32),
        ("dtype", c_int32),
    ]


_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GT4PY_B200_LIB") or os.path.join(_PKG, "lib", "libgt4py_b200.so")
_lib = None

_VP = c_void_p
_PROTOTYPES = {
    "b200_abi_version": ([], c_int),
    "b200_last_error": ([], c_char_p),
    "b200_last_kernel": ([], c_char_p),
    "b200_device_count": ([POINTER(c_int)], c_int),
    "b200_set_device": ([c_int], c_int),
    "b200_get_device": ([POINTER(c_int)], c_int),
    "b200_device_sync": ([], c_int),
    "b200_device_props": ([POINTER(c_int64)], c_int),
    "b200_malloc": ([POINTER(_VP), c_size_t], c_int),
    "b200_free": ([_VP], c_int),
    "b200_host_alloc": ([POINTER(_VP), c_size_t], c_int),
    "b200_host_free": ([_VP], c_int),
    "b200_memcpy_h2d": ([_VP, _VP, c_size_t, _VP], c_int),
    "b200_memcpy_d2h": ([_VP, _VP, c_size_t, _VP], c_int),
    "b200_memcpy_d2d": ([_VP, _VP, c_size_t, _VP], c_int),
    "b200_memcpy2d_h2d": ([_VP, c_size_t, _VP, c_size_t, c_size_t, c_size_t, _VP], c_int),
    "b200_memcpy2d_d2h": ([_VP, c_size_t, _VP, c_size_t, c_size_t, c_size_t, _VP], c_int),
    "b200_memset": ([_VP, c_int, c_size_t, _VP], c_int),
    "b200_fill_synthetic": ([POINTER(Field), POINTER(Field), ctypes.c_uint64, c_int64, c_int64, c_double, c_double, _VP], c_int),
    "b200_stream_create": ([POINTER(_VP)], c_int),
    "b200_stream_destroy": ([_VP], c_int),
    "b200_stream_sync": ([_VP], c_int),
    "b200_stream_l2_window": ([_VP, _VP, ctypes.c_size_t, ctypes.c_double, ctypes.POINTER(ctypes.c_size_t)], c_int),
    "b200_event_create": ([POINTER(_VP)], c_int),
    "b200_event_destroy": ([_VP], c_int),
    "b200_event_record": ([_VP, _VP], c_int),
    "b200_event_sync": ([_VP], c_int),
    "b200_event_elapsed_ms": ([_VP, _VP, POINTER(c_float)], c_int),
    "b200_tmap_encode": ([POINTER(Field), POINTER(c_uint32), _VP, POINTER(c_int32)], c_int),
    "b200_graph_begin": ([_VP], c_int),
    "b200_graph_end": ([_VP, POINTER(_VP)], c_int),
    "b200_graph_launch": ([_VP, _VP], c_int),
    "b200_graph_destroy": ([_VP], c_int),
    "b200_stream_wait_event": ([_VP, _VP], c_int),
    "b200_hdiff": ([POINTER(Field), POINTER(Field), POINTER(Field), POINTER(c_int64), c_int, c_int, _VP], c_int),
    "b200_vadv": ([POINTER(Field)] * 5 + [c_double, POINTER(c_int64), c_int, _VP], c_int),
    "b200_vadv_noalloc": ([POINTER(Field)] * 5 + [c_double, POINTER(c_int64), _VP], c_int),
    "b200_tridiag": ([POINTER(Field)] * 5 + [POINTER(c_int64), c_int, _VP], c_int),
    "b200_ipc_get_mem_handle": ([_VP, _VP], c_int),
    "b200_ipc_open_mem_handle": ([_VP, POINTER(_VP)], c_int),
    "b200_ipc_close_mem_handle": ([_VP], c_int),
    "b200_halo_refresh": ([_VP, c_int, c_int, POINTER(HaloSync), ctypes.c_uint64, _VP, c_int, _VP], c_int),
    "b200_halo_wait_done": ([POINTER(HaloSync), ctypes.c_uint64, _VP], c_int),
    "b200_module_load": ([_VP, c_size_t, POINTER(_VP)], c_int),
    "b200_module_unload": ([_VP], c_int),
    "b200_module_get_kernel": ([_VP, c_char_p, POINTER(_VP)], c_int),
    "b200_kernel_set_smem": ([_VP, c_int], c_int),
    "b200_launch": ([_VP, POINTER(c_uint32), POINTER(c_uint32), c_uint32, _VP, POINTER(_VP), c_int], c_int),
    "b200_plan_create": ([c_char_p, c_size_t, _VP, c_size_t, c_int, POINTER(_VP)], c_int),
    "b200_plan_launch": ([_VP, POINTER(c_int64), POINTER(Field), c_int, POINTER(Scalar), c_int, _VP, c_int], c_int),
    "b200_plan_destroy": ([_VP], c_int),
    "b200_plan_last_entries": ([_VP, c_char_p, c_size_t], c_int),
}


def declared_symbols():
    """Every function include/gt4py_b200.h declares (kept in sync by tests/test_abi.py)."""
    return sorted(_PROTOTYPES)


def lib():
    """Load (building first if the sources are newer) and return the ctypes library."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or os.environ.get("GT4PY_B200_REBUILD"):
            from . import build as _build

            _build.build()
        try:
            handle = ctypes.CDLL(LIB_PATH)
        except OSError as exc:  # pragma: no cover
            raise RuntimeError(
                f"gt4py_b200: cannot load {LIB_PATH} ({exc}); run `python -m gt4py_b200.build`. "
                "There is no fallback compute path."
            ) from exc
        for name, (argtypes, restype) in _PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.argtypes = argtypes
            fn.restype = restype
        if handle.b200_abi_version() != 1:
            raise RuntimeError("gt4py_b200: ABI version mismatch between Python host and library")
        _lib = handle
    return _lib


def check(code: int):
    if code != 0:
        msg = lib().b200_last_error()
        raise B200Error(code, msg.decode("utf-8", "replace") if msg else "")


def last_kernel() -> str:
    """Name of the kernel the last library-stencil call of this thread launched."""
    return (lib().b200_last_kernel() or b"").decode()


def device_count() -> int:
    n = c_int(0)
    try:
        check(lib().b200_device_count(byref(n)))
    except B200Error:
        return 0
    return n.value


def device_props() -> dict:
    arr = (c_int64 * 8)()
    check(lib().b200_device_props(arr))
    return {"sm_count": arr[0], "max_smem_optin": arr[1], "l2_bytes": arr[2],
            "cc": (arr[3], arr[4]), "global_mem_mib": arr[5]}


def synchronize():
    check(lib().b200_device_sync())


def set_device(dev: int):
    check(lib().b200_set_device(int(dev)))


class Stream:
    """A non-blocking CUDA stream owned by the library."""

    def __init__(self):
        h = _VP()
        check(lib().b200_stream_create(byref(h)))
        self.handle = h

    def synchronize(self):
        check(lib().b200_stream_sync(self.handle))

    def l2_window(self, ptr: int = 0, nbytes: int = 0, hit_ratio: float = 1.0) -> int:
        """Make accesses of this stream's work to [ptr, ptr + nbytes) persisting in L2 (b200_stream_l2_window; a field
        one stencil writes and the next one reads); no arguments: remove the window.  Returns the bytes set aside."""
        granted = ctypes.c_size_t(0)
        check(lib().b200_stream_l2_window(self.handle, _VP(ptr or None), int(nbytes), float(hit_ratio), byref(granted)))
        return int(granted.value)

    def wait_event(self, event: "Event"):
        check(lib().b200_stream_wait_event(self.handle, event.handle))

    def __del__(self):
        try:
            if self.handle:
                lib().b200_stream_destroy(self.handle)
        except Exception:
            pass


class Event:
    def __init__(self):
        h = _VP()
        check(lib().b200_event_create(byref(h)))
        self.handle = h

    def record(self, stream=None):
        check(lib().b200_event_record(self.handle, stream_handle(stream)))

    def synchronize(self):
        check(lib().b200_event_sync(self.handle))

    def elapsed_ms(self, stop: "Event") -> float:
        ms = c_float(0.0)
        check(lib().b200_event_elapsed_ms(self.handle, stop.handle, byref(ms)))
        return ms.value

    def __del__(self):
        try:
            if self.handle:
                lib().b200_event_destroy(self.handle)
        except Exception:
            pass


def stream_handle(stream) -> c_void_p:
    """Accept None (legacy default stream), a Stream, an int handle, or an object with .cuda_stream
    (e.g. torch.cuda.Stream) / .ptr (cupy)."""
    if stream is None:
        return _VP(None)
    if isinstance(stream, Stream):
        return stream.handle
    if isinstance(stream, _VP):
        return stream
    if isinstance(stream, int):
        return _VP(stream)
    for attr in ("cuda_stream", "ptr", "handle"):
        if hasattr(stream, attr):
            return _VP(int(getattr(stream, attr)))
    raise TypeError(f"cannot interpret {stream!r} as a CUDA stream")
