"""Run-time call path of the b200 backend: the host-side mirror of the reference's StencilObject.

Mirrors /root/reference/src/gt4py/cartesian/stencil_object.py:
  * ``_call_run`` (argument normalisation, (domain, origin) memoisation, validation)  :534-605
  * ``_normalize_origins``                                                            :499-532
  * ``_get_max_domain``                                                               :298-343
  * ``_validate_args`` (same checks, same exception types and messages)               :345-497
  * ``FrozenStencil`` / ``freeze``                                                    :94-128, 607-636
and the ``__call__`` / ``run`` / exec_info bookkeeping of the generated module
(/root/reference/src/gt4py/cartesian/backend/templates/stencil_module.py.in:91-169).

Differences that are the point of this backend ("StencilObject call fast path"):
  * fields are anything exposing ``__cuda_array_interface__`` (our DeviceArray, cupy, torch, numba);
    no cupy import, no ``asarray`` copy; host arrays are rejected with a TypeError instead of being
    silently copied to a temporary device array;
  * the memo key is a plain tuple of shapes/origins (no ``pickle.dumps``), and the ctypes field
    descriptors handed to the C ABI are cached per argument and reused while the pointer, shape,
    strides and origin are unchanged.

This module does not import gt4py: it is what runs on the GPU box.  When gt4py *is* importable,
``gt4py_b200.backend`` registers a real ``gt4py.cartesian.backend.Backend`` whose generated stencil
classes reuse the launch helpers below from within the reference's own StencilObject.
"""
from __future__ import annotations

import enum
import sys
import time
import warnings
from dataclasses import dataclass
from typing import Any, Dict, Optional, Sequence, Tuple

import numpy as np

from . import _lib, storage

BACKEND_NAME = storage.BACKEND_NAME


class AccessKind(enum.IntFlag):
    NONE = 0
    READ = 1
    WRITE = 2
    READ_WRITE = 3


class Boundary(tuple):
    """((lo_i, hi_i), (lo_j, hi_j), (lo_k, hi_k)) -- gtc/definitions.py:235-330 in the reference."""

    def __new__(cls, pairs):
        return super().__new__(cls, tuple((int(a), int(b)) for a, b in pairs))

    @property
    def lower_indices(self):
        return tuple(p[0] for p in self)

    @property
    def upper_indices(self):
        return tuple(p[1] for p in self)


@dataclass(frozen=True)
class FieldInfo:
    """cartesian/definitions.py:77-118 of the reference."""

    access: AccessKind
    boundary: Boundary
    axes: Tuple[str, ...]
    data_dims: Tuple[int, ...]
    dtype: np.dtype

    @property
    def domain_mask(self):
        return tuple(a in self.axes for a in "IJK")

    @property
    def domain_ndim(self):
        return len(self.axes)

    @property
    def mask(self):
        return (*self.domain_mask, *((True,) * len(self.data_dims)))

    @property
    def ndim(self):
        return len(self.axes) + len(self.data_dims)


@dataclass(frozen=True)
class ParameterInfo:
    access: AccessKind
    dtype: np.dtype


@dataclass(frozen=True)
class DomainInfo:
    parallel_axes: Tuple[str, ...]
    sequential_axis: str
    min_sequential_axis_size: int
    ndim: int


class ArrayView:
    """What the launcher needs from a field argument, read once from __cuda_array_interface__."""

    __slots__ = ("ptr", "shape", "strides", "dtype", "obj")

    def __init__(self, ptr, shape, strides, dtype, obj=None):
        self.ptr, self.shape, self.strides, self.dtype, self.obj = ptr, shape, strides, dtype, obj

    @property
    def ndim(self):
        return len(self.shape)

    def transpose(self, axes):
        return ArrayView(self.ptr, tuple(self.shape[a] for a in axes),
                         tuple(self.strides[a] for a in axes), self.dtype, self.obj)


def as_view(arg, name="<field>") -> ArrayView:
    """Device view of `arg`; the counterpart of storage_utils.asarray(arg, device="gpu")
    (storage/cartesian/utils.py:176-229) without any copy."""
    if isinstance(arg, storage.DeviceArray):
        return ArrayView(arg.ptr, arg.shape, arg.strides, arg.dtype, arg)
    cai = getattr(arg, "__cuda_array_interface__", None)
    if cai is None:
        if hasattr(arg, "__array_interface__") or isinstance(arg, np.ndarray):
            raise TypeError(
                f"Field '{name}' is a host array; the '{BACKEND_NAME}' backend needs device memory "
                f"(an object exposing __cuda_array_interface__, e.g. gt4py_b200.storage.from_array)."
            )
        raise TypeError(f"Field '{name}' of type {type(arg)} does not expose __cuda_array_interface__")
    shape = tuple(int(s) for s in cai["shape"])
    dtype = np.dtype(cai["typestr"])
    strides = cai.get("strides")
    if strides is None:
        strides, acc = [0] * len(shape), dtype.itemsize
        for d in range(len(shape) - 1, -1, -1):
            strides[d] = acc
            acc *= shape[d]
    return ArrayView(int(cai["data"][0] or 0), shape, tuple(int(s) for s in strides), dtype, arg)


def _get_dims(arg):
    dims = getattr(arg, "__gt_dims__", None)
    return tuple(str(d) for d in dims) if dims is not None else None


def _get_origin(arg):
    origin = getattr(arg, "__gt_origin__", None)
    return tuple(int(o) for o in origin) if origin is not None else None


def _filter_mask(seq, mask):
    return tuple(s for s, m in zip(seq, mask) if m)


class StencilObject:
    """Base class of b200 stencils; one immutable singleton per subclass (stencil_object.py:146-204).

    Subclasses define the class attributes ``_gt_backend_/_gt_source_/_gt_domain_info_/
    _gt_field_info_/_gt_parameter_info_/_gt_constants_/_gt_options_`` (same names as the reference's
    generated modules, stencil_module.py.in:49-61), ``_arg_names_`` (call signature order) and
    ``_launch(self, domain, fields: dict[str, Field struct], params, stream)``.
    """

    _gt_backend_ = BACKEND_NAME
    _gt_source_: Dict[str, str] = {}
    _gt_constants_: Dict[str, Any] = {}
    _gt_options_: Dict[str, Any] = {}
    _arg_names_: Tuple[str, ...] = ()
    _arg_defaults_: Dict[str, Any] = {}

    def __new__(cls, *args, **kwargs):
        if cls.__dict__.get("_instance") is None:
            cls._instance = object.__new__(cls)
            cls._domain_origin_cache = {}
            cls._field_desc_cache = {}
        return cls._instance

    def __setattr__(self, key, value):
        raise AttributeError("Attempting a modification of an attribute in a frozen class")

    def __delattr__(self, item):
        raise AttributeError("Attempting a deletion of an attribute in a frozen class")

    def __eq__(self, other):
        return type(self) is type(other)

    def __hash__(self):
        return hash(type(self).__qualname__)

    backend = property(lambda self: type(self)._gt_backend_)
    source = property(lambda self: type(self)._gt_source_)
    domain_info = property(lambda self: type(self)._gt_domain_info_)
    field_info = property(lambda self: type(self)._gt_field_info_)
    parameter_info = property(lambda self: type(self)._gt_parameter_info_)
    constants = property(lambda self: type(self)._gt_constants_)
    options = property(lambda self: type(self)._gt_options_)

    # ---- the user-facing call (stencil_module.py.in:91-158) ------------------------------------
    def __call__(self, *args, domain=None, origin=None, validate_args=True, exec_info=None, **kwargs):
        if exec_info is not None:
            exec_info["call_start_time"] = time.perf_counter()
        names = type(self)._arg_names_
        if len(args) > len(names):
            raise TypeError(f"{type(self).__name__}() takes {len(names)} positional arguments but "
                            f"{len(args)} were given")
        bound = dict(zip(names, args))
        for k, v in kwargs.items():
            if k not in names:
                raise TypeError(f"{type(self).__name__}() got an unexpected keyword argument '{k}'")
            if k in bound:
                raise TypeError(f"{type(self).__name__}() got multiple values for argument '{k}'")
            bound[k] = v
        for k in names:
            if k not in bound:
                if k in type(self)._arg_defaults_:
                    bound[k] = type(self)._arg_defaults_[k]
                else:
                    raise TypeError(f"{type(self).__name__}() missing required argument: '{k}'")
        finfo, pinfo = self.field_info, self.parameter_info
        field_args = {n: bound[n] for n in names if n in finfo}
        parameter_args = {n: bound[n] for n in names if n in pinfo}
        self._call_run(field_args, parameter_args, domain, origin,
                       validate_args=validate_args, exec_info=exec_info)
        if exec_info is not None:
            exec_info["call_end_time"] = time.perf_counter()
            if exec_info.setdefault("__aggregate_data", False):
                info = exec_info.setdefault(type(self).__name__, {})
                info["call_start_time"] = exec_info["call_start_time"]
                info["call_end_time"] = exec_info["call_end_time"]
                info["call_time"] = info["call_end_time"] - info["call_start_time"]
                info["total_call_time"] = info.get("total_call_time", 0.0) + info["call_time"]
                info["ncalls"] = info.get("ncalls", 0) + 1
                info["run_time"] = exec_info["run_end_time"] - exec_info["run_start_time"]
                info["total_run_time"] = info.get("total_run_time", 0.0) + info["run_time"]

    # ---- origin / domain handling ----------------------------------------------------------------
    @staticmethod
    def _make_origin_dict(origin):
        try:
            if isinstance(origin, dict):
                return {str(k): v for k, v in origin.items()}
            if origin is None:
                return {}
            if isinstance(origin, (str, bytes)):
                raise TypeError
            if hasattr(origin, "__iter__"):  # Index.from_value: any length, integers only (gtc/definitions.py:78-100)
                vals = list(origin)
                if not vals or not all(isinstance(v, (int, np.integer)) and not isinstance(v, bool) for v in vals):
                    raise TypeError
                return {"_all_": tuple(int(v) for v in vals)}
            if isinstance(origin, (int, np.integer)):  # Index.from_k: the same value on every axis
                return {"_all_": (int(origin),) * 3}
        except Exception:
            pass
        raise ValueError("Invalid 'origin' value ({})".format(origin))

    @staticmethod
    def _normalize_origins(views, dims_origins, field_infos, origin):
        origin = StencilObject._make_origin_dict(origin)
        all_origin = origin.get("_all_")
        for name, fi in field_infos.items():
            assert name in views, f"Missing value for '{name}' field."
            fo = origin.get(name)
            if fo is not None:
                if len(fo) != fi.ndim:
                    assert len(fo) == fi.domain_ndim, (
                        f"Invalid origin specification ({fo}) for '{name}' field.")
                    origin[name] = (*fo, *((0,) * len(fi.data_dims)))
                else:
                    origin[name] = tuple(fo)
            elif all_origin is not None:
                origin[name] = (*_filter_mask(all_origin, fi.domain_mask), *((0,) * len(fi.data_dims)))
            elif dims_origins.get(name) is not None:
                origin[name] = dims_origins[name]
            else:
                origin[name] = (0,) * fi.ndim
        return origin

    @staticmethod
    def _get_max_domain(views, domain_info, field_infos, origin, *, squeeze=True):
        big = sys.maxsize
        max_domain = [big] * domain_info.ndim
        for name, fi in field_infos.items():
            if fi.access != AccessKind.NONE:
                view = views.get(name)
                assert view is not None, f"Invalid value for '{name}' field."
                upper = _filter_mask(fi.boundary.upper_indices, fi.domain_mask)
                fo = origin[name]
                sizes = iter(view.shape[i] - (fo[i] + upper[i]) for i in range(fi.domain_ndim))
                for d, m in enumerate(fi.domain_mask):
                    if m:
                        max_domain[d] = min(max_domain[d], next(sizes))
        if squeeze:
            return tuple(v if v != big else 1 for v in max_domain)
        return tuple(max_domain)

    # ---- validation (stencil_object.py:345-497: same checks, exception types and messages) --------
    def _validate_args(self, views, dims, param_args, domain, origin):
        domain_ndim = self.domain_info.ndim
        if len(domain) != domain_ndim:
            raise ValueError(f"Invalid 'domain' value '{domain}'")
        try:
            domain = tuple(int(d) for d in domain)
        except Exception as ex:
            raise ValueError("Invalid 'domain' value ({})".format(domain)) from ex
        if not all(d > 0 for d in domain):
            raise ValueError(f"Compute domain contains zero sizes '{domain}')")
        max_domain = self._get_max_domain(views, self.domain_info, self.field_info, origin, squeeze=False)
        if not all(d <= m for d, m in zip(domain, max_domain)):
            offending = []
            for name, info in self.field_info.items():
                used = self._get_max_domain(views, self.domain_info, {name: info}, origin, squeeze=False)
                if any(u < d for u, d in zip(used, domain)):
                    offending.append((name, used))
            raise ValueError(
                f"Compute domain too large for stencil {self.options.get('name', type(self).__name__)}: \n"
                f"  Stencil domain is {domain} but field indexation leads to read outside of bounds.\n"
                f"  Check region/horizontal offsets or interval/vertical offsets, or stencil domain.\n"
                f"  Offending fields (name, size with offset removed): {offending}"
            )
        if domain[2] < self.domain_info.min_sequential_axis_size:
            raise ValueError(
                f"Compute domain too small. Sequential axis is {domain[2]}, but must be at least "
                f"{self.domain_info.min_sequential_axis_size}."
            )
        for name, fi in self.field_info.items():
            if fi.access == AccessKind.NONE:
                continue
            if name not in views or views[name] is None:
                raise ValueError(f"Missing value for '{name}' field.")
            view = views[name]
            full_dims = tuple(list(fi.axes) + [str(d) for d in range(len(fi.data_dims))])
            if not storage.is_optimal_layout(view, full_dims):
                warnings.warn(
                    f"The layout of the field '{name}' is not recommended for this backend."
                    f"This may lead to performance degradation. Please consider using the"
                    f"provided allocators in `gt4py.storage`.",
                    stacklevel=4,
                )
            if not view.dtype == fi.dtype:
                raise TypeError(f"The dtype of field '{name}' is '{view.dtype}' instead of '{fi.dtype}'")
            mask = fi.domain_mask
            if view.ndim != fi.domain_ndim + len(fi.data_dims):
                raise ValueError(
                    f"Storage for '{name}' has {view.ndim} dimensions but the API signature "
                    f"expects {fi.domain_ndim + len(fi.data_dims)} ('{fi.axes}[{fi.data_dims}]')"
                )
            if dims.get(name) is not None and full_dims != dims[name]:
                raise ValueError(
                    f"Storage for '{name}' has dimensions '{dims[name]}' but the API signature "
                    f"expects '[{', '.join(fi.axes)}]'"
                    + (f" and {len(fi.data_dims)}" if fi.data_dims else "")
                )
            if tuple(view.shape[fi.domain_ndim:]) != tuple(fi.data_dims):
                raise ValueError(
                    f"Field '{name}' expects data dimensions {fi.data_dims} but got "
                    f"{view.shape[fi.domain_ndim:]}"
                )
            lower = _filter_mask(fi.boundary.lower_indices, mask)
            upper = _filter_mask(fi.boundary.upper_indices, mask)
            field_origin = tuple(origin[name][: fi.domain_ndim])
            if any(o < lo for o, lo in zip(field_origin, lower)):
                raise ValueError(
                    f"Origin for field {name} too small. Must be at least {lower}, is {field_origin}"
                )
            spatial = _filter_mask(domain, mask)
            min_shape = tuple(lb + d + ub for lb, d, ub in zip(lower, spatial, upper))
            if any(m > s for m, s in zip(min_shape, view.shape)):
                raise ValueError(
                    f"Shape of field {name} is {view.shape} but must be at least {min_shape} for "
                    f"given domain and origin."
                )
        for name, pi in self.parameter_info.items():
            if pi.access != AccessKind.NONE:
                if name not in param_args:
                    raise ValueError(f"Missing value for '{name}' parameter.")
                if np.dtype(type(param_args[name])) != pi.dtype:
                    raise TypeError(
                        f"The type of parameter '{name}' is '{type(param_args[name])}' instead of '{pi.dtype}'"
                    )

    # ---- _call_run (stencil_object.py:534-605) ---------------------------------------------------
    def _call_run(self, field_args, parameter_args, domain, origin, *, validate_args=True, exec_info=None):
        if exec_info is not None:
            exec_info["call_run_start_time"] = time.perf_counter()
        views, dims, dims_origins = {}, {}, {}
        for name, arg in field_args.items():
            if arg is None:
                views[name] = None
                continue
            view = as_view(arg, name)
            d = _get_dims(arg)
            if d is not None:
                sorted_dims = [a for a in "IJK" if a in d]
                sorted_dims += [str(x) for x in sorted(int(x) for x in d if str(x).isdigit())]
                view = view.transpose([d.index(sd) for sd in sorted_dims])
                d = tuple(sorted_dims)
            views[name], dims[name], dims_origins[name] = view, d, _get_origin(arg)
        key = (
            tuple((n, v.shape, v.dtype.str, dims_origins.get(n) or (0, 0, 0)) for n, v in views.items() if v is not None),
            tuple(parameter_args),
            None if domain is None else tuple(domain),
            _freeze_origin(origin),
        )
        cache = type(self)._domain_origin_cache
        hit = cache.get(key)
        if hit is None:
            origin = self._normalize_origins(views, dims_origins, self.field_info, origin)
            if domain is None:
                domain = self._get_max_domain(views, self.domain_info, self.field_info, origin)
            if validate_args:
                self._validate_args(views, dims, parameter_args, domain, origin)
            domain = tuple(int(d) for d in domain)
            cache[key] = (domain, origin)
        else:
            domain, origin = hit
        self.run(_domain_=domain, _origin_=origin, exec_info=exec_info, **views, **parameter_args)
        if exec_info is not None:
            exec_info["call_run_end_time"] = time.perf_counter()

    def clean_call_args_cache(self):
        type(self)._domain_origin_cache.clear()
        type(self)._field_desc_cache.clear()

    # ---- run: build the C-ABI descriptors and launch (stencil_module.py.in:160-169) ---------------
    _device_sync_ = True  # like gt:gpu's default (gtc_common.py:288-296); see configure()

    @classmethod
    def configure(cls, *, device_sync: Optional[bool] = None):
        """Backend options that the reference passes at build time (`device_sync`,
        gtcpp_backend.py:176-179)."""
        if device_sync is not None:
            cls._device_sync_ = bool(device_sync)
        return cls

    def run(self, _domain_, _origin_, exec_info, **kwargs):
        if exec_info is not None:
            exec_info["domain"] = _domain_
            exec_info["origin"] = _origin_
            exec_info["run_start_time"] = time.perf_counter()
        cls = type(self)
        descs = {}
        cache = cls._field_desc_cache
        for name in cls._gt_field_info_:
            arg = kwargs.get(name)
            if arg is None:
                descs[name] = None
                continue
            view = arg if type(arg) is ArrayView else as_view(arg, name)
            org = _origin_[name]
            sig = (view.ptr, view.shape, view.strides, org)
            memo = cache.get(name)
            if memo is None or memo[0] != sig:
                memo = (sig, make_field_desc(view, org))
                cache[name] = memo
            descs[name] = memo[1]
        params = {name: kwargs[name] for name in cls._gt_parameter_info_}
        stream = current_stream()
        self._launch(_domain_, descs, params, stream)
        if cls._device_sync_ and not getattr(_tls, "capturing", False):
            if stream is None:
                _lib.synchronize()
            else:
                _lib.check(_lib.lib().b200_stream_sync(_lib.stream_handle(stream)))
        if exec_info is not None:
            exec_info["run_end_time"] = time.perf_counter()

    def _launch(self, domain, fields, params, stream):  # pragma: no cover - abstract
        raise NotImplementedError

    def freeze(self, *, origin, domain):
        return FrozenStencil(self, origin, tuple(domain))


def _freeze_origin(origin):
    if origin is None or isinstance(origin, int):
        return origin
    if isinstance(origin, dict):
        return tuple((str(k), tuple(v)) for k, v in origin.items())
    return tuple(origin)


def make_field_desc(view: ArrayView, origin) -> "_lib.Field":
    """Fill a struct b200_field from a device view (strides converted from bytes to elements)."""
    if view.ndim > _lib.MAX_DIMS:
        raise ValueError(f"fields with more than {_lib.MAX_DIMS} dimensions are not supported")
    code = _lib.DTYPE_CODES.get(view.dtype)
    if code is None:
        raise TypeError(f"unsupported field dtype {view.dtype}")
    f = _lib.Field()
    f.ptr = view.ptr
    item = view.dtype.itemsize
    for d in range(view.ndim):
        if view.strides[d] % item:
            raise ValueError("field strides must be multiples of the item size")
        f.stride[d] = view.strides[d] // item
        f.shape[d] = view.shape[d]
        f.origin[d] = origin[d]
    f.ndim = view.ndim
    f.dtype = code
    return f


# ---- current stream (thread-local), the hook for overlap with other work ----------------------------
import threading  # noqa: E402

_tls = threading.local()


def current_stream():
    return getattr(_tls, "stream", None)


class use_stream:
    """``with use_stream(s): stencil(...)`` launches on `s` (a gt4py_b200 Stream, a torch/cupy stream
    or a raw handle) instead of the legacy default stream."""

    def __init__(self, stream):
        self.stream = stream

    def __enter__(self):
        self.prev = current_stream()
        _tls.stream = self.stream
        return self.stream

    def __exit__(self, *exc):
        _tls.stream = self.prev
        return False


class Graph:
    """A sequence of stencil calls captured once into a CUDA graph and replayed with a single launch:

        g = Graph()
        with g:                              # nothing executes here: the launches are recorded
            advect(phi, u, v, tmp, origin=..., domain=...)
            diffuse(tmp, out, alpha=0.1, origin=..., domain=...)
        for _ in range(steps):
            g.launch()                       # one cudaGraphLaunch instead of two Python call paths
        g.synchronize()

    Field pointers, origins, domains and scalar parameters are baked in at capture time (call the stencils once
    before capturing: their first call may allocate scratch and build caches).  Inside the capture the stencils'
    `device_sync` is ignored.  The reference's callers pay `StencilObject.__call__` per stencil per step
    (stencil_object.py:534-605); this is the call-path fusion of SURVEY.md 8(f) rank 2."""

    def __init__(self, stream=None):
        self.stream = stream if stream is not None else _lib.Stream()
        self._exec = None
        self._ctx = None
        self._refs = []  # device buffers (scratch, TMA descriptors) whose addresses the captured launches carry

    def __enter__(self):
        if self._exec is not None:
            raise RuntimeError("Graph: already captured")
        self._ctx = use_stream(self.stream)
        self._ctx.__enter__()
        _lib.check(_lib.lib().b200_graph_begin(_lib.stream_handle(self.stream)))
        _tls.capturing = True
        _tls.graph_refs = self._refs
        return self

    def __exit__(self, exc_type, exc, tb):
        from ctypes import byref, c_void_p

        _tls.capturing = False
        _tls.graph_refs = None
        handle = c_void_p()
        rc = _lib.lib().b200_graph_end(_lib.stream_handle(self.stream), byref(handle))
        self._ctx.__exit__(exc_type, exc, tb)
        if exc_type is None:
            _lib.check(rc)
            self._exec = handle
        return False

    def launch(self, stream=None):
        if self._exec is None:
            raise RuntimeError("Graph: nothing captured")
        _lib.check(_lib.lib().b200_graph_launch(self._exec, _lib.stream_handle(stream if stream is not None else self.stream)))

    def synchronize(self):
        _lib.check(_lib.lib().b200_stream_sync(_lib.stream_handle(self.stream)))

    def __del__(self):
        try:
            if self._exec is not None:
                _lib.lib().b200_graph_destroy(self._exec)
        except Exception:
            pass


@dataclass(frozen=True)
class FrozenStencil:
    """Stencil with pre-computed domain and origin (stencil_object.py:94-128): no validation, no
    normalisation -- arguments go straight to ``run``."""

    stencil_object: StencilObject
    origin: Dict[str, Tuple[int, ...]]
    domain: Tuple[int, ...]

    def __post_init__(self):
        for name, fi in self.stencil_object.field_info.items():
            if name not in self.origin or len(self.origin[name]) != fi.ndim:
                raise ValueError(
                    f"'{name}' origin {self.origin.get(name)} is not a {fi.ndim}-dimensional integer tuple"
                )

    def __call__(self, **kwargs):
        assert "origin" not in kwargs and "domain" not in kwargs
        exec_info = kwargs.get("exec_info")
        if exec_info is not None:
            exec_info["call_run_start_time"] = time.perf_counter()
        so = self.stencil_object
        args = {name: kwargs[name] for name in so.field_info}
        args.update({name: kwargs[name] for name in so.parameter_info})
        so.run(_domain_=self.domain, _origin_=self.origin, exec_info=exec_info, **args)
        if exec_info is not None:
            exec_info["call_run_end_time"] = time.perf_counter()
