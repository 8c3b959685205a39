"""Host-buffer call path: run a stencil whose fields live in (pinned) HOST memory.

The reference's CPU backends take numpy arrays; a user switching from ``gt:cpu_ifirst`` keeps them on the
host.  ``HostPipeline`` makes that a first-class call: the compute domain is cut into slabs along an axis
the stencil is parallel in (K for hdiff, J for the K-recurrent vadv), and per slab the host->device copy of
the inputs, the kernel and the device->host copy of the outputs run on three streams, so that PCIe traffic
in both directions and the kernels overlap.  The device arrays are ordinary ``DeviceArray`` mirrors of the
host arrays (same shape and strides, ``storage.pinned_like``); a K slab of an I-contiguous field is one
contiguous copy, a J slab one pitched copy.
"""
from __future__ import annotations

from ctypes import c_void_p
from typing import Dict, Sequence

import numpy as np

from . import _lib
from .stencil_object import AccessKind, use_stream


def _slab_copy(dev, host, axis: int, lo: int, hi: int, to_device: bool, stream):
    """Copy the slab [lo, hi) along `axis` (all other indices) between a DeviceArray and its host mirror."""
    if host.shape != dev.shape or host.strides != dev.strides or host.dtype != dev.dtype:
        raise ValueError("host mirror must have the shape, strides and dtype of the device array")
    lib, st = _lib.lib(), _lib.stream_handle(stream)
    strides, shape, item = dev.strides, dev.shape, dev.dtype.itemsize
    order = sorted(range(dev.ndim), key=lambda d: strides[d])  # fastest ... slowest
    dense = strides[order[0]] == item and all(
        strides[order[n + 1]] >= strides[order[n]] * shape[order[n]] for n in range(dev.ndim - 1))
    if not dense or hi <= lo:
        if hi > lo:
            raise ValueError("slab copies need a dense layout")
        return 0
    off = lo * strides[axis]
    d, h = c_void_p(dev.ptr + off), c_void_p(host.ctypes.data + off)
    if axis == order[-1]:  # slowest axis: one contiguous range
        nbytes = (hi - lo - 1) * strides[axis] + _extent_below(dev, order, len(order) - 1)
        fn = lib.b200_memcpy_h2d if to_device else lib.b200_memcpy_d2h
        _lib.check(fn(d, h, nbytes, st) if to_device else fn(h, d, nbytes, st))
        return nbytes
    if dev.ndim >= 2 and axis == order[-2]:  # second slowest: one pitched copy, a row per outer index
        outer = order[-1]
        width = (hi - lo - 1) * strides[axis] + _extent_below(dev, order, len(order) - 2)
        pitch, height = strides[outer], shape[outer]
        if to_device:
            _lib.check(lib.b200_memcpy2d_h2d(d, pitch, h, pitch, width, height, st))
        else:
            _lib.check(lib.b200_memcpy2d_d2h(h, pitch, d, pitch, width, height, st))
        return width * height
    raise ValueError("slab axis must be one of the two slowest axes of the layout")


def _extent_below(dev, order, level):
    """Bytes spanned by one index of order[level] (all faster axes complete)."""
    if level == 0:
        return dev.dtype.itemsize
    d = order[level - 1]
    return (dev.shape[d] - 1) * dev.strides[d] + _extent_below(dev, order, level - 1)


class HostPipeline:
    """``HostPipeline(stencil, device_fields, origin=, domain=, axis="K", chunks=8)(host_fields, **params)``

    `device_fields`: name -> DeviceArray scratch mirrors; `host_fields`: name -> numpy arrays with identical
    shape / strides (pinned for asynchronous copies).  READ fields are uploaded, WRITE / READ_WRITE fields
    downloaded; returns (h2d_bytes, d2h_bytes).  The call returns when the results are in host memory.
    """

    def __init__(self, stencil, device_fields: Dict[str, object], *, origin: Sequence[int], domain: Sequence[int],
                 axis: str = "K", chunks: int = 8):
        from . import Event, Stream

        self.stencil, self.dev = stencil, device_fields
        self.origin, self.domain = tuple(int(o) for o in origin), tuple(int(d) for d in domain)
        self.axis = "IJK".index(axis)
        if self.axis == 0:
            raise ValueError("slabs along the contiguous I axis are not supported")
        n = self.domain[self.axis]
        chunks = max(1, min(int(chunks), n))
        if self.axis == 2 and n // chunks < stencil.domain_info.min_sequential_axis_size:
            chunks = max(1, n // max(1, stencil.domain_info.min_sequential_axis_size))
        bounds = [self.origin[self.axis] + (n * c) // chunks for c in range(chunks + 1)]
        self.slabs = [(bounds[c], bounds[c + 1]) for c in range(chunks) if bounds[c + 1] > bounds[c]]
        self.s_in, self.s_run, self.s_out = Stream(), Stream(), Stream()
        self.ev_in = [Event() for _ in self.slabs]
        self.ev_run = [Event() for _ in self.slabs]
        finfo = stencil.field_info
        self.reads = [k for k, fi in finfo.items() if fi.access & AccessKind.READ]
        self.writes = [k for k, fi in finfo.items() if fi.access & AccessKind.WRITE]
        # a field's slab must cover what the slab's compute domain reads along the slab axis
        self.halo = {k: finfo[k].boundary[self.axis] for k in finfo}

    def __call__(self, host_fields: Dict[str, np.ndarray], preloaded: Sequence[str] = (), **params):
        """`preloaded`: READ fields that are already on the device (e.g. uploaded whole so that their halo
        could be refreshed from neighbour GPUs) and must not be uploaded again."""
        ax = self.axis
        h2d = d2h = 0
        first = self.slabs[0][0]
        cls = type(self.stencil)
        sync_before, cls._device_sync_ = cls._device_sync_, False  # the pipeline orders everything with events
        try:
            h2d, d2h = self._run(host_fields, params, ax, first, set(preloaded))
        finally:
            cls._device_sync_ = sync_before
        self.s_out.synchronize()
        return h2d, d2h

    def _run(self, host_fields, params, ax, first, preloaded):
        h2d = d2h = 0
        for c, (lo, hi) in enumerate(self.slabs):
            for k in self.reads:
                if k in preloaded:
                    continue
                lo_h, hi_h = self.halo[k]
                a = lo - lo_h if lo == first else lo + hi_h  # rows [.., lo + hi_h) came with the previous slab
                b = hi + hi_h
                h2d += _slab_copy(self.dev[k], host_fields[k], ax, max(a, 0), min(b, self.dev[k].shape[ax]), True, self.s_in)
            self.ev_in[c].record(self.s_in)
            self.s_run.wait_event(self.ev_in[c])
            origin = list(self.origin)
            domain = list(self.domain)
            origin[ax], domain[ax] = lo, hi - lo
            with use_stream(self.s_run):
                self.stencil(*[self.dev[n] if n in self.dev else params[n] for n in type(self.stencil)._arg_names_],
                             origin=tuple(origin), domain=tuple(domain))
            self.ev_run[c].record(self.s_run)
            self.s_out.wait_event(self.ev_run[c])
            for k in self.writes:
                d2h += _slab_copy(self.dev[k], host_fields[k], ax, lo, hi, False, self.s_out)
        return h2d, d2h
