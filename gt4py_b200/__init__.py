"""gt4py_b200 -- a B200 (sm_100a) execution backend for GT4Py cartesian stencils.

Layers:
  csrc/ + lib/libgt4py_b200.so  hand-written CUDA kernels behind a C ABI (include/gt4py_b200.h)
  _lib                          ctypes binding of that ABI
  storage                       device allocation mirroring gt4py.storage (no cupy / torch)
  stencil_object                the StencilObject call path (validation, origin/domain, fast path)
  stencils                      hdiff / vadv / tridiagonal as ready-made stencil objects
  backend                       registration in gt4py.cartesian's backend registry + OIR lowering
                                (imported lazily: needs gt4py)

There is no CPU or PyTorch fallback anywhere in this package.
"""
from . import _lib, storage, stencil_object  # noqa: F401
from ._lib import B200Error, Event, Stream, device_count, synchronize  # noqa: F401
from .stencil_object import FrozenStencil, Graph, StencilObject, use_stream  # noqa: F401
from .hostio import HostPipeline  # noqa: F401,E402

__version__ = "0.1.0"
BACKEND_NAME = storage.BACKEND_NAME
