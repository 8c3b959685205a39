"""The library's hand-written hot stencils as callable stencil objects.

Each class has the call signature, ``field_info`` / ``parameter_info`` / ``domain_info`` and
validation behaviour that the reference produces for the GTScript definition of the same name in
/root/reference/tests/cartesian_tests/integration_tests/multi_feature_tests/stencil_definitions.py
(probe of the reference's ``make_args_data_from_gtir``, backend/module_generator.py:56-106), and
launches the matching sm_100a kernel through the C ABI (include/gt4py_b200.h).

    from gt4py_b200 import stencils, storage
    stencils.horizontal_diffusion(in_field, out_field, coeff, origin=(2, 2, 0), domain=(ni, nj, nk))
"""
from __future__ import annotations

from ctypes import byref, c_double, c_int64

import numpy as np

from . import _lib
from .stencil_object import (AccessKind, Boundary, DomainInfo, FieldInfo, ParameterInfo,
                             StencilObject)

_F64 = np.dtype("float64")
_IJK = ("I", "J", "K")
_B0 = Boundary(((0, 0), (0, 0), (0, 0)))


def _fi(access, boundary=_B0):
    return FieldInfo(access=access, boundary=boundary, axes=_IJK, data_dims=(), dtype=_F64)


def _dom(domain):
    return (c_int64 * 3)(*domain)


class _LibraryStencil(StencilObject):
    _variant_ = 0  # 0 = auto; tests/bench select specific kernels with `with_variant`

    @classmethod
    def with_variant(cls, variant: int):
        """A sibling stencil class pinned to one kernel variant (see include/gt4py_b200.h)."""
        name = f"{cls.__name__}_v{variant}"
        sub = type(name, (cls,), {"_variant_": int(variant)})
        return sub()


class HorizontalDiffusion(_LibraryStencil):
    """horizontal_diffusion(in_field, out_field, coeff) -- stencil_definitions.py:316-328."""

    _arg_names_ = ("in_field", "out_field", "coeff")
    _gt_domain_info_ = DomainInfo(("I", "J"), "K", 0, 3)
    _gt_field_info_ = {
        "in_field": _fi(AccessKind.READ, Boundary(((2, 2), (2, 2), (0, 0)))),
        "out_field": _fi(AccessKind.WRITE),
        "coeff": _fi(AccessKind.READ),
    }
    _gt_parameter_info_ = {}
    _gt_options_ = {"name": "horizontal_diffusion", "module": __name__}
    _limiter_ = 1

    def _launch(self, domain, f, params, stream):
        _lib.check(_lib.lib().b200_hdiff(byref(f["in_field"]), byref(f["out_field"]),
                                         byref(f["coeff"]), _dom(domain), self._limiter_,
                                         self._variant_, _lib.stream_handle(stream)))


class HorizontalDiffusionFloat32(HorizontalDiffusion):
    """The same definition built with float32 fields and ``literal_float_precision=32`` (every literal and
    temporary float32): csrc/hdiff.cu with T = float, variants 1-3."""

    _gt_field_info_ = {
        "in_field": FieldInfo(access=AccessKind.READ, boundary=Boundary(((2, 2), (2, 2), (0, 0))), axes=_IJK,
                              data_dims=(), dtype=np.dtype("float32")),
        "out_field": FieldInfo(access=AccessKind.WRITE, boundary=_B0, axes=_IJK, data_dims=(),
                               dtype=np.dtype("float32")),
        "coeff": FieldInfo(access=AccessKind.READ, boundary=_B0, axes=_IJK, data_dims=(), dtype=np.dtype("float32")),
    }
    _gt_options_ = {"name": "horizontal_diffusion_float32", "module": __name__}


class SimpleHorizontalDiffusion(HorizontalDiffusion):
    """simple_horizontal_diffusion(in_field, coeff, out_field) -- stencil_definitions.py:206-216."""

    _arg_names_ = ("in_field", "coeff", "out_field")
    _gt_field_info_ = {
        "in_field": _fi(AccessKind.READ, Boundary(((2, 2), (2, 2), (0, 0)))),
        "coeff": _fi(AccessKind.READ),
        "out_field": _fi(AccessKind.WRITE),
    }
    _gt_options_ = {"name": "simple_horizontal_diffusion", "module": __name__}
    _limiter_ = 0


class VerticalAdvectionDycore(_LibraryStencil):
    """vertical_advection_dycore(utens_stage, u_stage, wcon, u_pos, utens, *, dtr_stage) with
    externals BET_M = BET_P = 0.5 -- stencil_definitions.py:235-313."""

    _arg_names_ = ("utens_stage", "u_stage", "wcon", "u_pos", "utens", "dtr_stage")
    _gt_domain_info_ = DomainInfo(("I", "J"), "K", 3, 3)
    _gt_field_info_ = {
        "utens_stage": _fi(AccessKind.READ_WRITE),
        "u_stage": _fi(AccessKind.READ),
        "wcon": _fi(AccessKind.READ, Boundary(((0, 1), (0, 0), (-1, 0)))),
        "u_pos": _fi(AccessKind.READ),
        "utens": _fi(AccessKind.READ),
    }
    _gt_parameter_info_ = {"dtr_stage": ParameterInfo(AccessKind.READ, _F64)}
    _gt_constants_ = {"BET_M": 0.5, "BET_P": 0.5}
    _gt_options_ = {"name": "vertical_advection_dycore", "module": __name__}

    def _launch(self, domain, f, params, stream):
        _lib.check(_lib.lib().b200_vadv(byref(f["utens_stage"]), byref(f["u_stage"]),
                                        byref(f["wcon"]), byref(f["u_pos"]), byref(f["utens"]),
                                        c_double(params["dtr_stage"]), _dom(domain),
                                        self._variant_, _lib.stream_handle(stream)))


class TridiagonalSolver(_LibraryStencil):
    """tridiagonal_solver(inf, diag, sup, rhs, out) -- stencil_definitions.py:219-232."""

    _arg_names_ = ("inf", "diag", "sup", "rhs", "out")
    _gt_domain_info_ = DomainInfo(("I", "J"), "K", 2, 3)
    _gt_field_info_ = {
        "inf": _fi(AccessKind.READ, Boundary(((0, 0), (0, 0), (-1, 0)))),
        "diag": _fi(AccessKind.READ),
        "sup": _fi(AccessKind.READ_WRITE),
        "rhs": _fi(AccessKind.READ_WRITE),
        "out": _fi(AccessKind.WRITE),
    }
    _gt_parameter_info_ = {}
    _gt_options_ = {"name": "tridiagonal_solver", "module": __name__}

    def _launch(self, domain, f, params, stream):
        _lib.check(_lib.lib().b200_tridiag(byref(f["inf"]), byref(f["diag"]), byref(f["sup"]),
                                           byref(f["rhs"]), byref(f["out"]), _dom(domain),
                                           self._variant_, _lib.stream_handle(stream)))


horizontal_diffusion = HorizontalDiffusion()
simple_horizontal_diffusion = SimpleHorizontalDiffusion()
vertical_advection_dycore = VerticalAdvectionDycore()
tridiagonal_solver = TridiagonalSolver()

REGISTRY = {
    "horizontal_diffusion": horizontal_diffusion,
    "simple_horizontal_diffusion": simple_horizontal_diffusion,
    "vertical_advection_dycore": vertical_advection_dycore,
    "tridiagonal_solver": tridiagonal_solver,
}
