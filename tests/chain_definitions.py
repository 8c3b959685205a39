"""GTScript definitions written for this repository (build container only: needs gt4py importable).

BASELINE.json config 4 is an "FV3-style multi-stage stencil chain (advection, Laplacian smoothing, with
horizontal regions and conditionals)".  The reference ships no such single stencil; this
one composes the constructs its tests exercise separately: ternary upwinding
(/root/reference/tests/cartesian_tests/integration_tests/feature_tests/test_exec_info.py:24-44), the double
Laplacian (same file :46-67), ``with horizontal(region[...])`` edge specialisations
(multi_feature_tests/stencil_definitions.py:450-473) and data-dependent ``if`` masks (:195-203).
tools/make_generated_fixtures.py builds it with the reference's numpy backend (golden) and with ``b200``.
"""
import numpy as np

from gt4py.cartesian.gtscript import (  # noqa: F401
    BACKWARD,
    FORWARD,
    PARALLEL,
    Field,
    I,
    J,
    computation,
    horizontal,
    interval,
    region,
)

F64 = Field[np.float64]


def fv3_chain(phi: F64, u: F64, v: F64, out: F64, *, dt: np.float64, alpha: np.float64):
    with computation(PARALLEL), interval(...):
        # stage 1: face velocities and first-order upwind fluxes
        uf = 0.5 * (u[-1, 0, 0] + u)
        vf = 0.5 * (v[0, -1, 0] + v)
        fx = uf * (phi[-1, 0, 0] if uf > 0 else phi)
        fy = vf * (phi[0, -1, 0] if vf > 0 else phi)
        # the domain's west / south faces are closed (tile edge)
        with horizontal(region[I[0], :]):
            fx = 0.0
        with horizontal(region[:, J[0]]):
            fy = 0.0
        # stage 2: advective update
        adv = phi - dt * ((fx[1, 0, 0] - fx) + (fy[0, 1, 0] - fy))
        # stage 3: Laplacian smoothing, switched off in the tile corners and relaxed on the east edge
        lap = -4.0 * adv + adv[-1, 0, 0] + adv[1, 0, 0] + adv[0, -1, 0] + adv[0, 1, 0]
        with horizontal(region[I[0], J[0]], region[I[-1], J[-1]], region[I[0], J[-1]], region[I[-1], J[0]]):
            lap = 0.0
        with horizontal(region[I[-1], J[0] + 1 : J[-1]]):
            lap = 0.5 * lap
        # stage 4: monotonic limiter: smooth only where it does not steepen the advected field
        gx = lap[1, 0, 0] - lap
        gy = lap[0, 1, 0] - lap
        dx = 0.0 if gx * (adv[1, 0, 0] - adv) > 0.0 else gx
        dy = 0.0 if gy * (adv[0, 1, 0] - adv) > 0.0 else gy
        # stage 5: damped result, clipped below where the scalar went negative
        res = adv - alpha * ((dx - dx[-1, 0, 0]) + (dy - dy[0, -1, 0]))
        if res < 0.0:
            out = 0.0
        else:
            out = res


def column_physics(t: F64, q: F64, w: F64, t_out: F64, *, dt: np.float64):
    """A FORWARD accumulation followed by a BACKWARD relaxation (column kernels with rolling state)."""
    with computation(FORWARD):
        with interval(0, 1):
            acc = q * dt
            t_out = t
        with interval(1, None):
            acc = acc[0, 0, -1] + (q * dt if w > 0.0 else 0.0)
            t_out = t + 0.5 * (acc + acc[0, 0, -1])
    with computation(BACKWARD):
        with interval(-1, None):
            t_out = t_out
        with interval(0, -1):
            t_out = 0.75 * t_out + 0.25 * t_out[0, 0, 1]


DEFINITIONS = {"fv3_chain": (fv3_chain, {}), "column_physics": (column_physics, {})}


def seq_exchange(a: F64, b: F64, out: F64):
    """A sequential loop whose stages exchange values horizontally inside a level: `tmp` is written, then read at
    I / J offsets by the next statement of the SAME level, while `out` of level k-1 feeds level k.  The reference
    evaluates `for k: for stage: all ij` (gtc/numpy/npir_codegen.py:281-296); the b200 lowering runs the stages as
    per-level launches.  (The frontend only admits the pattern for a temporary that lives in ONE interval block and
    does not depend on a written API field: gtir.py:224-240, "non-zero read extent on written fields".)"""
    with computation(FORWARD):
        with interval(0, 1):
            out = b
        with interval(1, None):
            tmp = a * b[0, 0, -1] + b
            out = out[0, 0, -1] + 0.25 * (tmp[1, 0, 0] + tmp[-1, 0, 0] + tmp[0, 1, 0] + tmp[0, -1, 0])


def seq_exchange_backward(a: F64, b: F64, out: F64):
    """The BACKWARD twin of seq_exchange, with a third stage that reads the second one's result at an offset."""
    with computation(BACKWARD):
        with interval(-1, None):
            out = b
        with interval(0, -1):
            tmp = a[0, 0, 1] * 0.5 + b
            mid = tmp[1, 0, 0] - tmp[0, -1, 0]
            out = out[0, 0, 1] * 0.5 + mid[-1, 0, 0] + mid[0, 1, 0] + a


DEFINITIONS.update({"seq_exchange": (seq_exchange, {}), "seq_exchange_backward": (seq_exchange_backward, {})})
