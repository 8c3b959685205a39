"""TEST INFRASTRUCTURE (checker only; never imported by gt4py_b200/): numpy restatement of the pure-float32
build of horizontal_diffusion -- the reference definition
/root/reference/tests/cartesian_tests/integration_tests/multi_feature_tests/stencil_definitions.py:316-328
compiled with float32 fields and ``literal_float_precision=32`` (gtscript.py:242-290), i.e. every literal and
every temporary is float32.  Whole-array float32 operations in the order the reference's numpy backend evaluates
them (npir_codegen.py:243-296).  Pinned bit-for-bit by tests/golden/generated/horizontal_diffusion__f32lit/golden.npz,
which the reference itself produced (tests/test_oracle.py::test_hdiff_f32_oracle_is_pinned_by_the_reference)."""
import numpy as np

F = np.float32


def hdiff_f32(in_field, out_field, coeff, *, domain, origin=(2, 2, 0), out_origin=None, coeff_origin=None,
              limiter=True):
    """in_field needs a halo of 2 around the domain at `origin`; out_field / coeff default to the same origin."""
    assert in_field.dtype == F and out_field.dtype == F and coeff.dtype == F
    oi, oj, ok = origin
    ni, nj, nk = domain
    a = in_field[oi - 2:oi + ni + 2, oj - 2:oj + nj + 2, ok:ok + nk]  # a[2, 2] is the first domain point
    m, n = ni + 2, nj + 2
    # lap on the domain grown by 1: lap[1, 1] is the first domain point
    lap = F(4.0) * a[1:1 + m, 1:1 + n] - (((a[2:2 + m, 1:1 + n] + a[0:m, 1:1 + n]) + a[1:1 + m, 2:2 + n]) + a[1:1 + m, 0:n])
    # flx on i in [-1, ni), j in [0, nj): res = lap[+1, 0] - lap
    res = lap[1:ni + 2, 1:1 + nj] - lap[0:ni + 1, 1:1 + nj]
    d = a[2:ni + 3, 2:2 + nj] - a[1:ni + 2, 2:2 + nj]
    flx = np.where(res * d > F(0.0), F(0.0), res) if limiter else res
    # fly on i in [0, ni), j in [-1, nj)
    res = lap[1:1 + ni, 1:nj + 2] - lap[1:1 + ni, 0:nj + 1]
    d = a[2:2 + ni, 2:nj + 3] - a[2:2 + ni, 1:nj + 2]
    fly = np.where(res * d > F(0.0), F(0.0), res) if limiter else res
    ci, cj, ck = coeff_origin if coeff_origin is not None else origin
    co = coeff[ci:ci + ni, cj:cj + nj, ck:ck + nk]
    out = a[2:2 + ni, 2:2 + nj] - co * (((flx[1:, :] - flx[:-1, :]) + fly[:, 1:]) - fly[:, :-1])
    assert out.dtype == F
    qi, qj, qk = out_origin if out_origin is not None else origin
    out_field[qi:qi + ni, qj:qj + nj, qk:qk + nk] = out
    return out_field
