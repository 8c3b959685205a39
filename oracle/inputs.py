"""Seeded synthetic inputs for hdiff / vadv / tridiagonal (TEST + BENCH INFRASTRUCTURE; numpy only).

The recipes are the ones SURVEY.md §8(d) / BASELINE.md §5 fix:
  * hdiff: analytic field of /root/reference/examples/cartesian/demo_horizontal_diffusion.ipynb
    (cell 9) generalised to any shape, plus default_rng(seed) * 1e-2 noise so that both limiter
    branches are exercised; coeff = 0.025; arrays are (ni+4, nj+4, nk), origin (2, 2, 0).
  * vadv: default_rng(seed): fields ~ U(-1, 1), wcon ~ U(-0.1, 0.1), dtr_stage = 3/20 (system is
    strictly diagonally dominant); arrays are (ni+1, nj, nk), origin (0, 0, 0).
  * tridiagonal: diagonally dominant random system.

Arrays are returned in *logical* (I, J, K) index order as plain C-ordered numpy arrays; callers that
need the device layout (I contiguous) transpose/copy them.
"""
from __future__ import annotations

import numpy as np

HDIFF_HALO = 2


def hdiff_inputs(ni: int, nj: int, nk: int, seed: int = 42, dtype=np.float64):
    h = HDIFF_HALO
    x = (np.arange(ni + 2 * h, dtype=np.float64) - h) / max(ni, 1)
    y = (np.arange(nj + 2 * h, dtype=np.float64) - h) / max(nj, 1)
    xy = x[:, None] + 1.5 * y[None, :]
    base = 5.0 + 8.0 * (2.0 + np.cos(np.pi * xy) + np.sin(2.0 * np.pi * xy)) / 4.0
    rng = np.random.default_rng(seed)
    noise = rng.standard_normal((ni + 2 * h, nj + 2 * h, nk)) * 1e-2
    in_field = (base[:, :, None] + noise).astype(dtype)
    coeff = np.full((ni + 2 * h, nj + 2 * h, nk), 0.025, dtype=dtype)
    out_field = np.zeros((ni + 2 * h, nj + 2 * h, nk), dtype=dtype)
    return {"in_field": in_field, "coeff": coeff, "out_field": out_field}


def vadv_inputs(ni: int, nj: int, nk: int, seed: int = 7, dtype=np.float64):
    rng = np.random.default_rng(seed)
    shape = (ni + 1, nj, nk)
    f = {}
    for name in ("utens_stage", "u_stage", "u_pos", "utens"):
        f[name] = rng.uniform(-1.0, 1.0, shape).astype(dtype)
    f["wcon"] = rng.uniform(-0.1, 0.1, shape).astype(dtype)
    f["dtr_stage"] = dtype(3.0 / 20.0)
    return f


def tridiag_inputs(ni: int, nj: int, nk: int, seed: int = 11, dtype=np.float64):
    rng = np.random.default_rng(seed)
    shape = (ni, nj, nk)
    f = {
        "inf": rng.uniform(-1.0, 1.0, shape).astype(dtype),
        "sup": rng.uniform(-1.0, 1.0, shape).astype(dtype),
        "rhs": rng.uniform(-1.0, 1.0, shape).astype(dtype),
        "out": np.zeros(shape, dtype=dtype),
    }
    f["diag"] = (3.0 + rng.uniform(0.0, 1.0, shape)).astype(dtype)
    return f
