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


# ---- counter-based fields in GLOBAL coordinates (bench.py, halo parity) ---------------------------------------------
# numpy restatement of gt4py_b200/csrc/synthetic.cuh (splitmix64 finaliser over the global point index): the device
# fills its fields with b200_fill_synthetic, the checker computes any window of the same global field here.
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def hashed_uniform(gi, gj, k, seed: int, lo: float, hi: float) -> np.ndarray:
    """lo + (hi - lo) * u(gi, gj, k, seed) for broadcastable integer index arrays; u in [0, 1)."""
    with np.errstate(over="ignore"):
        off = np.int64(1 << 20)
        x = (np.uint64(seed)
             + (np.asarray(gi, dtype=np.int64) + off).astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
             + (np.asarray(gj, dtype=np.int64) + off).astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)
             + (np.asarray(k, dtype=np.int64) + off).astype(np.uint64) * np.uint64(0x165667B19E3779F9))
        x = x ^ (x >> np.uint64(30))
        x = x * np.uint64(0xBF58476D1CE4E5B9)
        x = x ^ (x >> np.uint64(27))
        x = x * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    u = (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    t = np.float64(hi - lo) * u
    return np.float64(lo) + t


def hashed_window(gi0: int, gj0: int, ni: int, nj: int, nk: int, seed: int, lo: float, hi: float, base2d=None):
    """(ni, nj, nk) window of the global field starting at global index (gi0, gj0), level 0."""
    gi = (gi0 + np.arange(ni, dtype=np.int64))[:, None, None]
    gj = (gj0 + np.arange(nj, dtype=np.int64))[None, :, None]
    k = np.arange(nk, dtype=np.int64)[None, None, :]
    v = hashed_uniform(gi, gj, k, seed, lo, hi)
    return v if base2d is None else base2d[:, :, None] + v


def hdiff_base2d(gi0: int, gj0: int, ni: int, nj: int, gni: int, gnj: int) -> np.ndarray:
    """The analytic field of the reference's hdiff notebook (demo_horizontal_diffusion.ipynb cell 9) over a window of
    a gni x gnj global domain: 5 + 8 (2 + cos(pi (x + 1.5 y)) + sin(2 pi (x + 1.5 y))) / 4,  x = gi / gni, y = gj / gnj.
    Computed on the host (numpy) by both sides -- the device only adds the hashed noise -- so libm never differs."""
    x = (gi0 + np.arange(ni, dtype=np.float64)) / max(gni, 1)
    y = (gj0 + np.arange(nj, dtype=np.float64)) / max(gnj, 1)
    xy = x[:, None] + 1.5 * y[None, :]
    return 5.0 + 8.0 * (2.0 + np.cos(np.pi * xy) + np.sin(2.0 * np.pi * xy)) / 4.0


#: (seed, lo, hi) of every synthetic field of the bench workload
SYNTH = {"in_field": (42, -0.02, 0.02), "coeff": (0, 0.025, 0.025), "out_field": (0, 0.0, 0.0),
         "utens_stage": (71, -1.0, 1.0), "u_stage": (72, -1.0, 1.0), "u_pos": (73, -1.0, 1.0), "utens": (74, -1.0, 1.0),
         "wcon": (75, -0.1, 0.1)}
DTR_STAGE = 3.0 / 20.0
