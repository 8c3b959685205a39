"""Single-domain expectation for bench.py's multi-step parity model (TEST / BENCH-CHECK INFRASTRUCTURE).

bench.py runs, after its timed region, T steps of a small "model" on the synthetic GLOBAL fields of oracle/inputs.py
(values are a function of the global point index, so any window can be computed anywhere):

  hdiff chain   X_0 = in_field recipe;  X_{t+1} = hdiff(X_t, coeff) on the global compute domain, the global halo
                frame keeps the recipe's values.  On the device the in / out arrays swap roles every step and the
                width-2 halo of the in-role array is refreshed from the neighbour ranks before each step.
  vadv chain    (U_0, W_0) = (utens_stage, wcon) recipes;  U'_t = vadv(U_t, u_stage, W_t, u_pos, utens);
                U_{t+1} = W_t,  W_{t+1} = U'_t  (the arrays swap roles; column GI of each array is never written and
                keeps its recipe value).  The +I halo column of the W-role array is refreshed before each step.

The exchanged values therefore CHANGE every step: a refresh that is late, early, missing or mis-addressed shows up as
a mismatch against the single-domain result computed here with the C oracle (oracle/stencils.c) on a window of the
global domain grown by the steps' dependency cone.  Only the core of the window is compared.
"""
from __future__ import annotations

import numpy as np

from . import cpu as oracle
from . import inputs as gen


def _f(a):
    return np.asfortranarray(a)  # I contiguous, like the device layout (the C oracle wants equal strides)


def hdiff_chain(window, steps: int, gdims, nk: int):
    """Expected X_steps on `window` = (i0, i1, j0, j1) in global compute coordinates (half-open)."""
    gi, gj = gdims
    i0, i1, j0, j1 = window
    g = 2 * steps + 2
    wi0, wi1 = max(i0 - g, -2), min(i1 + g, gi + 2)
    wj0, wj1 = max(j0 - g, -2), min(j1 + g, gj + 2)
    ni, nj = wi1 - wi0, wj1 - wj0
    base = gen.hdiff_base2d(wi0, wj0, ni, nj, gi, gj)
    seed, lo, hi = gen.SYNTH["in_field"]
    cur = _f(gen.hashed_window(wi0, wj0, ni, nj, nk, seed, lo, hi, base))
    coeff = _f(np.full(cur.shape, gen.SYNTH["coeff"][1]))
    # cells the window can update: inside the global compute domain and two cells away from the window's edge
    a0, a1 = max(wi0 + 2, 0), min(wi1 - 2, gi)
    b0, b1 = max(wj0 + 2, 0), min(wj1 - 2, gj)
    for _ in range(steps):
        new = cur.copy(order="F")
        oracle.hdiff(cur, new, coeff, origin=(a0 - wi0, b0 - wj0, 0), domain=(a1 - a0, b1 - b0, nk))
        cur = new
    return cur[i0 - wi0:i1 - wi0, j0 - wj0:j1 - wj0, :]


def vadv_chain(window, steps: int, gdims, nk: int):
    """Expected (U_steps-role array, W_steps-role array) contents on `window`; returns the array that was written
    last (the U-role array of the final step) restricted to the window."""
    gi, gj = gdims
    i0, i1, j0, j1 = window
    wi0, wi1 = i0, min(i1 + steps, gi + 1)  # wcon is read at i + 1: one more column per step
    ni, nj = wi1 - wi0, j1 - j0

    def rec(name):
        seed, lo, hi = gen.SYNTH[name]
        return _f(gen.hashed_window(wi0, j0, ni, nj, nk, seed, lo, hi))

    p, q = rec("utens_stage"), rec("wcon")
    u_stage, u_pos, utens = rec("u_stage"), rec("u_pos"), rec("utens")
    dom_i = min(wi1 - 1, gi) - wi0  # needs column i + 1 inside the window
    for t in range(steps):
        u, w = (p, q) if t % 2 == 0 else (q, p)
        oracle.vadv(u, u_stage, w, u_pos, utens, dtr_stage=gen.DTR_STAGE, origin=(0, 0, 0), domain=(dom_i, nj, nk))
    last = p if (steps - 1) % 2 == 0 else q
    return last[: i1 - i0, :, :]


def windows(goff, dom, width: int = 8, length: int = 96):
    """Check windows of one rank's block (global offset `goff`, extent `dom`): a strip along each of its four edges
    (where the halo cells matter) and one interior patch; (name, (i0, i1, j0, j1)) in global coordinates."""
    gi0, gj0 = goff
    ni, nj = int(dom[0]), int(dom[1])
    w, ln = min(width, ni, nj), min(length, ni, nj)
    mi, mj = gi0 + (ni - ln) // 2, gj0 + (nj - ln) // 2
    return [
        ("west", (gi0, gi0 + w, mj, mj + ln)),
        ("east", (gi0 + ni - w, gi0 + ni, mj, mj + ln)),
        ("south", (mi, mi + ln, gj0, gj0 + w)),
        ("north", (mi, mi + ln, gj0 + nj - w, gj0 + nj)),
        ("corner_ne", (gi0 + ni - w, gi0 + ni, gj0 + nj - w, gj0 + nj)),
        ("corner_sw", (gi0, gi0 + w, gj0, gj0 + w)),
        ("interior", (gi0 + ni // 2, gi0 + ni // 2 + w, gj0 + nj // 2, gj0 + nj // 2 + w)),
    ]
