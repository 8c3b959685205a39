"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the config-4 chain (tests/chain_definitions.py:fv3_chain).

Pinned against the reference's numpy backend through tests/golden/generated/fv3_chain/golden.npz (made by
tools/make_generated_fixtures.py, which RUNS /root/reference; tests/test_oracle.py checks this file against
it bit for bit).  Semantics followed: extent analysis (gtc/passes/oir_optimizations + compute_extents,
/root/reference/src/gt4py/cartesian/gtc/passes/gtir_pipeline / oir_access_kinds) means every temporary is
evaluated on the compute domain grown by the offsets of its readers; `horizontal(region[...])` masks are
relative to the COMPUTE DOMAIN bounds (oir_to_npir.py:187-196), also on the grown extents.
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import numpy as np

HALO = 3  # phi is read up to 3 points outside the compute domain


def fv3_chain(phi, u, v, dt, alpha):
    """phi, u, v: arrays with HALO points on every IJ side (shape (ni+6, nj+6, nk)); returns out (ni, nj, nk)."""
    h = HALO
    ni, nj = phi.shape[0] - 2 * h, phi.shape[1] - 2 * h

    def win(a, lo_i, hi_i, lo_j, hi_j, di=0, dj=0):
        """a on the domain grown by [lo_i, hi_i) x [lo_j, hi_j) shifted by (di, dj); a carries HALO."""
        return a[h + lo_i + di:h + ni + hi_i + di, h + lo_j + dj:h + nj + hi_j + dj]

    # stage 1 on the extent [-2, +3) x [-2, +3): everything downstream is sliced out of these
    e = (-2, 3, -2, 3)
    uf = 0.5 * (win(u, *e, di=-1) + win(u, *e))
    vf = 0.5 * (win(v, *e, dj=-1) + win(v, *e))
    fx = uf * np.where(uf > 0, win(phi, *e, di=-1), win(phi, *e))
    fy = vf * np.where(vf > 0, win(phi, *e, dj=-1), win(phi, *e))
    fx[2, :] = 0.0  # region[I[0], :]: compute-domain column 0 sits at index 2 of the grown window
    fy[:, 2] = 0.0  # region[:, J[0]]

    def sub(a, lo_i, hi_i, lo_j, hi_j, di=0, dj=0):
        """slice of a stage-1 array (extent e) on a smaller extent, shifted"""
        return a[lo_i - e[0] + di:a.shape[0] - (e[1] - hi_i) + di, lo_j - e[2] + dj:a.shape[1] - (e[3] - hi_j) + dj]

    # stage 2 on [-2, +2): adv is read at +-1 by lap, which is read at +-1 by the limiter
    a2 = (-2, 2, -2, 2)
    adv = win(phi, *a2) - dt * ((sub(fx, *a2, di=1) - sub(fx, *a2)) + (sub(fy, *a2, dj=1) - sub(fy, *a2)))

    def suba(lo_i, hi_i, lo_j, hi_j, di=0, dj=0):
        return adv[lo_i - a2[0] + di:adv.shape[0] - (a2[1] - hi_i) + di, lo_j - a2[2] + dj:adv.shape[1] - (a2[3] - hi_j) + dj]

    # stage 3 on [-1, +1)
    l3 = (-1, 1, -1, 1)
    lap = ((((-4.0) * suba(*l3) + suba(*l3, di=-1)) + suba(*l3, di=1)) + suba(*l3, dj=-1)) + suba(*l3, dj=1)
    i0, i1, j0, j1 = 1, ni, 1, nj  # compute-domain first / last points inside the [-1, +1) window
    for ii, jj in ((i0, j0), (i1, j1), (i0, j1), (i1, j0)):
        lap[ii, jj] = 0.0
    lap[i1, j0 + 1:j1] = 0.5 * lap[i1, j0 + 1:j1]  # region[I[-1], J[0]+1 : J[-1]]

    def subl(lo_i, hi_i, lo_j, hi_j, di=0, dj=0):
        return lap[lo_i - l3[0] + di:lap.shape[0] - (l3[1] - hi_i) + di, lo_j - l3[2] + dj:lap.shape[1] - (l3[3] - hi_j) + dj]

    # stage 4: dx on [-1, 0) x [0, 0), dy on [0, 0) x [-1, 0)
    def limited(ext, di, dj):
        g = subl(*ext, di=di, dj=dj) - subl(*ext)
        return np.where(g * (suba(*ext, di=di, dj=dj) - suba(*ext)) > 0.0, 0.0, g)

    dx_c, dx_w = limited((0, 0, 0, 0), 1, 0), limited((-1, -1, 0, 0), 1, 0)
    dy_c, dy_s = limited((0, 0, 0, 0), 0, 1), limited((0, 0, -1, -1), 0, 1)
    # stage 5
    res = suba(0, 0, 0, 0) - alpha * ((dx_c - dx_w) + (dy_c - dy_s))
    return np.where(res < 0.0, 0.0, res)


def column_physics(t, q, w, dt):
    """tests/chain_definitions.py:column_physics on (ni, nj, nk) arrays without halo."""
    nk = t.shape[2]
    acc = np.empty_like(t)
    out = np.empty_like(t)
    acc[:, :, 0] = q[:, :, 0] * dt
    out[:, :, 0] = t[:, :, 0]
    for k in range(1, nk):
        acc[:, :, k] = acc[:, :, k - 1] + np.where(w[:, :, k] > 0.0, q[:, :, k] * dt, 0.0)
        out[:, :, k] = t[:, :, k] + 0.5 * (acc[:, :, k] + acc[:, :, k - 1])
    for k in range(nk - 2, -1, -1):
        out[:, :, k] = 0.75 * out[:, :, k] + 0.25 * out[:, :, k + 1]
    return out
