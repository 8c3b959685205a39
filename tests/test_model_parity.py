"""The windowed single-domain expectation bench.py checks every rank against (oracle/model_parity.py) equals the
plain full-domain run of the same model -- CPU only; also pins the numpy restatement of the synthetic-field hash
against the C++ header the CUDA kernel is built from (through the host double of the C ABI)."""
import numpy as np
import pytest

from oracle import cpu as oracle
from oracle import inputs as gen
from oracle import model_parity as mp


def full_hdiff_chain(gdims, nk, steps):
    gi, gj = gdims
    base = gen.hdiff_base2d(-2, -2, gi + 4, gj + 4, gi, gj)
    seed, lo, hi = gen.SYNTH["in_field"]
    cur = np.asfortranarray(gen.hashed_window(-2, -2, gi + 4, gj + 4, nk, seed, lo, hi, base))
    coeff = np.asfortranarray(np.full(cur.shape, 0.025))
    for _ in range(steps):
        new = cur.copy(order="F")
        oracle.hdiff(cur, new, coeff, origin=(2, 2, 0), domain=(gi, gj, nk))
        cur = new
    return cur[2:-2, 2:-2, :]


def full_vadv_chain(gdims, nk, steps):
    gi, gj = gdims
    f = {k: np.asfortranarray(gen.hashed_window(0, 0, gi + 1, gj, nk, *gen.SYNTH[k]))
         for k in ("utens_stage", "wcon", "u_stage", "u_pos", "utens")}
    p, q = f["utens_stage"], f["wcon"]
    for t in range(steps):
        u, w = (p, q) if t % 2 == 0 else (q, p)
        oracle.vadv(u, f["u_stage"], w, f["u_pos"], f["utens"], dtr_stage=gen.DTR_STAGE, origin=(0, 0, 0), domain=(gi, gj, nk))
    return (p if (steps - 1) % 2 == 0 else q)[:gi]


@pytest.mark.parametrize("grid,dom", [((2, 2), (24, 20, 7)), ((1, 1), (30, 26, 5)), ((4, 2), (12, 14, 6))])
def test_windows_equal_the_full_domain_run(grid, dom):
    steps, nk = 2, dom[2]
    gdims = (grid[0] * dom[0], grid[1] * dom[1])
    want_h, want_v = full_hdiff_chain(gdims, nk, steps), full_vadv_chain(gdims, nk, steps)
    assert np.isfinite(want_h).all() and np.isfinite(want_v).all()
    for px in range(grid[0]):
        for py in range(grid[1]):
            goff = (px * dom[0], py * dom[1])
            for name, (i0, i1, j0, j1) in mp.windows(goff, dom, width=4, length=10):
                assert np.array_equal(mp.hdiff_chain((i0, i1, j0, j1), steps, gdims, nk), want_h[i0:i1, j0:j1]), (name, goff)
                assert np.array_equal(mp.vadv_chain((i0, i1, j0, j1), steps, gdims, nk), want_v[i0:i1, j0:j1]), (name, goff)


def test_model_really_changes_the_exchanged_values():
    gdims, nk = (24, 20), 5
    a, b = full_hdiff_chain(gdims, nk, 1), full_hdiff_chain(gdims, nk, 2)
    assert not np.array_equal(a, b)
    assert not np.array_equal(full_vadv_chain(gdims, nk, 1)[:, :, 1], full_vadv_chain(gdims, nk, 2)[:, :, 1])


def test_numpy_hash_matches_the_c_header():
    """oracle/inputs.py:hashed_uniform == csrc/synthetic.cuh (compiled into the host double of the C ABI)."""
    import subprocess
    import sys
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from tests import emu_abi; emu_abi.activate()\n"
        "import numpy as np\n"
        "from gt4py_b200 import storage\n"
        "from oracle import inputs as gen\n"
        "base = gen.hdiff_base2d(98, -7, 37, 21, 400, 300)\n"
        "a = storage.empty((37, 21, 9), aligned_index=(2, 2, 0))\n"
        "storage.fill_synthetic(a, 42, -0.02, 0.02, global_offset=(98, -7), base2d=storage.from_array(base, dimensions=['I', 'J']))\n"
        "assert np.array_equal(a.asnumpy(), gen.hashed_window(98, -7, 37, 21, 9, 42, -0.02, 0.02, base))\n"
        "b = storage.empty((5, 6, 4))\n"
        "storage.fill_synthetic(b, 75, -0.1, 0.1, global_offset=(1000000, 2000000))\n"
        "w = gen.hashed_window(1000000, 2000000, 5, 6, 4, 75, -0.1, 0.1)\n"
        "assert np.array_equal(b.asnumpy(), w) and w.min() >= -0.1 and w.max() < 0.1 and len(np.unique(w)) == w.size\n"
    ) % root
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
