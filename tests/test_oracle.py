"""Pin the CPU oracle (oracle/stencils.c) bit-for-bit against the reference's numpy backend.

Golden data: tests/golden/*.npz and digests.json, produced by oracle/make_golden.py which RUNS the
reference (gt4py 1.1.11, numpy backend) on the seeded inputs of oracle/inputs.py.
"""
import glob
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import cpu as oracle
from oracle import inputs as gen

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _cases(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "_[0-9]*.npz")))


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("path", _cases("hdiff"), ids=os.path.basename)
def test_hdiff_bit_exact(path):
    g = np.load(path)
    ni, nj, nk = (int(x) for x in g["shape"])
    f = gen.hdiff_inputs(ni, nj, nk, seed=int(g["seed"]))
    out = oracle.hdiff(f["in_field"], f["out_field"], f["coeff"], domain=(ni, nj, nk))
    assert np.array_equal(out, g["out_field"])  # includes untouched (zero) halo


@pytest.mark.parametrize("path", _cases("simple_hdiff"), ids=os.path.basename)
def test_simple_hdiff_bit_exact(path):
    g = np.load(path)
    ni, nj, nk = (int(x) for x in g["shape"])
    f = gen.hdiff_inputs(ni, nj, nk, seed=int(g["seed"]))
    out = oracle.simple_hdiff(f["in_field"], f["coeff"], f["out_field"], domain=(ni, nj, nk))
    assert np.array_equal(out, g["out_field"])


@pytest.mark.parametrize("path", _cases("vadv"), ids=os.path.basename)
def test_vadv_bit_exact(path):
    g = np.load(path)
    ni, nj, nk = (int(x) for x in g["shape"])
    f = gen.vadv_inputs(ni, nj, nk, seed=int(g["seed"]))
    ut = oracle.vadv(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                     dtr_stage=f["dtr_stage"], domain=(ni, nj, nk))
    assert np.array_equal(ut, g["utens_stage"])


@pytest.mark.parametrize("path", _cases("tridiag"), ids=os.path.basename)
def test_tridiag_bit_exact(path):
    g = np.load(path)
    ni, nj, nk = (int(x) for x in g["shape"])
    f = gen.tridiag_inputs(ni, nj, nk, seed=int(g["seed"]))
    oracle.tridiag(f["inf"], f["diag"], f["sup"], f["rhs"], f["out"], domain=(ni, nj, nk))
    for name in ("sup", "rhs", "out"):
        assert np.array_equal(f[name], g[name]), name


def test_config_size_digests():
    """BASELINE configs 1 and 2 (hdiff 128x128x80, vadv 256x256x80): SHA-256 of the reference output."""
    d = json.load(open(os.path.join(GOLDEN, "digests.json")))["cases"]
    f = gen.hdiff_inputs(128, 128, 80, seed=42)
    out = oracle.hdiff(f["in_field"], f["out_field"], f["coeff"], domain=(128, 128, 80))
    assert _sha(out[2:-2, 2:-2, :]) == d["hdiff_128x128x80"]["out_field_sha256"]
    f = gen.vadv_inputs(256, 256, 80, seed=7)
    ut = oracle.vadv(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                     dtr_stage=f["dtr_stage"], domain=(256, 256, 80))
    assert _sha(ut[:256]) == d["vadv_256x256x80"]["utens_stage_sha256"]
    f = gen.tridiag_inputs(64, 64, 80, seed=11)
    oracle.tridiag(f["inf"], f["diag"], f["sup"], f["rhs"], f["out"], domain=(64, 64, 80))
    for name in ("sup", "rhs", "out"):
        assert _sha(f[name]) == d["tridiag_64x64x80"][name + "_sha256"], name


def test_strided_layout_matches_c_order():
    """The oracle addresses through strides: I-contiguous (device layout) == C-order result."""
    f = gen.hdiff_inputs(13, 9, 4)
    ref = oracle.hdiff(f["in_field"], f["out_field"].copy(), f["coeff"], domain=(13, 9, 4))
    fi = {k: np.asfortranarray(v) for k, v in f.items()}
    out = oracle.hdiff(fi["in_field"], fi["out_field"], fi["coeff"], domain=(13, 9, 4))
    assert np.array_equal(out, ref)


def test_hdiff_f32_oracle_is_pinned_by_the_reference():
    """oracle/hdiff_f32.py (numpy float32 restatement) against the golden outputs the reference's numpy backend
    produced for horizontal_diffusion built with float32 fields and literal_float_precision=32."""
    import json

    from oracle.hdiff_f32 import hdiff_f32

    d = os.path.join(os.path.dirname(__file__), "golden", "generated", "horizontal_diffusion__f32lit")
    g = np.load(os.path.join(d, "golden.npz"))
    with open(os.path.join(d, "case.json")) as fh:
        case = json.load(fh)
    out = g["in__out_field"].copy()
    hdiff_f32(g["in__in_field"], out, g["in__coeff"], domain=tuple(case["domain"]),
              origin=tuple(case["origins"]["in_field"]), out_origin=tuple(case["origins"]["out_field"]),
              coeff_origin=tuple(case["origins"]["coeff"]))
    assert out.dtype == np.float32 and np.array_equal(out, g["out__out_field"])
