#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE ITSELF.

TEST INFRASTRUCTURE ONLY.  This script imports the read-only reference from
/root/reference/src (through the import shims in oracle/ref_shims, which contain no stencil
arithmetic) and executes the reference's own ``numpy`` backend
(/root/reference/src/gt4py/cartesian/backend/numpy_backend.py:28-84) on the reference's own stencil
definitions (/root/reference/tests/cartesian_tests/integration_tests/multi_feature_tests/
stencil_definitions.py: hdiff :316-328, simple hdiff :206-216, tridiagonal :219-232, vadv :235-313).

It only runs in the build container (the reference does not travel to the GPU box); its outputs
(small .npz files + a JSON of SHA-256 digests for the BASELINE config sizes) are committed.

Usage:  python oracle/make_golden.py            # writes tests/golden/*
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import tempfile
import time
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path[:0] = [os.path.join(HERE, "ref_shims"), "/root/reference/src", "/root/reference/tests"]
os.environ.setdefault("GT_CACHE_ROOT", tempfile.mkdtemp(prefix="gtcache_"))
warnings.filterwarnings("ignore")

import numpy as np  # noqa: E402

from oracle import inputs as gen  # noqa: E402  (seeded input recipes shared with tests/bench)


def _ref_stencil(name):
    from gt4py.cartesian import gtscript
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd

    return gtscript.stencil(
        "numpy",
        sd.REGISTRY[name],
        externals=sd.EXTERNALS_REGISTRY[name],
        format_source=False,
        rebuild=True,
    )


def _digest(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_hdiff(st, ni, nj, nk, seed=42, dtype=np.float64):
    f = gen.hdiff_inputs(ni, nj, nk, seed=seed, dtype=dtype)
    out = f["out_field"].copy()
    t0 = time.perf_counter()
    st(f["in_field"], out, f["coeff"], origin=(2, 2, 0), domain=(ni, nj, nk))
    return f, out, time.perf_counter() - t0


def run_simple_hdiff(st, ni, nj, nk, seed=43):
    f = gen.hdiff_inputs(ni, nj, nk, seed=seed)
    out = f["out_field"].copy()
    st(f["in_field"], f["coeff"], out, origin=(2, 2, 0), domain=(ni, nj, nk))
    return f, out


def run_vadv(st, ni, nj, nk, seed=7):
    f = gen.vadv_inputs(ni, nj, nk, seed=seed)
    ut = f["utens_stage"].copy()
    t0 = time.perf_counter()
    st(ut, f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=f["dtr_stage"],
       origin=(0, 0, 0), domain=(ni, nj, nk))
    return f, ut, time.perf_counter() - t0


def run_tridiag(st, ni, nj, nk, seed=11):
    f = gen.tridiag_inputs(ni, nj, nk, seed=seed)
    sup, rhs, out = f["sup"].copy(), f["rhs"].copy(), f["out"].copy()
    st(f["inf"], f["diag"], sup, rhs, out, origin=(0, 0, 0), domain=(ni, nj, nk))
    return f, sup, rhs, out


SMALL = {  # odd, ragged sizes on purpose (not multiples of any tile)
    "hdiff": [(24, 20, 6), (5, 3, 2), (1, 1, 1), (67, 33, 3)],
    "simple_hdiff": [(19, 23, 4)],
    "vadv": [(17, 13, 12), (3, 2, 3), (40, 9, 80)],
    "tridiag": [(11, 7, 9), (2, 3, 2)],
}
CONFIG = {  # BASELINE.json configs that the numpy backend finishes in seconds: digests only
    "hdiff": [(128, 128, 80)],
    "vadv": [(256, 256, 80)],
    "tridiag": [(64, 64, 80)],
}


def main():
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    digests = {"generator": "oracle/make_golden.py", "reference": "gt4py 1.1.11 numpy backend",
               "numpy": np.__version__, "cases": {}}

    st = _ref_stencil("horizontal_diffusion")
    for (ni, nj, nk) in SMALL["hdiff"]:
        f, out, _ = run_hdiff(st, ni, nj, nk)
        np.savez_compressed(os.path.join(gold, f"hdiff_{ni}x{nj}x{nk}.npz"), out_field=out,
                            shape=np.array([ni, nj, nk]), seed=np.array(42))
    for (ni, nj, nk) in CONFIG["hdiff"]:
        run_hdiff(st, ni, nj, nk)  # warm-up
        f, out, dt = run_hdiff(st, ni, nj, nk)
        digests["cases"][f"hdiff_{ni}x{nj}x{nk}"] = {
            "out_field_sha256": _digest(out[2:-2, 2:-2, :]), "seed": 42,
            "sum": float(out.sum()), "ref_numpy_seconds_build_container": dt}

    st = _ref_stencil("simple_horizontal_diffusion")
    for (ni, nj, nk) in SMALL["simple_hdiff"]:
        f, out = run_simple_hdiff(st, ni, nj, nk)
        np.savez_compressed(os.path.join(gold, f"simple_hdiff_{ni}x{nj}x{nk}.npz"), out_field=out,
                            shape=np.array([ni, nj, nk]), seed=np.array(43))

    st = _ref_stencil("vertical_advection_dycore")
    for (ni, nj, nk) in SMALL["vadv"]:
        f, ut, _ = run_vadv(st, ni, nj, nk)
        np.savez_compressed(os.path.join(gold, f"vadv_{ni}x{nj}x{nk}.npz"), utens_stage=ut,
                            shape=np.array([ni, nj, nk]), seed=np.array(7))
    for (ni, nj, nk) in CONFIG["vadv"]:
        f, ut, dt = run_vadv(st, ni, nj, nk)
        digests["cases"][f"vadv_{ni}x{nj}x{nk}"] = {
            "utens_stage_sha256": _digest(ut[:ni]), "seed": 7, "sum": float(ut[:ni].sum()),
            "ref_numpy_seconds_build_container": dt}

    st = _ref_stencil("tridiagonal_solver")
    for (ni, nj, nk) in SMALL["tridiag"]:
        f, sup, rhs, out = run_tridiag(st, ni, nj, nk)
        np.savez_compressed(os.path.join(gold, f"tridiag_{ni}x{nj}x{nk}.npz"), sup=sup, rhs=rhs,
                            out=out, shape=np.array([ni, nj, nk]), seed=np.array(11))
    for (ni, nj, nk) in CONFIG["tridiag"]:
        f, sup, rhs, out = run_tridiag(st, ni, nj, nk)
        digests["cases"][f"tridiag_{ni}x{nj}x{nk}"] = {
            "sup_sha256": _digest(sup), "rhs_sha256": _digest(rhs), "out_sha256": _digest(out),
            "seed": 11}

    with open(os.path.join(gold, "digests.json"), "w") as fh:
        json.dump(digests, fh, indent=1, sort_keys=True)
    print(json.dumps(digests, indent=1))


if __name__ == "__main__":
    main()
