#!/usr/bin/env python
"""Kernel micro-benchmark: time the hdiff / vadv / tridiag kernel variants with CUDA events.

    python tools/kbench.py --stencil hdiff --shape 1024 1024 80 --variants 1 2 3 --iters 20

Inputs are resident in HBM; three buffer sets are rotated so no field can stay in the 126 MB L2.
Prints one JSON line per variant with ms, grid-points/s and effective (algorithmic) GB/s.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gt4py_b200 import Event, stencils, storage, synchronize  # noqa: E402
from oracle import inputs as gen  # noqa: E402

ALG_BYTES = {"hdiff": 24, "vadv": 48, "tridiag": 56}


def make_sets(stencil, ni, nj, nk, nsets, dtype="float64"):
    sets = []
    if stencil == "hdiff":
        f = {k: v.astype(dtype) for k, v in gen.hdiff_inputs(ni, nj, nk).items()}
        for _ in range(nsets):
            sets.append({k: storage.from_array(v, backend="b200", aligned_index=(2, 2, 0)) for k, v in f.items()})
    elif stencil == "vadv":
        f = gen.vadv_inputs(ni, nj, nk)
        for _ in range(nsets):
            sets.append({k: (storage.from_array(v, backend="b200") if k != "dtr_stage" else v) for k, v in f.items()})
    else:
        f = gen.tridiag_inputs(ni, nj, nk)
        for _ in range(nsets):
            sets.append({k: storage.from_array(v, backend="b200") for k, v in f.items()})
    return sets


def call(stencil, st, d, dom):
    if stencil == "hdiff":
        st(d["in_field"], d["out_field"], d["coeff"], origin=(2, 2, 0), domain=dom)
    elif stencil == "vadv":
        st(d["utens_stage"], d["u_stage"], d["wcon"], d["u_pos"], d["utens"], dtr_stage=d["dtr_stage"],
           origin=(0, 0, 0), domain=dom)
    else:
        st(d["inf"], d["diag"], d["sup"], d["rhs"], d["out"], origin=(0, 0, 0), domain=dom)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stencil", default="hdiff")
    ap.add_argument("--shape", type=int, nargs=3, default=[1024, 1024, 80])
    ap.add_argument("--variants", type=int, nargs="+", default=[0])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--sets", type=int, default=3)
    ap.add_argument("--dtype", default="float64", choices=["float64", "float32"],
                    help="float32: hdiff only (the pure-float32 build; 12 B/point algorithmic)")
    ap.add_argument("--sweep", default="", help="environment sweep, e.g. 'B200_HDIFF_NW=2,3;B200_HDIFF_NS=3,4' "
                    "(cartesian product; the launchers read these hooks at every call)")
    ap.add_argument("--check", type=int, default=-1, help="compare each run's output of set 0 bit-for-bit with "
                    "this variant's (hdiff: out_field, vadv: utens_stage needs fresh inputs -> hdiff only)")
    a = ap.parse_args()
    ni, nj, nk = a.shape
    dom = (ni, nj, nk)
    cls = {"hdiff": stencils.HorizontalDiffusion, "vadv": stencils.VerticalAdvectionDycore,
           "tridiag": stencils.TridiagonalSolver}[a.stencil]
    if a.dtype == "float32":
        assert a.stencil == "hdiff", "float32 library kernel: hdiff only"
        cls = stencils.HorizontalDiffusionFloat32
    alg_bytes = ALG_BYTES[a.stencil] // (2 if a.dtype == "float32" else 1)
    sets = make_sets(a.stencil, ni, nj, nk, a.sets, a.dtype)
    n = ni * nj * nk
    import itertools

    axes = [(kv.split("=")[0], kv.split("=")[1].split(",")) for kv in a.sweep.split(";") if kv]
    combos = [dict(zip([k for k, _ in axes], vals)) for vals in itertools.product(*[v for _, v in axes])] or [{}]
    ref_out = None
    if a.check >= 0 and a.stencil == "hdiff":
        st = cls.with_variant(a.check)
        call(a.stencil, st, sets[0], dom)
        synchronize()
        ref_out = sets[0]["out_field"].asnumpy().copy()
    for v, env in itertools.product(a.variants, combos):
        os.environ.update(env)
        st = cls.with_variant(v)
        type(st).configure(device_sync=False)
        for it in range(a.warmup):
            call(a.stencil, st, sets[it % a.sets], dom)
        synchronize()
        evs = [(Event(), Event()) for _ in range(a.iters)]
        for it in range(a.iters):
            evs[it][0].record()
            call(a.stencil, st, sets[it % a.sets], dom)
            evs[it][1].record()
        synchronize()
        ms = np.array([s.elapsed_ms(e) for s, e in evs])
        best, med = float(ms.min()), float(np.median(ms))
        extra = dict(env)
        if ref_out is not None:
            extra["bit_exact"] = bool(np.array_equal(sets[0]["out_field"].asnumpy(), ref_out))
        print(json.dumps({"stencil": a.stencil, "shape": dom, "variant": v, **extra, "ms_min": round(best, 4),
                          "ms_median": round(med, 4), "gpts_s_median": round(n / med / 1e6, 2),
                          "alg_GBs_median": round(n * alg_bytes / med / 1e6, 1),
                          "alg_GBs_best": round(n * alg_bytes / best / 1e6, 1), "dtype": a.dtype}), flush=True)


if __name__ == "__main__":
    main()
