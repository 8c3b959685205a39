#!/usr/bin/env python
"""Sweep the tuning hooks of the library vadv (csrc/vadv_fused.cu reads them at every launch) at --shape, fields made
on the device:  B200_VADV_STAGES x B200_VADV_BSTAGES x B200_VADV_POL_{STREAM,KEEP,BWD}.  One JSON line per combination,
best first at the end."""
import argparse
import itertools
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gt4py_b200 as gb  # noqa: E402
from gt4py_b200 import _lib, stencils, storage  # noqa: E402
from oracle import inputs as gen  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs=3, default=[2048, 2048, 80])
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    ni, nj, nk = args.shape
    sets = []
    for n in range(2):
        f = {}
        for name in ("utens_stage", "u_stage", "wcon", "u_pos", "utens"):
            seed, lo, hi = gen.SYNTH[name]
            f[name] = storage.empty((ni + 1, nj, nk))
            storage.fill_synthetic(f[name], seed + 10 * n, lo, hi)
        sets.append(f)
    va = stencils.VerticalAdvectionDycore()

    def timed():
        def call(n):
            f = sets[n % 2]
            va(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=gen.DTR_STAGE,
               domain=(ni, nj, nk), validate_args=False)
        for n in range(2):
            call(n)
        gb.synchronize()
        e0, e1 = _lib.Event(), _lib.Event()
        e0.record(None)
        for n in range(args.reps):
            call(n)
        e1.record(None)
        e1.synchronize()
        return e0.elapsed_ms(e1) / args.reps

    rows = []
    base = timed()
    print(json.dumps({"defaults_ms": round(base, 4)}), flush=True)
    for s, sb, ps, pk, pb in itertools.product((4, 5, 6, 7), (4, 6, 8), (0, 1), (0, 2), (0, 1)):
        env = {"B200_VADV_STAGES": s, "B200_VADV_BSTAGES": sb, "B200_VADV_POL_STREAM": ps, "B200_VADV_POL_KEEP": pk,
               "B200_VADV_POL_BWD": pb}
        for k, v in env.items():
            os.environ[k] = str(v)
        rows.append({**env, "ms": round(timed(), 4)})
    for k in list(env):
        os.environ.pop(k)
    rows.sort(key=lambda r: r["ms"])
    for r in rows[:12]:
        print(json.dumps(r), flush=True)
    print(json.dumps({"worst": rows[-1], "defaults_again_ms": round(timed(), 4)}), flush=True)


if __name__ == "__main__":
    main()
