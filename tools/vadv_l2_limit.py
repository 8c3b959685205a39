#!/usr/bin/env python
"""Does the L2 set-aside (cudaLimitPersistingL2CacheSize, 0 by default) change what the evict_last hint on vadv's
forward u_pos loads achieves?  Times the library vadv at --shape with the default limit, then with the set-aside grown
through b200_stream_l2_window on an unused stream (the limit is device-wide), then once more.

    python tools/vadv_l2_limit.py [--shape 2048 2048 80] [--mb 16 32 64 96]
"""
import argparse
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
    ap.add_argument("--mb", type=int, nargs="+", default=[16, 32, 64, 96])
    ap.add_argument("--reps", type=int, default=20)
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
        for n in range(3):
            f = sets[n % 2]
            va(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=gen.DTR_STAGE, domain=(ni, nj, nk),
               validate_args=False)
        gb.synchronize()
        e0, e1 = _lib.Event(), _lib.Event()
        e0.record(None)
        for n in range(args.reps):
            f = sets[n % 2]
            va(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=gen.DTR_STAGE, domain=(ni, nj, nk),
               validate_args=False)
        e1.record(None)
        e1.synchronize()
        return e0.elapsed_ms(e1) / args.reps

    type(va).configure(device_sync=False) if hasattr(type(va), "configure") else None
    print(json.dumps({"set_aside_mb": 0, "ms": round(timed(), 4), "kernel": _lib.last_kernel()}), flush=True)
    dummy = storage.empty((1024, 1024, 16))
    side = _lib.Stream()
    for mb in args.mb:
        granted = side.l2_window(dummy.ptr, mb << 20)
        print(json.dumps({"set_aside_mb": mb, "granted": granted, "ms": round(timed(), 4)}), flush=True)
    side.l2_window()
    print(json.dumps({"set_aside_mb": "after switching the window off (set-aside returned)", "ms": round(timed(), 4)}), flush=True)


if __name__ == "__main__":
    main()
