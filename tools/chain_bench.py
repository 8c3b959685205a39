#!/usr/bin/env python
"""Time generated stencils (plan directories) on the device: CUDA events around `reps` calls after warm-up.

    python tools/chain_bench.py [--domain 512 512 128] [--reps 20] plan_dir [plan_dir ...]

Inputs are synthetic (seeded uniform); fields are allocated with a uniform halo of 3.  Prints one JSON line per
plan: ms per call, grid points/s and the algorithmic bandwidth (one read per input field, one write per output
field, 8 B each -- DESIGN.md §8) so that variants of the lowering can be compared on equal terms.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import gt4py_b200 as gb  # noqa: E402
from gt4py_b200 import _lib, generated, storage  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--domain", type=int, nargs=3, default=[512, 512, 128])
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--no-library", action="store_true")
    ap.add_argument("plans", nargs="+")
    args = ap.parse_args()
    ni, nj, nk = args.domain
    h = 3
    rng = np.random.default_rng(0)
    for d in args.plans:
        st = generated.load_plan(d, use_library=not args.no_library)
        type(st).configure(device_sync=False)  # time the kernels, not a host round trip per call
        plan = type(st)._plan_
        fields, nbytes = {}, 0
        for f in plan["fields"]:
            info = plan["info"]["field_info"][f["name"]]
            if info is None:
                fields[f["name"]] = None
                continue
            shape = [n + 2 * h if a in "IJ" else n for a, n in zip("IJK", (ni, nj, nk)) if a in info["axes"]]
            dims = list(info["axes"])
            host = rng.uniform(-1.0, 1.5, shape).astype(f["dtype"])
            fields[f["name"]] = storage.from_array(host, aligned_index=[h if a in "IJ" else 0 for a in dims], dimensions=dims)
            nbytes += int(np.prod([n for a, n in zip("IJK", (ni, nj, nk)) if a in info["axes"]])) * host.dtype.itemsize * (
                (1 if info["access"] & 1 else 0) + (1 if info["access"] & 2 else 0))
        params = {s["name"]: np.dtype(s["dtype"]).type(0.05) for s in plan["scalars"]}
        origin = {"_all_": (h, h, 0)}
        call = lambda: st(**fields, **params, origin=origin, domain=(ni, nj, nk), validate_args=False)  # noqa: E731
        for _ in range(3):
            call()
        gb.synchronize()
        e0, e1 = _lib.Event(), _lib.Event()
        e0.record(None)
        for _ in range(args.reps):
            call()
        e1.record(None)
        e1.synchronize()
        ms = e0.elapsed_ms(e1) / args.reps
        print(json.dumps({"plan": d, "domain": [ni, nj, nk], "ms": round(ms, 4), "points_per_s": ni * nj * nk / ms * 1e3,
                          "algorithmic_GBps": round(nbytes / ms / 1e6, 1),
                          "kernels": [(k["kind"], k.get("tile"), k.get("smem_bytes")) for k in plan["kernels"]]}), flush=True)


if __name__ == "__main__":
    main()
