#!/usr/bin/env python
"""L2 residency of an inter-stencil field: two generated stencils in a row,
A: hdiff(in, coeff) -> mid and B: hdiff(mid, coeff) -> out, where `mid` is written by A and read by B.

  whole         A over the whole domain, then B over the whole domain: `mid` (ni x nj x nk x 8 B, larger than L2) makes
                a round trip through HBM;
  slab          A then B on K slabs of --slab levels (the slab of `mid` fits L2): B should find it in L2, unless the
                streaming traffic of the other fields evicted it;
  slab_persist  the same with the stream's access-policy window on the slab of `mid` (b200_stream_l2_window:
                accesses to it persist in the set-aside part of L2, everything else streams);
  slab_overlap  slabs again, B of slab s on a second stream next to A of slab s + 1 (a slab's launches alone do not
                fill the GPU).

Prints one JSON line per mode (ms per pass, CUDA events around --reps passes).  With --once every mode runs exactly one
pass without warm-up: that is the form to run under `ncu --cache-control none --metrics dram__bytes_read.sum,
dram__bytes_write.sum` (the kernels of the modes follow each other in the order above: 2, 2 n, 2 n, 2 n launches for n slabs).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import gt4py_b200 as gb  # noqa: E402
from gt4py_b200 import _lib, generated, storage  # noqa: E402
from gt4py_b200.stencil_object import Graph, use_stream  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--domain", type=int, nargs=3, default=[512, 512, 128])
    ap.add_argument("--slab", type=int, default=16)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--library", action="store_true", help="the hand-written hdiff kernel (HBM-bound) instead of the "
                    "generated one")
    ap.add_argument("--graph", action="store_true", help="capture a pass in a CUDA graph and time its replays: no "
                    "host time between the small launches of the slab modes")
    ap.add_argument("--modes", nargs="+", default=["whole", "slab", "slab_persist", "slab_overlap"])
    args = ap.parse_args()
    ni, nj, nk = args.domain
    h = 2
    st = generated.load_plan(os.path.join(ROOT, "tests", "golden", "generated", "horizontal_diffusion"), use_library=args.library)
    type(st).configure(device_sync=False)
    shape = (ni + 2 * h, nj + 2 * h, nk)
    f = {}
    for n, (name, seed) in enumerate((("inp", 11), ("coeff", 12), ("mid", 13), ("out", 14))):
        f[name] = storage.empty(shape, aligned_index=(h, h, 0))
        storage.fill_synthetic(f[name], seed, -1.0, 1.5)
    stream = _lib.Stream()
    level_bytes = f["mid"].strides[2]
    info = {"domain": [ni, nj, nk], "slab_levels": args.slab, "slab_bytes_of_mid": level_bytes * args.slab,
            "kernel": "library hdiff" if args.library else [k["kind"] for k in type(st)._plan_["kernels"]]}

    def stencil_pair(k0, ks):
        # A on the full slab, B two points inside it (it reads mid at +-2: A's results only)
        st(f["inp"], f["mid"], f["coeff"], origin=(h, h, k0), domain=(ni, nj, ks), validate_args=False)
        st(f["mid"], f["out"], f["coeff"], origin=(2 * h, 2 * h, k0), domain=(ni - 2 * h, nj - 2 * h, ks), validate_args=False)

    stream_b = _lib.Stream()
    events = [_lib.Event() for _ in range((nk + args.slab - 1) // args.slab + 1)]

    def one_pass(mode):
        if mode == "whole":
            stencil_pair(0, nk)
            return
        if mode == "slab_overlap":
            # A of slab s + 1 next to B of slab s on a second stream: the small launches of one slab do not fill the
            # GPU on their own.  (B of slab s waits for A of slab s; the slabs of mid are distinct.)
            n = 0
            for k0 in range(0, nk, args.slab):
                ks = min(args.slab, nk - k0)
                with use_stream(stream):
                    st(f["inp"], f["mid"], f["coeff"], origin=(h, h, k0), domain=(ni, nj, ks), validate_args=False)
                events[n].record(stream)
                stream_b.wait_event(events[n])
                with use_stream(stream_b):
                    st(f["mid"], f["out"], f["coeff"], origin=(2 * h, 2 * h, k0), domain=(ni - 2 * h, nj - 2 * h, ks),
                       validate_args=False)
                n += 1
            events[n].record(stream_b)
            stream.wait_event(events[n])
            return
        for k0 in range(0, nk, args.slab):
            ks = min(args.slab, nk - k0)
            if mode == "slab_persist":
                granted = stream.l2_window(f["mid"].ptr + k0 * level_bytes, ks * level_bytes)
                info["persisting_bytes_granted"] = granted
            stencil_pair(k0, ks)

    if args.graph:
        for mode in [m for m in args.modes if m != "slab_overlap"]:
            with use_stream(stream):
                one_pass(mode)  # first call: scratch, caches, the L2 set-aside (nothing of that inside a capture)
            stream.synchronize()
            g = Graph(stream)
            with g:
                one_pass(mode)
            for _ in range(2):
                g.launch()
            g.synchronize()
            e0, e1 = _lib.Event(), _lib.Event()
            e0.record(stream)
            for _ in range(args.reps):
                g.launch()
            e1.record(stream)
            e1.synchronize()
            print(json.dumps({"mode": mode + "+graph", "ms_per_pass": round(e0.elapsed_ms(e1) / args.reps, 4), **info,
                              "algorithmic_bytes_whole": 2 * 3 * 8 * ni * nj * nk}), flush=True)
            if mode == "slab_persist":
                stream.l2_window()
        return
    with use_stream(stream):
        for mode in args.modes:
            if not args.once:
                for _ in range(2):
                    one_pass(mode)
            stream.synchronize()
            e0, e1 = _lib.Event(), _lib.Event()
            e0.record(stream)
            reps = 1 if args.once else args.reps
            for _ in range(reps):
                one_pass(mode)
            e1.record(stream)
            e1.synchronize()
            ms = e0.elapsed_ms(e1) / reps
            if mode == "slab_persist":
                stream.l2_window()  # window off, persisting lines back to normal
            print(json.dumps({"mode": mode, "ms_per_pass": round(ms, 4), **info,
                              "algorithmic_bytes_whole": 2 * 3 * 8 * ni * nj * nk}), flush=True)
    gb.synchronize()


if __name__ == "__main__":
    main()
