#!/usr/bin/env python
"""Host cost of one stencil call (the `StencilObject.__call__` fast path) on a tiny domain, where
the kernel itself is a few microseconds:  normal call / validate_args=False / FrozenStencil / CUDA-graph replay.

    python tools/call_overhead.py [--domain 128 128 80] [--calls 2000]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import gt4py_b200 as gb  # noqa: E402
from gt4py_b200 import stencils, storage  # noqa: E402
from oracle import inputs as gen  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--domain", type=int, nargs=3, default=[128, 128, 80])
    ap.add_argument("--calls", type=int, default=2000)
    args = ap.parse_args()
    dom = tuple(args.domain)
    hf = gen.hdiff_inputs(*dom, seed=1)
    h = {k: storage.from_array(v, aligned_index=(2, 2, 0)) for k, v in hf.items()}
    hd = stencils.HorizontalDiffusion()
    type(hd).configure(device_sync=False)
    stream = gb.Stream()
    out = {"domain": list(dom), "calls": args.calls, "stencil": "horizontal_diffusion (library kernel)"}

    def timed(fn, n):
        fn()
        stream.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        t_issue = time.perf_counter() - t0
        stream.synchronize()
        return t_issue / n * 1e6, (time.perf_counter() - t0) / n * 1e6

    with gb.use_stream(stream):
        full = lambda: hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)  # noqa: E731
        out["call_us"] = timed(full, args.calls)
        fast = lambda: hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom, validate_args=False)  # noqa: E731
        out["call_novalidate_us"] = timed(fast, args.calls)
        frozen = hd.freeze(origin={k: (2, 2, 0) for k in h}, domain=dom)
        fr = lambda: frozen(in_field=h["in_field"], out_field=h["out_field"], coeff=h["coeff"])  # noqa: E731
        out["frozen_us"] = timed(fr, args.calls)
    # a generated stencil (the config-4 chain): argument packing / entry choice / grids in C (b200_plan_launch), or,
    # with B200_PY_LAUNCH=1 in the environment, interpreted in Python as in round 1
    from gt4py_b200 import generated

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    st = generated.load_plan(os.path.join(root, "tests", "golden", "generated", "fv3_chain"))
    type(st).configure(device_sync=False)
    cdom, hh = (64, 64, 8), 3
    rng = np.random.default_rng(0)
    cf = {k: storage.from_array(rng.uniform(-1, 1.5, (cdom[0] + 2 * hh, cdom[1] + 2 * hh, cdom[2])), aligned_index=(hh, hh, 0))
          for k in ("phi", "u", "v")}
    cf["out"] = storage.zeros(cdom)
    org = {"_all_": (hh, hh, 0), "out": (0, 0, 0)}
    out["generated"] = {"stencil": "fv3_chain (generated tile kernel)", "domain": list(cdom),
                        "launch_path": "python" if os.environ.get("B200_PY_LAUNCH") == "1" else "b200_plan_launch"}
    with gb.use_stream(stream):
        full = lambda: st(cf["phi"], cf["u"], cf["v"], cf["out"], dt=0.05, alpha=0.03, origin=org, domain=cdom)  # noqa: E731
        out["generated"]["call_us"] = timed(full, args.calls)
        fast = lambda: st(cf["phi"], cf["u"], cf["v"], cf["out"], dt=0.05, alpha=0.03, origin=org, domain=cdom,  # noqa: E731
                          validate_args=False)
        out["generated"]["call_novalidate_us"] = timed(fast, args.calls)
        fz = st.freeze(origin={"phi": (hh, hh, 0), "u": (hh, hh, 0), "v": (hh, hh, 0), "out": (0, 0, 0)}, domain=cdom)
        fr2 = lambda: fz(phi=cf["phi"], u=cf["u"], v=cf["v"], out=cf["out"], dt=0.05, alpha=0.03)  # noqa: E731
        out["generated"]["frozen_us"] = timed(fr2, args.calls)
    g = gb.Graph(stream)
    per_graph = 20
    with g:
        for _ in range(per_graph):
            hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)
    a, b = timed(g.launch, max(args.calls // per_graph, 10))
    out["graph_replay_us_per_stencil"] = (a / per_graph, b / per_graph)
    e0, e1 = gb.Event(), gb.Event()
    e0.record(stream)
    for _ in range(50):
        g.launch()
    e1.record(stream)
    e1.synchronize()
    out["device_us_per_stencil_in_graph"] = e0.elapsed_ms(e1) * 1e3 / (50 * per_graph)
    out["columns"] = "(host issue time, wall time incl. final sync) per call, microseconds"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
