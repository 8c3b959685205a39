#!/usr/bin/env python
"""Build variants of the warp-march kernels (lowering_march.py knobs) under tools/_variants/ for A/B timing on the
GPU box (build container only; the directory is git-ignored but ships with gpurun).  The cubins are compiled here too
(content-addressed cache under gt4py_b200/lib/cubin_cache), so the box only times.

    python tools/make_march_variants.py [C:STRIP:LEAD:WARPS[:MIN_BLOCKS] ...]      tags: march_c<C>_s<STRIP>_p<P>_w<W>, smem
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
NAMES = ("fv3_chain", "exec_info_diffusion", "horizontal_diffusion", "exec_info_advection")
CHILD = r"""
import os, sys, shutil, tempfile, warnings
sys.path[:0] = [%(root)r, os.path.join(%(root)r, "tools")]
warnings.filterwarnings("ignore")
from oracle import ref
ref.activate()
import gt4py_b200.backend
from gt4py_b200 import generated
import make_generated_fixtures as mk
from gt4py.cartesian import gtscript
defs = mk.definitions()
for name in %(names)r:
    fn, ext = defs[name][:2]
    st = gtscript.stencil("b200", fn, externals=ext, rebuild=True, name=name, use_library_kernels=False,
                          fuse_temporaries=True, %(opts)s)
    src = type(st)._b200()._dir_
    dst = os.path.join(%(root)r, "tools", "_variants", "%(tag)s", name)
    shutil.rmtree(dst, ignore_errors=True)
    os.makedirs(dst)
    for f in ("plan.json", "kernel.cu"):
        shutil.copy(os.path.join(src, f), os.path.join(dst, f))
    generated.compile_cubin(os.path.join(dst, "kernel.cu"))
    print("made", dst, [k["kind"] for k in type(st)._b200()._plan_["kernels"]])
"""


def main():
    specs = sys.argv[1:] or ["2:64:2:4", "1:64:2:4", "2:128:2:4", "2:64:1:4", "2:64:3:4", "2:64:2:2", "2:64:2:8", "2:32:2:4"]
    jobs = [("smem", {}, "warp_march=False")]
    for spec in specs:
        c, strip, pre, warps, *rest = spec.split(":")
        minb = rest[0] if rest else "0"
        jobs.append((f"march_c{c}_s{strip}_p{pre}_w{warps}" + (f"_b{minb}" if minb != "0" else ""),
                     {"B200_MARCH_COLUMNS": c, "B200_MARCH_STRIP": strip, "B200_MARCH_LEAD": pre,
                      "B200_MARCH_WARPS": warps, "B200_MARCH_MIN_BLOCKS": minb}, ""))
    procs = []
    for tag, env, opts in jobs:
        code = CHILD % {"root": ROOT, "tag": tag, "opts": opts, "names": NAMES}
        procs.append(subprocess.Popen([sys.executable, "-c", code], env=dict(os.environ, **env)))
    for p in procs:
        if p.wait():
            raise SystemExit(f"variant build failed: {p.args[-1][:80]}")


if __name__ == "__main__":
    main()
