#!/usr/bin/env python
"""Build variants of the fused column kernels (prefetch lead) under tools/_variants/ for A/B timing on the GPU box
(build container only; the directory is git-ignored but ships with gpurun).

    python tools/make_colfuse_variants.py [lead ...]
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CHILD = r"""
import os, sys, shutil, tempfile, warnings
sys.path[:0] = [%(root)r, os.path.join(%(root)r, "tools")]
warnings.filterwarnings("ignore")
from oracle import ref
ref.activate()
import gt4py_b200.backend
import make_generated_fixtures as mk
from gt4py.cartesian import gtscript
defs = mk.definitions()
for name in ("vertical_advection_dycore", "tridiagonal_solver", "column_physics"):
    fn, ext = defs[name][:2]
    st = gtscript.stencil("b200", fn, externals=ext, rebuild=True, name=name, use_library_kernels=False, %(opts)s)
    src = type(st)._b200()._dir_
    dst = os.path.join(%(root)r, "tools", "_variants", "%(tag)s", name)
    shutil.rmtree(dst, ignore_errors=True)
    os.makedirs(dst)
    for f in ("plan.json", "kernel.cu"):
        shutil.copy(os.path.join(src, f), os.path.join(dst, f))
    print("made", dst)
"""


def main():
    """tags: twokernel, and b<BX>x<BY>_l<LEAD> for every "BXxBY:LEAD" argument (default: a sweep)."""
    specs = sys.argv[1:] or ["128x2:6", "128x2:4", "192x2:3", "192x2:4", "128x3:4", "256x2:2", "256x2:3", "128x4:3"]
    jobs = [("twokernel", {}, "fuse_columns=False")]
    for spec in specs:
        block, lead = spec.split(":")
        jobs.append((f"b{block}_l{lead}", {"B200_COLFUSE_BLOCK": block, "B200_COLFUSE_LEAD": lead}, ""))
    for tag, env, opts in jobs:
        code = CHILD % {"root": ROOT, "tag": tag, "opts": opts}
        subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), check=True)


if __name__ == "__main__":
    main()
