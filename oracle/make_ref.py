#!/usr/bin/env python
"""Stage the UNMODIFIED reference under oracle/_ref/ so that it travels to the GPU box.

TEST INFRASTRUCTURE ONLY.  The reference (romanc/gt4py, GT4Py 1.1.11) is pure Python; /root/reference exists
only in the build container.  This recipe copies, byte for byte and without editing anything,

    /root/reference/src/gt4py                  -> oracle/_ref/src/gt4py
    /root/reference/tests/cartesian_tests      -> oracle/_ref/tests/cartesian_tests
    /root/reference/tests/storage_tests        -> oracle/_ref/tests/storage_tests
    /root/reference/tests/conftest.py          -> oracle/_ref/tests/conftest.py

oracle/_ref/ is git-ignored (no reference source ever enters the history) but not gpurun-ignored, so the
snapshot that goes to the B200 box carries it, exactly like the built .so files.  There it serves three purposes
(all of them the checker / the timed CPU baseline, never the product path):

  * `bench.py --impl reference` and the `cpu_baseline` leg time the reference's own ``numpy`` backend
    (src/gt4py/cartesian/backend/numpy_backend.py:28-84) on the box's host cores;
  * tests/test_seam_gpu.py builds stencils with ``@gtscript.stencil(backend="b200")`` ON the GPU box and compares
    them with ``backend="numpy"`` in the same process;
  * tests/test_reference_suites_gpu.py runs the reference's own test modules for the registered name.

Usage:  python oracle/make_ref.py [--force]      (called by __graft_entry__.build() when /root/reference exists)
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")
ITEMS = [("src/gt4py", "src/gt4py"), ("tests/cartesian_tests", "tests/cartesian_tests"),
         ("tests/storage_tests", "tests/storage_tests"), ("tests/conftest.py", "tests/conftest.py")]
IGNORE = shutil.ignore_patterns("__pycache__", "*.pyc", ".gt_cache*", ".pytest_cache", ".hypothesis")


def tree_digest(path: str) -> str:
    h = hashlib.sha256()
    if os.path.isfile(path):
        with open(path, "rb") as fh:
            h.update(fh.read())
        return h.hexdigest()
    for root, dirs, files in os.walk(path):
        dirs[:] = sorted(d for d in dirs if d != "__pycache__" and not d.startswith(".gt_cache"))
        for f in sorted(files):
            if f.endswith(".pyc"):
                continue
            p = os.path.join(root, f)
            h.update(os.path.relpath(p, path).encode())
            with open(p, "rb") as fh:
                h.update(fh.read())
    return h.hexdigest()


def stage(force: bool = False) -> str:
    if not os.path.isdir(os.path.join(SRC, "src", "gt4py")):
        raise RuntimeError(f"{SRC} is not present: the reference can only be staged in the build container")
    manifest_path = os.path.join(DST, "MANIFEST.json")
    want = {dst: tree_digest(os.path.join(SRC, src)) for src, dst in ITEMS}
    if not force and os.path.exists(manifest_path):
        with open(manifest_path) as fh:
            have = json.load(fh)
        if have.get("items") == want and all(os.path.exists(os.path.join(DST, d)) for d in want):
            return DST
    for src, dst in ITEMS:
        s, d = os.path.join(SRC, src), os.path.join(DST, dst)
        if os.path.isdir(d):
            shutil.rmtree(d)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if os.path.isdir(s):
            shutil.copytree(s, d, ignore=IGNORE)
        else:
            shutil.copy2(s, d)
    with open(manifest_path, "w") as fh:
        json.dump({"source": SRC, "version": "gt4py 1.1.11 (romanc/gt4py)", "items": want,
                   "note": "verbatim copy made by oracle/make_ref.py; git-ignored"}, fh, indent=1)
    return DST


if __name__ == "__main__":
    print(stage(force="--force" in sys.argv))
