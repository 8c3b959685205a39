"""Make the reference importable (TEST INFRASTRUCTURE: tests/, bench.py's CPU legs, tools/ only).

    from oracle import ref
    if ref.activate():            # False when neither oracle/_ref nor /root/reference exists
        from gt4py.cartesian import gtscript

Search order: the staged copy oracle/_ref/ (made by oracle/make_ref.py; it travels to the GPU box), then
/root/reference (build container).  The import shims of oracle/ref_shims (stand-ins for the reference's absent
third-party dependencies: boltons, toolz, ... -- none contains stencil arithmetic) go in
front.  With ``gpu=True`` and a CUDA device present, oracle/ref_shims_gpu is added as well: a stand-in for the
``cupy`` module (absent from this image) built on gt4py_b200's DeviceArray, which the reference's own tests use
as the array library of every GPU backend (tests/cartesian_tests/definitions.py:57-65).
"""
from __future__ import annotations

import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
STAGED = os.path.join(HERE, "_ref")


def location():
    """(src dir, tests dir) of the reference, or None."""
    for base in (STAGED, "/root/reference"):
        if os.path.isdir(os.path.join(base, "src", "gt4py", "cartesian")):
            return os.path.join(base, "src"), os.path.join(base, "tests")
    return None


def available() -> bool:
    return location() is not None


def activate(gpu: bool = False, cache_root: str | None = None) -> bool:
    loc = location()
    if loc is None:
        return False
    paths = [os.path.join(HERE, "ref_shims"), loc[0], loc[1]]
    if gpu:
        import gt4py_b200

        if gt4py_b200.device_count() > 0:
            paths.insert(0, os.path.join(HERE, "ref_shims_gpu"))
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    for p in reversed(paths):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    if cache_root is not None:
        os.environ["GT_CACHE_ROOT"] = cache_root
    elif "GT_CACHE_ROOT" not in os.environ:
        os.environ["GT_CACHE_ROOT"] = tempfile.mkdtemp(prefix="gtcache_")
    return True


def env(gpu: bool = False) -> dict:
    """Environment for a subprocess that should see the reference (PYTHONPATH in the same order)."""
    loc = location()
    if loc is None:
        raise RuntimeError("the reference is not available (no oracle/_ref, no /root/reference)")
    paths = [os.path.join(HERE, "ref_shims"), loc[0], loc[1], ROOT]
    if gpu:
        paths.insert(0, os.path.join(HERE, "ref_shims_gpu"))
    e = dict(os.environ)
    e["PYTHONPATH"] = os.pathsep.join(paths + ([e["PYTHONPATH"]] if e.get("PYTHONPATH") else []))
    return e
