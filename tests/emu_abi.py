"""TEST INFRASTRUCTURE: run the Python host side of the backend against a host-memory double of the C ABI
(tests/emu_abi.cpp) -- for machines without a GPU.

    from tests import emu_abi
    emu_abi.activate()          # before the first gt4py_b200 call of the process

After that ``gt4py_b200._lib`` loads the double instead of libgt4py_b200.so, "compiling" a generated kernel.cu means
building its host form with tests/host_emulation.py, and stencils recognised as library kernels run their lowered
kernels instead (the hand-written CUDA kernels have no double).  The product never does any of this by itself: the
library path is only ever overridden by this explicit call (or the GT4PY_B200_LIB variable a test sets).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def build() -> str:
    src = os.path.join(HERE, "emu_abi.cpp")
    deps = [src, os.path.join(ROOT, "include", "gt4py_b200.h"), os.path.join(ROOT, "gt4py_b200", "csrc", "synthetic.cuh"),
            os.path.join(ROOT, "gt4py_b200", "csrc", "plan.inl")]
    tag = hashlib.sha256(b"".join(open(d, "rb").read() for d in deps)).hexdigest()[:16]
    cache = os.path.join(tempfile.gettempdir(), "gt4py_b200_host_emulation")
    os.makedirs(cache, exist_ok=True)
    lib = os.path.join(cache, f"libgt4py_b200_emu_{tag}.so")
    if not os.path.exists(lib):
        tmp = f"{lib}.{os.getpid()}.tmp"
        subprocess.run(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-w", src, "-o", tmp, "-ldl"], check=True)
        os.replace(tmp, lib)
    return lib


_active = [False]


def activate() -> None:
    if _active[0]:
        return
    from gt4py_b200 import _lib, generated

    from tests import host_emulation

    if _lib._lib is not None and _lib.LIB_PATH != build():
        raise RuntimeError("emu_abi.activate(): the CUDA library is already loaded in this process")
    _lib.LIB_PATH = build()
    os.environ["GT4PY_B200_LIB"] = _lib.LIB_PATH  # subprocesses inherit the choice

    def compile_host(src_path: str) -> str:
        plan, lib = host_emulation.build_lib(os.path.dirname(src_path))
        out = os.path.splitext(src_path)[0] + ".emu.path"
        with open(out, "w") as fh:
            fh.write(lib)
        return out

    real_load_plan = generated.load_plan

    def load_plan(directory, use_library=True):
        return real_load_plan(directory, use_library=False)

    generated.compile_cubin = compile_host
    generated.load_plan = load_plan
    _active[0] = True


def active() -> bool:
    return _active[0]


if os.environ.get("GT4PY_B200_EMU") == "1":  # a subprocess started by a test that wants the double
    activate()
