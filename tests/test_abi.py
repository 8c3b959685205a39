"""The C-ABI library loads on a CPU-only box and exports every symbol include/gt4py_b200.h declares."""
import ctypes
import os
import re

import pytest

from gt4py_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "gt4py_b200.h")).read()
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", text)))


def test_library_builds_and_loads():
    path = build.build()
    assert os.path.exists(path)
    handle = ctypes.CDLL(path)
    assert handle.b200_abi_version() == 1


def test_every_declared_symbol_is_exported():
    syms = _header_symbols()
    assert len(syms) >= 30
    handle = ctypes.CDLL(build.build())
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in the header but not exported"


def test_python_prototypes_match_header():
    assert _header_symbols() == _lib.declared_symbols()


def test_no_silent_fallback_without_gpu():
    """Compute entry points must fail loudly (negative status + message) when there is no device."""
    if _lib.device_count() > 0:
        pytest.skip("a GPU is present")
    from gt4py_b200.stencil_object import ArrayView, make_field_desc
    import numpy as np

    v = ArrayView(0x1000, (8, 8, 4), (8, 64, 512), np.dtype("float64"))
    f = make_field_desc(v, (2, 2, 0))
    dom = (ctypes.c_int64 * 3)(4, 4, 4)
    rc = _lib.lib().b200_hdiff(ctypes.byref(f), ctypes.byref(f), ctypes.byref(f), dom, 1, 0, None)
    assert rc < 0
    assert _lib.lib().b200_last_error()


def test_argument_validation_in_c():
    from gt4py_b200.stencil_object import ArrayView, make_field_desc
    import numpy as np

    v = ArrayView(0x1000, (8, 8, 4), (8, 64, 512), np.dtype("float64"))
    f = make_field_desc(v, (1, 2, 0))  # halo of 2 required in I
    dom = (ctypes.c_int64 * 3)(4, 4, 4)
    rc = _lib.lib().b200_hdiff(ctypes.byref(f), ctypes.byref(f), ctypes.byref(f), dom, 1, 0, None)
    assert rc == _lib.EINVAL
    assert b"origin" in _lib.lib().b200_last_error()
    g = make_field_desc(v, (2, 2, 0))
    dom = (ctypes.c_int64 * 3)(5, 4, 4)  # 2 + 5 + 2 > 8
    rc = _lib.lib().b200_hdiff(ctypes.byref(g), ctypes.byref(g), ctypes.byref(g), dom, 1, 0, None)
    assert rc == _lib.EINVAL and b"too small" in _lib.lib().b200_last_error()


# ---- the plan-level C ABI driven from plain C (tests/abi_driver.c): b200_plan_create / launch / destroy ---------------
def _run_c_driver(tmp_path, library, image_path, dom, fixture="fv3_chain"):
    import subprocess

    import numpy as np

    from gt4py_b200 import generated
    from oracle import chain
    from oracle import inputs as gen

    gen_dir = os.path.join(ROOT, "tests", "golden", "generated", fixture)
    import json

    with open(os.path.join(gen_dir, "plan.json")) as fh:
        plan = json.load(fh)
    desc = tmp_path / "plan.desc"
    desc.write_text(generated.plan_descriptor(plan))
    exe = tmp_path / "abi_driver"
    subprocess.run(["gcc", "-std=c11", "-O1", "-o", str(exe), os.path.join(ROOT, "tests", "abi_driver.c"), "-ldl"], check=True)
    out = tmp_path / "out.bin"
    r = subprocess.run([str(exe), library, str(desc), image_path, *[str(d) for d in dom], str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    got = np.fromfile(out, dtype=np.float64).reshape(dom[2], dom[1], dom[0]).transpose(2, 1, 0)
    h = chain.HALO
    host = {n: gen.hashed_window(-h, -h, dom[0] + 2 * h, dom[1] + 2 * h, dom[2], seed, -1.0, 1.5)
            for n, seed in (("phi", 91), ("u", 92), ("v", 93))}
    want = chain.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03)
    assert np.array_equal(got, want)
    return r.stdout


def test_plan_abi_from_c_on_the_host_double(tmp_path):
    """A C program (no Python in the process) creates the plan of fv3_chain from plan.desc, launches it twice and
    destroys it -- against the host double of the C ABI, which compiles the same csrc/plan.inl."""
    from tests import emu_abi, host_emulation

    _, hostlib = host_emulation.build_lib(os.path.join(ROOT, "tests", "golden", "generated", "fv3_chain"))
    image = tmp_path / "image.path"
    image.write_text(hostlib)
    out = _run_c_driver(tmp_path, emu_abi.build(), str(image), (37, 21, 4))
    assert "entries k0_0_u" in out  # the warp-march kernel (shuffles emulated), unit-stride entry point


def test_plan_abi_from_c_on_the_host_double_tile_kernel(tmp_path):
    """... and on the shared-memory tile form of the chain (no TMA in the double: the unit-stride entry point)."""
    from tests import emu_abi, host_emulation

    _, hostlib = host_emulation.build_lib(os.path.join(ROOT, "tests", "golden", "generated", "fv3_chain__smem"))
    image = tmp_path / "image.path"
    image.write_text(hostlib)
    out = _run_c_driver(tmp_path, emu_abi.build(), str(image), (37, 21, 4), fixture="fv3_chain__smem")
    assert "entries k0_0_u" in out


@pytest.mark.gpu
def test_plan_abi_from_c_on_the_gpu(tmp_path):
    from gt4py_b200 import generated

    cubin = generated.compile_cubin(os.path.join(ROOT, "tests", "golden", "generated", "fv3_chain", "kernel.cu"))
    out = _run_c_driver(tmp_path, build.build(), cubin, (200, 130, 9))
    assert "entries k0_0_u" in out  # the warp-march kernel, unit-stride entry point


@pytest.mark.gpu
def test_plan_abi_from_c_on_the_gpu_tile_kernel(tmp_path):
    """The same C program on the shared-memory tile form of the chain (backend option warp_march=False)."""
    from gt4py_b200 import generated

    cubin = generated.compile_cubin(os.path.join(ROOT, "tests", "golden", "generated", "fv3_chain__smem", "kernel.cu"))
    out = _run_c_driver(tmp_path, build.build(), cubin, (200, 130, 9), fixture="fv3_chain__smem")
    assert "entries k0_0_t" in out  # TMA-staged entry point chosen by the plan
