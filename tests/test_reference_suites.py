"""The reference's OWN acceptance tests, unmodified, run for the backend name ``b200``.

``tests/cartesian_tests/definitions.py:34-55`` derives ALL_BACKENDS / GPU_BACKENDS from the registry at import time, so
a backend registered before collection is swept by every parametrised test.  oracle/ref_plugin.py registers the name
(and adds the value check the reference's harness skips for non-numpy arrays); the modules run from the staged copy
oracle/_ref (made by oracle/make_ref.py, travels to the GPU box), selected with ``-k b200``:

  unit_tests/backend_tests/test_backend.py            args data vs GTIR, device_sync option, build_info, bad name
  feature_tests/test_exec_info.py                     exec_info keys / ordering / aggregation
  feature_tests/test_field_layouts.py                 C / F ordered arrays, layout warning, data-dimension strides
  feature_tests/test_call_interface.py                data dimensions, origin handling (+ CPU_BACKENDS tests, pass 2)
  storage_tests/unit_tests/test_interface.py          gt4py.storage.* for the name
  multi_feature_tests/test_code_generation.py         70 feature stencils built and run
  multi_feature_tests/test_suites.py                  40 StencilTestSuite classes through the Hypothesis harness
  feature_tests/test_stencil_object.py                stencil object attributes / repeated calls for the name
  multi_feature_tests/test_math_functions.py          every unary / binary native function against NumPy
  unit_tests/test_lazy_stencil.py                     LazyStencil: build on first use, field_info / call through it

``-m gpu``: on the B200 through libgt4py_b200.so (cupy is absent from the image: oracle/ref_shims_gpu holds a stand-in
built on DeviceArray that the TESTS use as their array library).  CPU suite: the same modules against the host
double of the C ABI (tests/emu_abi.py), which covers the Python half of the seam without a GPU.
"""
import os
import subprocess
import sys
import xml.etree.ElementTree as ET

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref  # noqa: E402

CT = "cartesian_tests"
MODULES = [
    f"{CT}/unit_tests/backend_tests/test_backend.py",
    f"{CT}/integration_tests/feature_tests/test_exec_info.py",
    f"{CT}/integration_tests/feature_tests/test_field_layouts.py",
    f"{CT}/integration_tests/feature_tests/test_call_interface.py",
    "storage_tests/unit_tests/test_interface.py",
    f"{CT}/integration_tests/multi_feature_tests/test_code_generation.py",
    f"{CT}/integration_tests/multi_feature_tests/test_suites.py",
    f"{CT}/integration_tests/feature_tests/test_stencil_object.py",
    f"{CT}/integration_tests/multi_feature_tests/test_math_functions.py",
    f"{CT}/unit_tests/test_lazy_stencil.py",
]
#: tests of pass 2: parametrised over CPU_BACKENDS in the reference, backend-agnostic in what they assert
AS_CPU_MODULES = [f"{CT}/integration_tests/feature_tests/test_call_interface.py"]

#: test id -> reason.  Every failure outside this table fails the run.  (INTEGRATION.md lists the same.)
EXPECTED_FAILURES = {
    # pass 2 only (the reference itself restricts this test to CPU backends): it doubles `arg3` through
    # `tmp = np.asarray(arg3); tmp *= 2`, i.e. relies on np.asarray ALIASING the storage, which no device array can
    "test_call_interface::test_default_arguments[b200]": "host aliasing of np.asarray(storage) (CPU-only idiom)",
}

#: what must at least have run (guards against a silently empty selection)
MINIMUM = {"pass1": 197, "pass2": 3}


def run_reference_tests(modules, tag, emu, as_cpu=False, workers=6, extra=()):
    loc = ref.location()
    if loc is None:
        pytest.fail("the reference is not staged (oracle/make_ref.py)") if not emu else pytest.skip("no reference here")
    tests_dir = loc[1]
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    xml = os.path.join(out_dir, f"reference_suites_{tag}.xml")
    env = dict(os.environ, PYTHONHASHSEED="0", PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    env.pop("GT_CACHE_PYTEST_DIR", None)
    if emu:
        env["GT4PY_B200_EMU"] = "1"
    else:
        env.pop("GT4PY_B200_EMU", None)
        env.pop("GT4PY_B200_LIB", None)
    if as_cpu:
        env["B200_REF_AS_CPU_BACKEND"] = "1"
    cmd = [sys.executable, "-m", "pytest", "-p", "oracle.ref_plugin", "--rootdir", tests_dir, "-c", os.devnull,
           *[os.path.join(tests_dir, m) for m in modules], "-k", "b200", "-q", "-p", "no:cacheprovider",
           "-n", str(workers), "--dist", "loadgroup", f"--junitxml={xml}", "-W", "ignore", *extra]
    r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True)
    with open(os.path.join(out_dir, f"reference_suites_{tag}.log"), "w") as fh:
        fh.write(r.stdout[-200000:] + "\n--- stderr ---\n" + r.stderr[-20000:])
    assert os.path.exists(xml), r.stdout[-3000:] + r.stderr[-3000:]
    passed, failed, skipped = [], {}, []
    for case in ET.parse(xml).getroot().iter("testcase"):
        tid = f"{case.get('classname')}::{case.get('name')}"
        bad = [c for c in case if c.tag in ("failure", "error")]
        if bad:
            failed[tid] = (bad[0].get("message") or "")[:300]
        elif any(c.tag == "skipped" for c in case):
            skipped.append(tid)
        else:
            passed.append(tid)
    return passed, failed, skipped, r


def check(passed, failed, skipped, r, minimum):
    unexpected = {k: v for k, v in failed.items() if not any(k.endswith(e) or e in k for e in EXPECTED_FAILURES)}
    summary = f"{len(passed)} passed, {len(failed)} failed, {len(skipped)} skipped"
    print("reference tests for 'b200':", summary)
    assert not unexpected, summary + "\n" + "\n".join(f"{k}: {v}" for k, v in list(unexpected.items())[:20]) + \
        "\n" + r.stdout[-2500:]
    assert len(passed) >= minimum, summary + "\n" + r.stdout[-2500:]


@pytest.mark.gpu
def test_reference_modules_for_b200_on_the_gpu():
    check(*run_reference_tests(MODULES, "gpu", emu=False, workers=8), MINIMUM["pass1"])


@pytest.mark.gpu
def test_reference_cpu_backend_tests_for_b200_on_the_gpu():
    check(*run_reference_tests(AS_CPU_MODULES, "gpu_as_cpu", emu=False, as_cpu=True, workers=4), MINIMUM["pass2"])


def test_reference_modules_for_b200_on_the_host_double():
    if not ref.available():
        pytest.skip("reference not available on this machine")
    check(*run_reference_tests(MODULES, "emu", emu=True), MINIMUM["pass1"])


def test_reference_cpu_backend_tests_for_b200_on_the_host_double():
    if not ref.available():
        pytest.skip("reference not available on this machine")
    check(*run_reference_tests(AS_CPU_MODULES, "emu_as_cpu", emu=True, as_cpu=True, workers=4), MINIMUM["pass2"])
