"""The real seam: ``@gtscript.stencil(backend="b200")`` built AND run in one process, next to the reference's own
``numpy`` backend on the same inputs -- SURVEY.md 8(b), VERDICT r1 "the drop-in path has never executed on a GPU".

The reference is the staged copy oracle/_ref (made by oracle/make_ref.py; it travels to the GPU box) or
/root/reference in the build container.  ``-m gpu``: on the B200 through libgt4py_b200.so.  The same test bodies also
run in the CPU suite (``*_on_the_host_double``) against the host double of the C ABI (tests/emu_abi.py), in a
subprocess, so that the Python half of the seam is covered where there is no GPU.

Bar: bit-exact (np.array_equal) against the reference numpy backend; transcendentals 1e-12.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import inputs as gen  # noqa: E402
from oracle import ref  # noqa: E402

EMU = os.environ.get("GT4PY_B200_EMU") == "1"


def _gt4py():
    """Import the reference + register the backend; returns (gtscript, gt_storage, stencil_definitions)."""
    if not ref.available():
        pytest.fail("the reference is not staged: run `python oracle/make_ref.py` in the build container "
                    "(__graft_entry__.build() does it)")
    if EMU:
        from tests import emu_abi

        emu_abi.activate()
    ref.activate(gpu=True)
    import gt4py_b200.backend  # noqa: F401
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from gt4py import storage as gt_storage
    from gt4py.cartesian import gtscript

    return gtscript, gt_storage, sd


def _both(gtscript, definition, externals=None, **opts):
    kw = dict(externals=externals or {}, rebuild=True)
    return (gtscript.stencil("b200", definition, **kw, **opts),
            gtscript.stencil("numpy", definition, format_source=False, **kw))


def _run_pair(gt_storage, st_dev, st_ref, fields, params, origin, domain, written, aligned=None):
    dev = {k: gt_storage.from_array(v, backend="b200", aligned_index=(aligned or {}).get(k, (0, 0, 0)))
           for k, v in fields.items()}
    host = {k: gt_storage.from_array(v, backend="numpy", aligned_index=(aligned or {}).get(k, (0, 0, 0)))
            for k, v in fields.items()}
    st_dev(**dev, **params, origin=origin, domain=domain)
    st_ref(**host, **params, origin=origin, domain=domain)
    for name in written:
        got, want = np.asarray(dev[name]), np.asarray(host[name])
        assert np.array_equal(got, want, equal_nan=True), (name, float(np.abs(got - want).max()))
    return dev


# sizes: BASELINE configs[0] (hdiff 128x128x80) and configs[1] (vadv 256x256x80) on the device; small under emulation
HD = (24, 20, 5) if EMU else (128, 128, 80)
VA = (12, 9, 11) if EMU else (256, 256, 80)
CH = (20, 18, 3) if EMU else (200, 130, 9)


def body_hdiff_config0():
    gtscript, gt_storage, sd = _gt4py()
    st_dev, st_ref = _both(gtscript, sd.REGISTRY["horizontal_diffusion"])
    assert type(st_dev).__mro__[1].__module__.startswith("gt4py.cartesian")  # the reference's own StencilObject
    f = gen.hdiff_inputs(*HD, seed=42)
    org = {k: (2, 2, 0) for k in f}
    _run_pair(gt_storage, st_dev, st_ref, f, {}, (2, 2, 0), HD, ["out_field"], aligned=org)
    impl = type(st_dev)._b200()
    if not EMU:
        assert impl._plan_["library"]["kernel"] == "hdiff"  # fingerprint-matched: the hand-written kernel ran


def body_vadv_config1():
    gtscript, gt_storage, sd = _gt4py()
    st_dev, st_ref = _both(gtscript, sd.REGISTRY["vertical_advection_dycore"], {"BET_M": 0.5, "BET_P": 0.5})
    f = gen.vadv_inputs(*VA, seed=7)
    dtr = f.pop("dtr_stage")
    _run_pair(gt_storage, st_dev, st_ref, f, {"dtr_stage": dtr}, (0, 0, 0), VA, ["utens_stage"])


def body_vadv_other_externals():
    """BET_M = 0.4: not the library's kernel -> the generated FORWARD / BACKWARD kernels."""
    gtscript, gt_storage, sd = _gt4py()
    import gt4py_b200.backend as b200_backend

    del b200_backend._WARNED[:]
    with pytest.warns(UserWarning, match="does not match a hand-written library kernel"):
        st_dev, st_ref = _both(gtscript, sd.REGISTRY["vertical_advection_dycore"], {"BET_M": 0.4, "BET_P": 0.6})
    assert "library" not in type(st_dev)._b200()._plan_
    dom = (12, 9, 11) if EMU else (96, 70, 40)
    f = gen.vadv_inputs(*dom, seed=8)
    dtr = f.pop("dtr_stage")
    _run_pair(gt_storage, st_dev, st_ref, f, {"dtr_stage": dtr}, (0, 0, 0), dom, ["utens_stage"])


def body_tridiagonal():
    gtscript, gt_storage, sd = _gt4py()
    st_dev, st_ref = _both(gtscript, sd.REGISTRY["tridiagonal_solver"])
    dom = (10, 7, 9) if EMU else (130, 67, 80)
    f = gen.tridiag_inputs(*dom, seed=3)
    _run_pair(gt_storage, st_dev, st_ref, f, {}, (0, 0, 0), dom, ["out", "sup", "rhs"])


def body_chain():
    gtscript, gt_storage, _ = _gt4py()
    from tests import chain_definitions as cd

    st_dev, st_ref = _both(gtscript, cd.fv3_chain)
    h = 3
    rng = np.random.default_rng(4)
    f = {k: rng.uniform(-1.0, 1.5, (CH[0] + 2 * h, CH[1] + 2 * h, CH[2])) for k in ("phi", "u", "v")}
    f["out"] = np.zeros(f["phi"].shape)
    _run_pair(gt_storage, st_dev, st_ref, f, {"dt": 0.05, "alpha": 0.03}, (h, h, 0), CH, ["out"],
              aligned={k: (h, h, 0) for k in f})
    kinds = [k["kind"] for k in type(st_dev)._b200()._plan_["kernels"]]
    assert kinds == ["march"]  # one fused kernel, temporaries in registers


def body_sequential_exchange():
    """FORWARD / BACKWARD loops whose stages exchange values horizontally inside a level (chain_definitions.py): by
    default loop fission turns them into column kernels (the exchanged temporary does not depend on the carried values);
    with `level_fission=False` they run as `level` kernels launched level by level by the C plan launcher.  Both
    bit-exact with the reference's numpy backend."""
    gtscript, gt_storage, _ = _gt4py()
    from tests import chain_definitions as cd

    dom = (14, 9, 7) if EMU else (150, 67, 33)
    for n, definition in enumerate((cd.seq_exchange, cd.seq_exchange_backward)):
        for fission in (True, False):
            st_dev, st_ref = _both(gtscript, definition, level_fission=fission)
            rng = np.random.default_rng(20 + n)
            f = {k: rng.uniform(-1.0, 1.5, (dom[0] + 2, dom[1] + 2, dom[2])) for k in ("a", "b")}
            f["out"] = np.zeros(f["a"].shape)
            _run_pair(gt_storage, st_dev, st_ref, f, {}, (1, 1, 0), dom, ["out"], aligned={k: (1, 1, 0) for k in f})
            impl = type(st_dev)._b200()
            kinds = [k["kind"] for k in impl._plan_["kernels"]]
            assert kinds == (["column"] if fission else ["level"]) * (2 + n)
            assert impl._last_entry_points == [k["name"] + "_u" for k in impl._plan_["kernels"]]


def body_storage_drop_in():
    """gt4py.storage.*(backend="b200") -> device arrays with the registered layout / alignment (interface.py:40-327)."""
    gtscript, gt_storage, sd = _gt4py()
    from gt4py_b200 import storage as b200_storage

    z = gt_storage.zeros((9, 7, 5), np.float64, backend="b200", aligned_index=(2, 1, 0))
    assert isinstance(z, b200_storage.DeviceArray) and z.shape == (9, 7, 5)
    assert z.strides[0] == 8 and z.strides[0] < z.strides[1] < z.strides[2]      # I contiguous: layout_map (2, 1, 0)
    assert (z.ptr + 2 * z.strides[0]) % (b200_storage.ALIGNMENT * 8) == 0         # the aligned index sits on 128 B
    assert z.strides[1] % (b200_storage.ALIGNMENT * 8) == 0                       # rows padded to whole blocks
    assert not np.asarray(z).any()
    o = gt_storage.ones((4, 3, 2), np.float32, backend="b200")
    assert np.asarray(o).dtype == np.float32 and (np.asarray(o) == 1).all()
    fl = gt_storage.full((4, 3), 2.5, np.float64, backend="b200", dimensions=["I", "J"])
    assert (np.asarray(fl) == 2.5).all()
    data = np.arange(2 * 3 * 4 * 2, dtype=np.float64).reshape(2, 3, 4, 2)
    fa = gt_storage.from_array(data, backend="b200", dtype=(np.float64, (2,)), aligned_index=(0, 0, 0))
    assert fa.shape == (2, 3, 4, 2) and np.array_equal(np.asarray(fa), data)
    assert fa.strides[3] > max(fa.strides[:3])                                    # data dimension outermost
    e = gt_storage.empty((3, 3, 3), backend="b200")
    assert e.__cuda_array_interface__["version"] == 3 and e.__cuda_array_interface__["data"][0] == e.ptr
    with pytest.raises(ValueError):
        gt_storage.empty((3, 3), backend="b200", aligned_index=(0, 0, 0))
    # a stencil takes them as they are
    st = gtscript.stencil("b200", sd.REGISTRY["copy_stencil"], rebuild=True)
    a = gt_storage.from_array(np.random.default_rng(0).standard_normal((6, 5, 4)), backend="b200")
    b = gt_storage.zeros((6, 5, 4), backend="b200")
    st(a, b)
    assert np.array_equal(np.asarray(a), np.asarray(b))


def body_lazy_stencil_and_exec_info():
    """lazy_stencil (lazy_stencil.py:24-92) builds on first use; exec_info carries the reference's keys."""
    gtscript, gt_storage, sd = _gt4py()
    lazy = gtscript.lazy_stencil(backend="b200", definition=sd.REGISTRY["copy_stencil"], rebuild=True)
    assert lazy.backend.name == "b200"
    a = gt_storage.from_array(np.random.default_rng(1).standard_normal((5, 4, 3)), backend="b200")
    b = gt_storage.zeros((5, 4, 3), backend="b200")
    info = {}
    lazy(a, b, exec_info=info)
    assert np.array_equal(np.asarray(a), np.asarray(b))
    for key in ("call_start_time", "call_end_time", "run_start_time", "run_end_time"):
        assert key in info
    assert info["call_start_time"] <= info["run_start_time"] <= info["run_end_time"] <= info["call_end_time"]
    frozen = lazy.implementation.freeze(origin={"field_a": (0, 0, 0), "field_b": (0, 0, 0)}, domain=(5, 4, 3))
    b2 = gt_storage.zeros((5, 4, 3), backend="b200")
    frozen(field_a=a, field_b=b2)
    assert np.array_equal(np.asarray(a), np.asarray(b2))


def body_default_arguments():
    """Parameters with default values (test_call_interface.py:166-216 form) through the b200 call path."""
    gtscript, gt_storage, _ = _gt4py()
    from gt4py.cartesian.gtscript import PARALLEL, Field, computation, interval

    def scaled(inp: Field[np.float64], out: Field[np.float64], factor: np.float64 = 3.0):
        with computation(PARALLEL), interval(...):
            out = inp * factor

    st = gtscript.stencil("b200", scaled, rebuild=True)
    a = gt_storage.from_array(np.full((4, 4, 2), 2.0), backend="b200")
    b = gt_storage.zeros((4, 4, 2), backend="b200")
    st(a, b)
    assert (np.asarray(b) == 6.0).all()
    st(a, b, factor=0.5)
    assert (np.asarray(b) == 1.0).all()


BODIES = {n[len("body_"):]: f for n, f in list(globals().items()) if n.startswith("body_")}


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(BODIES))
def test_seam_on_the_gpu(name):
    BODIES[name]()


@pytest.mark.parametrize("name", sorted(BODIES))
def test_seam_on_the_host_double(name):
    """Same bodies in a subprocess whose C ABI is the host double (the CUDA library cannot be unloaded)."""
    if not ref.available():
        pytest.skip("reference not available on this machine")
    env = dict(os.environ, GT4PY_B200_EMU="1", PYTHONHASHSEED="0")
    code = f"import sys; sys.path.insert(0, {ROOT!r}); from tests import test_seam as t; t.BODIES[{name!r}]()"
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
