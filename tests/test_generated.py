"""Generated stencils: plans + CUDA sources lowered from the reference's OIR by the ``b200`` backend
(gt4py_b200/backend.py, lowering.py), loaded WITHOUT gt4py by gt4py_b200/generated.py.

Fixtures: tests/golden/generated/<name>/{plan.json, kernel.cu, case.json, golden.npz}, one per stencil of the
reference's registry (multi_feature_tests/stencil_definitions.py), of feature_tests/test_exec_info.py and of
tests/chain_definitions.py (config 4); golden.npz holds the seeded inputs and the outputs of the reference's
numpy backend (tools/make_generated_fixtures.py).

CPU tests: fixture integrity, ctypes <-> C layout of the argument block, nvcc compiles the sources, the
numpy restatement of the chain (oracle/chain.py) is pinned by the golden outputs, and -- where gt4py is
importable (build container) -- registration, determinism and library-kernel matching of the backend.
GPU tests (-m gpu): run every plan on the device through b200_module_load / b200_launch and compare with the
reference outputs: bit-exact, except ``native_functions`` (transcendentals: CUDA libm vs numpy, 1e-12).
"""
import glob
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from gt4py_b200 import generated

HERE = os.path.dirname(os.path.abspath(__file__))
GEN = os.path.join(HERE, "golden", "generated")
NAMES = sorted(os.path.basename(os.path.dirname(p)) for p in glob.glob(os.path.join(GEN, "*", "plan.json")))
LIBRARY_MATCHED = ["horizontal_diffusion", "simple_horizontal_diffusion", "tridiagonal_solver",
                   "vertical_advection_dycore"]
TOLERANT = {"native_functions": 1e-12, "native_functions__tile": 1e-12}


def _plan(name):
    with open(os.path.join(GEN, name, "plan.json")) as fh:
        return json.load(fh)


def _case(name):
    with open(os.path.join(GEN, name, "case.json")) as fh:
        case = json.load(fh)
    return case, np.load(os.path.join(GEN, name, "golden.npz"))


# ---- CPU ----------------------------------------------------------------------------------------------------
def test_fixture_inventory():
    assert len(NAMES) >= 30
    for must in LIBRARY_MATCHED + ["fv3_chain", "column_physics", "horizontal_regions", "runtime_if",
                                   "exec_info_advection", "exec_info_diffusion", "data_types", "while_stencil"]:
        assert must in NAMES
    with open(os.path.join(GEN, "REPORT.json")) as fh:
        report = json.load(fh)
    assert all(v.startswith("ok") for v in report.values()), report  # nothing of the registry is unsupported
    with open(os.path.join(GEN, "REPORT_suites.json")) as fh:
        suites = json.load(fh)  # the reference's 40 StencilTestSuite definitions (test_suites.py)
    assert len(suites) == 40 and all(v.startswith("ok") for v in suites.values()), suites


@pytest.mark.parametrize("name", NAMES)
def test_plan_is_loadable_without_gt4py(name):
    plan = _plan(name)
    cls = generated.stencil_class(plan, os.path.join(GEN, name))
    assert set(cls._gt_field_info_) | set(cls._gt_parameter_info_) == set(plan["info"]["arg_names"])
    kinds = {k["kind"] for k in plan["kernels"]}
    assert kinds <= {"point", "column", "tile", "colfuse", "level", "march"}
    src = open(os.path.join(GEN, name, "kernel.cu")).read()
    for k in plan["kernels"]:
        assert f" {k['name']}(const Args a)" in src and f" {k['name']}_u(const Args a)" in src
    # float32 re-annotations must NOT match the float64 library kernels; the pure-float32 build of hdiff (float32
    # fields and literals) has a library kernel of its own (csrc/hdiff.cu, T = float), the others do not
    expected = name.split("__")[0] in LIBRARY_MATCHED and not name.endswith(("__f32", "__f32lit", "__kcache"))
    expected = expected or name == "horizontal_diffusion__f32lit"
    assert expected == ("library" in plan)


def test_argument_block_layout_matches_c(tmp_path):
    """ctypes packs `struct Args` exactly as nvcc/gcc lay it out (mixed bool/int8/.../double members)."""
    lines = ["#include <cstdint>", "#include <cstdio>", "#include <cstddef>"]
    calls = []
    for n, name in enumerate(NAMES):
        src = open(os.path.join(GEN, name, "kernel.cu")).read()
        struct = src[src.index("struct Args {"):]
        struct = struct[:struct.index("};") + 2]
        lines.append(f"namespace s{n} {{ {struct} }}")
        plan = _plan(name)
        last = plan["args"][-1][0]
        calls.append(f'printf("%zu %zu %zu\\n", sizeof(s{n}::Args), offsetof(s{n}::Args, {last}), offsetof(s{n}::Args, ni));')
    lines.append("int main() { " + " ".join(calls) + " return 0; }")
    (tmp_path / "layout.cpp").write_text("\n".join(lines))
    subprocess.run(["g++", "-std=c++17", "-Wno-invalid-offsetof", "-o", str(tmp_path / "layout"),
                    str(tmp_path / "layout.cpp")], check=True)
    out = subprocess.run([str(tmp_path / "layout")], check=True, capture_output=True, text=True).stdout.split("\n")
    for name, line in zip(NAMES, out):
        plan = _plan(name)
        A = generated._args_struct(plan)
        size, off_last, off_ni = (int(x) for x in line.split())
        assert (size, off_last, off_ni) == (__import__("ctypes").sizeof(A), getattr(A, plan["args"][-1][0]).offset,
                                            A.ni.offset), name


@pytest.mark.parametrize("name", ["fv3_chain", "column_physics", "data_types", "native_functions"])
def test_generated_source_compiles_for_sm100a(name, tmp_path):
    import shutil

    shutil.copy(os.path.join(GEN, name, "kernel.cu"), tmp_path / "kernel.cu")
    cubin = generated.compile_cubin(str(tmp_path / "kernel.cu"))
    r = subprocess.run(["cuobjdump", "-elf", cubin], capture_output=True, text=True)
    assert "sm_100" in r.stdout or os.path.getsize(cubin) > 1000


def test_chain_oracle_is_pinned_by_the_reference():
    """oracle/chain.py == the reference numpy backend's output, bit for bit."""
    from oracle import chain

    case, g = _case("fv3_chain")
    u = np.pad(g["in__u"], ((0, 0), (1, 1), (0, 0)))
    v = np.pad(g["in__v"], ((1, 1), (0, 0), (0, 0)))
    out = chain.fv3_chain(g["in__phi"], u, v, float(g["param__dt"]), float(g["param__alpha"]))
    assert np.array_equal(out, g["out__out"])
    assert 0.05 < (out == 0).mean() < 0.95  # both branches of the final conditional are exercised
    case, g = _case("column_physics")
    out = chain.column_physics(g["in__t"], g["in__q"], g["in__w"], float(g["param__dt"]))
    assert np.array_equal(out, g["out__t_out"])


# ---- CPU, needs the reference importable (build container only) ---------------------------------------------
def _gt4py():
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("/root/reference is not present on this machine")
    root = os.path.dirname(HERE)
    for p in ("/root/reference/tests", "/root/reference/src", os.path.join(root, "oracle", "ref_shims")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import tempfile

    os.environ.setdefault("GT_CACHE_ROOT", tempfile.mkdtemp(prefix="gtcache_"))
    try:
        import gt4py.cartesian  # noqa: F401
    except Exception as exc:  # pragma: no cover
        pytest.skip(f"gt4py not importable: {exc}")
    import gt4py_b200.backend as b200_backend

    return b200_backend


def test_backend_is_registered_in_the_reference_registry():
    b200_backend = _gt4py()
    from gt4py.cartesian import backend as gt_backend
    from gt4py.storage.cartesian import layout_registry as layout

    assert gt_backend.from_name("b200") is b200_backend.B200Backend
    assert "b200" in layout.REGISTRY
    info = layout.from_name("b200")
    assert info["device"] == "gpu" and info["alignment"] == 16
    assert info["layout_map"](("I", "J", "K")) == (2, 1, 0)  # I contiguous


def test_gtscript_stencil_builds_and_refuses_host_arrays():
    _gt4py()
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from gt4py.cartesian import gtscript
    from gt4py.cartesian.stencil_object import StencilObject

    info = {}
    st = gtscript.stencil("b200", sd.REGISTRY["horizontal_diffusion"], rebuild=True, build_info=info)
    assert isinstance(st, StencilObject) and st.backend == "b200"
    assert st.field_info["in_field"].boundary == ((2, 2), (2, 2), (0, 0))
    impl = type(st)._b200()
    assert impl._plan_["library"]["kernel"] == "hdiff"
    host = np.zeros((12, 12, 4))
    with pytest.raises(TypeError, match="host array"):  # no CPU fallback
        st(host, host.copy(), host.copy(), origin=(2, 2, 0), domain=(8, 8, 4))


@pytest.mark.parametrize("name", ["fv3_chain", "horizontal_region_with_conditional", "vertical_advection_dycore",
                                  "horizontal_diffusion__tile", "exec_info_diffusion__point"])
def test_lowering_is_deterministic_and_matches_the_fixture(name):
    _gt4py()
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
    import make_generated_fixtures as mk
    from gt4py.cartesian import gtscript

    fn, ext, *opts = mk.definitions()[name]
    st = gtscript.stencil("b200", fn, externals=ext, rebuild=True, name=name, **(opts[0] if opts else {}))
    plan_dir = type(st)._b200()._dir_
    # the reference's OIR pipeline orders independent statements by set iteration (PYTHONHASHSEED), so the
    # text is compared as a multiset of lines; the plan (kernels, extents, temporaries, fingerprint of the
    # unoptimised OIR) is compared exactly
    import re

    def lines(path):  # (hoisted loads and merged locals are numbered in statement order)
        return sorted(re.sub(r"\bld\d+\b|_gen_\d+\b", "#", ln) for ln in open(path).read().splitlines())

    assert lines(os.path.join(plan_dir, "kernel.cu")) == lines(os.path.join(GEN, name, "kernel.cu"))
    with open(os.path.join(plan_dir, "plan.json")) as fh:
        got, want = json.load(fh), _plan(name)
    got.pop("fingerprint"), want.pop("fingerprint")
    assert got == want


def test_library_matching_ignores_names_and_respects_externals():
    b200_backend = _gt4py()
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from gt4py.cartesian import gtscript

    from tests import renamed_definitions as rd

    st = gtscript.stencil("b200", rd.smoothing, rebuild=True)  # simple hdiff under other names
    lib = type(st)._b200()._plan_["library"]
    assert lib["kernel"] == "hdiff" and lib["limiter"] == 0
    assert lib["args"] == {"in_field": "p", "coeff": "q", "out_field": "r"}
    st = gtscript.stencil("b200", sd.REGISTRY["vertical_advection_dycore"], externals={"BET_M": 0.4, "BET_P": 0.6},
                          rebuild=True)
    assert "library" not in type(st)._b200()._plan_  # other externals: not the library's kernel
    st = gtscript.stencil("b200", sd.REGISTRY["horizontal_diffusion"], rebuild=True,
                          use_library_kernels=False)
    assert "library" not in type(st)._b200()._plan_
    table = b200_backend.library_table().values()
    assert sorted(e["reference"] for e in table if e.get("dtype", "float64") == "float64") == sorted(LIBRARY_MATCHED)
    assert sorted(e["reference"] for e in table if e.get("dtype") == "float32") == [
        "horizontal_diffusion", "simple_horizontal_diffusion"]  # csrc/hdiff.cu with T = float


# ---- GPU -------------------------------------------------------------------------------------------------------
def _run(name, use_library):
    from gt4py_b200 import storage

    case, g = _case(name)
    plan = _plan(name)
    st = generated.load_plan(os.path.join(GEN, name), use_library=use_library)
    dims = {f["name"]: [a for a, m in zip("IJK", f["dims"]) if m] + [str(n) for n in range(len(f["data_dims"]))]
            for f in plan["fields"]}
    fields = {}
    for f in plan["fields"]:
        n = f["name"]
        if n in case["none_fields"]:
            fields[n] = None
            continue
        fields[n] = storage.from_array(g["in__" + n], aligned_index=case["origins"][n], dimensions=dims[n])
    params = {k[len("param__"):]: g[k][()] for k in g.files if k.startswith("param__")}
    fields.update({n: None for n in case["none_fields"]})  # optional fields the definition does not reference
    st(**fields, **params, origin={k: tuple(v) for k, v in case["origins"].items()}, domain=tuple(case["domain"]))
    for n in case["written"]:
        got, want = fields[n].asnumpy(), g["out__" + n]
        if name in TOLERANT:
            np.testing.assert_allclose(got, want, rtol=TOLERANT[name], atol=1e-300, equal_nan=True, err_msg=n)
        else:
            assert np.array_equal(got, want, equal_nan=got.dtype.kind == "f"), (name, n)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_generated_kernels_match_the_reference_numpy_backend(name):
    _run(name, use_library=False)


@pytest.mark.gpu
@pytest.mark.parametrize("name", LIBRARY_MATCHED)
def test_library_kernel_selected_by_fingerprint_matches_too(name):
    _run(name, use_library=True)


@pytest.mark.gpu
@pytest.mark.parametrize("fixture", ["fv3_chain", "fv3_chain__smem"])  # warp-march form / shared-memory tile form
@pytest.mark.parametrize("domain", [(64, 48, 5), (130, 67, 9), (257, 129, 3)])
def test_fv3_chain_against_the_oracle_at_other_sizes(domain, fixture):
    from gt4py_b200 import storage
    from oracle import chain

    ni, nj, nk = domain
    rng = np.random.default_rng(ni * 7 + nj)
    h = chain.HALO
    phi = rng.uniform(-0.2, 1.5, (ni + 2 * h, nj + 2 * h, nk))
    u = rng.uniform(-1.0, 1.0, phi.shape)
    v = rng.uniform(-1.0, 1.0, phi.shape)
    want = chain.fv3_chain(phi, u, v, 0.05, 0.03)
    st = generated.load_plan(os.path.join(GEN, fixture))
    d = {k: storage.from_array(a, aligned_index=(h, h, 0)) for k, a in (("phi", phi), ("u", u), ("v", v))}
    out = storage.zeros((ni, nj, nk))
    st(d["phi"], d["u"], d["v"], out, dt=0.05, alpha=0.03, origin={"_all_": (h, h, 0), "out": (0, 0, 0)}, domain=domain)
    assert np.array_equal(out.asnumpy(), want)
    # tile form: the TMA-staged entry point (odd halo: base 8 bytes off); warp-march form: unit stride
    assert type(st)._last_entry_points == ["k0_0_t" if fixture.endswith("__smem") else "k0_0_u"]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["tma", "unit", "generic", "march_unit", "march_generic"])
def test_fv3_chain_entry_points_agree(mode, monkeypatch):
    """the three entry points of the tile kernel (<k>_t TMA-staged, <k>_u unit stride, <k> any strides) and the two of
    the warp-march kernel"""
    from gt4py_b200 import storage
    from oracle import chain

    ni, nj, nk = domain = (70, 37, 4)
    rng = np.random.default_rng(3)
    h = chain.HALO
    host = {k: rng.uniform(-1.0, 1.5, (ni + 2 * h, nj + 2 * h, nk)) for k in ("phi", "u", "v")}
    want = chain.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03)
    if mode == "unit":
        monkeypatch.setenv("B200_GENERATED_TMA", "0")
    st = generated.load_plan(os.path.join(GEN, "fv3_chain" if mode.startswith("march") else "fv3_chain__smem"))
    if mode.endswith("generic"):  # K-fastest storage viewed as (I, J, K): stride_I != 1
        d = {k: storage.from_array(np.ascontiguousarray(a.transpose(2, 1, 0))).transpose(2, 1, 0) for k, a in host.items()}
    else:
        d = {k: storage.from_array(a, aligned_index=(h, h, 0)) for k, a in host.items()}
    out = storage.zeros((ni, nj, nk))
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # non-optimal layout warning for the transposed fields
        st(d["phi"], d["u"], d["v"], out, dt=0.05, alpha=0.03, origin={"_all_": (h, h, 0), "out": (0, 0, 0)}, domain=domain)
    assert np.array_equal(out.asnumpy(), want)
    assert type(st)._last_entry_points == [{"tma": "k0_0_t", "unit": "k0_0_u", "generic": "k0_0", "march_unit": "k0_0_u",
                                            "march_generic": "k0_0"}[mode]]


# ---- CPU: the point / column lowerings executed on the host against the reference-made golden outputs --------------
def _host_emulable():
    from tests import host_emulation

    return [n for n in NAMES if host_emulation.emulable(_plan(n))]


@pytest.fixture(scope="module")
def _host_libraries():
    """g++ over all fixture kernels up front, in parallel (each is cached by source hash under the temp dir)."""
    from concurrent.futures import ThreadPoolExecutor

    from tests import host_emulation

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        list(pool.map(lambda n: host_emulation.build(os.path.join(GEN, n)), _host_emulable()))


@pytest.mark.parametrize("name", _host_emulable())
def test_lowered_kernels_on_the_host(name, _host_libraries):
    """The committed kernel.cu compiled as plain C++ (tests/host_emulation.py) and run thread by thread must
    reproduce the outputs of the reference's numpy backend -- the same bar as the GPU test, reachable without
    a GPU.  Checks the lowering (indexing, guards, row marches, two-level column trips, K register cache, the
    shared-memory stages of the tile kernels through their plain entry point), not the device; the TMA-staged
    entry points are GPU-only."""
    from tests import host_emulation

    case, g = _case(name)
    fields = {}
    for f in _plan(name)["fields"]:
        n = f["name"]
        fields[n] = None if n in case["none_fields"] else np.array(g["in__" + n], order="F")
    params = {k[len("param__"):]: g[k][()] for k in g.files if k.startswith("param__")}
    host_emulation.run(os.path.join(GEN, name), fields, params, case["origins"], case["domain"])
    for n in case["written"]:
        got, want = fields[n], g["out__" + n]
        if name in TOLERANT:
            np.testing.assert_allclose(got, want, rtol=TOLERANT[name], atol=1e-300, equal_nan=True, err_msg=n)
        else:
            assert np.array_equal(got, want, equal_nan=got.dtype.kind == "f"), (name, n)


_SIZES = [(5, 3, 7), (66, 19, 5), (130, 35, 4), (64, 16, 6), (1, 1, 5)]
_SIZED = ["fv3_chain", "horizontal_diffusion", "horizontal_diffusion__tile", "vertical_advection_dycore",
          "tridiagonal_solver", "column_physics", "column_physics__kcache", "exec_info_advection",
          "exec_info_diffusion", "horizontal_regions", "horizontal_region_with_conditional__tile", "fv3_chain__f32lit"]
# ... and every other definition of the registry fixtures once, at a ragged size
_CASES = [(n, d) for n in _SIZED for d in _SIZES] + [(n, (67, 9, 7)) for n in NAMES
                                                      if not n.startswith("suite_") and n not in _SIZED]
# fused FORWARD / BACKWARD column kernels (lowering_colfuse.py) with columns deep enough that the value stack leaves
# Tensor Memory (128 values per thread) for shared memory and the spill area, and with more tiles than CTAs
_CASES += [("vertical_advection_dycore", (9, 5, 80)), ("vertical_advection_dycore", (131, 3, 150)),
           ("vertical_advection_dycore", (300, 7, 66)), ("tridiagonal_solver", (7, 4, 100)),
           ("tridiagonal_solver", (260, 3, 70)), ("column_physics", (5, 6, 200)),
           ("vertical_advection_dycore__f32", (40, 3, 90))]


@pytest.mark.parametrize("name,domain,knobs", [
    ("fv3_chain", (130, 37, 2), {"B200_MARCH_STRIP": "8"}),                              # 5 warps x 5 strips
    ("fv3_chain", (130, 21, 3), {"B200_MARCH_STRIP": "5", "B200_MARCH_COLUMNS": "1", "B200_MARCH_WARPS": "4"}),   # one column per lane
    ("fv3_chain", (61, 19, 2), {"B200_MARCH_STRIP": "16", "B200_MARCH_LEAD": "1", "B200_MARCH_WARPS": "1"}),
    ("horizontal_diffusion__tile", (120, 40, 2), {"B200_MARCH_STRIP": "8", "B200_MARCH_COLUMNS": "1"}),
    ("horizontal_diffusion__tile", (121, 40, 2), {"B200_MARCH_STRIP": "8", "B200_MARCH_PAIRS": "0"}),  # a ring per column
    ("exec_info_diffusion", (59, 33, 3), {"B200_MARCH_STRIP": "4", "B200_MARCH_LEAD": "7"}),
    ("simple_horizontal_diffusion__tile", (63, 30, 2), {"B200_MARCH_STRIP": "8"}),
    ("native_functions__tile", (40, 13, 2), {"B200_MARCH_STRIP": "8", "B200_MARCH": "force"}),      # two stages: forced
    ("exec_info_advection__tile", (70, 23, 2), {"B200_MARCH_STRIP": "8", "B200_MARCH": "force", "B200_MARCH_COLUMNS": "2"}),
    ("fv3_chain__f32", (70, 23, 2), {"B200_MARCH_STRIP": "8"}),
])
def test_warp_march_kernels_over_strips_and_knobs(name, domain, knobs, monkeypatch):
    """The warp-march form (lowering_march.py) with short strips (several strips per column block, prologue rows of one
    strip overlapping the previous one), one and two columns per lane (paired ring elements or a ring per column),
    other prefetch distances: bit-exact against the
    reference's numpy backend on the host-emulated kernels."""
    for k, v in knobs.items():
        monkeypatch.setenv(k, v)
    test_lowering_against_the_reference_at_other_sizes(name, domain, rebuild=True, expect_kind="march")


def test_tma_fed_column_kernel_has_a_producer_row_and_compiles(monkeypatch):
    """The `<k>_t` entry of the fused column template (lowering_colfuse.TMA_RINGS, off by default): one more row of
    threads than the consumers' block, whose first lane issues the CTA's whole level stream; the plan says so, the
    host emulation leaves the entry out, and nvcc accepts it for sm_100a.  (Values: tests/test_large_sizes_gpu.py.)"""
    _gt4py()
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
    import make_generated_fixtures as mk
    from gt4py.cartesian import gtscript

    from gt4py_b200 import lowering_colfuse

    monkeypatch.setattr(lowering_colfuse, "TMA_RINGS", True)
    fn, ext = mk.definitions()["vertical_advection_dycore"][:2]
    st = gtscript.stencil("b200", fn, externals=ext, rebuild=True, name="vadv_tma_cpu", use_library_kernels=False)
    impl = type(st)._b200()
    (k,) = impl._plan_["kernels"]
    assert k["kind"] == "colfuse" and k["tma"]["producer_rows"] == 1
    bx, by = k["block"]
    src = open(os.path.join(impl._dir_, "kernel.cu")).read()
    assert f"__launch_bounds__({bx * (by + 1)}, 1) {k['name']}_t(" in src
    assert f"if (tid == {bx * by}) {{" in src and "b200_tma_load_3d(" in src
    assert os.path.getsize(generated.compile_cubin(os.path.join(impl._dir_, "kernel.cu"))) > 0


@pytest.mark.parametrize("name,domain", [("vertical_advection_dycore", (300, 7, 66)), ("tridiagonal_solver", (260, 5, 70)),
                                         ("column_physics", (131, 6, 9)), ("vertical_advection_dycore__f32", (40, 3, 90))])
def test_pipelined_column_kernels(name, domain, monkeypatch):
    """The pipelined schedule of the fused column template (lowering_colfuse.PIPELINE, off by default): loop B of the
    previous tile interleaved with loop A of the next one, stack direction alternating from tile to tile, unguarded
    middle part for tiles inside the domain.  More tiles than CTAs, ragged edges; bit-exact against the reference."""
    _gt4py()
    from gt4py_b200 import lowering_colfuse

    monkeypatch.setattr(lowering_colfuse, "PIPELINE", True)
    test_lowering_against_the_reference_at_other_sizes(name, domain, rebuild=True, expect_kind="colfuse")


@pytest.mark.parametrize("name,domain", _CASES)
def test_lowering_against_the_reference_at_other_sizes(name, domain, rebuild=False, expect_kind=None):
    """Differential test in the build container: the reference's numpy backend and the host-emulated kernels of a
    FRESH lowering on the same seeded inputs, at sizes that cross tile / block / row-march boundaries (the
    committed golden files are all 13 x 11 x 10).  Bit-exact."""
    _gt4py()
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
    import make_generated_fixtures as mk
    from gt4py.cartesian import gtscript

    from tests import host_emulation

    fn, ext, *rest = mk.definitions()[name]
    opts = rest[0] if rest else {}
    common = rest[1] if len(rest) > 1 else {}
    ref = gtscript.stencil("numpy", fn, externals=ext, format_source=False, name=name, **common)
    b2 = gtscript.stencil("b200", fn, externals=ext, format_source=False, name=name, rebuild=rebuild, **common, **opts)
    if expect_kind is not None:
        assert expect_kind in [k["kind"] for k in type(b2)._b200()._plan_["kernels"]]
    domain = (domain[0], domain[1], max(domain[2], ref.domain_info.min_sequential_axis_size))
    fields, params, origins = mk.make_inputs(ref, domain, seed=sum(domain) + len(name))
    want = {k: (None if v is None else v.copy()) for k, v in fields.items()}
    ref(**want, **params, origin=origins, domain=domain)
    got = {k: (None if v is None else np.array(v, order="F")) for k, v in fields.items()}
    # tile kernels: shared-memory window against the front guard page for even sizes, the back one for odd sizes
    host_emulation.run(type(b2)._b200()._dir_, got, params, origins, domain, window_mode=domain[0] & 1)
    for n, fi in ref.field_info.items():
        if fi is not None and int(fi.access) & 2:
            assert np.array_equal(got[n], want[n], equal_nan=True), (name, domain, n)
