#!/usr/bin/env python
"""Build the fixtures of the generated-stencil GPU tests by RUNNING THE REFERENCE (build container only).

For every stencil of the reference's own registry
(/root/reference/tests/cartesian_tests/integration_tests/multi_feature_tests/stencil_definitions.py), the two
stencils of feature_tests/test_exec_info.py and the config-4 chain of tests/chain_definitions.py:
  1. build it with the reference's ``numpy`` backend and run it on seeded inputs  -> golden outputs;
  2. build it with the ``b200`` backend (gt4py_b200.backend: OIR -> CUDA lowering)  -> plan.json + kernel.cu.
Both land in tests/golden/generated/<name>/ and are committed; the -m gpu tests load the plan WITHOUT gt4py
(gt4py_b200.generated), run it on the B200 and compare with the golden outputs.

    python tools/make_generated_fixtures.py [name ...]
"""
from __future__ import annotations

import json
import os
import shutil
import sys
import tempfile
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "ref_shims"), "/root/reference/src", "/root/reference/tests"]
os.environ.setdefault("GT_CACHE_ROOT", tempfile.mkdtemp(prefix="gtcache_"))
warnings.filterwarnings("ignore")

import numpy as np  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "generated")
DOMAIN = (13, 11, 10)
SKIP = {"native_functions_unsupported"}


def make_inputs(st, domain, seed):
    """Seeded inputs sized domain + boundary; returns (fields, params, origins)."""
    rng = np.random.default_rng(seed)
    fields, origins, params = {}, {}, {}
    for name, fi in st.field_info.items():
        if fi is None or int(fi.access) == 0:
            fields[name] = None
            continue
        mask = [a in fi.axes for a in "IJK"]
        lower = [max(b[0], 0) for b, m in zip(fi.boundary, mask) if m]
        upper = [max(b[1], 0) for b, m in zip(fi.boundary, mask) if m]
        dom = [d for d, m in zip(domain, mask) if m]
        shape = [lo + d + hi for lo, d, hi in zip(lower, dom, upper)] + list(fi.data_dims)
        if fi.dtype.kind == "f":
            arr = rng.uniform(0.5, 1.5, shape).astype(fi.dtype)
            arr *= rng.choice([-1.0, 1.0], shape).astype(fi.dtype) if "land" not in st.options["name"] else 1
        elif fi.dtype.kind == "b":
            arr = rng.integers(0, 2, shape).astype(bool)
        else:
            arr = rng.integers(-3, 9, shape).astype(fi.dtype)
        fields[name] = arr
        origins[name] = tuple(lower) + (0,) * len(fi.data_dims)
    for name, pi in st.parameter_info.items():
        if pi is None or int(pi.access) == 0:
            params[name] = pi.dtype.type(0) if pi is not None else None
            continue
        params[name] = pi.dtype.type(rng.uniform(0.5, 1.5)) if pi.dtype.kind == "f" else pi.dtype.type(rng.integers(1, 4))
    return fields, params, origins


def definitions():
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from cartesian_tests.integration_tests.feature_tests import test_exec_info as tei

    from tests import chain_definitions as chain

    defs = {name: (sd.REGISTRY[name], sd.EXTERNALS_REGISTRY[name]) for name in sd.REGISTRY.names}
    defs["exec_info_advection"] = (tei.TestExecInfo.advection_def, {})
    defs["exec_info_diffusion"] = (tei.TestExecInfo.diffusion_def, {})
    for name, (fn, ext) in chain.DEFINITIONS.items():
        defs[name] = (fn, ext)
    # the same stencils forced through the other lowering than the backend's own choice, to cover both
    for name in ("horizontal_diffusion", "simple_horizontal_diffusion", "exec_info_advection", "native_functions",
                 "horizontal_region_with_conditional", "form_land_mask", "data_types"):
        defs[name + "__tile"] = (*defs[name][:2], {"fuse_temporaries": True})
    for name in ("fv3_chain", "exec_info_diffusion"):
        defs[name + "__point"] = (*defs[name][:2], {"fuse_temporaries": False})
    # ... and through the shared-memory tile kernel instead of the warp-march form (lowering_march.py)
    for name in ("fv3_chain", "exec_info_diffusion", "horizontal_diffusion", "exec_info_advection"):
        defs[name + "__smem"] = (*defs[name][:2], {"fuse_temporaries": True, "warp_march": False})
    # sequential loops whose stages exchange values horizontally, run level by level (loop fission switched off)
    for name in ("seq_exchange", "seq_exchange_backward"):
        defs[name + "__level"] = (*defs[name][:2], {"level_fission": False})
    # column kernels with the K register cache (rolling buffer) switched on; library matching off so that the
    # lowered kernels are what runs
    for name in ("vertical_advection_dycore", "tridiagonal_solver", "column_physics", "while_stencil"):
        defs[name + "__kcache"] = (*defs[name][:2], {"k_register_cache": True, "use_library_kernels": False})
    # float32 variants of the headline stencils (north_star tolerance for float32: 1e-5; these are bit-exact):
    # the same definitions re-annotated the way the reference's suites do it (testing/utils.py:24-30)
    import inspect

    import numpy as np
    from gt4py.cartesian import gtscript
    from gt4py.cartesian.testing.utils import annotate_function

    for name in ("horizontal_diffusion", "vertical_advection_dycore", "tridiagonal_solver", "fv3_chain"):
        fn, ext = defs[name][:2]
        spec = inspect.getfullargspec(fn)
        dtypes = {a: gtscript.Field[np.float32] for a in spec.args}
        dtypes.update({a: np.float32 for a in spec.kwonlyargs})
        defs[name + "__f32"] = (annotate_function(fn, dtypes), ext)
        # pure float32: float32 fields AND float32 literals (`literal_float_precision=32`, gtscript.py:242-290,
        # ADR docs/development/ADRs/cartesian/frontend/literal-precision.md) -- SURVEY.md 8(f) rank 4
        defs[name + "__f32lit"] = (annotate_function(fn, dtypes), ext, {}, {"literal_float_precision": 32})
    return defs


def main():
    from gt4py.cartesian import gtscript

    import gt4py_b200.backend  # noqa: F401  registers "b200"

    defs = definitions()
    names = sys.argv[1:] or sorted(defs)
    report_path = os.path.join(OUT, "REPORT.json")
    report = json.load(open(report_path)) if sys.argv[1:] and os.path.exists(report_path) else {}
    for name in names:
        fn, ext, *rest = defs[name]
        opts = rest[0] if rest else {}             # b200 backend options
        common = rest[1] if len(rest) > 1 else {}  # frontend options, given to both builds
        try:
            ref = gtscript.stencil("numpy", fn, externals=ext, format_source=False, rebuild=True, name=name, **common)
            b2 = gtscript.stencil("b200", fn, externals=ext, format_source=False, rebuild=True, name=name, **common,
                                  **opts)
        except Exception as exc:  # unsupported feature: recorded, not hidden
            report[name] = f"{type(exc).__name__}: {str(exc)[:200]}"
            print("SKIP", name, report[name])
            continue
        domain = (DOMAIN[0], DOMAIN[1], max(DOMAIN[2], ref.domain_info.min_sequential_axis_size + 3))
        fields, params, origins = make_inputs(ref, domain, seed=sum(map(ord, name)))
        work = {k: (None if v is None else v.copy()) for k, v in fields.items()}
        try:
            ref(**work, **{k: v for k, v in params.items()}, origin=origins, domain=domain)
        except Exception as exc:
            report[name] = f"reference run failed: {type(exc).__name__}: {str(exc)[:200]}"
            print("SKIP", name, report[name])
            continue
        dst = os.path.join(OUT, name)
        shutil.rmtree(dst, ignore_errors=True)
        os.makedirs(dst)
        plan_dir = type(b2)._b200()._dir_ if hasattr(type(b2), "_b200") else None
        for fname in ("plan.json", "kernel.cu"):
            shutil.copy(os.path.join(plan_dir, fname), os.path.join(dst, fname))
        written = [k for k, fi in ref.field_info.items() if fi is not None and int(fi.access) & 2]
        np.savez_compressed(
            os.path.join(dst, "golden.npz"),
            **{"in__" + k: v for k, v in fields.items() if v is not None},
            **{"out__" + k: work[k] for k in written},
            **{"param__" + k: np.asarray(v) for k, v in params.items() if v is not None},
        )
        with open(os.path.join(dst, "case.json"), "w") as fh:
            json.dump({"domain": list(domain), "origins": {k: list(v) for k, v in origins.items()},
                       "none_fields": [k for k, v in fields.items() if v is None], "written": written}, fh, indent=1)
        with open(os.path.join(dst, "plan.json")) as fh:
            plan = json.load(fh)
        report[name] = "ok: " + ", ".join(f"{k['name']}:{k['kind']}" for k in plan["kernels"])
        print("ok  ", name, report[name])
    with open(report_path, "w") as fh:
        json.dump(report, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
