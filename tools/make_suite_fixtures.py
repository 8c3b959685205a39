#!/usr/bin/env python
"""Fixtures from the reference's own stencil test suites (build container only; RUNS /root/reference).

/root/reference/tests/cartesian_tests/integration_tests/multi_feature_tests/test_suites.py defines 40
`StencilTestSuite` classes (data dimensions, IJ / K fields, variable K offsets, `while`, typed temporaries,
horizontal regions, ...).  For each suite this script takes the annotated definition of its first float64 test
case (suites.py:163-195), draws the externals the suite declares, builds the stencil with the reference's numpy
backend and with ``b200``, runs numpy on seeded inputs drawn inside the ranges the suite's `symbols` give, and
writes tests/golden/generated/suite_<name>/ in the same format as tools/make_generated_fixtures.py.

    python tools/make_suite_fixtures.py [SuiteName ...]
"""
from __future__ import annotations

import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_generated_fixtures as base  # noqa: E402  (sets sys.path / cache root for the reference)

import numpy as np  # noqa: E402

OUT = base.OUT


def closure_vars(fn):
    return dict(zip(fn.__code__.co_freevars, (c.cell_contents for c in (fn.__closure__ or ()))))


def symbol_range(symbol):
    rng = closure_vars(symbol.value_st_factory).get("in_range") if getattr(symbol, "value_st_factory", None) else None
    return rng


def draw(rng, dtype, lohi, shape):
    lo, hi = lohi if lohi is not None else (0.5, 1.5)
    dtype = np.dtype(dtype)
    if dtype.kind == "f":
        return rng.uniform(lo, hi, shape).astype(dtype)
    if dtype.kind == "b":
        return rng.integers(0, 2, shape).astype(bool)
    return rng.integers(int(np.ceil(lo)), int(np.floor(hi)) + 1, shape).astype(dtype)


def pick_test(cls):
    tests = [t for t in cls.tests if t["backend"] == "numpy"] or cls.tests
    for t in tests:
        if all(np.dtype(d) != np.float32 for d in t["dtypes"].values()):
            return t
    return tests[0]


def main():
    from cartesian_tests.integration_tests.multi_feature_tests import test_suites as ts
    from gt4py.cartesian import gtscript
    from gt4py.cartesian.testing import StencilTestSuite
    from gt4py.cartesian.testing.input_strategies import SymbolKind

    import gt4py_b200.backend  # noqa: F401

    suites = {n: c for n, c in vars(ts).items()
              if isinstance(c, type) and issubclass(c, StencilTestSuite) and getattr(c, "tests", None)}
    names = sys.argv[1:] or sorted(suites)
    report_path = os.path.join(OUT, "REPORT_suites.json")
    report = json.load(open(report_path)) if os.path.exists(report_path) and sys.argv[1:] else {}
    for name in names:
        cls = suites[name]
        test = pick_test(cls)
        seed = sum(map(ord, name))
        rng = np.random.default_rng(seed)
        externals = dict(cls.singletons)
        for gname, const in test["constants"].items():
            externals[gname] = const
        for gname, symbol in cls.symbols.items():  # globals given as a range: draw once
            if symbol.kind == SymbolKind.GLOBAL_STRATEGY and gname not in externals:
                externals[gname] = draw(rng, test["dtypes"][gname], symbol_range(symbol), ())[()]
        tag = "suite_" + name[4:]
        try:
            ref = gtscript.stencil("numpy", test["definition"], externals=externals, rebuild=True, name=tag)
            b2 = gtscript.stencil("b200", test["definition"], externals=externals, rebuild=True, name=tag)
        except Exception as exc:
            report[tag] = f"{type(exc).__name__}: {str(exc)[:300]}"
            print("SKIP", tag, report[tag])
            continue
        lo_dom = [max(r[0], 1) for r in cls.domain_range]
        hi_dom = [r[1] for r in cls.domain_range]
        domain = [min(max(lo, want), hi) for lo, hi, want in zip(lo_dom, hi_dom, (13, 11, 10))]
        domain = tuple(domain + [1] * (3 - len(domain)))
        domain = (domain[0], domain[1], max(domain[2], ref.domain_info.min_sequential_axis_size))
        fields, params, origins = {}, {}, {}
        for fname, fi in ref.field_info.items():
            if fi is None or int(fi.access) == 0:
                fields[fname] = None
                continue
            symbol = cls.symbols[fname]
            mask = [a in fi.axes for a in "IJK"]
            bnd = [b for b, m in zip(symbol.boundary or fi.boundary, mask) if m] if symbol.boundary and len(symbol.boundary) == 3 \
                else list(symbol.boundary or [b for b, m in zip(fi.boundary, mask) if m])
            bnd = [(max(int(a), 0), max(int(b), 0)) for a, b in bnd]
            need = [(max(b[0], 0), max(b[1], 0)) for b, m in zip(fi.boundary, mask) if m]
            bnd = [(max(a[0], n[0]), max(a[1], n[1])) for a, n in zip(bnd, need)]
            dom = [d for d, m in zip(domain, mask) if m]
            shape = [lo + d + hi for (lo, hi), d in zip(bnd, dom)] + list(fi.data_dims)
            fields[fname] = draw(rng, fi.dtype, symbol_range(symbol), shape)
            origins[fname] = tuple(b[0] for b in bnd) + (0,) * len(fi.data_dims)
        for pname, pi in ref.parameter_info.items():
            if pi is None or int(pi.access) == 0:
                params[pname] = None if pi is None else pi.dtype.type(0)
                continue
            symbol = cls.symbols[pname]
            if symbol.kind == SymbolKind.PARAMETER and symbol_range(symbol) is None:
                vals = closure_vars(symbol.value_st_factory).get("one_of") or (1,)
                params[pname] = pi.dtype.type(vals[0])
            else:
                params[pname] = draw(rng, pi.dtype, symbol_range(symbol), ())[()]
        work = {k: (None if v is None else v.copy()) for k, v in fields.items()}
        try:
            with np.errstate(all="ignore"):
                ref(**work, **params, origin=origins, domain=domain)
        except Exception as exc:
            report[tag] = f"reference run failed: {type(exc).__name__}: {str(exc)[:300]}"
            print("SKIP", tag, report[tag])
            continue
        dst = os.path.join(OUT, tag)
        shutil.rmtree(dst, ignore_errors=True)
        os.makedirs(dst)
        plan_dir = type(b2)._b200()._dir_
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
                       "none_fields": [k for k, v in fields.items() if v is None], "written": written,
                       "externals": {k: (v.item() if hasattr(v, "item") else v) for k, v in externals.items()},
                       "suite": name}, fh, indent=1, default=str)
        with open(os.path.join(dst, "plan.json")) as fh:
            plan = json.load(fh)
        report[tag] = "ok: " + ", ".join(f"{k['name']}:{k['kind']}" for k in plan["kernels"])
        print("ok  ", tag, report[tag])
    with open(report_path, "w") as fh:
        json.dump(report, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
