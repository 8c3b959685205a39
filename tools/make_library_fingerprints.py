#!/usr/bin/env python
"""Write gt4py_b200/library_kernels.json: the structural fingerprints (gt4py_b200.lowering.fingerprint, taken
on the unoptimised OIR) of the reference stencils the library has hand-written kernels for.  Build container
only (imports /root/reference).  The table holds digests and argument roles, no reference code.

    python tools/make_library_fingerprints.py
"""
import json
import os
import sys
import tempfile
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle", "ref_shims"), "/root/reference/src", "/root/reference/tests"]
os.environ.setdefault("GT_CACHE_ROOT", tempfile.mkdtemp(prefix="gtcache_"))
warnings.filterwarnings("ignore")

# reference definition -> (library kernel, extra, roles of the API arguments in signature order, min nk)
LIBRARY = {
    "horizontal_diffusion": ("hdiff", {"limiter": 1}, ["in_field", "out_field", "coeff"], 1),
    "simple_horizontal_diffusion": ("hdiff", {"limiter": 0}, ["in_field", "coeff", "out_field"], 1),
    "vertical_advection_dycore": ("vadv", {}, ["utens_stage", "u_stage", "wcon", "u_pos", "utens", "dtr_stage"], 3),
    "tridiagonal_solver": ("tridiag", {}, ["inf", "diag", "sup", "rhs", "out"], 2),
}


FLOAT32_KERNELS = {"hdiff"}  # library kernels that also exist for all-float32 fields (csrc/hdiff.cu)


def main():
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from gt4py.cartesian.backend import from_name
    from gt4py.cartesian.definitions import BuildOptions
    from gt4py.cartesian.gtc.gtir_to_oir import GTIRToOIR
    from gt4py.cartesian.stencil_builder import StencilBuilder

    from gt4py_b200 import lowering

    import inspect

    import numpy as np
    from gt4py.cartesian import gtscript
    from gt4py.cartesian.testing.utils import annotate_function

    table = {}
    for name, (kernel, extra, roles, min_k) in LIBRARY.items():
        variants = [("float64", sd.REGISTRY[name], {})]
        if kernel in FLOAT32_KERNELS:
            # the pure-float32 build: float32 fields and parameters, `literal_float_precision=32`
            spec = inspect.getfullargspec(sd.REGISTRY[name])
            dtypes = {a: gtscript.Field[np.float32] for a in spec.args}
            dtypes.update({a: np.float32 for a in spec.kwonlyargs})
            variants.append(("float32", annotate_function(sd.REGISTRY[name], dtypes), {"literal_float_precision": 32}))
        for dtype, fn, opts in variants:
            builder = StencilBuilder(fn, backend=from_name("numpy"),
                                     options=BuildOptions(name=name, module="m", format_source=False, rebuild=True,
                                                          **opts)
                                     ).with_externals(sd.EXTERNALS_REGISTRY[name])
            plan, _ = lowering.lower(GTIRToOIR().visit(builder.gtir), name, fuse=False)
            fp = plan["fingerprint"]
            assert fp["api_order"] == roles, (name, fp["api_order"])
            assert fp["digest"] not in table, (name, dtype)
            table[fp["digest"]] = {"kernel": kernel, **extra, "roles": roles, "min_k": min_k, "reference": name,
                                   "dtype": dtype}
            print(name, dtype, fp["digest"])
    out = os.path.join(ROOT, "gt4py_b200", "library_kernels.json")
    with open(out, "w") as fh:
        json.dump({"comment": "made by tools/make_library_fingerprints.py", "kernels": table}, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
