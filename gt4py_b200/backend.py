"""``b200``: a backend in ``gt4py.cartesian``'s backend registry (needs ``gt4py`` importable).

    import gt4py_b200.backend                      # registers the name
    @gtscript.stencil(backend="b200") ...          # gtscript.py:171 -> StencilBuilder -> B200Backend

The seam is the reference's own: ``backend.register(cls)`` (backend/base.py:142-152, which also registers
``storage_info`` in the storage layout registry), ``Backend.generate() / load()`` called by
``StencilBuilder.build`` (stencil_builder.py:71-81) and the module template
(backend/templates/stencil_module.py.in).  ``generate()`` runs GTIR -> OIR (gtir_to_oir.py:255-266), the OIR
pipeline, then either selects one of the library's hand-written kernels (hdiff / vadv / tridiagonal, matched
by a structural fingerprint of the OIR) or lowers the OIR to CUDA with ``gt4py_b200.lowering``; it writes
``plan.json`` + ``kernel.cu`` next to the generated module.  At run time the generated class delegates
``_call_run`` / ``run`` to ``gt4py_b200.generated`` -- the gt4py-free call path (no cupy, no pybind11),
which performs the same validation with the same exception types and messages
(gt4py_b200/stencil_object.py).
"""
from __future__ import annotations

import hashlib
import json
import time
import warnings
from typing import Any, ClassVar

from gt4py.cartesian import backend as gt_backend
from gt4py.cartesian.backend.base import BaseBackend
from gt4py.cartesian.backend.module_generator import BaseModuleGenerator, ModuleData, make_args_data_from_gtir
from gt4py.cartesian.gtc import gtir
from gt4py.cartesian.gtc.gtir_to_oir import GTIRToOIR
from gt4py.cartesian.gtc.passes.oir_optimizations.horizontal_execution_merging import OnTheFlyMerging
from gt4py.cartesian.gtc.passes.oir_pipeline import DefaultPipeline, OirPipeline
from gt4py.storage.cartesian import layout

from . import lowering, storage

NAME = storage.BACKEND_NAME  # "b200"


def serialize_args_data(args_data: ModuleData, gtir_stencil: gtir.Stencil) -> dict:
    """The reference's field / parameter / domain info (module_generator.py:56-106) as plain JSON."""
    finfo = {}
    for name, fi in args_data.field_info.items():
        finfo[name] = None if fi is None else {
            "access": int(fi.access), "boundary": [[int(a), int(b)] for a, b in fi.boundary],
            "axes": list(fi.axes), "data_dims": [int(d) for d in fi.data_dims], "dtype": str(fi.dtype)}
    pinfo = {name: (None if pi is None else {"access": int(pi.access), "dtype": str(pi.dtype)})
             for name, pi in args_data.parameter_info.items()}
    di = args_data.domain_info
    api = [a.name for a in gtir_stencil.api_signature]
    defaults = {}
    return {"field_info": finfo, "parameter_info": pinfo, "arg_names": api, "arg_defaults": defaults,
            "unreferenced": list(args_data.unreferenced),
            "domain_info": {"parallel_axes": list(di.parallel_axes), "sequential_axis": di.sequential_axis,
                            "min_sequential_axis_size": int(di.min_sequential_axis_size), "ndim": int(di.ndim)}}


_WARNED: list = []  # the fingerprint-miss warning of make_plan is shown once per process


def library_table() -> dict:
    import os

    with open(os.path.join(os.path.dirname(__file__), "library_kernels.json")) as fh:
        return json.load(fh)["kernels"]


class B200ModuleGenerator(BaseModuleGenerator):
    def generate_imports(self) -> str:
        return ("from gt4py.cartesian.stencil_object import StencilObject\n"
                "from gt4py_b200 import generated as _b200_generated")

    def generate_module_members(self) -> str:
        return f"_B200_PLAN_DIR = pathlib.Path(__file__).parent / {self.builder.backend.plan_dir_name!r}"

    def generate_class_members(self) -> str:
        return '''
_b200_impl_ = None

@classmethod
def _b200(cls):
    """The gt4py-free stencil object (gt4py_b200.generated) that owns the kernels."""
    if cls._b200_impl_ is None:
        cls._b200_impl_ = _b200_generated.load_plan(str(_B200_PLAN_DIR))
        type(cls._b200_impl_).configure(device_sync=cls._gt_options_.get("backend_opts", {}).get("device_sync", True))
    return cls._b200_impl_

def _call_run(self, field_args, parameter_args, domain, origin, *, validate_args=True, exec_info=None):
    # same normalisation / validation / memoisation as StencilObject._call_run (stencil_object.py:534-605),
    # on raw device pointers instead of cupy arrays
    return type(self)._b200()._call_run(field_args, parameter_args, domain, origin,
                                        validate_args=validate_args, exec_info=exec_info)
'''

    def generate_implementation(self) -> str:
        names = [n for n in self.args_data.field_info] + [n for n in self.args_data.parameter_info]
        kwargs = ", ".join(f"{n}={n}" for n in names)
        sync = self.builder.options.backend_opts.get("device_sync", True)
        note = ("# device_sync=True: run() returns after the stream has been synchronised (b200_stream_sync / "
                "b200_device_sync) --\n# this backend's counterpart of cupy.cuda.Device(0).synchronize() "
                "(gtc_common.py:288-296)\n" if sync else "# device_sync=False: run() returns after the launch\n")
        return note + f"type(self)._b200().run(_domain_=_domain_, _origin_=_origin_, exec_info=None, {kwargs})"


class B200Backend(BaseBackend):
    """CUDA (sm_100a) backend: OIR lowered onto hand-written kernel templates, launched through
    libgt4py_b200.so with raw device pointers."""

    name = NAME
    options: ClassVar[dict[str, Any]] = {
        "oir_pipeline": {"versioning": True, "type": OirPipeline},
        "device_sync": {"versioning": True, "type": bool},
        "use_library_kernels": {"versioning": True, "type": bool},
        "fuse_temporaries": {"versioning": True, "type": bool},
        "k_register_cache": {"versioning": True, "type": bool},
        "fuse_columns": {"versioning": True, "type": bool},
        "warp_march": {"versioning": True, "type": bool},
        "level_fission": {"versioning": True, "type": bool},
    }
    storage_info: ClassVar[layout.LayoutInfo] = {
        "alignment": storage.ALIGNMENT,
        "device": "gpu",
        "layout_map": layout.layout_maker_factory(storage.BASE_LAYOUT),
        "is_optimal_layout": layout.layout_checker_factory(layout.layout_maker_factory(storage.BASE_LAYOUT)),
    }
    languages: ClassVar[dict] = {"computation": "cuda", "bindings": ["python"]}
    MODULE_GENERATOR_CLASS = B200ModuleGenerator

    @property
    def plan_dir_name(self) -> str:
        caching = self.builder.caching
        return f"{caching.module_prefix}b200{caching.module_postfix}"

    #: fused kernels (temporaries in registers / shared memory) win over recompute-on-the-fly point kernels from three
    #: dependent stages on (measured at 512x512x128 on B200, ms: hdiff point 0.227 / tile 0.240 / warp-march 0.203;
    #: lap-of-lap diffusion 0.37 / 0.195 / 0.166; config-4 chain 0.96 / 0.393 / 0.397; two-stage upwind advection
    #: point 0.214 / tile 0.166 when forced -- left to the point kernel by this rule, as before)
    AUTO_FUSE_MIN_STAGES = 3

    @property
    def fuse(self):
        """backend option `fuse_temporaries`: True / False / None (default: decide per stencil)."""
        return self.builder.options.backend_opts.get("fuse_temporaries", None)

    def default_pipeline(self, fuse: bool):
        """The reference's OIR pipeline; for tile kernels without OnTheFlyMerging: that pass removes temporaries
        by RECOMPUTING them at every offset they are read at (horizontal_execution_merging.py:OnTheFlyMerging),
        which the tile kernels make unnecessary -- they keep the temporaries in shared memory."""
        return DefaultPipeline(skip=[OnTheFlyMerging]) if fuse else DefaultPipeline()

    def oir(self, fuse: bool = False):
        key = f"b200:oir:{int(bool(fuse))}"
        if key not in self.builder.backend_data:
            base = GTIRToOIR().visit(self.builder.gtir)
            pipeline = self.builder.options.backend_opts.get("oir_pipeline", self.default_pipeline(fuse))
            self.builder.with_backend_data({key: pipeline.run(base)})
        return self.builder.backend_data[key]

    def library_match(self):
        """The hand-written kernel this stencil is, if any: the UNOPTIMISED OIR (program order, independent of
        the naming-dependent merge order of the OIR pipeline) is fingerprinted structurally and looked up in
        gt4py_b200/library_kernels.json (made by tools/make_library_fingerprints.py)."""
        if not self.builder.options.backend_opts.get("use_library_kernels", True):
            return None
        raw_plan, _ = lowering.lower(GTIRToOIR().visit(self.builder.gtir), self.builder.options.name, fuse=False,
                                     fuse_columns=False)
        fp = raw_plan["fingerprint"]
        entry = library_table().get(fp["digest"])
        if entry is None:
            return None
        return {"kernel": entry["kernel"], "limiter": entry.get("limiter", 0), "min_k": entry.get("min_k", 1),
                "args": dict(zip(entry["roles"], fp["api_order"]))}

    def make_plan(self):
        name, fuse = self.builder.options.name, self.fuse
        kc = bool(self.builder.options.backend_opts.get("k_register_cache", False))
        fc = bool(self.builder.options.backend_opts.get("fuse_columns", True))
        wm = bool(self.builder.options.backend_opts.get("warp_march", True))
        lf = bool(self.builder.options.backend_opts.get("level_fission", True))
        if fuse is None or fuse:
            plan, source = lowering.lower(self.oir(True), name, fuse=True, k_register_cache=kc, fuse_columns=fc, warp_march=wm, level_fission=lf)
            deepest = max([k.get("stages", 0) for k in plan["kernels"]], default=0)
            if fuse is None and deepest < self.AUTO_FUSE_MIN_STAGES:
                fuse = False
        if fuse is not None and not fuse:
            plan, source = lowering.lower(self.oir(False), name, fuse=False, k_register_cache=kc, fuse_columns=fc, warp_march=wm, level_fission=lf)
        match = self.library_match()
        if match is not None:
            plan["library"] = match
        elif (self.builder.options.backend_opts.get("use_library_kernels", True) and not _WARNED
              and any(k["kind"] in ("colfuse", "column") for k in plan["kernels"])):
            # once per process: a vadv / Thomas-shaped stencil that differs from the library's fingerprint (other
            # externals, an extra statement) still runs on the device, but on the generated column templates
            _WARNED.append(name)
            warnings.warn(
                f"b200: stencil '{name}' has FORWARD/BACKWARD loops and does not match a hand-written library kernel "
                "(gt4py_b200/library_kernels.json); it runs on the generated column templates (measured 0.7-0.8 of "
                "the HBM roofline against 0.9 for the library's vadv / tridiagonal kernels).  Shown once.",
                UserWarning, stacklevel=2)
        plan["info"] = serialize_args_data(make_args_data_from_gtir(self.builder.gtir_pipeline), self.builder.gtir)
        plan["source"] = "kernel.cu"
        return plan, source

    def generate(self):
        self.check_options(self.builder.options)
        build_info = self.builder.options.build_info
        if not self.builder.options._impl_opts.get("disable-code-generation", False):
            t0 = time.perf_counter()
            plan, source = self.make_plan()
            if build_info is not None:
                build_info["codegen_time"] = time.perf_counter() - t0
            t0 = time.perf_counter()
            out = self.builder.module_path.parent / self.plan_dir_name
            out.mkdir(parents=True, exist_ok=True)
            (out / "kernel.cu").write_text(source, encoding="utf-8")
            (out / "plan.json").write_text(json.dumps(plan, indent=1), encoding="utf-8")
            from . import generated as _generated  # what a host without Python hands to b200_plan_create

            (out / "plan.desc").write_text(_generated.plan_descriptor(plan), encoding="utf-8")
            self.builder.with_backend_data({"b200:plan_md5": hashlib.md5(source.encode()).hexdigest()})
            from . import generated  # nvcc -cubin now, so that the first call does not pay for it

            generated.compile_cubin(str(out / "kernel.cu"))
            if build_info is not None:
                build_info["build_time"] = time.perf_counter() - t0
        return self.make_module()

    @property
    def extra_cache_info(self) -> dict[str, Any]:
        return {**super().extra_cache_info, "plan_md5": self.builder.backend_data.get("b200:plan_md5", "")}


def install_storage_hooks() -> bool:
    """``gt4py.storage.{empty, zeros, ones, full, from_array}(backend="b200")`` without cupy.

    The reference allocates every ``device == "gpu"`` storage through cupy (storage/cartesian/utils.py:258-276:
    ``allocate_gpu`` asserts it, ``asarray(..., device="gpu")`` is ``cupy.asarray``, interface.py:81-86, :325).  When
    the reference found no cupy at import, this installs cupy-free replacements for exactly those two functions:
    the same padded / aligned allocation algorithm (gt4py_b200.storage.allocate, after allocators.py:189-275) over
    cudaMalloc, returning ``gt4py_b200.storage.DeviceArray`` (``__cuda_array_interface__`` v3).  With a real cupy the
    reference's own path is left alone (cupy arrays expose ``__cuda_array_interface__``, which is all the kernels need).
    """
    import numpy as np
    from gt4py.storage.cartesian import utils as storage_utils

    cp = getattr(storage_utils, "cp", None)
    if cp is not None and not getattr(cp, "__gt4py_b200_standin__", False):
        return False
    if getattr(storage_utils.allocate_gpu, "__gt4py_b200__", False):
        return True

    def allocate_gpu(shape, layout_map, dtype, alignment_bytes, aligned_index):
        shape = tuple(int(x) for x in shape)
        arr = storage.allocate(shape, tuple(int(x) for x in layout_map), np.dtype(dtype), int(alignment_bytes),
                               tuple(int(x) for x in (aligned_index or (0,) * len(shape))))
        return arr._buf, arr

    reference_asarray = storage_utils.asarray

    def asarray(array, *, device=None):
        if hasattr(array, "ndarray"):
            array = array.ndarray
        if isinstance(array, storage.DeviceArray):
            return array.asnumpy() if device == "cpu" else array
        if device == "gpu" or (device is None and hasattr(array, "__cuda_array_interface__")
                               and not hasattr(array, "__array_interface__")):
            return storage.as_device_array(array)
        return reference_asarray(array, device=device)

    allocate_gpu.__gt4py_b200__ = True
    storage_utils.allocate_gpu = allocate_gpu
    storage_utils.asarray = asarray
    return True


if NAME not in gt_backend.REGISTRY:
    gt_backend.register(B200Backend)
install_storage_hooks()
