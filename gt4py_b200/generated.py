"""Run-time half of the OIR lowering: load a *plan* (+ CUDA source) produced by ``gt4py_b200.lowering`` and
expose it as a stencil object.  Does not import gt4py -- this is what runs on the GPU box.

    st = generated.load_plan("tests/golden/generated/horizontal_regions")   # dir with plan.json + kernel.cu
    st(in_f, out_f, origin=(1, 1, 0), domain=(ni, nj, nk))

The kernels are compiled once with ``nvcc -cubin`` for sm_100a (cached next to the source), loaded through
``b200_module_load`` and launched with ``b200_launch`` (include/gt4py_b200.h); fields cross as raw pointers
and element strides packed into the generated ``struct Args``.  There is no CPU fallback: without nvcc or
without a device, loading / calling raises.
"""
from __future__ import annotations

import ctypes
import hashlib
import json
import os
import subprocess
import threading
from ctypes import byref, c_void_p
from typing import Any, Dict

import numpy as np

from . import _lib, build as _build, storage
from .stencil_object import AccessKind, Boundary, DomainInfo, FieldInfo, ParameterInfo, StencilObject

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
              "-prec-div=true", "-prec-sqrt=true"]
_CT = {"bool": ctypes.c_bool, "int8": ctypes.c_int8, "int16": ctypes.c_int16, "int32": ctypes.c_int32,
       "int64": ctypes.c_int64, "float32": ctypes.c_float, "float64": ctypes.c_double}


def cubin_cache_dir() -> str:
    """Content-addressed cubin store shared by every build of this installation (GT4PY_B200_CUBIN_CACHE overrides):
    the reference's JIT cache keeps one extension module per stencil id (caching.py:159-342), so ``rebuild=True`` or
    a new stencil id recompiles identical code; here identical generated source is compiled once."""
    d = os.environ.get("GT4PY_B200_CUBIN_CACHE") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "cubin_cache")
    os.makedirs(d, exist_ok=True)
    return d


def compile_cubin(src_path: str) -> str:
    """nvcc -cubin for sm_100a, cached by the hash of (source, flags)."""
    with open(src_path, "rb") as fh:
        digest = hashlib.sha256(fh.read() + " ".join(NVCC_FLAGS).encode()).hexdigest()[:24]
    out = os.path.join(cubin_cache_dir(), f"{digest}.cubin")
    if not os.path.exists(out):
        # unique per process AND thread: two plans with identical kernels (same digest) may compile at the same time
        tmp = f"{out}.{os.getpid()}.{threading.get_ident()}.tmp"
        cmd = [_build.nvcc(), *NVCC_FLAGS, "-cubin", "-o", tmp, src_path]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"nvcc failed on generated source {src_path}:\n{r.stderr}")
        os.replace(tmp, out)
    return out


def plan_descriptor(plan: Dict[str, Any]) -> str:
    """The plan as the line-based text `b200_plan_create` parses (include/gt4py_b200.h): what a host without Python
    needs besides the cubin.  The backend writes it as `plan.desc` next to `plan.json`."""
    code = {k.name: v for k, v in _lib.DTYPE_CODES.items()}
    L = [f"b200plan 1 {plan['name'].replace(' ', '_')}"]
    for f in plan["fields"]:
        d = f["dims"]
        L.append(f"field {f['name']} {code[f['dtype']]} {int(d[0])} {int(d[1])} {int(d[2])} {len(f['data_dims'])}")
    for s in plan["scalars"]:
        L.append(f"scalar {s['name']} {code[s['dtype']]}")
    for t in plan["temporaries"]:
        d, (pi, pj) = t["dims"], t["pad"]
        L.append(f"temp {t['name']} {code[t['dtype']]} {int(d[0])} {int(d[1])} {int(d[2])} {len(t['data_dims'])} "
                 + "".join(f"{int(x)} " for x in t["data_dims"]) + f"{pi[0]} {pi[1]} {pj[0]} {pj[1]}")
    for k in plan["kernels"]:
        head = f"kernel {k['name']} {k['kind']} {k['i_ext'][0]} {k['i_ext'][1]} {k['j_ext'][0]} {k['j_ext'][1]}"
        if k["kind"] == "point":
            L.append(f"{head} {int(k.get('rows_per_thread', 1))}")
        elif k["kind"] == "level":  # rows, loop id, direction, the K intervals (0 = START / 1 = END, offset) x 2
            iv = "".join(f" {a} {b} {c} {d}" for a, b, c, d in k["intervals"])
            L.append(f"{head} {int(k.get('rows_per_thread', 1))} {k['loop']} {int(k['backward'])} {len(k['intervals'])}{iv}")
        elif k["kind"] == "march":  # result columns per warp, warps per CTA, rows per strip, dynamic shared memory
            L.append(f"{head} {k['result_columns']} {k['warps']} {k['strip']} {k.get('smem_bytes', 0)}")
        elif k["kind"] == "tile":
            tma = k.get("tma")
            L.append(f"{head} {k['tile'][0]} {k['tile'][1]} {k['block'][0]} {k['block'][1]} {k['smem_bytes']} "
                     f"{1 if tma else 0} {tma['smem_bytes'] if tma else 0}")
        elif k["kind"] == "colfuse":
            tma = k.get("tma")
            L.append(f"{head} {k['block'][0]} {k['block'][1]} {k['ring_slots']} {k['stack_values_per_level']} "
                     f"{k.get('stack_tmem', 128)} {1 if tma else 0} {tma['smem_bytes'] if tma else 0}")
        else:
            L.append(head)
    for name, kind, ref, extra in plan["args"]:
        if kind == "field":
            L.append(f"member field {name} {int(extra)}")
        elif kind == "tmap":
            L.append(f"member tmap {name} {ref} {extra['kernel']} {extra['box'][0]} {extra['box'][1]} {extra['box'][2]}")
        elif kind == "origin":
            L.append(f"member origin {name} {ref}")
        elif kind in ("cf_spill", "cf_smem"):
            L.append(f"member {kind} {name} {ref}")
        elif kind == "level":
            L.append(f"member level {name}")
        else:
            L.append(f"member scalar {name} {code[ref]}")
    lib = plan.get("library")
    if lib is not None:
        roles = " ".join(f"{r}={a}" for r, a in lib["args"].items())
        L.append(f"library {lib['kernel']} {int(lib.get('limiter', 0))} {int(lib.get('min_k', 1))} {len(lib['args'])} {roles}")
    L.append("end")
    return "\n".join(L) + "\n"


class _Module:
    """A loaded cubin and its kernels."""

    def __init__(self, cubin_path: str, names):
        with open(cubin_path, "rb") as fh:
            self.image = fh.read()
        self.handle = c_void_p()
        buf = ctypes.create_string_buffer(self.image, len(self.image))
        _lib.check(_lib.lib().b200_module_load(buf, len(self.image), byref(self.handle)))
        self.kernels = {}
        for n in names:
            k = c_void_p()
            _lib.check(_lib.lib().b200_module_get_kernel(self.handle, n.encode(), byref(k)))
            self.kernels[n] = k


def _args_struct(plan):
    """ctypes mirror of the generated `struct Args`."""
    fields = []
    for name, kind, dtype, nstr in plan["args"]:
        if kind == "field":
            fields.append((name, c_void_p))
            fields.append((name + "_s", ctypes.c_int64 * nstr))
        elif kind == "tmap":
            fields.append((name, c_void_p))  # device address of a CUtensorMap
        elif kind == "origin":
            fields.append((name, ctypes.c_int32 * 3))
        elif kind == "cf_spill":
            fields.append((name, c_void_p))
        elif kind in ("cf_smem", "level"):
            fields.append((name, ctypes.c_int32))
        else:
            fields.append((name, _CT[dtype]))
    fields += [("ni", ctypes.c_int32), ("nj", ctypes.c_int32), ("nk", ctypes.c_int32)]
    return type("Args", (ctypes.Structure,), {"_fields_": fields})


class GeneratedStencil(StencilObject):
    """Base of plan-driven stencils; subclasses are created by `load_plan`."""

    _plan_: Dict[str, Any] = {}
    _dir_: str = ""
    _use_library_ = True  # False: always run the lowered kernels (tests compare the two paths)

    @classmethod
    def _module(cls):
        if cls.__dict__.get("_mod") is None:
            cubin = compile_cubin(os.path.join(cls._dir_, cls._plan_.get("source", "kernel.cu")))
            names = [k["name"] + sfx for k in cls._plan_["kernels"] for sfx in ("", "_u") + (("_t",) if "tma" in k else ())]
            cls._mod = _Module(cubin, names)
            for k in cls._plan_["kernels"]:
                if k["kind"] == "colfuse":
                    for sfx in ("", "_u") + (("_t",) if "tma" in k else ()):
                        _lib.check(_lib.lib().b200_kernel_set_smem(cls._mod.kernels[k["name"] + sfx], _COLFUSE_SMEM))
                    continue
                for sfx, nbytes in (("", k.get("smem_bytes", 0)), ("_u", k.get("smem_bytes", 0)),
                                    ("_t", k.get("tma", {}).get("smem_bytes", 0))):
                    if nbytes > 48 * 1024:  # opt in to more than the default dynamic shared memory
                        _lib.check(_lib.lib().b200_kernel_set_smem(cls._mod.kernels[k["name"] + sfx], int(nbytes)))
            cls._tmaps = {}
            cls._Args = _args_struct(cls._plan_)
            cls._scratch = {}
        return cls._mod

    #: scratch sets (one per (domain, stream)) and TMA descriptors kept per class; the least recently used go first
    SCRATCH_SETS = 8
    TMAP_ENTRIES = 256

    @classmethod
    def _cplan(cls):
        """The stencil as ONE C-ABI object (b200_plan_create, csrc/plan.inl): the C side packs the argument block,
        picks entry points and grids and owns scratch, TMA descriptors and spill areas."""
        if cls.__dict__.get("_cplan_handle") is None:
            cubin = compile_cubin(os.path.join(cls._dir_, cls._plan_.get("source", "kernel.cu")))
            with open(cubin, "rb") as fh:
                image = fh.read()
            desc = plan_descriptor(cls._plan_).encode()
            handle = c_void_p()
            buf = ctypes.create_string_buffer(image, len(image))
            flags = 1 | (2 if os.environ.get("B200_GENERATED_TMA", "1") == "0" else 0)
            _lib.check(_lib.lib().b200_plan_create(desc, len(desc), buf, len(image), flags, byref(handle)))
            cls._cplan_handle = handle
            cls._cplan_image = buf
            cls._cplan_arrays = {}
            cls._code = {f["name"]: _lib.DTYPE_CODES[np.dtype(f["dtype"])] for f in cls._plan_["fields"]}
            cls._entries_buf = ctypes.create_string_buffer(512)
        return cls._cplan_handle

    def _launch_cplan(self, domain, fields, params, stream):
        cls = type(self)
        handle = cls._cplan()
        specs = cls._plan_["fields"]
        key = tuple(id(fields.get(f["name"])) for f in specs)
        arr = cls._cplan_arrays.get(key)
        if arr is None:  # the b200_field structs are cached per argument by StencilObject.run: identity is stable
            arr = (_lib.Field * max(len(specs), 1))()
            for n, f in enumerate(specs):
                src = fields.get(f["name"])
                if src is None:
                    arr[n].ptr = None
                    arr[n].dtype = cls._code[f["name"]]
                    arr[n].ndim = f["ndim"]
                else:
                    ctypes.memmove(ctypes.byref(arr[n]), ctypes.byref(src), ctypes.sizeof(_lib.Field))
            if len(cls._cplan_arrays) > 64:
                cls._cplan_arrays.clear()
            cls._cplan_arrays[key] = (arr, [fields.get(f["name"]) for f in specs])  # keeps the ids alive
        else:
            arr = arr[0]
        scal = cls._plan_["scalars"]
        sarr = (_lib.Scalar * max(len(scal), 1))()
        for n, s in enumerate(scal):
            c = _lib.DTYPE_CODES[np.dtype(s["dtype"])]
            sarr[n].dtype = c
            setattr(sarr[n].value, _lib.SCALAR_MEMBER[c], params[s["name"]])
        _lib.check(_lib.lib().b200_plan_launch(handle, (ctypes.c_int64 * 3)(*domain), arr, len(specs), sarr, len(scal),
                                               _lib.stream_handle(stream), 0))
        _lib.lib().b200_plan_last_entries(handle, cls._entries_buf, 512)
        cls._last_entry_points = [e for e in cls._entries_buf.value.decode().split(",") if e]

    def _temporaries(self, domain, stream=None):
        """Global scratch for the OIR temporaries, owned by the stencil class and cached per (domain, stream): two
        streams (or a captured graph and a live stream) running the same stencil never share temporaries, and a
        call with another domain does not free what an earlier capture still points at -- a capture in progress
        takes references to the buffers it bakes in (stencil_object.Graph keeps them alive for its lifetime).
        OIR temporaries have no initial-value semantics: plain `empty`, no memset."""
        cls = type(self)
        key = (tuple(domain), _stream_key(stream))
        hit = cls._scratch.pop(key, None)
        if hit is None:
            hit = {}
            ni, nj, nk = domain
            for t in cls._plan_["temporaries"]:
                (li, hi), (lj, hj) = t["pad"]
                shape = [ni + li + hi if t["dims"][0] else 1, nj + lj + hj if t["dims"][1] else 1,
                         nk if t["dims"][2] else 1] + list(t["data_dims"])
                arr = storage.empty(shape, np.dtype(t["dtype"]), aligned_index=(li, lj, 0) + (0,) * len(t["data_dims"]),
                                    dimensions=["I", "J", "K"] + [str(d) for d in range(len(t["data_dims"]))])
                item = arr.dtype.itemsize
                strides = [s // item for s in arr.strides]
                for d in range(3):
                    if not t["dims"][d]:
                        strides[d] = 0
                origin_ptr = arr.ptr + (li * arr.strides[0] if t["dims"][0] else 0) + (lj * arr.strides[1] if t["dims"][1] else 0)
                hit[t["name"]] = (arr, origin_ptr, strides)
            while len(cls._scratch) >= cls.SCRATCH_SETS:
                cls._scratch.pop(next(iter(cls._scratch)))
        cls._scratch[key] = hit  # most recently used last
        _hold_for_capture(hit)
        return hit

    def _cf_spill(self, kname, nbytes, stream):
        """Spill area of a fused column kernel's value stack (per stream; grown on demand, never shrunk)."""
        cls = type(self)
        key = ("cf_spill", kname, _stream_key(stream))
        buf = cls._scratch.get(key)
        if buf is None or buf.nbytes < max(nbytes, 256):
            buf = storage._Buffer(max(nbytes, 256))
            cls._scratch[key] = buf
        _hold_for_capture(buf)
        return buf.ptr

    def _launch_library(self, lib_plan, domain, fields, params, stream):
        """The stencil was recognised as one the library has a hand-written kernel for (backend.library_match)."""
        L, arg, sh = _lib.lib(), lib_plan["args"], _lib.stream_handle(stream)
        dom = (ctypes.c_int64 * 3)(*domain)
        f = {role: byref(fields[name]) for role, name in arg.items() if name in fields}
        variant = int(os.environ.get("B200_LIBRARY_VARIANT", "0"))
        if lib_plan["kernel"] == "hdiff":
            rc = L.b200_hdiff(f["in_field"], f["out_field"], f["coeff"], dom, int(lib_plan["limiter"]), variant, sh)
        elif lib_plan["kernel"] == "vadv":
            rc = L.b200_vadv(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                             ctypes.c_double(params[arg["dtr_stage"]]), dom, variant, sh)
        elif lib_plan["kernel"] == "tridiag":
            rc = L.b200_tridiag(f["inf"], f["diag"], f["sup"], f["rhs"], f["out"], dom, variant, sh)
        else:
            raise RuntimeError(f"plan names an unknown library kernel {lib_plan['kernel']!r}")
        _lib.check(rc)

    def _launch(self, domain, fields, params, stream):
        cls = type(self)
        lib_plan = cls._plan_.get("library")
        if lib_plan is not None and cls._use_library_ and domain[2] >= lib_plan.get("min_k", 1):
            return self._launch_library(lib_plan, domain, fields, params, stream)
        if not _PY_LAUNCH:
            return self._launch_cplan(domain, fields, params, stream)
        # the same launch interpreted in Python (B200_PY_LAUNCH=1: kept for A/B runs of the call overhead)
        mod = cls._module()
        a = cls._Args()
        a.ni, a.nj, a.nk = domain
        unit = True  # every field I-contiguous -> the <kernel>_u entry points (stride folded to 1)
        for spec in cls._plan_["fields"]:
            name = spec["name"]
            f = fields.get(name)
            if f is None:  # unreferenced / optional field passed as None
                setattr(a, name, None)
                continue
            item = np.dtype(spec["dtype"]).itemsize
            # the launcher trusts nothing the validation may have skipped (validate_args=False, or the memoised
            # (shapes, origins) key, which like the reference's carries no dtype): the reference's bindings reject a
            # wrong buffer format at the C++ boundary (gtc_common.py:30-62), so does this
            if int(f.dtype) != _lib.DTYPE_CODES[np.dtype(spec["dtype"])]:
                raise TypeError(f"The dtype of field '{name}' does not match the stencil's ('{spec['dtype']}')")
            if int(f.ndim) != spec["ndim"]:
                raise ValueError(f"Field '{name}' has {int(f.ndim)} dimensions, the stencil expects {spec['ndim']}")
            # the descriptor's axes are the field's own (masked) axes; the kernel indexes I, J, K, data...
            strides, ptr, pos = [0, 0, 0], f.ptr, 0
            for d in range(3):
                if spec["dims"][d]:
                    strides[d] = f.stride[pos]
                    ptr += f.origin[pos] * f.stride[pos] * item
                    pos += 1
            strides += [f.stride[pos + n] for n in range(len(spec["data_dims"]))]
            span = sum((int(f.shape[n]) - 1) * abs(int(f.stride[n])) for n in range(int(f.ndim)))
            unit = unit and (not spec["dims"][0] or strides[0] == 1) and span < (1 << 31) - (1 << 20)
            setattr(a, name, ptr)
            getattr(a, name + "_s")[:] = strides
        for name, (arr, origin_ptr, strides) in self._temporaries(tuple(domain), stream).items():
            setattr(a, name, origin_ptr)
            getattr(a, name + "_s")[:] = strides
            unit = unit and (strides[0] in (0, 1))
        for s in cls._plan_["scalars"]:
            setattr(a, s["name"], params[s["name"]])
        # TMA descriptors of the staged inputs (cached per field layout); a field that is not TMA-compatible
        # (alignment, strides) sends its kernel to the plain entry point instead
        tma_ok = {}
        if unit and os.environ.get("B200_GENERATED_TMA", "1") != "0":
            shifts = {}
            for name, kind, fname, extra in cls._plan_["args"]:
                if kind != "tmap":
                    continue
                f = fields[fname]
                key = (name, f.ptr, tuple(f.shape[:3]), tuple(f.stride[:3]))
                hit = cls._tmaps.get(key)
                if hit is None:
                    blob, shift = (ctypes.c_uint8 * 128)(), ctypes.c_int32(0)
                    rc = _lib.lib().b200_tmap_encode(byref(f), (ctypes.c_uint32 * 3)(*extra["box"]), blob, byref(shift))
                    if rc == -3:  # B200_EUNSUP: layout not TMA-compatible
                        hit = False
                    else:
                        _lib.check(rc)
                        dev = storage._Buffer(128)  # the descriptor lives in global memory (256-byte aligned)
                        _lib.check(_lib.lib().b200_memcpy_h2d(c_void_p(dev.ptr), blob, 128, None))
                        hit = (dev, int(shift.value))
                    while len(cls._tmaps) >= cls.TMAP_ENTRIES:  # least recently made first; a capture holds its own
                        cls._tmaps.pop(next(iter(cls._tmaps)))
                    cls._tmaps[key] = hit
                if hit is not False:
                    _hold_for_capture(hit[0])
                kname = extra["kernel"]
                if hit is False:
                    tma_ok[kname] = False
                else:
                    setattr(a, name, hit[0].ptr)
                    shifts[fname] = hit[1]
                    tma_ok.setdefault(kname, True)
            for name, kind, fname, extra in cls._plan_["args"]:
                if kind == "origin":  # array index of the compute origin, in the descriptor's coordinates
                    org = [int(x) for x in fields[fname].origin[:3]]
                    org[0] += shifts.get(fname, 0)
                    getattr(a, name)[:] = org
        ni, nj, nk = domain
        cls._last_entry_points = []  # what the last call launched (tests check that the intended variant ran)
        argp = (c_void_p * 1)(ctypes.cast(byref(a), c_void_p))
        sh = _lib.stream_handle(stream)
        kernels = cls._plan_["kernels"]
        for n, k in enumerate(kernels):
            wi = ni + k["i_ext"][1] - k["i_ext"][0]
            wj = nj + k["j_ext"][1] - k["j_ext"][0]
            if k["kind"] == "level":
                # a sequential loop run level by level: its stages are consecutive `level` kernels of one loop id
                if n and kernels[n - 1]["kind"] == "level" and kernels[n - 1]["loop"] == k["loop"]:
                    continue  # launched with the head of its group
                group = [g for g in kernels[n:] if g["kind"] == "level" and g["loop"] == k["loop"]]
                sfx = "_u" if unit else ""
                for lvl in level_order(nk, k["backward"]):
                    for g in group:
                        gi = ni + g["i_ext"][1] - g["i_ext"][0]
                        gj = nj + g["j_ext"][1] - g["j_ext"][0]
                        if gi <= 0 or gj <= 0 or not level_in(g["intervals"], lvl, nk):
                            continue
                        rows = 4 * int(g.get("rows_per_thread", 1))
                        a.lvl = lvl
                        if g["name"] + sfx not in cls._last_entry_points:
                            cls._last_entry_points.append(g["name"] + sfx)
                        _lib.check(_lib.lib().b200_launch(
                            mod.kernels[g["name"] + sfx], (ctypes.c_uint32 * 3)((gi + 63) // 64, (gj + rows - 1) // rows, 1),
                            (ctypes.c_uint32 * 3)(64, 4, 1), 0, sh, argp, 1))
                continue
            if wi <= 0 or wj <= 0:
                continue
            smem, sfx = 0, ("_u" if unit else "")
            if k["kind"] == "point":
                rows = 4 * int(k.get("rows_per_thread", 1))  # rows of the compute block a CTA covers
                block, grid = (64, 4, 1), ((wi + 63) // 64, (wj + rows - 1) // rows, nk)
            elif k["kind"] == "march":  # a warp per (result columns x strip of rows) of a level (lowering_march.py)
                per_cta = k["result_columns"] * k["warps"]
                block, grid = (32 * k["warps"], 1, 1), ((ni + per_cta - 1) // per_cta, (nj + k["strip"] - 1) // k["strip"], nk)
                smem = k.get("smem_bytes", 0)  # the lanes' input rings
            elif k["kind"] == "tile":  # a CTA per TI x TJ tile of the compute domain and level; halo recomputed
                ti, tj = k["tile"]
                block, grid, smem = (*k["block"], 1), ((ni + ti - 1) // ti, (nj + tj - 1) // tj, nk), k["smem_bytes"]
                if tma_ok.get(k["name"]):
                    sfx, smem = "_t", k["tma"]["smem_bytes"]
            elif k["kind"] == "colfuse":
                # persistent fused FORWARD / BACKWARD kernel (lowering_colfuse.py): a CTA per SM walks 256-column tiles;
                # the value stack beyond Tensor Memory goes to shared memory, then to an L2-resident spill area
                bx, by = k["block"]
                ntiles = ((wi + bx - 1) // bx) * ((wj + by - 1) // by)
                ctas = max(1, min(ntiles, _sm_count()))
                slot = bx * by * 8
                need = max(0, nk * k["stack_values_per_level"] - k.get("stack_tmem", 128))
                ring = k["ring_slots"] * slot
                if tma_ok.get(k["name"]):  # the <k>_t entry: rings fed by TMA tile loads
                    sfx, ring = "_t", k["tma"]["smem_bytes"]
                n_smem = max(0, min(need, (_COLFUSE_SMEM - ring) // slot))
                smem = ring + n_smem * slot
                spill = self._cf_spill(k["name"], (need - n_smem) * ctas * bx * by * 8, stream)
                setattr(a, f"cf_{k['name']}_smem", n_smem)
                setattr(a, f"cf_{k['name']}_spill", spill)
                # (the <k>_t entry runs with one more row of threads: the producer of its TMA rings)
                block, grid = (bx, by + (1 if sfx == "_t" else 0), 1), (ctas, 1, 1)
            else:
                bx, by = _COLUMN_BLOCK
                block, grid = (bx, by, 1), ((wi + bx - 1) // bx, (wj + by - 1) // by, 1)
            if nk <= 0 and k["kind"] not in ("column", "colfuse"):
                continue
            cls._last_entry_points.append(k["name"] + sfx)
            _lib.check(_lib.lib().b200_launch(mod.kernels[k["name"] + sfx], (ctypes.c_uint32 * 3)(*grid),
                                              (ctypes.c_uint32 * 3)(*block), smem, sh, argp, 1))


def level_order(nk: int, backward: bool):
    return range(nk - 1, -1, -1) if backward else range(nk)


def level_in(intervals, lvl: int, nk: int) -> bool:
    """Is level `lvl` inside one of a `level` kernel's K intervals ([start level, offset, end level, offset]; level
    0 = START, 1 = END)?  (The kernel checks again; this only saves the launch.)"""
    return any((nk * sl + so) <= lvl < (nk * el + eo) for sl, so, el, eo in intervals)


_PY_LAUNCH = os.environ.get("B200_PY_LAUNCH", "0") == "1"
_COLFUSE_SMEM = 226 * 1024  # dynamic shared memory a fused column kernel may use (227 KB minus its static part)
_SM_COUNT = []


def _sm_count() -> int:
    if not _SM_COUNT:
        _SM_COUNT.append(int(_lib.device_props()["sm_count"]))
    return _SM_COUNT[0]


def _stream_key(stream):
    if stream is None:
        return 0
    h = _lib.stream_handle(stream)
    return int(h.value or 0)


def _hold_for_capture(obj):
    """During a CUDA-graph capture the graph being built keeps `obj` (device buffers whose addresses the captured
    launches carry) alive: stencil_object.Graph collects the list on exit."""
    from . import stencil_object as so

    refs = getattr(so._tls, "graph_refs", None)
    if refs is not None:
        refs.append(obj)


# threads of a column-kernel CTA along I x J (a thread per column); tuning hook B200_COLUMN_BLOCK=IxJ
_COLUMN_BLOCK = tuple(int(x) for x in os.environ.get("B200_COLUMN_BLOCK", "32x4").split("x"))


def stencil_class(plan: Dict[str, Any], directory: str):
    """Build the stencil class described by `plan` (plan["info"] holds the reference's args data)."""
    info = plan["info"]
    finfo = {}
    for name, fi in info["field_info"].items():
        if fi is None:
            finfo[name] = FieldInfo(AccessKind.NONE, Boundary(((0, 0), (0, 0), (0, 0))), ("I", "J", "K"), (), np.dtype("float64"))
            continue
        finfo[name] = FieldInfo(AccessKind(fi["access"]), Boundary(fi["boundary"]), tuple(fi["axes"]),
                                tuple(fi["data_dims"]), np.dtype(fi["dtype"]))
    pinfo = {name: (ParameterInfo(AccessKind(pi["access"]), np.dtype(pi["dtype"])) if pi is not None
                    else ParameterInfo(AccessKind.NONE, np.dtype("float64")))
             for name, pi in info["parameter_info"].items()}
    di = info["domain_info"]
    attrs = {
        "_plan_": plan, "_dir_": directory, "_arg_names_": tuple(info["arg_names"]),
        "_arg_defaults_": dict(info.get("arg_defaults", {})),
        "_gt_domain_info_": DomainInfo(tuple(di["parallel_axes"]), di["sequential_axis"],
                                       di["min_sequential_axis_size"], di["ndim"]),
        "_gt_field_info_": finfo, "_gt_parameter_info_": pinfo,
        "_gt_options_": {"name": plan["name"], "module": __name__},
    }
    return type(f"{plan['name']}__b200", (GeneratedStencil,), attrs)


def load_plan(directory: str, use_library: bool = True):
    """Stencil object for the plan in `directory` (plan.json + the CUDA source it names)."""
    with open(os.path.join(directory, "plan.json")) as fh:
        plan = json.load(fh)
    cls = stencil_class(plan, directory)
    cls._use_library_ = bool(use_library)
    return cls()
