#!/usr/bin/env python
"""bench.py -- the measured hot path: hdiff + vadv on synthetic float64 fields (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--domain NI NJ NK] [--no-e2e] [--no-cpu-baseline] [--no-configs] [--no-parity]

Workload (config.workload): BASELINE.json configs[4], "hdiff + vadv weak-scaling sweep at 2048x2048x80 per GPU with IJ
decomposition and halo exchange at 1/2/4/8 B200" -- for every N, N = 1 included.  One *step* = one
horizontal_diffusion call + one vertical_advection_dycore call over one set of fields (2 kernel launches,
2 * NI*NJ*NK grid-point updates).  With N > 1 (torchrun, one rank per GPU) every rank owns an NI x NJ x NK block of a
Px x Py IJ decomposition of a global (Px*NI) x (Py*NJ) x NK domain (weak scaling): each step first
refreshes the width-2 halo of in_field and the +I halo column of wcon from the neighbour ranks over NVLink.
The fields are synthetic (b200_fill_synthetic: a counter-based hash of the GLOBAL point index, restated in numpy in
oracle/inputs.py), so every rank -- and the checker -- can compute any window of the same global field.

Numbers on the JSON line
  value        grid-point updates/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks.
  roofline     dominant kernel: algorithmic bytes (vadv 48 B/point, hdiff 24) / its mean launch time
               from per-launch CUDA events in the same timed region, vs the measured HBM copy bandwidth in
               MEASURED_PEAKS.json; the kernel NAME is what the library reports it launched (b200_last_kernel).
  kernels      the same for both kernels of the step.
  configs      (N = 1) driver-visible timings + rooflines of the other BASELINE configs: [2] hdiff + vadv at
               1024x1024x80 (the north star's single-GPU target), [3] the FV3-style chain at 512x512x128 (generated
               fused kernel, 32 B/point), [1] vadv 256x256x80, [0] hdiff 128x128x80, and the Thomas solver
               (a3, 56 B/point) at 1024x1024x80.
  parity       after the timed region: two steps of a small model whose exchanged fields CHANGE every step (hdiff
               feeds its own output back, vadv's result becomes the next wcon), every received halo cell poisoned
               with NaN before each refresh, each rank's block compared bit-for-bit with the single-domain C oracle
               on strips along its four edges, two corners and an interior patch (oracle/model_parity.py).
               `halo_parity` (N > 1) is the AND over ranks; a false value makes the run exit non-zero.
  e2e          same metric through the public host-buffer API (gt4py_b200.HostPipeline) at configs[2]'s size per
               rank: every step copies that step's inputs from pinned host memory, runs the kernels and copies the
               results back.
  cpu_baseline the REFERENCE's own numpy backend (oracle/_ref, staged by oracle/make_ref.py) on a bounded sample of
               the workload on the box's host cores (kind "reference", 1 core: numpy's elementwise loops are
               single-threaded); the oracle's C/OpenMP restatement on all cores is reported beside it (`cpu_port`).

--impl reference times the reference numpy backend alone (rank 0 only).
The oracle / the reference are used here ONLY as checker and CPU baseline; the measured path is the CUDA library.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_BYTES = {"hdiff": 24, "vadv": 48, "tridiag": 56, "chain": 32}  # fp64 bytes per grid point
METRIC = "grid-points/s (hdiff+vadv, float64)"
UNIT = "grid-points/s"
DEFAULT_DOMAIN = (2048, 2048, 80)   # BASELINE configs[4], per GPU
CONFIG2_DOMAIN = (1024, 1024, 80)   # BASELINE configs[2]
CPU_SAMPLE_DOMAIN = (256, 256, 80)  # sub-block of the workload the reference numpy backend is timed on


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--domain", type=int, nargs=3, default=list(DEFAULT_DOMAIN), help="per-GPU compute domain NI NJ NK")
    ap.add_argument("--sets", type=int, default=2, help="field sets rotated between steps")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the `configs` block (N = 1)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--halo-mode", default="push", choices=["push", "push2", "pull"],
                    help="N>1: halo refresh by pushing faces into the neighbours' staging buffers "
                         "or by pulling boxes out of the neighbours' fields; both over NVLink peer memory")
    ap.add_argument("--split", action="store_true",
                    help="N>1: run both stencils through halo.DistributedStencil (refresh on its own stream, interior of "
                         "the domain meanwhile, boundary strips after it) instead of hiding one stencil's refresh behind "
                         "the other stencil")
    ap.add_argument("--no-overlap", action="store_true",
                    help="N>1: run the halo exchange on the compute stream instead of beside it")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def workload_name(dom):
    tag = {DEFAULT_DOMAIN: " (BASELINE configs[4])", CONFIG2_DOMAIN: " (BASELINE configs[2])"}.get(tuple(dom), "")
    return f"hdiff + vadv at {dom[0]}x{dom[1]}x{dom[2]} float64 per GPU{tag}"


def decomposition(world):
    return {1: "1x1", 2: "2x1", 4: "2x2", 8: "4x2"}.get(world, f"{world}x1") + " (IxJ ranks)"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---- clocks sampler --------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        rows = [r for (t, r) in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1 + 0.02)]
        window = "timed region"
        if not rows:  # timed region shorter than the sampling period: fall back to warm-up + timed region
            rows = [r for _, r in self.rows]
            window = "warm-up + timed region"
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1])); pw.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "power_w_max": float(max(pw)) if pw else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


# ---- CPU legs (the reference's numpy backend; the oracle's OpenMP port beside it) ----------------------------------
def host_inputs(dom, goff=(0, 0), gdims=None):
    """The workload's fields for a block at global offset `goff`, as host arrays in the device layout (I contiguous)."""
    from oracle import inputs as gen  # input recipes only (numpy); no oracle arithmetic

    ni, nj, nk = dom
    gdims = gdims or (ni, nj)
    base = gen.hdiff_base2d(goff[0] - 2, goff[1] - 2, ni + 4, nj + 4, gdims[0], gdims[1])
    h, v = {}, {}
    for name in ("in_field", "coeff", "out_field"):
        seed, lo, hi = gen.SYNTH[name]
        h[name] = np.asfortranarray(gen.hashed_window(goff[0] - 2, goff[1] - 2, ni + 4, nj + 4, nk, seed, lo, hi,
                                                      base if name == "in_field" else None))
    for name in ("utens_stage", "u_stage", "wcon", "u_pos", "utens"):
        seed, lo, hi = gen.SYNTH[name]
        v[name] = np.asfortranarray(gen.hashed_window(goff[0], goff[1], ni + 1, nj, nk, seed, lo, hi))
    return h, v, gen.DTR_STAGE


def reference_numpy_leg(steps, warmup, budget_s):
    """Time the reference's own numpy backend (src/gt4py/cartesian/backend/numpy_backend.py:28-84, staged under
    oracle/_ref) on CPU_SAMPLE_DOMAIN: hdiff + vadv per step.  Returns (ms_per_step, steps, n_points) or None."""
    from oracle import ref

    if not ref.activate():
        return None
    import warnings

    warnings.filterwarnings("ignore")
    from cartesian_tests.integration_tests.multi_feature_tests import stencil_definitions as sd
    from gt4py import storage as gt_storage
    from gt4py.cartesian import gtscript

    dom = CPU_SAMPLE_DOMAIN
    hd = gtscript.stencil("numpy", sd.REGISTRY["horizontal_diffusion"], format_source=False)
    va = gtscript.stencil("numpy", sd.REGISTRY["vertical_advection_dycore"], externals={"BET_M": 0.5, "BET_P": 0.5},
                          format_source=False)
    h, v, dtr = host_inputs(dom)
    hs = {k: gt_storage.from_array(a, backend="numpy", aligned_index=(2, 2, 0)) for k, a in h.items()}
    vs = {k: gt_storage.from_array(a, backend="numpy") for k, a in v.items()}

    def step():
        hd(hs["in_field"], hs["out_field"], hs["coeff"], origin=(2, 2, 0), domain=dom, validate_args=False)
        va(vs["utens_stage"], vs["u_stage"], vs["wcon"], vs["u_pos"], vs["utens"], dtr_stage=np.float64(dtr),
           origin=(0, 0, 0), domain=dom, validate_args=False)

    t = time.perf_counter()
    step()
    one = time.perf_counter() - t
    for _ in range(max(0, min(warmup - 1, int(budget_s / 4 / max(one, 1e-9))))):
        step()
    steps = max(1, min(steps, int(budget_s / max(one, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    return (time.perf_counter() - t0) / steps * 1e3, steps, dom[0] * dom[1] * dom[2]


def port_leg(dom, steps, budget_s):
    """The oracle's C/OpenMP restatement on all host threads, same fields.  Returns (ms_per_step, steps)."""
    from oracle import cpu as oracle

    h, v, dtr = host_inputs(dom)

    def step():
        oracle.hdiff(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)
        oracle.vadv(v["utens_stage"], v["u_stage"], v["wcon"], v["u_pos"], v["utens"], dtr_stage=dtr,
                    origin=(0, 0, 0), domain=dom)

    t = time.perf_counter()
    step()
    one = time.perf_counter() - t
    steps = max(1, min(steps, int(budget_s / max(one, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    return (time.perf_counter() - t0) / steps * 1e3, steps


def cpu_baseline_block(steps, warmup, budget_ref, budget_port):
    cores = host_threads()
    out = {}
    ref_leg = None
    try:
        ref_leg = reference_numpy_leg(steps, warmup, budget_ref)
    except Exception as exc:  # the reference could not be imported / built: say so, fall back to the port
        out["reference_error"] = f"{type(exc).__name__}: {exc}"[:300]
    d = CPU_SAMPLE_DOMAIN
    os.environ["OMP_NUM_THREADS"] = str(cores)  # torchrun pins it to 1; the port uses every core
    ms_p, steps_p = port_leg(CONFIG2_DOMAIN, steps, budget_port)
    n2 = CONFIG2_DOMAIN[0] * CONFIG2_DOMAIN[1] * CONFIG2_DOMAIN[2]
    port = {"value": 2 * n2 / (ms_p * 1e-3), "unit": UNIT, "cores": cores, "kind": "port", "ms_per_step": ms_p,
            "steps": steps_p,
            "sample": f"{steps_p} steps of hdiff + vadv at {CONFIG2_DOMAIN} with the oracle's C/OpenMP restatement "
                      f"(oracle/stencils.c), {cores} threads"}
    if ref_leg is None:
        out.update(port)
        out["note"] = "reference numpy backend unavailable here (oracle/_ref missing); the port is the baseline"
        return out, port
    ms, n_steps, n = ref_leg
    out.update({"value": 2 * n / (ms * 1e-3), "unit": UNIT, "cores": 1, "kind": "reference", "ms_per_step": ms,
                "steps": n_steps,
                "sample": f"{n_steps} steps of hdiff + vadv on a {d[0]}x{d[1]}x{d[2]} block of the workload with the "
                          f"reference's numpy backend (gt4py 1.1.11 staged under oracle/_ref; numpy {np.__version__}, "
                          "single-threaded elementwise loops: 1 core)",
                "host_cores_available": cores})
    return out, port


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    dom = tuple(args.domain)
    cpu, port = cpu_baseline_block(args.steps, args.warmup, budget_ref=90.0, budget_port=15.0)
    value = cpu["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": cpu.get("steps", args.steps), "warmup": args.warmup, "ms_per_step": cpu["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(dom), "domain_per_gpu": list(dom), "device": "host cpu",
                   "sample_domain": list(CPU_SAMPLE_DOMAIN)},
        "cpu_baseline": cpu, "cpu_port": port,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def bind_near_gpu(index: int):
    """Pin this rank to the host cores NVML reports as closest to its GPU (intersected with the cores this
    process may use), so that the pinned host buffers of the e2e path are first-touched on the GPU's NUMA node.
    Returns a short description for the JSON line; never fatal."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        allowed = os.sched_getaffinity(0)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(allowed) // 64) + 1)
        near = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        pick = sorted(near & allowed)
        if pick and len(pick) < len(allowed):
            os.sched_setaffinity(0, pick)
            return f"bound to {len(pick)} of {len(allowed)} allowed cores near GPU {index}"
        return f"{len(allowed)} allowed cores, all equally near GPU {index}" if pick else "GPU-near cores not allowed"
    except Exception as exc:  # NVML missing / restricted container
        return f"unbound ({type(exc).__name__})"


# ---- device fields -----------------------------------------------------------------------------------------------
def make_device_fields(storage, dom, goff, gdims, stream=None):
    """One set of the workload's eight fields on the device, filled by b200_fill_synthetic in global coordinates."""
    from oracle import inputs as gen  # recipes (seeds / ranges) only

    ni, nj, nk = dom
    base = storage.from_array(gen.hdiff_base2d(goff[0] - 2, goff[1] - 2, ni + 4, nj + 4, gdims[0], gdims[1]),
                              dimensions=["I", "J"])
    h, v = {}, {}
    for name in ("in_field", "coeff", "out_field"):
        seed, lo, hi = gen.SYNTH[name]
        h[name] = storage.empty((ni + 4, nj + 4, nk), aligned_index=(2, 2, 0))
        storage.fill_synthetic(h[name], seed, lo, hi, global_offset=(goff[0] - 2, goff[1] - 2),
                               base2d=base if name == "in_field" else None, stream=stream)
    for name in ("utens_stage", "u_stage", "wcon", "u_pos", "utens"):
        seed, lo, hi = gen.SYNTH[name]
        v[name] = storage.empty((ni + 1, nj, nk))
        storage.fill_synthetic(v[name], seed, lo, hi, global_offset=goff, stream=stream)
    return h, v, base


def refill(storage, arr, recipe, goff_arr, base=None):
    from oracle import inputs as gen

    seed, lo, hi = gen.SYNTH[recipe]
    storage.fill_synthetic(arr, seed, lo, hi, global_offset=goff_arr, base2d=base)


def halo_leg(halo, groups, stream, gb, barrier, allmax, reps=50):
    """The halo exchange alone: `reps` x (wcon refresh + in_field refresh) back to back on one stream, CUDA
    events, max over ranks.  Bytes are what one GPU receives over NVLink per step."""
    nbytes = sum(halo.bytes_received(g) for g in groups[0])
    for i in range(5):
        for g in groups[i % len(groups)]:
            halo.refresh(g, stream)
    barrier()
    e0, e1 = gb.Event(), gb.Event()
    e0.record(stream)
    for i in range(reps):
        for g in groups[i % len(groups)]:
            halo.refresh(g, stream)
    e1.record(stream)
    stream.synchronize()
    us = allmax(e0.elapsed_ms(e1)) * 1e3 / reps
    barrier()
    gbs = nbytes / us / 1e3
    return {"bytes_received_per_gpu_per_step": int(nbytes), "us_per_step": round(us, 2), "achieved": round(gbs, 1),
            "unit": "GB/s", "peak": 900.0, "frac": round(gbs / 900.0, 4),
            "peak_source": "NVLink 5 nominal, per direction per GPU (B200_PROFILING.md; measured peer copy 770 GB/s)",
            "note": "faces of 1.3-2.6 MB and 2.5 KB corners, one launch per refresh (push + unpack fused): launch / flag "
                    "latency, not link bandwidth, bounds it",
            "launches_counted": (5 + reps) * len(groups[0]) * halo.launches_per_refresh}


# ---- the multi-step parity model (see oracle/model_parity.py) -----------------------------------------------------
def model_parity(gb, storage, hd, va, hset, vset, base, dom, goff, gdims, halo, pgroups, stream, steps=2, dstencils=None):
    """Run `steps` steps of the evolving model on the device and compare this rank's block with the single-domain
    oracle on the check windows.  Returns (ok, report)."""
    from oracle import model_parity as mp

    ni, nj, nk = dom
    A, B = hset["in_field"], hset["out_field"]
    P, Q = vset["utens_stage"], vset["wcon"]
    # initial state: both arrays of each chain hold their recipe everywhere (halo frame / column GI included)
    refill(storage, A, "in_field", (goff[0] - 2, goff[1] - 2), base)
    refill(storage, B, "in_field", (goff[0] - 2, goff[1] - 2), base)
    refill(storage, hset["coeff"], "coeff", (goff[0] - 2, goff[1] - 2))
    refill(storage, P, "utens_stage", goff)
    refill(storage, Q, "wcon", goff)
    gb.synchronize()
    with gb.use_stream(stream):
        for t in range(steps):
            x_in, x_out = (A, B) if t % 2 == 0 else (B, A)
            u, w = (P, Q) if t % 2 == 0 else (Q, P)
            if halo is not None:
                g_in, g_w = pgroups[("in", id(x_in))], pgroups[("w", id(w))]
                halo.poison(g_in, stream)
                halo.poison(g_w, stream)
                if dstencils:  # --split: refresh + interior + boundary strips inside the distributed stencils
                    dstencils[("in", id(x_in))](x_in, x_out, hset["coeff"])
                    dstencils[("w", id(w))](u, vset["u_stage"], w, vset["u_pos"], vset["utens"], dtr_stage=gen_dtr())
                    continue
                halo.refresh(g_w, stream)
                halo.refresh(g_in, stream)
            hd(x_in, x_out, hset["coeff"], origin=(2, 2, 0), domain=dom)
            va(u, vset["u_stage"], w, vset["u_pos"], vset["utens"], dtr_stage=gen_dtr(), origin=(0, 0, 0), domain=dom)
        stream.synchronize()
    x_last = B if (steps - 1) % 2 == 0 else A
    u_last = P if (steps - 1) % 2 == 0 else Q
    report, ok = {}, True
    for name, (i0, i1, j0, j1) in mp.windows(goff, dom):
        want_h = mp.hdiff_chain((i0, i1, j0, j1), steps, gdims, nk)
        want_v = mp.vadv_chain((i0, i1, j0, j1), steps, gdims, nk)
        li, lj = i0 - goff[0], j0 - goff[1]
        got_h = x_last[2 + li:2 + li + (i1 - i0), 2 + lj:2 + lj + (j1 - j0), :].asnumpy()
        got_v = u_last[li:li + (i1 - i0), lj:lj + (j1 - j0), :].asnumpy()
        eq_h = bool(np.array_equal(got_h, want_h, equal_nan=True)) and not np.isnan(got_h).any()
        eq_v = bool(np.array_equal(got_v, want_v, equal_nan=True)) and not np.isnan(got_v).any()
        report[name] = {"hdiff": eq_h, "vadv": eq_v}
        ok = ok and eq_h and eq_v
    return ok, report


def gen_dtr():
    from oracle import inputs as gen

    return gen.DTR_STAGE


# ---- the b200 arm ------------------------------------------------------------------------------------
def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"),
                   os.path.abspath(__file__)] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")

    # libraries (NCCL prints its version banner) must not write into the one-JSON-line stdout contract:
    # everything written to fd 1 from here on goes to stderr, the result line goes to the saved descriptor
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)

    import gt4py_b200 as gb
    from gt4py_b200 import _lib, stencils, storage

    if gb.device_count() <= local:
        raise SystemExit("bench.py: no CUDA device for this rank; the b200 path has no CPU fallback")
    _lib.set_device(local)
    affinity = bind_near_gpu(local) if world > 1 else None

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_

        torch.cuda.set_device(local)
        dist_.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_

    def barrier():
        if dist is not None:
            dist.barrier()
        _lib.synchronize()

    def allmax(x: float) -> float:
        if dist is None:
            return x
        import torch

        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allmin(x: float) -> float:
        return -allmax(-x)

    dom = tuple(args.domain)
    ni, nj, nk = dom
    n = ni * nj * nk
    from gt4py_b200 import halo as halo_mod

    decomp = halo_mod.Decomposition.create(world, rank)
    goff = decomp.global_offset(dom)
    gdims = (decomp.px_count * ni, decomp.py_count * nj)

    # device field sets (set 0 filled by the synthetic-data kernel, the others are device copies of it)
    t_setup = time.perf_counter()
    h0, v0, base = make_device_fields(storage, dom, goff, gdims)
    hsets, vsets = [h0], [v0]
    for _ in range(1, max(1, args.sets)):
        hsets.append({k: a.copy() for k, a in h0.items()})
        vsets.append({k: a.copy() for k, a in v0.items()})
    gb.synchronize()
    setup_s = time.perf_counter() - t_setup
    dtr = gen_dtr()

    hd = stencils.HorizontalDiffusion()
    va = stencils.VerticalAdvectionDycore()
    type(hd).configure(device_sync=False)
    type(va).configure(device_sync=False)

    halo, groups, pgroups, dstencils = None, [], {}, {}
    if world > 1:
        halo = halo_mod.HaloExchanger(halo_mod.TorchComm(dist), decomp, mode=args.halo_mode)
        h_in = halo_mod.halo_from_field_info(hd, "in_field")   # ((2, 2), (2, 2)) from field_info.boundary
        h_w = halo_mod.halo_from_field_info(va, "wcon")        # ((0, 1), (0, 0))
        def register(arr, role):
            """A halo group for `arr` in the in_field / wcon role (with --split: inside a DistributedStencil)."""
            if args.split:
                ds = (halo_mod.DistributedStencil(hd, halo, {"in_field": arr}, origin=(2, 2, 0), domain=dom) if role == "in"
                      else halo_mod.DistributedStencil(va, halo, {"wcon": arr}, origin=(0, 0, 0), domain=dom))
                dstencils[(role, id(arr))] = ds
                g = ds.group
            else:
                g = halo.add_group([(arr, (2, 2, 0), dom, h_in)] if role == "in" else [(arr, (0, 0, 0), dom, h_w)])
            pgroups[(role, id(arr))] = g
            return g

        for h, v in zip(hsets, vsets):
            groups.append((register(v["wcon"], "w"), register(h["in_field"], "in")))
        # the parity model swaps array roles: set 0's out_field takes the in-role, its utens_stage the wcon-role
        register(h0["out_field"], "in")
        register(v0["utens_stage"], "w")
        halo.enable_multi_stream()  # the timed loop refreshes on the halo stream, the parity model on the compute stream
        halo.finalize()

    stream = gb.Stream()
    overlap = halo is not None and not args.no_overlap and not args.split
    stream_h = gb.Stream() if overlap else None          # halo exchange runs beside the compute stream
    ev_step, ev_wcon, ev_in = (gb.Event(), gb.Event(), gb.Event()) if overlap else (None, None, None)
    names = {}

    def run_hdiff(h):
        hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)

    def run_vadv(v):
        va(v["utens_stage"], v["u_stage"], v["wcon"], v["u_pos"], v["utens"], dtr_stage=dtr,
           origin=(0, 0, 0), domain=dom)

    def step(i, ev=None, refresh=True):
        """hdiff + vadv on field set i.  N>1: the +I halo column of wcon and the width-2 halo of in_field are
        refreshed from the neighbour GPUs first (every rank calls refresh the same number of times).  With
        overlap the exchange runs on its own stream: vadv waits only for wcon's (one column, one neighbour),
        and in_field's (8 neighbours) is hidden behind vadv; ev = (before first, between, after second kernel)."""
        s = i % len(hsets)
        h, v = hsets[s], vsets[s]
        if args.split and halo is not None and refresh:
            if ev is not None:
                ev[0].record(stream)
            dstencils[("in", id(h["in_field"]))](h["in_field"], h["out_field"], h["coeff"])
            if ev is not None:
                ev[1].record(stream)
            dstencils[("w", id(v["wcon"]))](v["utens_stage"], v["u_stage"], v["wcon"], v["u_pos"], v["utens"], dtr_stage=dtr)
            if ev is not None:
                ev[2].record(stream)
            return
        if overlap and refresh:
            g_w, g_in = groups[s]
            ev_step.record(stream)
            stream_h.wait_event(ev_step)                 # not before the previous users of these fields
            halo.refresh(g_w, stream_h)                  # wcon
            ev_wcon.record(stream_h)
            halo.refresh(g_in, stream_h)                 # in_field
            ev_in.record(stream_h)
            stream.wait_event(ev_wcon)
            if ev is not None:
                ev[0].record(stream)
            run_vadv(v)
            if ev is not None:
                ev[1].record(stream)
            stream.wait_event(ev_in)
            run_hdiff(h)
            if ev is not None:
                ev[2].record(stream)
            return
        if halo is not None and refresh:
            for g in groups[s]:
                halo.refresh(g, stream)
        if ev is not None:
            ev[0].record(stream)
        run_hdiff(h)
        if ev is not None:
            ev[1].record(stream)
        run_vadv(v)
        if ev is not None:
            ev[2].record(stream)

    sampler = ClockSampler(local) if rank == 0 else None
    with gb.use_stream(stream):
        run_hdiff(hsets[0])
        names["hdiff"] = _lib.last_kernel()              # what `variant = 0` resolved to for these fields
        run_vadv(vsets[0])
        names["vadv"] = _lib.last_kernel()
        for i in range(max(args.warmup, 3)):
            step(i)
        barrier()
        if sampler is not None:  # nvidia-smi needs a moment to start; keep the GPU busy meanwhile
            t_s = time.perf_counter()
            while not sampler.rows and time.perf_counter() - t_s < 3.0:
                step(0, refresh=False)
                stream.synchronize()
        evs = [(gb.Event(), gb.Event(), gb.Event()) for _ in range(args.steps)]
        e0, e1 = gb.Event(), gb.Event()
        barrier()
        t0 = time.perf_counter()
        e0.record(stream)
        for i in range(args.steps):
            step(i, evs[i])
        e1.record(stream)
        stream.synchronize()
        t1 = time.perf_counter()
        barrier()
    total_ms = allmax(e0.elapsed_ms(e1))
    clocks = sampler.stop(t0, t1) if sampler is not None else None
    ms_step = total_ms / args.steps
    value = 2 * n * world / (ms_step * 1e-3)
    t_first = float(np.mean([a.elapsed_ms(b) for a, b, _ in evs]))
    t_second = float(np.mean([b.elapsed_ms(c) for _, b, c in evs]))
    t_h, t_v = (t_second, t_first) if overlap else (t_first, t_second)  # overlap runs vadv first
    peak, peak_src = peaks()

    def roof(name, t_ms, kernel, npts=n):
        ach = npts * ALG_BYTES[name] / (t_ms * 1e-3) / 1e9
        return {"kernel": kernel, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "frac_of_nominal_8TBs": round(ach / 8000.0, 4),
                "traffic": traffic_for(kernel, npts), "ms_per_launch": round(t_ms, 4),
                "alg_bytes_per_launch": npts * ALG_BYTES[name], "peak_source": peak_src}

    kernels = {"hdiff": roof("hdiff", t_h, names["hdiff"]), "vadv": roof("vadv", t_v, names["vadv"])}
    dominant = "vadv" if t_v >= t_h else "hdiff"
    halo_line = halo_leg(halo, groups, stream, gb, barrier, allmax) if halo is not None else None

    # ---- parity of the step (and of the halo exchange) against the single-domain oracle ----------------
    parity = None
    if not args.no_parity:
        t_p = time.perf_counter()
        ok, report = model_parity(gb, storage, hd, va, hsets[0], vsets[0], base, dom, goff, gdims, halo, pgroups, stream,
                                  dstencils=dstencils)
        all_ok = allmin(1.0 if ok else 0.0) == 1.0
        parity = {"ok": all_ok, "rank0_windows": report, "steps": 2, "seconds": round(time.perf_counter() - t_p, 2),
                  "how": "2 model steps with changing exchanged fields, received halo cells poisoned with NaN before "
                         "each refresh; strips along the 4 block edges, 2 corners and an interior patch of every rank "
                         "compared bit-for-bit with the single-domain C oracle (oracle/model_parity.py)"}

    # free the big sets before the side legs
    del evs
    hsets, vsets = hsets[:1], vsets[:1]

    # ---- the other BASELINE configs (N = 1) -----------------------------------------------------------
    configs = None
    if world == 1 and not args.no_configs:
        configs = run_configs(gb, _lib, storage, stencils, stream, roof, args)

    # ---- e2e: host buffers in, host buffers out, every step (public API) ----------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(gb, storage, hd, va, args, barrier, allmax, world, decomp)

    # ---- CPU baseline (rank 0, N == 1 only) ---------------------------------------------------------
    cpu, port = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, port = cpu_baseline_block(args.steps, 2, budget_ref=20.0, budget_port=8.0)

    # launches per step: the two stencils (x their pieces with --split) + the refresh launches
    launches_per_step = 2
    if halo is not None:
        refreshes = 2 * (1 if args.halo_mode != "push2" else 2)  # 'push' may take 2 launches for large faces: upper bound below
        if args.halo_mode == "push":
            refreshes = sum(1 if halo.bytes_received(g) <= halo.FUSED_MAX_BYTES else 2 for g in groups[0])
        pieces = 0
        if args.split:
            for ds in (dstencils[("in", id(hsets[0]["in_field"]))], dstencils[("w", id(vsets[0]["wcon"]))]):
                pieces += (1 + len(ds.strips)) if (ds.interior is not None and ds.strips) else 1
        launches_per_step = (pieces if args.split else 2) + refreshes
    rc = 0
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(dom), "domain_per_gpu": list(dom),
                       "global_domain": [gdims[0], gdims[1], nk],
                       "decomposition": decomposition(world), "field_sets_rotated": args.sets,
                       "l2_policy": f"inputs larger than L2 ({n * 8 / 1e6:.0f} MB per field, {args.sets} sets rotated); "
                                    "no flush",
                       "data_generation": "b200_fill_synthetic on the device (counter-based hash of the global point "
                                          f"index), {setup_s:.1f} s",
                       "halo_mode": args.halo_mode if world > 1 else None,
                       "halo_overlap": ("exchange on its own stream: wcon's exposed, in_field's hidden behind vadv"
                                        if overlap else ("DistributedStencil: refresh beside the interior, boundary "
                                                         "strips after it" if (args.split and halo is not None) else None))},
            "effective_GBs": round(value / 2 * (ALG_BYTES["hdiff"] + ALG_BYTES["vadv"]) / 1e9, 1),
            "roofline": kernels[dominant], "kernels": kernels,
            "parity": parity, "configs": configs,
            "e2e": e2e, "cpu_baseline": cpu, "cpu_port": port, "clocks": clocks,
            "gpu_launches": launches_per_step * args.steps,
        }
        if halo is not None:
            line["halo_parity"] = None if parity is None else parity["ok"]
        if halo_line is not None:
            line["halo_exchange"] = halo_line
            line["config"]["host_affinity"] = affinity
        os.write(result_fd, (json.dumps(line) + "\n").encode())
        if parity is not None and not parity["ok"]:
            sys.stderr.write("bench.py: PARITY FAILED against the single-domain oracle\n")
            rc = 3
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return rc


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures (profiles/):
# {kernel name: {"bytes": per launch, "points": grid points of the captured launch}}; None until measured.
def traffic_for(kernel, npts):
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            table = json.load(fh)
    except Exception:
        return None
    for key in (f"{kernel}@{npts}", kernel):  # one entry per (kernel, size)
        e = table.get(key)
        if isinstance(e, dict) and e.get("points") == npts:
            return e.get("bytes")
    return None


def time_calls(gb, stream, call, reps, warm=3):
    for _ in range(warm):
        call()
    stream.synchronize()
    e0, e1 = gb.Event(), gb.Event()
    e0.record(stream)
    for _ in range(reps):
        call()
    e1.record(stream)
    stream.synchronize()
    return e0.elapsed_ms(e1) / reps


def run_configs(gb, _lib, storage, stencils, stream, roof, args):
    """Driver-visible lines for the other BASELINE configs (single GPU): CUDA-event timed on `stream`, fields far
    larger than the L2 where the config's size allows it, each with a parity check against the oracle."""
    from gt4py_b200 import generated
    from oracle import chain as chain_oracle
    from oracle import cpu as oracle
    from oracle import inputs as gen

    out = {}
    hd, va, tri = stencils.HorizontalDiffusion(), stencils.VerticalAdvectionDycore(), stencils.TridiagonalSolver()
    for cls in (type(hd), type(va), type(tri)):
        cls.configure(device_sync=False)
    with gb.use_stream(stream):
        # configs[2]: hdiff + vadv at 1024x1024x80 (north star: >= 75 % of HBM peak on one GPU), 3 sets rotated
        dom = CONFIG2_DOMAIN
        npts = dom[0] * dom[1] * dom[2]
        sets = []
        h0, v0, _ = make_device_fields(storage, dom, (0, 0), dom[:2], stream)
        sets.append((h0, v0))
        for _ in range(2):
            sets.append(({k: a.copy() for k, a in h0.items()}, {k: a.copy() for k, a in v0.items()}))
        it = [0]

        def call_h():
            h = sets[it[0] % 3][0]
            it[0] += 1
            hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)

        def call_v():
            v = sets[it[0] % 3][1]
            it[0] += 1
            va(v["utens_stage"], v["u_stage"], v["wcon"], v["u_pos"], v["utens"], dtr_stage=gen.DTR_STAGE,
               origin=(0, 0, 0), domain=dom)

        t_h = time_calls(gb, stream, call_h, 60)
        k_h = _lib.last_kernel()
        t_v = time_calls(gb, stream, call_v, 60)
        k_v = _lib.last_kernel()
        out["configs[2] hdiff + vadv 1024x1024x80"] = {
            "ms_per_step": round(t_h + t_v, 4), "grid_points_per_s": 2 * npts / ((t_h + t_v) * 1e-3),
            "hdiff": roof("hdiff", t_h, k_h, npts), "vadv": roof("vadv", t_v, k_v, npts),
            "target": ">= 0.75 of nominal 8 TB/s for both kernels (BASELINE.json north_star)"}
        # a3: the Thomas solver at the same size (sup / rhs are overwritten: refill not needed for timing)
        tf = {k: storage.empty(dom) for k in ("inf", "diag", "sup", "rhs", "out")}
        for k, (seed, lo, hi) in {"inf": (81, -1.0, 1.0), "diag": (82, 3.0, 4.0), "sup": (83, -1.0, 1.0),
                                  "rhs": (84, -1.0, 1.0), "out": (0, 0.0, 0.0)}.items():
            storage.fill_synthetic(tf[k], seed, lo, hi, stream=stream)
        t_t = time_calls(gb, stream, lambda: tri(tf["inf"], tf["diag"], tf["sup"], tf["rhs"], tf["out"], domain=dom), 40)
        out["a3 tridiagonal_solver 1024x1024x80"] = roof("tridiag", t_t, _lib.last_kernel(), npts)
        # parity of the solver on a J strip (columns are independent), fresh inputs
        for k, (seed, lo, hi) in {"sup": (83, -1.0, 1.0), "rhs": (84, -1.0, 1.0)}.items():
            storage.fill_synthetic(tf[k], seed, lo, hi, stream=stream)
        tri(tf["inf"], tf["diag"], tf["sup"], tf["rhs"], tf["out"], domain=dom)
        stream.synchronize()
        js = slice(500, 504)
        hostf = {k: np.asfortranarray(gen.hashed_window(0, 500, dom[0], 4, dom[2], *{"inf": (81, -1.0, 1.0),
                 "diag": (82, 3.0, 4.0), "sup": (83, -1.0, 1.0), "rhs": (84, -1.0, 1.0)}[k])) for k in ("inf", "diag", "sup", "rhs")}
        hostf["out"] = np.zeros_like(hostf["inf"])
        oracle.tridiag(hostf["inf"], hostf["diag"], hostf["sup"], hostf["rhs"], hostf["out"], domain=(dom[0], 4, dom[2]))
        out["a3 tridiagonal_solver 1024x1024x80"]["parity"] = bool(
            np.array_equal(tf["out"][:, js, :].asnumpy(), hostf["out"]))
        del tf, sets, h0, v0

        # configs[3]: the FV3-style chain at 512x512x128, one generated fused kernel (warp-march form)
        cdom, hgt = (512, 512, 128), chain_oracle.HALO
        cn = cdom[0] * cdom[1] * cdom[2]
        st = generated.load_plan(os.path.join(ROOT, "tests", "golden", "generated", "fv3_chain"))
        type(st).configure(device_sync=False)
        csets = []
        for s in range(3):
            f = {}
            for name, seed in (("phi", 91), ("u", 92), ("v", 93)):
                f[name] = storage.empty((cdom[0] + 2 * hgt, cdom[1] + 2 * hgt, cdom[2]), aligned_index=(hgt, hgt, 0))
                storage.fill_synthetic(f[name], seed + 10 * s, -1.0, 1.5, global_offset=(-hgt, -hgt), stream=stream)
            f["out"] = storage.zeros(cdom)
            csets.append(f)
        org = {"_all_": (hgt, hgt, 0), "out": (0, 0, 0)}

        def call_c():
            f = csets[it[0] % 3]
            it[0] += 1
            st(f["phi"], f["u"], f["v"], f["out"], dt=0.05, alpha=0.03, origin=org, domain=cdom)

        t_c = time_calls(gb, stream, call_c, 40)
        entry = "+".join(type(st)._last_entry_points)
        r = roof("chain", t_c, f"generated fv3_chain: {entry}", cn)
        # parity on two levels (the chain is PARALLEL in K; its horizontal regions need the whole IJ plane)
        f = csets[0]
        st(f["phi"], f["u"], f["v"], f["out"], dt=0.05, alpha=0.03, origin=org, domain=cdom)
        stream.synchronize()
        lv = [0, cdom[2] - 1]
        host = {name: gen.hashed_window(-hgt, -hgt, cdom[0] + 2 * hgt, cdom[1] + 2 * hgt, cdom[2], seed, -1.0, 1.5)[:, :, lv]
                for name, seed in (("phi", 91), ("u", 92), ("v", 93))}
        want = chain_oracle.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03)
        r["parity"] = bool(np.array_equal(f["out"].asnumpy()[:, :, lv], want))
        out["configs[3] fv3_chain 512x512x128"] = r
        del csets

        # configs[1] / configs[0]: the small single-stencil cases (launch-latency territory; L2 resident)
        for label, stencil_name, d in (("configs[1] vadv 256x256x80", "vadv", (256, 256, 80)),
                                       ("configs[0] hdiff 128x128x80", "hdiff", (128, 128, 80))):
            h1, v1, _ = make_device_fields(storage, d, (0, 0), d[:2], stream)
            if stencil_name == "vadv":
                t = time_calls(gb, stream, lambda: va(v1["utens_stage"], v1["u_stage"], v1["wcon"], v1["u_pos"],
                                                      v1["utens"], dtr_stage=gen.DTR_STAGE, origin=(0, 0, 0), domain=d), 200)
            else:
                t = time_calls(gb, stream, lambda: hd(h1["in_field"], h1["out_field"], h1["coeff"], origin=(2, 2, 0),
                                                      domain=d), 200)
            r = roof(stencil_name, t, _lib.last_kernel(), d[0] * d[1] * d[2])
            r["note"] = "fields fit the 126 MB L2: back-to-back calls are served from L2, not HBM"
            out[label] = r
    return out


def run_e2e(gb, storage, hd, va, args, barrier, allmax, world, decomp):
    """Host buffers -> device -> kernels -> host buffers, all inside the timed region, through the public
    host-buffer call path (gt4py_b200.HostPipeline: slab-wise H2D / kernel / D2H on three streams), at
    configs[2]'s size per rank (the pinned mirrors of configs[4]'s 21.5 GB per rank would not fit 8 ranks' host)."""
    dom = CONFIG2_DOMAIN
    n = dom[0] * dom[1] * dom[2]
    goff = decomp.global_offset(dom)
    h, v, _ = make_device_fields(storage, dom, goff, (decomp.px_count * dom[0], decomp.py_count * dom[1]))
    gb.synchronize()
    pin_h = {k: storage.pinned_like(a) for k, a in h.items()}
    pin_v = {k: storage.pinned_like(a) for k, a in v.items()}
    for k, a in h.items():
        a.copy_to_host(pin_h[k])
    for k, a in v.items():
        a.copy_to_host(pin_v[k])
    gb.synchronize()
    pipe_h = gb.HostPipeline(hd, h, origin=(2, 2, 0), domain=dom, axis="K", chunks=args.e2e_chunks)
    pipe_v = gb.HostPipeline(va, v, origin=(0, 0, 0), domain=dom, axis="J", chunks=args.e2e_chunks)
    h2d = d2h = 0
    dtr = gen_dtr()

    def step():
        nonlocal h2d, d2h
        a_in, a_out = pipe_h(pin_h)
        b_in, b_out = pipe_v(pin_v, dtr_stage=dtr)
        h2d, d2h = a_in + b_in, a_out + b_out

    steps = max(1, min(args.steps, 10))
    for _ in range(2):
        step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()  # returns when the results are in host memory
    gb.synchronize()
    ms = allmax((time.perf_counter() - t0) * 1e3) / steps
    barrier()
    for p in list(pin_h.values()) + list(pin_v.values()):
        storage.free_pinned(p)
    pcie = (h2d + d2h) / (ms * 1e-3) / 1e9
    return {"value": 2 * n * world / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": ms, "steps": steps,
            "workload": workload_name(dom) + " (host-resident fields; no halo exchange on this leg)",
            "pcie": {"achieved_GBs_both_directions": round(pcie, 1), "h2d_GBs": round(h2d / (ms * 1e-3) / 1e9, 1),
                     "peak_note": "PCIe Gen5 x16: ~55-57 GB/s per direction in practice; the H2D direction (5 of 6 "
                                  "field transfers) is the bound"},
            "path": f"pinned host -> gt4py_b200.HostPipeline ({args.e2e_chunks} slabs: H2D / kernel / D2H on 3 streams) "
                    "-> pinned host; wall clock around the calls, which return with results on the host"}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
