#!/usr/bin/env python
"""bench.py -- the measured hot path: hdiff + vadv on synthetic float64 fields (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--domain NI NJ NK] [--no-e2e] [--no-cpu-baseline]

Workload (config.workload): BASELINE.json configs[2], "hdiff + vadv at 1024x1024x80 float64 on
1xB200 vs HBM roofline" -- per GPU.  One *step* = one horizontal_diffusion call + one
vertical_advection_dycore call over one set of fields (2 kernel launches, 2 * NI*NJ*NK grid-point
updates).  With N > 1 (torchrun, one rank per GPU) every rank owns an NI x NJ x NK block of a
Px x Py IJ decomposition of a global (Px*NI) x (Py*NJ) x NK domain (weak scaling):
each step first refreshes the width-2 halo of in_field and the +I halo column of wcon from the
neighbour ranks over NVLink, then runs the two kernels.

Numbers on the JSON line
  value       grid-point updates/s over all ranks, inputs resident in HBM, CUDA-event timed, max
              over ranks.  Three field sets are rotated (each field is 671 MB >> 126 MB L2).
  roofline    dominant kernel (vadv): algorithmic bytes (48 B/point) / its mean
              launch time from per-launch CUDA events in the same timed region, vs the measured
              HBM copy bandwidth in MEASURED_PEAKS.json.  `kernels` carries the same for hdiff.
  e2e         same metric through the public host-buffer API: every step copies that step's inputs
              from pinned host memory, runs the kernels and copies the results back.
  cpu_baseline  the oracle's C/OpenMP restatement of the reference (kind "port": the reference's
              numpy backend cannot travel to the GPU box) on the box's host cores.

--impl reference times that CPU port alone (all host threads) on the same workload.
The oracle is used here ONLY as the CPU baseline; the measured path is the CUDA library.
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

ALG_BYTES = {"hdiff": 24, "vadv": 48}  # fp64 bytes per grid point
METRIC = "grid-points/s (hdiff+vadv, float64)"
UNIT = "grid-points/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--domain", type=int, nargs=3, default=[1024, 1024, 80],
                    help="per-GPU compute domain NI NJ NK")
    ap.add_argument("--sets", type=int, default=3, help="field sets rotated between steps")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--halo-mode", default="push", choices=["push", "pull"],
                    help="N>1: halo refresh by pushing faces into the neighbours' staging buffers (2 kernels) "
                         "or by pulling boxes out of the neighbours' fields (1 kernel); both over NVLink peer memory")
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
    tag = " (BASELINE configs[2])" if tuple(dom) == (1024, 1024, 80) else ""
    return f"hdiff + vadv at {dom[0]}x{dom[1]}x{dom[2]} float64 per GPU{tag}"


# ---- host-side synthetic inputs in the device layout (I contiguous) --------------------------------
def ifirst(a: np.ndarray) -> np.ndarray:
    """Same logical (I, J, K) array with I-contiguous strides (layout_map (2, 1, 0))."""
    return np.ascontiguousarray(a.transpose(2, 1, 0)).transpose(2, 1, 0)


def make_host_inputs(dom):
    from oracle import inputs as gen  # input recipes only (numpy); no oracle arithmetic

    ni, nj, nk = dom
    h = {k: ifirst(v) for k, v in gen.hdiff_inputs(ni, nj, nk, seed=42).items()}
    v = {k: (ifirst(a) if k != "dtr_stage" else float(a)) for k, a in gen.vadv_inputs(ni, nj, nk, seed=7).items()}
    return h, v


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


# ---- CPU leg (oracle port; the reference's numpy backend cannot travel to the GPU box) ---------------
def cpu_leg(dom, hin, vin, steps, warmup, budget_s=25.0):
    """Time the OpenMP C restatement of the reference on host arrays.  Returns (ms_per_step, steps)."""
    from oracle import cpu as oracle

    def step():
        oracle.hdiff(hin["in_field"], hin["out_field"], hin["coeff"], origin=(2, 2, 0), domain=dom)
        oracle.vadv(vin["utens_stage"], vin["u_stage"], vin["wcon"], vin["u_pos"], vin["utens"],
                    dtr_stage=vin["dtr_stage"], origin=(0, 0, 0), domain=dom)

    t = time.perf_counter()
    step()
    one = time.perf_counter() - t
    for _ in range(max(0, min(warmup - 1, int(budget_s / 4 / max(one, 1e-9))))):
        step()
    steps = max(1, min(steps, int(budget_s / max(one, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    return dt / steps * 1e3, steps


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    os.environ["OMP_NUM_THREADS"] = str(host_threads())  # torchrun pins it to 1; the CPU arm uses every core
    dom = tuple(args.domain)
    n = dom[0] * dom[1] * dom[2]
    hin, vin = make_host_inputs(dom)
    ms, steps = cpu_leg(dom, hin, vin, args.steps, args.warmup, budget_s=90.0)
    value = 2 * n / (ms * 1e-3)
    cores = host_threads()
    sample = (f"{steps} full steps of the workload (hdiff {dom} + vadv {dom}), oracle C/OpenMP port of the "
              f"reference semantics, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(dom), "domain_per_gpu": list(dom), "device": "host cpu"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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


def halo_leg(halo, sets, stream, gb, barrier, allmax, reps=50):
    """The halo exchange alone: `reps` x (wcon refresh + in_field refresh) back to back on one stream, CUDA
    events, max over ranks.  Bytes are what one GPU receives over NVLink per step."""
    nbytes = halo.bytes_received(0) + halo.bytes_received(1)
    for i in range(5):
        halo.refresh(2 * (i % sets) + 1, stream)
        halo.refresh(2 * (i % sets), stream)
    barrier()
    e0, e1 = gb.Event(), gb.Event()
    e0.record(stream)
    for i in range(reps):
        halo.refresh(2 * (i % sets) + 1, stream)
        halo.refresh(2 * (i % sets), stream)
    e1.record(stream)
    stream.synchronize()
    us = allmax(e0.elapsed_ms(e1)) * 1e3 / reps
    barrier()
    gbs = nbytes / us / 1e3
    return {"bytes_received_per_gpu_per_step": int(nbytes), "us_per_step": round(us, 2), "achieved": round(gbs, 1),
            "unit": "GB/s", "peak": 900.0, "frac": round(gbs / 900.0, 4),
            "peak_source": "NVLink 5 nominal, per direction per GPU (B200_PROFILING.md; measured peer copy 770 GB/s)",
            "note": "faces of 1.3-2.6 MB and 2.5 KB corners: launch/flag latency, not link bandwidth, bounds it",
            "launches_counted": (5 + reps) * 2 * halo.launches_per_refresh}


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

    dom = tuple(args.domain)
    ni, nj, nk = dom
    n = ni * nj * nk
    hin, vin = make_host_inputs(dom)

    # device field sets
    hsets, vsets = [], []
    for s in range(args.sets):
        hsets.append({k: storage.from_array(v, aligned_index=(2, 2, 0)) for k, v in hin.items()})
        vsets.append({k: storage.from_array(v) for k, v in vin.items() if k != "dtr_stage"})
    dtr = vin["dtr_stage"]

    hd = stencils.HorizontalDiffusion()
    va = stencils.VerticalAdvectionDycore()
    type(hd).configure(device_sync=False)
    type(va).configure(device_sync=False)

    halo = None
    if world > 1:
        from gt4py_b200 import halo as halo_mod

        halo = halo_mod.HaloExchanger.for_bench(dist, rank, world, local, hsets, vsets, dom, mode=args.halo_mode)

    stream = gb.Stream()
    overlap = halo is not None and not args.no_overlap
    stream_h = gb.Stream() if overlap else None          # halo exchange runs beside the compute stream
    ev_step, ev_wcon, ev_in = (gb.Event(), gb.Event(), gb.Event()) if overlap else (None, None, None)

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
        s = i % args.sets
        h, v = hsets[s], vsets[s]
        if overlap and refresh:
            ev_step.record(stream)
            stream_h.wait_event(ev_step)                 # not before the previous users of these fields
            halo.refresh(2 * s + 1, stream_h)            # wcon
            ev_wcon.record(stream_h)
            halo.refresh(2 * s, stream_h)                # in_field
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
            halo.refresh(2 * s + 1, stream)
            halo.refresh(2 * s, stream)
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

    def roof(name, t_ms, kernel):
        ach = n * ALG_BYTES[name] / (t_ms * 1e-3) / 1e9
        return {"kernel": kernel, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "frac_of_nominal_8TBs": round(ach / 8000.0, 4),
                "traffic": TRAFFIC.get(name), "ms_per_launch": round(t_ms, 4),
                "alg_bytes_per_launch": n * ALG_BYTES[name], "peak_source": peak_src}

    kernels = {"hdiff": roof("hdiff", t_h, "hdiff_tma_kernel"), "vadv": roof("vadv", t_v, "vadv_fused2_kernel")}
    dominant = "vadv" if t_v >= t_h else "hdiff"
    halo_line = halo_leg(halo, args.sets, stream, gb, barrier, allmax) if halo is not None else None

    # ---- e2e: host buffers in, host buffers out, every step (public API) ----------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(gb, storage, hd, va, hsets[0], vsets[0], hin, vin, dom, dtr, args, barrier, allmax, world, halo)

    # ---- CPU baseline (rank 0, N == 1 only) ---------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ms_cpu, steps_cpu = cpu_leg(dom, hin, vin, args.steps, 2, budget_s=20.0)
        cores = host_threads()
        cpu = {"value": 2 * n / (ms_cpu * 1e-3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{steps_cpu} full steps of the same workload with the oracle C/OpenMP port "
                         f"({cores} threads); the reference numpy backend cannot travel to the GPU box",
               "ms_per_step": ms_cpu}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(dom), "domain_per_gpu": list(dom),
                       "decomposition": decomposition(world), "field_sets_rotated": args.sets,
                       "l2_policy": "inputs larger than L2 (671 MB per field, 3 sets rotated); no flush",
                       "halo_mode": args.halo_mode if world > 1 else None,
                       "halo_overlap": ("exchange on its own stream: wcon's exposed, in_field's hidden behind vadv"
                                        if overlap else None)},
            "effective_GBs": round(value / 2 * (ALG_BYTES["hdiff"] + ALG_BYTES["vadv"]) / 1e9, 1),
            "roofline": kernels[dominant], "kernels": kernels,
            "e2e": e2e, "cpu_baseline": cpu, "clocks": clocks,
            "gpu_launches": 2 * args.steps + (2 * halo.launches_per_refresh * args.steps if halo else 0),
        }
        if halo_line is not None:
            line["halo_exchange"] = halo_line
            line["config"]["host_affinity"] = affinity
        os.write(result_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
# (profiles/): filled in once measured, None until then.
TRAFFIC = {"hdiff": None, "vadv": None}
try:
    with open(os.path.join(ROOT, "profiles", "traffic.json")) as _fh:
        TRAFFIC.update({k: v for k, v in json.load(_fh).items() if k in TRAFFIC})
except Exception:
    pass


def decomposition(world):
    return {1: "1x1", 2: "2x1", 4: "2x2", 8: "4x2"}.get(world, f"{world}x1") + " (IxJ ranks)"


def run_e2e(gb, storage, hd, va, h, v, hin, vin, dom, dtr, args, barrier, allmax, world, halo=None):
    """Host buffers -> device -> kernels -> host buffers, all inside the timed region, through the public
    host-buffer call path (gt4py_b200.HostPipeline: slab-wise H2D / kernel / D2H on three streams)."""
    n = dom[0] * dom[1] * dom[2]
    pin_h = {k: storage.pinned_like(h[k], hin[k]) for k in h}
    pin_v = {k: storage.pinned_like(v[k], vin[k]) for k in v}
    pipe_h = gb.HostPipeline(hd, h, origin=(2, 2, 0), domain=dom, axis="K", chunks=args.e2e_chunks)
    pipe_v = gb.HostPipeline(va, v, origin=(0, 0, 0), domain=dom, axis="J", chunks=args.e2e_chunks)
    h2d = d2h = 0

    def step():
        nonlocal h2d, d2h
        if halo is not None:
            # decomposed run: the halo cells of in_field / wcon come from the neighbour GPUs, which need this
            # rank's boundary values on the device first
            h["in_field"].copy_from_host(pin_h["in_field"], pipe_h.s_in)
            v["wcon"].copy_from_host(pin_v["wcon"], pipe_h.s_in)
            halo.refresh(1, pipe_h.s_in)   # groups of field set 0: wcon, in_field
            halo.refresh(0, pipe_h.s_in)
            pipe_h.s_in.synchronize()
            pre_h, pre_v = ("in_field",), ("wcon",)
            extra = pin_h["in_field"].nbytes + pin_v["wcon"].nbytes
        else:
            pre_h = pre_v = ()
            extra = 0
        a_in, a_out = pipe_h(pin_h, preloaded=pre_h)
        b_in, b_out = pipe_v(pin_v, preloaded=pre_v, dtr_stage=dtr)
        h2d, d2h = a_in + b_in + extra, a_out + b_out

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
    return {"value": 2 * n * world / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": ms, "steps": steps,
            "path": f"pinned host -> gt4py_b200.HostPipeline ({args.e2e_chunks} slabs: H2D / kernel / D2H on 3 streams) "
                    "-> pinned host; wall clock around the calls, which return with results on the host"}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
