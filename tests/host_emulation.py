"""TEST INFRASTRUCTURE: run the generated `point` / `column` kernels of a plan ON THE HOST, to check the
OIR -> CUDA lowering (gt4py_b200/lowering.py) against the reference-made golden outputs without a GPU.

The committed `kernel.cu` is compiled as plain C++ with g++: the CUDA qualifiers are defined away, blockIdx /
blockDim / threadIdx become thread-local globals, and a launcher per kernel loops over the grid and the block
sequentially (legal for these two templates: their threads never communicate).  `tile` kernels run through their
plain entry point (no TMA): one CTA at a time, one OS thread per CUDA thread, `__syncthreads()` = a std::barrier,
the dynamic shared memory a static buffer.  Arithmetic is IEEE on both sides
(-ffp-contract=off here, -fmad=false on the device), so the comparison is bit-exact like the GPU tests.

Nothing here is used by the product: gt4py_b200 has no CPU path.
"""
from __future__ import annotations

import ctypes
import hashlib
import json
import os
import re
import subprocess
import tempfile
import threading

import numpy as np

from gt4py_b200 import generated

SHIM = r"""
#include <cstdint>
#include <cmath>
#include <cstdlib>
#include <barrier>
#include <thread>
#include <vector>
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__ __restrict
#define __launch_bounds__(...)
#include <algorithm>
using std::isnan; using std::isinf; using std::isfinite; using std::min; using std::max;
struct emu_dim3 { unsigned x, y, z; };
static thread_local emu_dim3 blockIdx, blockDim, threadIdx, gridDim;
#define __shared__ static
#define __align__(x) alignas(x)
#include <sys/mman.h>
static unsigned char *emu_smem = nullptr;  // the CTA's dynamic shared-memory window, fenced by guard pages
static std::barrier<> *emu_barrier = nullptr;
// Map `bytes` of shared memory between two inaccessible pages; mode 0: the window starts right after the front
// guard, mode 1: it ends (to 16 bytes) right before the back guard -- a read outside the window the plan asked
// for (plan["kernels"][..]["smem_bytes"]) is a segmentation fault here instead of a silent stray read.
static unsigned char *emu_map_window(size_t bytes, int mode, void **base, size_t *total)
{
    const size_t page = 4096, body = (bytes + page - 1) / page * page;
    *total = body + 2 * page;
    unsigned char *p = (unsigned char *)mmap(nullptr, *total, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) abort();
    mprotect(p, page, PROT_NONE);
    mprotect(p + page + body, page, PROT_NONE);
    *base = p;
    return p + page + (mode ? (body - bytes) / 16 * 16 : 0);
}
#define __syncthreads() do { if (emu_barrier) emu_barrier->arrive_and_wait(); } while (0)
#define __syncwarp()
// host forms of the fused column kernels' helpers (lowering.PRELUDE): cp.async is an immediate copy, the value
// stack's Tensor-Memory part a per-thread array.  Threads of these kernels never communicate, so the launcher runs
// them one after the other, each to completion.
#include <cstring>
// every prefetch must read inside one of the arrays the launch was given (registered by host_emulation.run): a
// prefetch that runs past a field's first or last level aborts here instead of silently reading a neighbour
#include <cstdio>
static std::vector<std::pair<const char *, const char *>> emu_ranges;
extern "C" void emu_clear_ranges() { emu_ranges.clear(); }
extern "C" void emu_add_range(const void *lo, size_t bytes) { emu_ranges.emplace_back((const char *)lo, (const char *)lo + bytes); }
static inline void emu_check_src(const void *src, size_t n)
{
    if (emu_ranges.empty()) return;
    for (auto &r : emu_ranges) if ((const char *)src >= r.first && (const char *)src + n <= r.second) return;
    fprintf(stderr, "host emulation: cp.async source %p outside every field of the launch\n", src);
    abort();
}
static inline void b200_cp_async8(void *dst, const void *src, bool valid) { if (valid) { emu_check_src(src, 8); memcpy(dst, src, 8); } else memset(dst, 0, 8); }
static inline void b200_cp_async4(void *dst, const void *src, bool valid) { if (valid) { emu_check_src(src, 4); memcpy(dst, src, 4); } else memset(dst, 0, 4); }
static inline void b200_cp_async16(void *dst, const void *src, bool valid) { if (valid) { emu_check_src(src, 16); memcpy(dst, src, 16); } else memset(dst, 0, 16); }
template <typename T> struct alignas(2 * sizeof(T)) b200_pair { T x, y; };
static inline void b200_cp_commit() {}
template <int N> static inline void b200_cp_wait() {}
static double emu_tmem[1024][256];
struct b200_stack { unsigned tmem; double *smem; double *spill; int n_tmem; int n_smem; int cta; long long nthreads; };
static inline unsigned b200_stack_setup(unsigned *slot, int n_tmem) { return threadIdx.y * blockDim.x + threadIdx.x; }
static inline void b200_stack_teardown(unsigned *) {}
static inline void b200_stack_push(const b200_stack &st, int s, double v)
{
    if (s < st.n_tmem) emu_tmem[st.tmem][s] = v;
    else if (s - st.n_tmem < st.n_smem) st.smem[(s - st.n_tmem) * st.cta] = v;
    else st.spill[(long long)(s - st.n_tmem - st.n_smem) * st.nthreads] = v;
}
static inline double b200_stack_pop(const b200_stack &st, int s)
{
    if (s < st.n_tmem) return emu_tmem[st.tmem][s];
    if (s - st.n_tmem < st.n_smem) return st.smem[(s - st.n_tmem) * st.cta];
    return st.spill[(long long)(s - st.n_tmem - st.n_smem) * st.nthreads];
}
static inline void b200_stack_pop2(const b200_stack &st, int s, double &v0, double &v1) { v0 = b200_stack_pop(st, s); v1 = b200_stack_pop(st, s + 1); }
static inline void b200_stack_push2(const b200_stack &st, int s, double v0, double v1) { b200_stack_push(st, s, v0); b200_stack_push(st, s + 1, v1); }
static inline void b200_stack_fence() {}
// Warp-march kernels: the 32 lanes of a warp are fibers of ONE OS thread, resumed round-robin; a shuffle publishes the
// lane's value and yields, so when the lane is resumed every lane has published (two buffers: a lane may already be
// at shuffle n + 1 while others still read the values of shuffle n).  Warps of a CTA never interact: run one by one.
#include <ucontext.h>
static ucontext_t emu_sched, emu_lane_ctx[32];
static int emu_cur = -1;
static bool emu_yielded = false;
static unsigned emu_shfl_count[32];
static unsigned long long emu_shfl_buf[2][32];
static inline void emu_yield() { const int me = emu_cur; emu_yielded = true; swapcontext(&emu_lane_ctx[me], &emu_sched); }
template <typename T> static inline T b200_shfl(T v, int delta)
{
    const int me = emu_cur;
    const unsigned ph = emu_shfl_count[me]++ & 1u;
    unsigned long long raw = 0;
    memcpy(&raw, &v, sizeof(T));
    emu_shfl_buf[ph][me] = raw;
    emu_yield();
    int src = me + delta;
    if (src < 0 || src > 31) src = me;
    raw = emu_shfl_buf[ph][src];
    T out;
    memcpy(&out, &raw, sizeof(T));
    return out;
}
static void emu_run_warp(void (*lane_fn)(), unsigned warp)
{
    const size_t STK = 512 * 1024;
    static char *stacks = (char *)malloc(32 * STK);
    bool alive[32];
    for (int l = 0; l < 32; ++l) {
        getcontext(&emu_lane_ctx[l]);
        emu_lane_ctx[l].uc_stack.ss_sp = stacks + (size_t)l * STK;
        emu_lane_ctx[l].uc_stack.ss_size = STK;
        emu_lane_ctx[l].uc_link = &emu_sched;
        makecontext(&emu_lane_ctx[l], lane_fn, 0);
        alive[l] = true;
        emu_shfl_count[l] = 0;
    }
    for (bool any = true; any;) {
        any = false;
        for (int l = 0; l < 32; ++l) {
            if (!alive[l]) continue;
            emu_cur = l;
            threadIdx = emu_dim3{warp * 32u + (unsigned)l, 0, 0};
            emu_yielded = false;
            swapcontext(&emu_sched, &emu_lane_ctx[l]);
            if (!emu_yielded) alive[l] = false; else any = true;
        }
    }
    emu_cur = -1;
}
"""

MARCH_LAUNCHER = r"""
static const Args *emu_args_{name};
static void emu_lane_{name}() {{ {name}(*emu_args_{name}); }}
extern "C" void emu_{name}(const Args *a, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz)
{{
    emu_args_{name} = a;
    static unsigned char *window = (unsigned char *)aligned_alloc(128, 256 * 1024);
    emu_smem = window;
    blockDim = emu_dim3{{bx, by, bz}};
    gridDim = emu_dim3{{gx, gy, gz}};
    for (unsigned Z = 0; Z < gz; ++Z) for (unsigned Y = 0; Y < gy; ++Y) for (unsigned X = 0; X < gx; ++X) {{
        blockIdx = emu_dim3{{X, Y, Z}};
        for (unsigned w = 0; w < bx / 32; ++w) emu_run_warp(emu_lane_{name}, w);
    }}
}}
"""

COLFUSE_LAUNCHER = r"""
extern "C" void emu_{name}(const Args *a, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz)
{{
    static unsigned char *window = (unsigned char *)aligned_alloc(128, 256 * 1024);
    emu_smem = window;
    emu_barrier = nullptr;
    blockDim = emu_dim3{{bx, by, bz}};
    gridDim = emu_dim3{{gx, gy, gz}};
    for (unsigned X = 0; X < gx; ++X) {{
        blockIdx = emu_dim3{{X, 0, 0}};
        for (unsigned y = 0; y < by; ++y) for (unsigned x = 0; x < bx; ++x) {{
            threadIdx = emu_dim3{{x, y, 0}};
            {name}(*a);
        }}
    }}
}}
"""

TILE_LAUNCHER = r"""
extern "C" {{ int emu_is_tile_{name} = 1; }}
extern "C" void emu_{name}(const Args *a, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz,
                           unsigned smem_bytes, int mode)
{{
    void *base;
    size_t total;
    emu_smem = emu_map_window(smem_bytes, mode, &base, &total);
    for (unsigned Z = 0; Z < gz; ++Z) for (unsigned Y = 0; Y < gy; ++Y) for (unsigned X = 0; X < gx; ++X) {{
        std::barrier<> bar((std::ptrdiff_t)(bx * by * bz));
        emu_barrier = &bar;
        std::vector<std::thread> pool;
        for (unsigned z = 0; z < bz; ++z) for (unsigned y = 0; y < by; ++y) for (unsigned x = 0; x < bx; ++x)
            pool.emplace_back([=] {{
                blockDim = emu_dim3{{bx, by, bz}};
                blockIdx = emu_dim3{{X, Y, Z}};
                threadIdx = emu_dim3{{x, y, z}};
                {name}(*a);
            }});
        for (auto &t : pool) t.join();
    }}
    munmap(base, total);
}}
"""

LAUNCHER = r"""
extern "C" void emu_{name}(const Args *a, unsigned gx, unsigned gy, unsigned gz, unsigned bx, unsigned by, unsigned bz)
{{
    blockDim = emu_dim3{{bx, by, bz}};
    for (unsigned Z = 0; Z < gz; ++Z) for (unsigned Y = 0; Y < gy; ++Y) for (unsigned X = 0; X < gx; ++X) {{
        blockIdx = emu_dim3{{X, Y, Z}};
        for (unsigned z = 0; z < bz; ++z) for (unsigned y = 0; y < by; ++y) for (unsigned x = 0; x < bx; ++x) {{
            threadIdx = emu_dim3{{x, y, z}};
            {name}(*a);
        }}
    }}
}}
"""


def emulable(plan) -> bool:
    return all(k["kind"] in ("point", "column", "tile", "colfuse", "level", "march") for k in plan["kernels"])


def _host_source(src: str, plan) -> str:
    # drop the TMA plumbing (inline PTX) that only the tile kernels' <k>_t entry points use
    a = src.index("// TMA plumbing")
    b = src.index("struct Args {")
    body = src[:a] + src[b:]
    body = body.replace("#include <cstdint>\n#include <cmath>\n", "", 1)
    # the <k>_t entry points (TMA-staged inputs) are device-only
    body = re.sub(r"__device__ __forceinline__ void (\w+)_tma_impl\(const Args &a\).*?\n"
                  r"extern \"C\" __global__ void __launch_bounds__\([\d, ]+\) \1_t\(const Args a\)[^\n]*\n?", "", body,
                  flags=re.S)
    body = re.sub(r"extern __shared__ __align__\(\d+\) unsigned char b200_smem\[\];",
                  "unsigned char *b200_smem = emu_smem;", body)
    out = SHIM + body
    for k in plan["kernels"]:  # the generic-stride entry point: host arrays keep numpy's strides
        out += {"tile": TILE_LAUNCHER, "colfuse": COLFUSE_LAUNCHER, "march": MARCH_LAUNCHER}.get(k["kind"], LAUNCHER).format(name=k["name"])
    return out


def build(plan_dir: str):
    plan, lib = build_lib(plan_dir)
    return plan, ctypes.CDLL(lib)


def build_lib(plan_dir: str):
    """(plan, path of the host shared library made from the plan's kernel.cu); cached by content."""
    with open(os.path.join(plan_dir, "plan.json")) as fh:
        plan = json.load(fh)
    with open(os.path.join(plan_dir, plan.get("source", "kernel.cu"))) as fh:
        text = _host_source(fh.read(), plan)
    cache = os.path.join(tempfile.gettempdir(), "gt4py_b200_host_emulation")
    os.makedirs(cache, exist_ok=True)
    tag = hashlib.sha256(text.encode()).hexdigest()[:16]
    lib = os.path.join(cache, f"emu_{tag}.so")
    if not os.path.exists(lib):
        uniq = f"{os.getpid()}_{threading.get_ident()}"  # two plans with identical kernels may build concurrently
        cpp = os.path.join(cache, f"emu_{tag}_{uniq}.cpp")
        tmp = os.path.join(cache, f"emu_{tag}_{uniq}.so")
        with open(cpp, "w") as fh:
            fh.write(text)
        cmd = ["g++", "-std=c++20", "-O1", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-pthread", "-w", cpp,
               "-o", tmp]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"g++ failed on the host form of {plan_dir}:\n{r.stderr[:4000]}")
        os.replace(tmp, lib)
        os.remove(cpp)
    return plan, lib


def run(plan_dir: str, fields: dict, params: dict, origins: dict, domain, window_mode: int = 0):
    """fields: name -> numpy array (modified in place) or None; origins: name -> index of the compute origin.
    window_mode: where the tile kernels' shared-memory window sits between its guard pages (0 front, 1 back)."""
    plan, lib = build(plan_dir)
    Args = generated._args_struct(plan)
    a = Args()
    ni, nj, nk = (int(x) for x in domain)
    a.ni, a.nj, a.nk = ni, nj, nk
    keep = []
    for spec in plan["fields"]:
        name = spec["name"]
        arr = fields.get(name)
        if arr is None:
            setattr(a, name, None)
            continue
        item = arr.dtype.itemsize
        strides, ptr, pos = [0, 0, 0], arr.ctypes.data, 0
        for d in range(3):
            if spec["dims"][d]:
                strides[d] = arr.strides[pos] // item
                ptr += int(origins[name][pos]) * arr.strides[pos]
                pos += 1
        strides += [arr.strides[pos + n] // item for n in range(len(spec["data_dims"]))]
        setattr(a, name, ptr)
        getattr(a, name + "_s")[:] = strides
    for t in plan["temporaries"]:  # same shapes / origins as generated.GeneratedStencil._temporaries
        (li, hi), (lj, hj) = t["pad"]
        shape = [ni + li + hi if t["dims"][0] else 1, nj + lj + hj if t["dims"][1] else 1,
                 nk if t["dims"][2] else 1] + list(t["data_dims"])
        arr = np.zeros(shape, dtype=np.dtype(t["dtype"]), order="F")
        keep.append(arr)
        item = arr.dtype.itemsize
        strides = [s // item for s in arr.strides]
        for d in range(3):
            if not t["dims"][d]:
                strides[d] = 0
        ptr = arr.ctypes.data + (li * arr.strides[0] if t["dims"][0] else 0) + (lj * arr.strides[1] if t["dims"][1] else 0)
        setattr(a, t["name"], ptr)
        getattr(a, t["name"] + "_s")[:] = strides
    for s in plan["scalars"]:
        setattr(a, s["name"], params[s["name"]])
    if any(k["kind"] in ("colfuse", "march") for k in plan["kernels"]):  # bounds of every array, for the prefetch check
        lib.emu_clear_ranges()
        lib.emu_add_range.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
        for arr in list(fields.values()) + keep:
            if arr is not None:
                base = arr.base if isinstance(arr.base, np.ndarray) else arr
                lo = min(arr.ctypes.data, base.ctypes.data)
                lib.emu_add_range(lo, max(base.nbytes, arr.nbytes))
    kernels = plan["kernels"]
    for n, k in enumerate(kernels):  # launch geometry of generated.GeneratedStencil._launch
        wi = ni + k["i_ext"][1] - k["i_ext"][0]
        wj = nj + k["j_ext"][1] - k["j_ext"][0]
        if k["kind"] == "level":  # a sequential loop run level by level: for k: for stage: all ij
            if n and kernels[n - 1]["kind"] == "level" and kernels[n - 1]["loop"] == k["loop"]:
                continue
            group = [g for g in kernels[n:] if g["kind"] == "level" and g["loop"] == k["loop"]]
            for lvl in generated.level_order(nk, k["backward"]):
                for g in group:
                    gi = ni + g["i_ext"][1] - g["i_ext"][0]
                    gj = nj + g["j_ext"][1] - g["j_ext"][0]
                    if gi <= 0 or gj <= 0 or not generated.level_in(g["intervals"], lvl, nk):
                        continue
                    rows = 4 * int(g.get("rows_per_thread", 1))
                    a.lvl = lvl
                    fn = getattr(lib, "emu_" + g["name"])
                    fn.restype = None
                    fn.argtypes = [ctypes.c_void_p] + [ctypes.c_uint] * 6
                    fn(ctypes.byref(a), (gi + 63) // 64, (gj + rows - 1) // rows, 1, 64, 4, 1)
            continue
        if wi <= 0 or wj <= 0:
            continue
        if k["kind"] == "point":
            rows = 4 * int(k.get("rows_per_thread", 1))
            block, grid = (64, 4, 1), ((wi + 63) // 64, (wj + rows - 1) // rows, nk)
            if nk <= 0:
                continue
        elif k["kind"] == "tile":
            ti, tj = k["tile"]
            block, grid = (*k["block"], 1), ((ni + ti - 1) // ti, (nj + tj - 1) // tj, nk)
            if nk <= 0:
                continue
        elif k["kind"] == "march":
            per_cta = k["result_columns"] * k["warps"]
            block, grid = (32 * k["warps"], 1, 1), ((ni + per_cta - 1) // per_cta, (nj + k["strip"] - 1) // k["strip"], nk)
            if nk <= 0:
                continue
        elif k["kind"] == "colfuse":  # persistent: fewer CTAs than tiles, so that the tile loop is exercised
            bx, by = k["block"]
            ntiles = ((wi + bx - 1) // bx) * ((wj + by - 1) // by)
            ctas = max(1, min(ntiles, 3))
            slot = bx * by * 8
            need = max(0, nk * k["stack_values_per_level"] - k.get("stack_tmem", 128))
            n_smem = max(0, min((need + 1) // 2, (250 * 1024 - k["ring_slots"] * slot) // slot))  # some in smem, the rest spilled
            spill = np.zeros(max(1, (need - n_smem) * ctas * bx * by), dtype=np.float64)
            keep.append(spill)
            setattr(a, f"cf_{k['name']}_smem", n_smem)
            setattr(a, f"cf_{k['name']}_spill", spill.ctypes.data)
            block, grid = (bx, by, 1), (ctas, 1, 1)
        else:
            block, grid = (32, 4, 1), ((wi + 31) // 32, (wj + 3) // 4, 1)
        fn = getattr(lib, "emu_" + k["name"])
        fn.restype = None
        if k["kind"] == "tile":
            fn.argtypes = [ctypes.c_void_p] + [ctypes.c_uint] * 7 + [ctypes.c_int]
            fn(ctypes.byref(a), *grid, *block, int(k["smem_bytes"]), int(window_mode))
        else:
            fn.argtypes = [ctypes.c_void_p] + [ctypes.c_uint] * 6
            fn(ctypes.byref(a), *grid, *block)
    return fields
