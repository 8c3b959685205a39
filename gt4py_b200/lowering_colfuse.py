"""Fused FORWARD / BACKWARD column kernels -- the generic form of csrc/vadv_fused.cu's machine mapping.

A run of consecutive sequential vertical loops whose cross-loop dependencies stay inside a column (vadv: Thomas
forward elimination + back substitution; the tridiagonal solver; column physics) is lowered to ONE persistent
kernel instead of one thread-per-column kernel per loop with the intermediate fields in HBM:

* a CTA of 256 threads walks 256-column tiles; a thread owns a column and runs all loops of the run on it;
* fields that carry values from level to level and from loop to loop -- the ones the reference's pipeline
  annotates ``KCache(fill, flush)`` (gtc/oir.py:310-320, passes/oir_optimizations/caches.py:44-253; GridTools
  realises them as ring buffers, gtcpp_codegen.py:227-232) -- live in registers (current and previous level) and,
  between loops, on a per-thread VALUE STACK: Tensor Memory for the first 128 values of a thread, shared memory
  next, an L2-resident scratch beyond (helpers in lowering.PRELUDE).  Temporaries that never leave the run cost no
  DRAM byte; API fields written in one loop and read in the next are stored once and never re-read;
* read-only inputs are prefetched `LEAD` levels ahead by cp.async into a per-thread shared-memory ring (a thread
  reads only what it copied itself: no barrier), which replaces occupancy as the latency-hiding mechanism -- with
  ~256 columns per SM in flight the stack stays on chip.

Semantics are those of the column kernels of lowering.py (sections in program order, one level at a time).
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Set, Tuple

from gt4py.cartesian.gtc import common, oir

# threads of a CTA along I x J (one column each) and the levels the prefetch runs ahead; tuning hooks for A/B builds
BX, BY = (int(x) for x in os.environ.get("B200_COLFUSE_BLOCK", "128x2").split("x"))
LEAD = int(os.environ.get("B200_COLFUSE_LEAD", "6"))
# two-loop runs (FORWARD then BACKWARD over all levels, or the reverse): the second loop of the previous tile runs
# interleaved, level by level, with the first loop of the next tile -- two independent dependency chains per thread,
# the schedule of the hand-written vadv.  OFF by default, by measurement (profiles/r2_colfuse_pipelined.jsonl, B200,
# 1024x1024x80): vadv 0.898 ms sequential / 0.996 pipelined / 0.912 pipelined with the unguarded middle part, Thomas
# solver 0.903 / 1.000 / 0.928 -- the generated kernels are bound by instruction issue (every thread issues its own
# cp.async copies), not by the latency the second chain would hide; kept for parts where that differs, with a test.
PIPELINE = os.environ.get("B200_COLFUSE_PIPELINE", "0") != "0"
MIDPATH = os.environ.get("B200_COLFUSE_MIDPATH", "1") != "0"
# third entry point <k>_t: prefetch rings fed by TMA tile loads issued by a producer lane (see _emit_tma).  OFF by default,
# by measurement (B200, 1024x1024x80, profiles/r2_colfuse_tma.jsonl): bit-exact, but vadv 1.25 ms against 0.90 ms with the
# per-thread cp.async rings, Thomas solver 1.01 against 0.90, column_physics 0.64-0.69 against 0.59, with 8 or 16 slots.
# (First form, thread 0 of the consumers issuing from inside the level loop: 2.04 / 1.42 / 1.00 ms -- the rings starved.)
# ncu on the producer-row form: the ring is not starved (9 % of the stall samples on try_wait results), but a level's
# barrier test + arrive cost the instructions its five cp.async did (538 M against 511 M warp instructions) and the
# try_wait round trip is on every warp's critical path once per level: 550 cycles per level and CTA against 392.
TMA_RINGS = os.environ.get("B200_COLFUSE_TMA", "0") != "0"
assert (BX * BY) % 128 == 0 and BX * BY <= 1024, "a fused column CTA is a whole number of 4-warp groups"


def tmem_values(threads: int) -> int:
    """Doubles per thread in Tensor Memory: the threads / 128 warps of a lane quarter split its 512 columns."""
    return (512 // (threads // 128)) // 2 // 2 * 2


def _walk(stmts, cond=False, restricted=False):
    """(kind, FieldAccess, conditional, restricted) in evaluation order; kind 'r' / 'w'."""
    for st in stmts:
        if isinstance(st, oir.AssignStmt):
            for n in st.right.walk_values():
                if isinstance(n, oir.FieldAccess):
                    yield "r", n, cond, restricted
            if isinstance(st.left, oir.FieldAccess):
                subs = list(st.left.data_index)
                if isinstance(st.left.offset, oir.VariableKOffset):
                    subs.append(st.left.offset.k)
                for sub in subs:
                    for n in sub.walk_values():
                        if isinstance(n, oir.FieldAccess):
                            yield "r", n, cond, restricted
                yield "w", st.left, cond, restricted
        elif isinstance(st, oir.MaskStmt):
            for n in st.mask.walk_values():
                if isinstance(n, oir.FieldAccess):
                    yield "r", n, cond, restricted
            yield from _walk(st.body, True, restricted)
        elif isinstance(st, oir.While):
            for n in st.cond.walk_values():
                if isinstance(n, oir.FieldAccess):
                    yield "r", n, True, restricted
            yield from _walk(st.body, True, restricted)
        elif isinstance(st, oir.HorizontalRestriction):
            yield from _walk(st.body, True, True)
        elif hasattr(st, "body"):
            yield from _walk(st.body, cond, restricted)


def _sections(unit):
    """[(interval, [hexec, ...])] in program order."""
    out = []
    for interval, h in unit:
        if out and out[-1][0] is interval:
            out[-1][1].append(h)
        else:
            out.append((interval, [h]))
    return out


def _same_bound(a: common.AxisBound, b: common.AxisBound) -> bool:
    return a.level == b.level and int(a.offset) == int(b.offset)


def _contiguous(sections, backward: bool) -> bool:
    for (a, _), (b, _) in zip(sections, sections[1:]):
        if backward and not _same_bound(b.end, a.start):
            return False
        if not backward and not _same_bound(a.end, b.start):
            return False
    return True


def find_runs(stencil: oir.Stencil, units_by_loop) -> List[List[int]]:
    """Maximal runs (>= 2 loops) of consecutive sequential vertical loops that can share one kernel: every loop is a
    single column kernel, its sections tile a contiguous K range in order, and every field written inside the run is
    only ever touched at horizontal offset (0, 0) inside the run."""
    ok = []
    for n_loop, loop, units in units_by_loop:
        seq = loop.loop_order != common.LoopOrder.PARALLEL
        good = seq and len(units) == 1 and not units[0][1]
        if good:
            secs = _sections(units[0][0])
            good = _contiguous(secs, loop.loop_order == common.LoopOrder.BACKWARD)
        ok.append(good)
    runs, cur = [], []
    for n, good in enumerate(ok):
        if good:
            cur.append(n)
        else:
            if len(cur) >= 2:
                runs.append(cur)
            cur = []
    if len(cur) >= 2:
        runs.append(cur)
    out = []
    for run in runs:
        while len(run) >= 2:  # shrink from the end until the column-locality condition holds
            written, offset = set(), set()
            for n in run:
                for _, h in units_by_loop[n][2][0][0]:
                    for kind, acc, _, _ in _walk(h.body):
                        name = str(acc.name)
                        if kind == "w":
                            written.add(name)
                        off = acc.offset
                        if isinstance(off, common.CartesianOffset) and (off.i or off.j):
                            offset.add(name)
            if written & offset:
                run = run[:-1]
            else:
                break
        if len(run) >= 2:
            out.append(run)
    return out


def _field(fields, name):
    return next(v for k, v in fields.items() if str(k) == name)


class Analysis:
    """Which fields of a run are cached in registers / pushed / popped / loaded / prefetched, per loop."""

    def __init__(self, stencil: oir.Stencil, run_loops, fields, accessed_outside: Set[str]):
        self.loops = run_loops            # [(loop, sections, backward)]
        temps = {str(t.name) for t in stencil.declarations}
        nl = len(run_loops)
        acc: List[List[List[Tuple]]] = []  # accesses per loop / section
        for loop, secs, _ in run_loops:
            acc.append([[a for h in hs for a in _walk(h.body)] for _, hs in secs])
        written_in = [set(str(a.name) for sec in la for k, a, _, _ in sec if k == "w") for la in acc]
        touched_in = [set(str(a.name) for sec in la for _, a, _, _ in sec) for la in acc]
        self.written = set().union(*written_in)
        # ---- cached fields: float64 / float32 IJK fields written in the run, touched only at (0, 0, 0) and (0, 0, prev) ------
        cached: Dict[str, Dict[str, Any]] = {}
        for name in sorted(self.written):
            f = _field(fields, name)
            if not all(f["dims"]) or f["data_dims"] or f["dtype"] not in (common.DataType.FLOAT64, common.DataType.FLOAT32):
                continue  # (a float32 value crosses the value stack as the double it converts to exactly)
            good = True
            for n, (loop, secs, backward) in enumerate(run_loops):
                prev = 1 if backward else -1
                reads_prev = False
                for s_idx, sec in enumerate(acc[n]):
                    for kind, a, _, _ in sec:
                        if str(a.name) != name:
                            continue
                        off = a.offset
                        if not isinstance(off, common.CartesianOffset) or off.i or off.j or a.data_index:
                            good = False
                        elif kind == "w" and off.k != 0:
                            good = False
                        elif kind == "r" and off.k not in (0, prev):
                            good = False
                        elif kind == "r" and off.k == prev:
                            reads_prev = True
                            if s_idx == 0:
                                good = False  # the loop's first level would read a level the loop never visits
                if reads_prev and not all(any(str(a.name) == name for _, a, _, _ in sec) for sec in acc[n]):
                    good = False
            if good:
                cached[name] = {"store": name not in temps or name in accessed_outside}
        self.cached = cached

        def first_is_write(sec, name):
            for kind, a, cond, _ in sec:
                if str(a.name) != name or int(a.offset.k) != 0:
                    continue
                return kind == "w" and not cond
            return False

        needs_initial = [{name: not all(first_is_write(sec, name) for sec in acc[n])
                          for name in cached if name in touched_in[n]} for n in range(nl)]
        self.pop = [{name for name, need in needs_initial[n].items()
                     if need and any(name in written_in[m] for m in range(n))} for n in range(nl)]
        self.load = [{name for name, need in needs_initial[n].items()
                      if need and name not in self.pop[n] and (name not in temps or name in accessed_outside)}
                     for n in range(nl)]
        self.push = [{name for name in cached if name in written_in[n]
                      and any(name in self.pop[m] for m in range(n + 1, nl))} for n in range(nl)]
        self.pushed = sorted(set().union(*self.push)) if self.push else []
        self.reads_prev = [{str(a.name) for sec in acc[n] for k, a, _, _ in sec
                            if k == "r" and str(a.name) in cached and int(a.offset.k) != 0} for n in range(nl)]
        self.touched_in = touched_in
        # ---- prefetch ring keys per loop: (field, di, dj) -> {"dk": offsets, "uses": [(section index, hexec)]} ------
        self.ring: List[Dict[Tuple[str, int, int], Dict[str, Any]]] = []
        ring_dtypes = (common.DataType.FLOAT64, common.DataType.FLOAT32, common.DataType.INT64, common.DataType.INT32)
        for n, (loop, secs, backward) in enumerate(run_loops):
            keys: Dict[Tuple[str, int, int], Dict[str, Any]] = {}
            for s_idx, (_, hs) in enumerate(secs):
                for h in hs:
                    for kind, a, _, restricted in _walk(h.body):
                        name = str(a.name)
                        if kind != "r" or name in cached or name in self.written:
                            continue
                        f = _field(fields, name)
                        if (not isinstance(a.offset, common.CartesianOffset) or a.data_index or not all(f["dims"])
                                or f["data_dims"] or f["dtype"] not in ring_dtypes):
                            continue
                        key = (name, int(a.offset.i), int(a.offset.j))
                        e = keys.setdefault(key, {"dk": set(), "uses": [], "free": True, "dk_by_sec": {}})
                        e["dk"].add(int(a.offset.k))
                        e["dk_by_sec"].setdefault(s_idx, set()).add(int(a.offset.k))
                        if restricted:  # read inside horizontal(region[...]): its halo only exists inside the region
                            e["free"] = False
                        if not any(x == s_idx and y is h for x, y in e["uses"]):
                            e["uses"].append((s_idx, h))
            for name in self.load[n]:  # the level-start value of a cached field nobody in the run wrote before
                e = keys.setdefault((name, 0, 0), {"dk": set(), "uses": [], "free": True, "dk_by_sec": {}})
                e["dk"].add(0)
                for s_idx in range(len(secs)):
                    e["dk_by_sec"].setdefault(s_idx, set()).add(0)
                e["uses"] = [(s_idx, h) for s_idx, (_, hs) in enumerate(secs) for h in hs]
            self.ring.append({k: v for k, v in keys.items() if v["free"] and v["uses"]})


def pow2ceil(n: int) -> int:
    p = 1
    while p < n:
        p *= 2
    return p


def emit(kname: str, stencil: oir.Stencil, run: List[int], units_by_loop, block_ext, fields, scalar_names,
         accessed_outside: Set[str], L) -> Tuple[str, Dict[str, Any]]:
    """CUDA source + launch metadata of the fused kernel for the vertical loops `run`.  `L` is the lowering module
    (ExprGen, CTYPES, _interval_cond)."""
    run_loops = []
    for n in run:
        _, loop, units = units_by_loop[n]
        run_loops.append((loop, _sections(units[0][0]), loop.loop_order == common.LoopOrder.BACKWARD))
    an = Analysis(stencil, run_loops, fields, accessed_outside)
    if PIPELINE and _pipelinable(run_loops):
        return _emit_pipelined(kname, run, run_loops, an, block_ext, fields, scalar_names, L)
    all_h = [h for _, secs, _ in run_loops for _, hs in secs for h in hs]
    exts = [block_ext[id(h)] for h in all_h]
    ilo, ihi = min(e[0][0] for e in exts), max(e[0][1] for e in exts)
    jlo, jhi = min(e[1][0] for e in exts), max(e[1][1] for e in exts)
    locals_ = {d.name for h in all_h for d in h.declarations}
    gen = L.ExprGen(fields, scalar_names, locals_, True)
    never_written = {str(f) for f in fields} - an.written
    read_anywhere = set().union(*an.touched_in)
    gen.readonly = {n for n in never_written if n in read_anywhere}
    gen.cf_cached = an.cached
    NP = len(an.pushed)
    pslot = {name: p for p, name in enumerate(an.pushed)}
    T, NT = BX * BY, tmem_values(BX * BY)

    # ---- ring layout: slots of 256 x 8 bytes; per loop the keys are laid out from slot 0 (loops reuse the region) ----
    ring_slots = 0
    ring_of_loop = []
    for keys in an.ring:
        off, table = 0, {}
        for key in sorted(keys):
            dks = keys[key]["dk"]
            depth = pow2ceil(LEAD + (max(dks) - min(dks)) + 1)
            table[key] = {"slot": off, "depth": depth, "dkmin": min(dks), "dkmax": max(dks), **keys[key]}
            off += depth
        ring_of_loop.append(table)
        ring_slots = max(ring_slots, off)
    span_max = max([t["dkmax"] - t["dkmin"] for table in ring_of_loop for t in table.values()], default=0)

    def ct_of(name):
        return L.CTYPES[_field(fields, name)["dtype"]]

    size_of = {"double": 8, "float": 4, "int64_t": 8, "int32_t": 4}

    def guard_text(h):
        e = block_ext[id(h)]
        return f"i >= {e[0][0]} && i < a.ni + ({e[0][1]}) && j >= {e[1][0]} && j < a.nj + ({e[1][1]})"

    guards: Dict[int, str] = {}
    by_text: Dict[str, str] = {}
    G = []  # guard declarations (per tile)
    for h in all_h:
        text = guard_text(h)
        if text not in by_text:
            by_text[text] = f"g{len(by_text)}"
            G.append(f"        const bool {by_text[text]} = live && {text};")
        guards[id(h)] = by_text[text]

    def index_at(name, di, dj, level):
        f = _field(fields, name)
        terms = []
        if f["dims"][0]:
            terms.append(f"(idx_t)(i + ({di})) * S0(a.{name}_s[0])")
        if f["dims"][1]:
            terms.append(f"(idx_t)(j + ({dj})) * (idx_t)a.{name}_s[1]")
        if f["dims"][2]:
            terms.append(f"(idx_t)({level}) * (idx_t)a.{name}_s[2]")
        return " + ".join(terms) if terms else "0"

    def fold(fn, items):
        items = sorted(set(items))
        text = items[0]
        for x in items[1:]:
            text = f"{fn}({text}, {x})"
        return text

    S = []  # kernel body
    S.append("    typedef typename b200_idx<U>::type idx_t;")
    for name in (str(f) for f in fields if str(f) in gen.readonly):
        S.append(f"    const {ct_of(name)} *__restrict__ g_{name} = a.{name};")
    S.append("    extern __shared__ __align__(16) unsigned char b200_smem[];")
    S.append("    __shared__ unsigned cf_tmem_slot;")
    S.append("    const int tid = (int)(threadIdx.y * blockDim.x + threadIdx.x);")
    S.append("    double *const cf_ring = reinterpret_cast<double *>(b200_smem) + tid;")
    S.append("    b200_stack cf_stk;")
    S.append(f"    cf_stk.n_tmem = {NT};")
    S.append(f"    cf_stk.cta = {T};")
    S.append(f"    cf_stk.tmem = b200_stack_setup(&cf_tmem_slot, {NT});")
    S.append(f"    cf_stk.smem = reinterpret_cast<double *>(b200_smem) + {ring_slots} * {T} + tid;")
    S.append(f"    cf_stk.n_smem = a.cf_{kname}_smem;")
    S.append(f"    cf_stk.nthreads = (long long)gridDim.x * {T};")
    S.append(f"    cf_stk.spill = a.cf_{kname}_spill + (long long)blockIdx.x * {T} + tid;")
    S.append(f"    const int cf_tiles_i = (a.ni + ({ihi - ilo}) + {BX - 1}) / {BX};")
    S.append(f"    const int cf_tiles = cf_tiles_i * ((a.nj + ({jhi - jlo}) + {BY - 1}) / {BY});")
    S.append("    for (int cf_tile = (int)blockIdx.x; cf_tile < cf_tiles; cf_tile += (int)gridDim.x) {")
    S.append(f"        const int i = ({ilo}) + (cf_tile % cf_tiles_i) * {BX} + (int)threadIdx.x;")
    S.append(f"        const int j = ({jlo}) + (cf_tile / cf_tiles_i) * {BY} + (int)threadIdx.y;")
    S.append(f"        const bool live = i < a.ni + ({ihi}) && j < a.nj + ({jhi});")
    S += G
    for name in sorted(an.cached):
        S.append(f"        {ct_of(name)} cur_{name} = 0, prev_{name} = 0;")

    for n, (loop, secs, backward) in enumerate(run_loops):
        step = -1 if backward else 1
        table = ring_of_loop[n]
        gen.cf_prev = 1 if backward else -1
        gen.cf_ring = {}
        S.append(f"        {{  // ---- vertical loop {run[n]}: {'BACKWARD' if backward else 'FORWARD'}")
        issue = []  # the prefetch of one level, per ring key
        order = sorted(table)
        for q, key in enumerate(order):
            t = table[key]
            name, di, dj = key
            ct = ct_of(name)
            per = 8 // size_of[ct]
            S.append(f"            {ct} *const rg{q} = reinterpret_cast<{ct} *>(cf_ring + {t['slot']} * {T});  // {name}[{di},{dj}]")
            gen.cf_ring[key] = (f"rg{q}", t["depth"] - 1, T * per)
            far = t["dkmax"] if step > 0 else t["dkmin"]
            use_guards = sorted({guards[id(h)] for _, h in t["uses"]})
            lows, highs = [], []
            # levels any statement reads: per section its own K offsets (a section that reads only k and k+1 must not
            # make the prefetch touch k-1 below its first level); the hull of the pieces lies inside the field
            for s_idx, dks in sorted(t["dk_by_sec"].items()):
                lo, hi = L._interval_cond(secs[s_idx][0])
                lows.append(f"({lo}) + ({min(dks)})")
                highs.append(f"({hi}) + ({max(dks)})")
            S.append(f"            const int lo{q} = {fold('min', lows)}, hi{q} = {fold('max', highs)};")
            fn = "b200_cp_async8" if size_of[ct] == 8 else "b200_cp_async4"
            issue.append(f"{{ const int lv = kv + ({step * LEAD + far}); const bool ok = ({' || '.join(use_guards)}) && "
                         f"lv >= lo{q} && lv < hi{q}; {fn}(rg{q} + (lv & {t['depth'] - 1}) * {T * per}, "
                         f"a.{name} + (ok ? {index_at(name, di, dj, 'lv')} : 0), ok); }}")
        lo0, hi0 = L._interval_cond(secs[0][0])
        first = f"({hi0}) - 1" if backward else f"({lo0})"
        if issue:
            S.append("            auto cf_issue = [&](const int kv) {  // copy one more level of every input into the ring")
            S += ["                " + x for x in issue]
            S.append("                b200_cp_commit();")
            S.append("            };")
            S.append(f"            for (int t = {LEAD + span_max}; t >= 1; --t) cf_issue({first} - ({step}) * t);  // prologue")
        pops, loads, pushes = sorted(an.pop[n]), sorted(an.load[n]), sorted(an.push[n])

        def paired(todo, fmt2, fmt1):
            out, todo = [], list(todo)
            while todo:
                a_name = todo.pop(0)
                pa = pslot[a_name]
                if (todo and pslot[todo[0]] == pa + 1 and pa % 2 == 0 and NP % 2 == 0
                        and ct_of(a_name) == "double" and ct_of(todo[0]) == "double"):
                    out.append(fmt2.format(s=f"k * {NP} + {pa}", a=a_name, b=todo.pop(0)))
                else:
                    out.append(fmt1.format(s=f"k * {NP} + {pa}", a=a_name))
            return out

        for interval, hs in secs:
            lo, hi = L._interval_cond(interval)
            if backward:
                S.append(f"            for (int k = ({hi}) - 1; k >= ({lo}); --k) {{")
            else:
                S.append(f"            for (int k = ({lo}); k < ({hi}); ++k) {{")
            if issue:
                S.append(f"                cf_issue(k); b200_cp_wait<{LEAD}>();")
            if pops:
                S.append("                __syncwarp();")
                S += paired(pops, "                b200_stack_pop2(cf_stk, {s}, cur_{a}, cur_{b});",
                            "                cur_{a} = b200_stack_pop(cf_stk, {s});")
            for name in loads:
                q = order.index((name, 0, 0))
                S.append(f"                cur_{name} = rg{q}[(k & {table[(name, 0, 0)]['depth'] - 1}) * {T * (8 // size_of[ct_of(name)])}];")
            for h in hs:
                S.append(f"                if ({guards[id(h)]}) {{")
                for d in L._declarations(h):
                    S.append(" " * 20 + f"{L.CTYPES[d.dtype]} {d.name} = 0;")
                S += gen.stmts(h.body, 20)
                S.append("                }")
            if pushes:
                S.append("                __syncwarp();")
                S += paired(pushes, "                b200_stack_push2(cf_stk, {s}, cur_{a}, cur_{b});",
                            "                b200_stack_push(cf_stk, {s}, cur_{a});")
            for name in sorted(an.reads_prev[n]):
                S.append(f"                prev_{name} = cur_{name};")
            S.append("            }")
        if issue:
            S.append("            b200_cp_wait<0>();")
        if pushes:
            S.append("            b200_stack_fence();")
        S.append("        }")
    S.append("    }")
    S.append("    b200_stack_teardown(&cf_tmem_slot);")
    gen.cf_ring, gen.cf_cached = {}, {}

    src = [f"template <bool U> __device__ __forceinline__ void {kname}_impl(const Args &a)", "{"] + S + ["}"]
    for suffix, flag in (("", "false"), ("_u", "true")):
        src.append(f'extern "C" __global__ void __launch_bounds__({BX * BY}, 1) {kname}{suffix}(const Args a) '
                   f"{{ {kname}_impl<{flag}>(a); }}")
    meta = {"name": kname, "kind": "colfuse", "block": [BX, BY], "i_ext": [int(ilo), int(ihi)], "j_ext": [int(jlo), int(jhi)],
            "ring_slots": int(ring_slots), "stack_values_per_level": NP, "stack_tmem": NT, "loops": [int(x) for x in run],
            "cached": sorted(an.cached), "pushed": list(an.pushed),
            "private": sorted(nm for nm, c in an.cached.items() if not c["store"])}
    if TMA_RINGS:
        made = _emit_tma(kname, run_loops, an, block_ext, fields, scalar_names, L)
        if made is not None:
            src.append(made[0])
            meta["tma"] = made[1]
    return "\n".join(src), meta


def _emit_tma(kname, run_loops, an, block_ext, fields, scalar_names, L):
    """Third entry point `<k>_t` of a fused column kernel: the prefetch rings are fed by TMA tile loads (one box per
    field and level, completion on an mbarrier per ring slot) instead of one cp.async per thread, field and level -- the
    copies are a quarter of the instructions of a level in the `_u` form.  Same persistent-tile / value-stack structure
    for the BX x BY consumer threads; the CTA is launched with ONE MORE ROW of threads (block BX x (BY + 1)), whose
    first lane is the producer: it walks the same tiles, loops and levels as the consumers and issues the boxes of the
    CTA's whole level stream, position after position, held back only by the slots' `empty` barriers -- so the ring
    stays full across loop and tile boundaries (the `_u` form drains and refills it at each).  Every consumer waits on
    a slot's `full` barrier before it reads a level; one lane per warp arrives on its `empty` barrier when the warp is
    done with it.  Levels and columns outside a field are zero-filled by the hardware (never read: extent analysis).
    Returns (source, tma meta) or None."""
    size_of = {"double": 8, "float": 4, "int64_t": 8, "int32_t": 4}

    def ct_of(name):
        return L.CTYPES[_field(fields, name)["dtype"]]

    if not any(an.ring):
        return None
    all_h = [h for _, secs, _ in run_loops for _, hs in secs for h in hs]
    exts = [block_ext[id(h)] for h in all_h]
    ilo, ihi = min(e[0][0] for e in exts), max(e[0][1] for e in exts)
    jlo, jhi = min(e[1][0] for e in exts), max(e[1][1] for e in exts)
    locals_ = {d.name for h in all_h for d in h.declarations}
    gen = L.ExprGen(fields, scalar_names, locals_, True)
    never_written = {str(f) for f in fields} - an.written
    read_anywhere = set().union(*an.touched_in)
    gen.readonly = {n for n in never_written if n in read_anywhere}
    gen.cf_cached = an.cached
    NP = len(an.pushed)
    pslot = {name: p for p, name in enumerate(an.pushed)}
    T, NT = BX * BY, tmem_values(BX * BY)
    # one box per field, the same for every loop of the run (one tensor map per field): the hull of its offsets
    hull: Dict[str, List[int]] = {}
    for keys in an.ring:
        for (name, di, dj) in keys:
            if ct_of(name) not in size_of:
                return None
            h_ = hull.setdefault(name, [di, di, dj, dj])
            h_[0], h_[1], h_[2], h_[3] = min(h_[0], di), max(h_[1], di), min(h_[2], dj), max(h_[3], dj)
    box, foff, slot_bytes = {}, {}, 0
    for name in sorted(hull):
        item = size_of[ct_of(name)]
        per16 = 16 // item
        w = (BX + hull[name][1] - hull[name][0] + (per16 - 1) + per16 - 1) // per16 * per16
        hgt = BY + hull[name][3] - hull[name][2]
        if w > 256 or hgt > 256:
            return None
        box[name] = (w, hgt, per16)
        foff[name] = slot_bytes
        slot_bytes += (w * hgt * item + 127) // 128 * 128
    # ring slots (a slot holds one level of EVERY field here): a power of two, at least the K span of the reads + 2
    SLOTS = int(os.environ.get("B200_COLFUSE_TMA_SLOTS", "8"))
    # per loop: the fields it streams, the K offsets, the ring depth
    loops = []
    ring_bytes = 0
    for n, (loop, secs, backward) in enumerate(run_loops):
        keys = an.ring[n]
        if not keys:
            loops.append(None)
            continue
        dkmin = min(min(v["dk"]) for v in keys.values())
        dkmax = max(max(v["dk"]) for v in keys.values())
        span = dkmax - dkmin
        depth = max(pow2ceil(span + 2), pow2ceil(SLOTS))
        if depth > 16:
            return None
        names = sorted({k_[0] for k_ in keys})
        loops.append({"dkmin": dkmin, "dkmax": dkmax, "span": span, "depth": depth, "fields": names,
                      "tx": sum(box[nm][0] * box[nm][1] * size_of[ct_of(nm)] for nm in names)})
    # ONE ring depth for every loop of the kernel: the barriers' phases follow the CTA-wide stream position cf_mb, so
    # position -> (slot, parity) must be the same function in every loop
    depth_all = max(lp["depth"] for lp in loops if lp is not None)
    for lp in loops:
        if lp is not None:
            lp["depth"] = depth_all
    ring_bytes = depth_all * slot_bytes

    def guard_text(h):
        e = block_ext[id(h)]
        return f"i >= {e[0][0]} && i < a.ni + ({e[0][1]}) && j >= {e[1][0]} && j < a.nj + ({e[1][1]})"

    guards: Dict[int, str] = {}
    by_text: Dict[str, str] = {}
    G = []
    for h in all_h:
        text = guard_text(h)
        if text not in by_text:
            by_text[text] = f"g{len(by_text)}"
            G.append(f"        const bool {by_text[text]} = live && {text};")
        guards[id(h)] = by_text[text]

    S = []
    S.append("    constexpr bool U = true;")
    S.append("    typedef int idx_t;")
    for name in (str(f) for f in fields if str(f) in gen.readonly):
        S.append(f"    const {ct_of(name)} *__restrict__ g_{name} = a.{name}; (void)g_{name};")
    S.append("    extern __shared__ __align__(1024) unsigned char b200_smem[];")
    S.append("    __shared__ unsigned cf_tmem_slot;")
    S.append("    __shared__ __align__(8) unsigned long long cf_full[16], cf_empty[16];")
    S.append("    const int tid = (int)(threadIdx.y * blockDim.x + threadIdx.x);")
    S.append(f"    if (tid == 0) {{ for (int s = 0; s < 16; ++s) {{ b200_mbar_init(&cf_full[s], 1); b200_mbar_init(&cf_empty[s], {T // 32}); }} }}")
    S.append("    b200_stack cf_stk;")
    S.append(f"    cf_stk.n_tmem = {NT};")
    S.append(f"    cf_stk.cta = {T};")
    S.append(f"    cf_stk.tmem = b200_stack_setup(&cf_tmem_slot, {NT});  // (its __syncthreads() also publishes the barriers)")
    S.append(f"    cf_stk.smem = reinterpret_cast<double *>(b200_smem + {ring_bytes}) + tid;")
    S.append(f"    cf_stk.n_smem = a.cf_{kname}_smem;")
    S.append(f"    cf_stk.nthreads = (long long)gridDim.x * {T};")
    S.append(f"    cf_stk.spill = a.cf_{kname}_spill + (long long)blockIdx.x * {T} + tid;")
    S.append(f"    const int cf_tiles_i = (a.ni + ({ihi - ilo}) + {BX - 1}) / {BX};")
    S.append(f"    const int cf_tiles = cf_tiles_i * ((a.nj + ({jhi - jlo}) + {BY - 1}) / {BY});")
    tile_head = ["    for (int cf_tile = (int)blockIdx.x; cf_tile < cf_tiles; cf_tile += (int)gridDim.x) {",
                 f"        const int ti0 = ({ilo}) + (cf_tile % cf_tiles_i) * {BX}, tj0 = ({jlo}) + (cf_tile / cf_tiles_i) * {BY};"]
    for name in sorted(hull):  # the box starts on a 16-byte boundary of global memory (see lowering._tile_kernel)
        tile_head.append(f"        const int adj_{name} = (a.org_{name}[0] + ti0 + ({hull[name][0]})) & {box[name][2] - 1};")
    # ---- the producer: thread (0, BY) of the block
    P = ["    unsigned cf_u = 0;  // position in the CTA's stream of ring slots"] + tile_head
    for n, (loop, secs, backward) in enumerate(run_loops):
        lp = loops[n]
        if lp is None:
            continue
        step = -1 if backward else 1
        lo_all = L._interval_cond(secs[-1][0] if backward else secs[0][0])[0]
        hi_all = L._interval_cond(secs[0][0] if backward else secs[-1][0])[1]
        D = lp["depth"]
        start = f"(({lo_all}) + ({lp['dkmin']}))" if not backward else f"(({hi_all}) - 1 + ({lp['dkmax']}))"
        P.append(f"        {{  // ---- vertical loop {n}: position m of its stream holds level cf_start {'-' if backward else '+'} m")
        P.append(f"            const int cf_start = {start}, cf_npos = ({hi_all}) - ({lo_all}) + {lp['span']};")
        P.append("            for (int m = 0; m < cf_npos; ++m, ++cf_u) {")
        P.append(f"                const unsigned s = cf_u & {D - 1}u;")
        P.append(f"                b200_mbar_wait(&cf_empty[s], ((cf_u / {D}u) & 1u) ^ 1u);")
        P.append(f"                b200_mbar_expect_tx(&cf_full[s], {lp['tx']});")
        P.append(f"                const int lv = cf_start + ({step}) * m;")
        for name in lp["fields"]:
            P.append(f"                b200_tma_load_3d(b200_smem + s * {slot_bytes} + {foff[name]}, a.tm_{kname}_{name}, "
                     f"a.org_{name}[0] + ti0 + ({hull[name][0]}) - adj_{name}, a.org_{name}[1] + tj0 + ({hull[name][2]}), "
                     f"a.org_{name}[2] + lv, &cf_full[s]);")
        P.append("            }")
        P.append("        }")
    P.append("    }")
    # ---- the consumers
    head_at = len(S)
    S.append("    unsigned cf_mb = 0;  // position of the next loop's first level in the CTA's stream of ring slots")
    S += tile_head
    S.append("        const int i = ti0 + (int)threadIdx.x, j = tj0 + (int)threadIdx.y;")
    S.append(f"        const bool live = i < a.ni + ({ihi}) && j < a.nj + ({jhi});")
    S += G
    for name in sorted(an.cached):
        S.append(f"        {ct_of(name)} cur_{name} = 0, prev_{name} = 0;")
    for n, (loop, secs, backward) in enumerate(run_loops):
        step = -1 if backward else 1
        lp = loops[n]
        gen.cf_prev = 1 if backward else -1
        gen.cf_ring = {}
        S.append(f"        {{  // ---- vertical loop {n}: {'BACKWARD' if backward else 'FORWARD'}")
        lo_all = L._interval_cond(secs[-1][0] if backward else secs[0][0])[0]
        hi_all = L._interval_cond(secs[0][0] if backward else secs[-1][0])[1]
        if lp is not None:
            D, span = lp["depth"], lp["span"]
            # the loop's level stream: position m holds level `start + step * m`, m = 0 .. nlev + span - 1
            start = f"(({lo_all}) + ({lp['dkmin']}))" if not backward else f"(({hi_all}) - 1 + ({lp['dkmax']}))"
            S.append(f"            const int cf_start = {start}, cf_nlev = ({hi_all}) - ({lo_all}), cf_npos = cf_nlev + {span};")
            for q, key in enumerate(sorted(an.ring[n])):
                name, di, dj = key
                ct = ct_of(name)
                w, hgt, per16 = box[name]
                S.append(f"            const {ct} *const rq{q} = reinterpret_cast<const {ct} *>(b200_smem + {foff[name]}) + "
                         f"((int)threadIdx.y + ({dj - hull[name][2]})) * {w} + ((int)threadIdx.x + ({di - hull[name][0]}) + adj_{name});"
                         f"  // {name}[{di},{dj}]")
                gen.cf_ring[key] = ((f"rq{q}", "cf_start", step), D - 1, slot_bytes // size_of[ct])
            S.append("            auto cf_wait = [&](const int m) {  // every thread: position m has landed")
            S.append(f"                const unsigned u = cf_mb + (unsigned)m; b200_mbar_wait(&cf_full[u & {D - 1}u], (u / {D}u) & 1u);")
            S.append("            };")
            S.append("            auto cf_done = [&](const int m) {  // this warp no longer reads position m")
            S.append(f"                __syncwarp(); if ((tid & 31) == 0) b200_mbar_arrive(&cf_empty[(cf_mb + (unsigned)m) & {D - 1}u]);")
            S.append("            };")
            S.append(f"            for (int m = 0; m < {span}; ++m) cf_wait(m);")
        pops, loads, pushes = sorted(an.pop[n]), sorted(an.load[n]), sorted(an.push[n])

        def paired(todo, fmt2, fmt1):
            out, todo = [], list(todo)
            while todo:
                a_name = todo.pop(0)
                pa = pslot[a_name]
                if (todo and pslot[todo[0]] == pa + 1 and pa % 2 == 0 and NP % 2 == 0
                        and ct_of(a_name) == "double" and ct_of(todo[0]) == "double"):
                    out.append(fmt2.format(s=f"k * {NP} + {pa}", a=a_name, b=todo.pop(0)))
                else:
                    out.append(fmt1.format(s=f"k * {NP} + {pa}", a=a_name))
            return out

        first = f"(({hi_all}) - 1)" if backward else f"({lo_all})"
        for interval, hs in secs:
            lo, hi = L._interval_cond(interval)
            if backward:
                S.append(f"            for (int k = ({hi}) - 1; k >= ({lo}); --k) {{")
            else:
                S.append(f"            for (int k = ({lo}); k < ({hi}); ++k) {{")
            if lp is not None:
                S.append(f"                const int cf_t = (k - {first}) * ({step});  // the level's number in the loop")
                S.append(f"                cf_wait(cf_t + {lp['span']});")
            if pops:
                S.append("                __syncwarp();")
                S += paired(pops, "                b200_stack_pop2(cf_stk, {s}, cur_{a}, cur_{b});",
                            "                cur_{a} = b200_stack_pop(cf_stk, {s});")
            for name in loads:
                (rq, start_, step_), mask_, stride_ = gen.cf_ring[(name, 0, 0)]  # (the cached name itself reads cur_<f>)
                S.append(f"                cur_{name} = {rq}[((cf_mb + ((k) - ({start_})) * ({step_})) & {mask_}) * {stride_}];")
            for h in hs:
                S.append(f"                if ({guards[id(h)]}) {{")
                for d in L._declarations(h):
                    S.append(" " * 20 + f"{L.CTYPES[d.dtype]} {d.name} = 0;")
                S += gen.stmts(h.body, 20)
                S.append("                }")
            if pushes:
                S.append("                __syncwarp();")
                S += paired(pushes, "                b200_stack_push2(cf_stk, {s}, cur_{a}, cur_{b});",
                            "                b200_stack_push(cf_stk, {s}, cur_{a});")
            for name in sorted(an.reads_prev[n]):
                S.append(f"                prev_{name} = cur_{name};")
            if lp is not None:
                S.append("                cf_done(cf_t);")
            S.append("            }")
        if lp is not None:
            S.append(f"            for (int m = 0; m < {lp['span']}; ++m) cf_done(cf_nlev + m);  // the positions only read ahead")
            S.append("            cf_mb += (unsigned)cf_npos;")
        if pushes:
            S.append("            b200_stack_fence();")
        S.append("        }")
    S.append("    }")
    body = S[head_at:]
    S = S[:head_at]
    S.append(f"    if (tid >= {T}) {{  // the extra row of threads: its first lane feeds the rings")
    S.append(f"        if (tid == {T}) {{")
    S += ["    " * 2 + line for line in P]
    S.append("        }")
    S.append("    } else {")
    S += ["    " + line for line in body]
    S.append("    }")
    S.append("    b200_stack_teardown(&cf_tmem_slot);")
    gen.cf_ring, gen.cf_cached = {}, {}
    src = [f"__device__ __forceinline__ void {kname}_tma_impl(const Args &a)", "{"] + S + ["}",
           f'extern "C" __global__ void __launch_bounds__({BX * (BY + 1)}, 1) {kname}_t(const Args a) {{ {kname}_tma_impl(a); }}']
    tma = {"smem_bytes": int(ring_bytes), "producer_rows": 1,
           "fields": {nm: [int(box[nm][0]), int(box[nm][1]), 1] for nm in sorted(hull)}}
    return "\n".join(src), tma


def _full_range(secs, backward: bool) -> bool:
    """Do the sections of a loop tile [START, END) exactly, in loop order?"""
    first, last = secs[0][0], secs[-1][0]
    lo, hi = (last.start, first.end) if backward else (first.start, last.end)
    return (lo.level == common.LevelMarker.START and int(lo.offset) == 0
            and hi.level == common.LevelMarker.END and int(hi.offset) == 0 and _contiguous(secs, backward))


def _section_weight(interval) -> int:
    """Levels of a section for a deep column (to pick a loop's longest section)."""
    nk = 1000

    def at(b):
        return (nk if b.level == common.LevelMarker.END else 0) + int(b.offset)

    return at(interval.end) - at(interval.start)


def _pipelinable(run_loops) -> bool:
    if len(run_loops) != 2:
        return False
    (_, secs_a, back_a), (_, secs_b, back_b) = run_loops
    return back_a != back_b and _full_range(secs_a, back_a) and _full_range(secs_b, back_b)


def _emit_pipelined(kname, run, run_loops, an, block_ext, fields, scalar_names, L):
    """The fused kernel of a two-loop run with the machine schedule of the hand-written vadv (csrc/vadv_fused.cu): per
    CTA, round r runs loop B (the second loop) of tile r - 1 and loop A (the first loop) of tile r level by level in ONE
    loop -- two independent dependency chains per thread, which is what hides the latency of the division chain with
    8 resident warps.  The value stack changes direction from tile to tile: the slot loop B pops at an iteration is
    the slot loop A refills in the same iteration (slot of level k: k for even tiles, nk - 1 - k for odd ones).  Each
    loop has its own prefetch ring and its own register set (nothing crosses from A to B except through the stack)."""
    all_h = [h for _, secs, _ in run_loops for _, hs in secs for h in hs]
    exts = [block_ext[id(h)] for h in all_h]
    ilo, ihi = min(e[0][0] for e in exts), max(e[0][1] for e in exts)
    jlo, jhi = min(e[1][0] for e in exts), max(e[1][1] for e in exts)
    locals_ = {d.name for h in all_h for d in h.declarations}
    gen = L.ExprGen(fields, scalar_names, locals_, True)
    never_written = {str(f) for f in fields} - an.written
    read_anywhere = set().union(*an.touched_in)
    gen.readonly = {n for n in never_written if n in read_anywhere}
    gen.cf_cached = an.cached
    NP = len(an.pushed)
    pslot = {name: p for p, name in enumerate(an.pushed)}
    T, NT = BX * BY, tmem_values(BX * BY)
    size_of = {"double": 8, "float": 4, "int64_t": 8, "int32_t": 4}

    def ct_of(name):
        return L.CTYPES[_field(fields, name)["dtype"]]

    # ring layout: both loops are live at once, so their rings lie side by side
    ring_of_loop, off = [], 0
    for keys in an.ring:
        table = {}
        for key in sorted(keys):
            dks = keys[key]["dk"]
            depth = pow2ceil(LEAD + (max(dks) - min(dks)) + 1)
            table[key] = {"slot": off, "depth": depth, "dkmin": min(dks), "dkmax": max(dks), **keys[key]}
            off += depth
        ring_of_loop.append(table)
    ring_slots = off

    def guard_text(h):
        e = block_ext[id(h)]
        return f"i >= {e[0][0]} && i < a.ni + ({e[0][1]}) && j >= {e[1][0]} && j < a.nj + ({e[1][1]})"

    def index_at(name, di, dj, level):
        f = _field(fields, name)
        terms = []
        if f["dims"][0]:
            terms.append(f"(idx_t)(i + ({di})) * S0(a.{name}_s[0])")
        if f["dims"][1]:
            terms.append(f"(idx_t)(j + ({dj})) * (idx_t)a.{name}_s[1]")
        if f["dims"][2]:
            terms.append(f"(idx_t)({level}) * (idx_t)a.{name}_s[2]")
        return " + ".join(terms) if terms else "0"

    def fold(fn, items):
        items = sorted(set(items))
        text = items[0]
        for x in items[1:]:
            text = f"{fn}({text}, {x})"
        return text

    S = []
    S.append("    typedef typename b200_idx<U>::type idx_t;")
    for name in (str(f) for f in fields if str(f) in gen.readonly):
        S.append(f"    const {ct_of(name)} *__restrict__ g_{name} = a.{name};")
    S.append("    extern __shared__ __align__(16) unsigned char b200_smem[];")
    S.append("    __shared__ unsigned cf_tmem_slot;")
    S.append("    const int tid = (int)(threadIdx.y * blockDim.x + threadIdx.x);")
    S.append("    double *const cf_ring = reinterpret_cast<double *>(b200_smem) + tid;")
    S.append("    b200_stack cf_stk;")
    S.append(f"    cf_stk.n_tmem = {NT};")
    S.append(f"    cf_stk.cta = {T};")
    S.append(f"    cf_stk.tmem = b200_stack_setup(&cf_tmem_slot, {NT});")
    S.append(f"    cf_stk.smem = reinterpret_cast<double *>(b200_smem) + {ring_slots} * {T} + tid;")
    S.append(f"    cf_stk.n_smem = a.cf_{kname}_smem;")
    S.append(f"    cf_stk.nthreads = (long long)gridDim.x * {T};")
    S.append(f"    cf_stk.spill = a.cf_{kname}_spill + (long long)blockIdx.x * {T} + tid;")
    S.append(f"    const int cf_tiles_i = (a.ni + ({ihi - ilo}) + {BX - 1}) / {BX};")
    S.append(f"    const int cf_tiles = cf_tiles_i * ((a.nj + ({jhi - jlo}) + {BY - 1}) / {BY});")
    S.append("    const int cf_mine = (int)blockIdx.x < cf_tiles ? (cf_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;")
    S.append("    const int nk = a.nk;")
    # per-loop contexts: tile coordinates, guards, register sets
    guard_names: List[Dict[int, str]] = []
    for n, (loop, secs, backward) in enumerate(run_loops):
        X = "AB"[n]
        S.append(f"    int i{X} = 0, j{X} = 0;  // the tile loop {X} works on")
        by_text: Dict[str, str] = {}
        names: Dict[int, str] = {}
        for _, hs in secs:
            for h in hs:
                text = guard_text(h)
                if text not in by_text:
                    by_text[text] = f"g{X}{len(by_text)}"
                    S.append(f"    bool {by_text[text]} = false;")
                names[id(h)] = by_text[text]
        guard_names.append(names)
        for name in sorted(an.cached):
            if name in an.touched_in[n]:
                S.append(f"    {ct_of(name)} cur{X}_{name} = 0, prev{X}_{name} = 0;")

    set_tile = []  # text that (re)computes a loop's tile context from a tile number
    for n, (loop, secs, backward) in enumerate(run_loops):
        X = "AB"[n]
        lines = [f"i{X} = ({ilo}) + (tile % cf_tiles_i) * {BX} + (int)threadIdx.x; "
                 f"j{X} = ({jlo}) + (tile / cf_tiles_i) * {BY} + (int)threadIdx.y;",
                 f"{{ const int i = i{X}, j = j{X}; const bool live = i < a.ni + ({ihi}) && j < a.nj + ({jhi});"]
        seen = set()
        for _, hs in secs:
            for h in hs:
                g = guard_names[n][id(h)]
                if g not in seen:
                    seen.add(g)
                    lines.append(f"  {g} = live && {guard_text(h)};")
        lines.append("}")
        for name in sorted(an.cached):
            if name in an.touched_in[n]:
                lines.append(f"cur{X}_{name} = 0; prev{X}_{name} = 0;")
        set_tile.append(lines)

    level_fn, issue_fn, prologue = [], [], []
    mid_range: List[Tuple[str, str]] = []
    inside: List[str] = []
    for n, (loop, secs, backward) in enumerate(run_loops):
        X = "AB"[n]
        step = -1 if backward else 1
        table = ring_of_loop[n]
        gen.cf_prev = 1 if backward else -1
        gen.cf_ring = {}
        order = sorted(table)
        issue = []
        span = max([t["dkmax"] - t["dkmin"] for t in table.values()], default=0)
        for q, key in enumerate(order):
            t = table[key]
            name, di, dj = key
            ct = ct_of(name)
            per = 8 // size_of[ct]
            S.append(f"    {ct} *const rg{X}{q} = reinterpret_cast<{ct} *>(cf_ring + {t['slot']} * {T});  // loop {X}: {name}[{di},{dj}]")
            gen.cf_ring[key] = (f"rg{X}{q}", t["depth"] - 1, T * per)
            far = t["dkmax"] if step > 0 else t["dkmin"]
            use_guards = sorted({guard_names[n][id(h)] for _, h in t["uses"]})
            lows, highs = [], []
            for s_idx, dks in sorted(t["dk_by_sec"].items()):
                lo, hi = L._interval_cond(secs[s_idx][0])
                lows.append(f"({lo}) + ({min(dks)})")
                highs.append(f"({hi}) + ({max(dks)})")
            S.append(f"    const int lo{X}{q} = {fold('min', lows)}, hi{X}{q} = {fold('max', highs)};")
            fn = "b200_cp_async8" if size_of[ct] == 8 else "b200_cp_async4"
            issue.append(f"{{ const int lv = kv + ({step * LEAD + far}); const bool ok = ({' || '.join(use_guards)}) && "
                         f"lv >= lo{X}{q} && lv < hi{X}{q}; {fn}(rg{X}{q} + (lv & {t['depth'] - 1}) * {T * per}, "
                         f"a.{name} + (ok ? {index_at(name, di, dj, 'lv')} : 0), ok); }}")
        # the copies of one more level of loop X's inputs (no commit: the caller commits once per iteration)
        F = [f"    auto cf_issue{X} = [&](const int kv) {{", f"        const int i = i{X}, j = j{X}; (void)i; (void)j;"]
        F += ["        " + x for x in issue] + ["    };"]
        issue_fn.append((F, bool(issue)))
        first = "nk - 1" if backward else "0"
        prologue.append(f"for (int t = {LEAD + span}; t >= 1; --t) {{ cf_issue{X}({first} - ({step}) * t); b200_cp_commit(); }}"
                        if issue else "")
        pops, loads, pushes = sorted(an.pop[n]), sorted(an.load[n]), sorted(an.push[n])

        def paired(todo, fmt2, fmt1):
            out, todo = [], list(todo)
            while todo:
                a_name = todo.pop(0)
                pa = pslot[a_name]
                if (todo and pslot[todo[0]] == pa + 1 and pa % 2 == 0 and NP % 2 == 0
                        and ct_of(a_name) == "double" and ct_of(todo[0]) == "double"):
                    out.append(fmt2.format(s=f"sl * {NP} + {pa}", a=a_name, b=todo.pop(0)))
                else:
                    out.append(fmt1.format(s=f"sl * {NP} + {pa}", a=a_name))
            return out

        # one level of loop X: k = the level, sl = its stack slot (the direction alternates from tile to tile)
        F = [f"    auto cf_level{X} = [&](const int k, const int sl) {{",
             f"        const int i = i{X}, j = j{X}; (void)i; (void)j; (void)sl;"]
        for name in sorted(an.cached):
            if name in an.touched_in[n]:
                F.append(f"        {ct_of(name)} &cur_{name} = cur{X}_{name}, &prev_{name} = prev{X}_{name}; (void)prev_{name};")
        for g in sorted(set(guard_names[n].values())):
            F.append(f"        const bool {g[:1] + g[2:]} = {g};")  # gA0 -> g0 (the name the bodies were written with)
        if pops:
            F.append("        __syncwarp();")
            F += paired(pops, "        b200_stack_pop2(cf_stk, {s}, cur_{a}, cur_{b});",
                        "        cur_{a} = b200_stack_pop(cf_stk, {s});")
        for name in loads:
            q = order.index((name, 0, 0))
            F.append(f"        cur_{name} = rg{X}{q}[(k & {table[(name, 0, 0)]['depth'] - 1}) * {T * (8 // size_of[ct_of(name)])}];")
        for s_idx, (interval, hs) in enumerate(secs):
            lo, hi = L._interval_cond(interval)
            F.append(f"        {'if' if s_idx == 0 else 'else if'} (k >= ({lo}) && k < ({hi})) {{")
            for h in hs:
                g = guard_names[n][id(h)]
                F.append(f"            if ({g[:1] + g[2:]}) {{")
                for d in L._declarations(h):
                    F.append(" " * 16 + f"{L.CTYPES[d.dtype]} {d.name} = 0;")
                F += gen.stmts(h.body, 16)
                F.append("            }")
            F.append("        }")
        if pushes:
            F.append("        __syncwarp();")
            F += paired(pushes, "        b200_stack_push2(cf_stk, {s}, cur_{a}, cur_{b});",
                        "        b200_stack_push(cf_stk, {s}, cur_{a});")
        for name in sorted(an.reads_prev[n]):
            F.append(f"        prev_{name} = cur_{name};")
        F.append("    };")
        level_fn.append(F)
        # the loop's longest section once more, without section dispatch and without guards: for iterations where the
        # tile is entirely inside the domain -- straight-line code that the compiler can interleave with the other loop's
        mid = max(range(len(secs)), key=lambda q: _section_weight(secs[q][0]))
        interval, hs = secs[mid]
        lo, hi = L._interval_cond(interval)
        mid_range.append((f"nk - ({hi})", f"nk - ({lo})") if backward else (lo, hi))  # in iteration space
        M = [f"    auto cf_mid{X} = [&](const int k, const int sl) {{",
             f"        const int i = i{X}, j = j{X}; (void)i; (void)j; (void)sl;"]
        for name in sorted(an.cached):
            if name in an.touched_in[n]:
                M.append(f"        {ct_of(name)} &cur_{name} = cur{X}_{name}, &prev_{name} = prev{X}_{name}; (void)prev_{name};")
        if pops:
            M.append("        __syncwarp();")
            M += paired(pops, "        b200_stack_pop2(cf_stk, {s}, cur_{a}, cur_{b});",
                        "        cur_{a} = b200_stack_pop(cf_stk, {s});")
        for name in loads:
            q = order.index((name, 0, 0))
            M.append(f"        cur_{name} = rg{X}{q}[(k & {table[(name, 0, 0)]['depth'] - 1}) * {T * (8 // size_of[ct_of(name)])}];")
        for h in hs:
            M.append("        {")
            for d in L._declarations(h):
                M.append(" " * 12 + f"{L.CTYPES[d.dtype]} {d.name} = 0;")
            M += gen.stmts(h.body, 12)
            M.append("        }")
        if pushes:
            M.append("        __syncwarp();")
            M += paired(pushes, "        b200_stack_push2(cf_stk, {s}, cur_{a}, cur_{b});",
                        "        b200_stack_push(cf_stk, {s}, cur_{a});")
        for name in sorted(an.reads_prev[n]):
            M.append(f"        prev_{name} = cur_{name};")
        M.append("    };")
        level_fn.append(M)
        # "every guard of the loop holds for every thread of the CTA": the tile lies inside the domain
        es = [block_ext[id(h)] for _, hs_ in secs for h in hs_]
        inside.append(f"(i{X} - (int)threadIdx.x) >= ({max(e[0][0] for e in es)}) && "
                      f"(i{X} - (int)threadIdx.x) + {BX} <= a.ni + ({min(e[0][1] for e in es)}) && "
                      f"(j{X} - (int)threadIdx.y) >= ({max(e[1][0] for e in es)}) && "
                      f"(j{X} - (int)threadIdx.y) + {BY} <= a.nj + ({min(e[1][1] for e in es)})")
    for F, _ in issue_fn:
        S += F
    for F in level_fn:
        S += F
    backA, backB = run_loops[0][2], run_loops[1][2]
    kA = "nk - 1 - it" if backA else "it"
    kB = "nk - 1 - it" if backB else "it"
    has_issue = [x[1] for x in issue_fn]

    def round_loop(doA: bool, doB: bool) -> List[str]:
        if doA and doB and MIDPATH:
            # iterations in which both loops are in their longest section, in a CTA whose two tiles lie inside the
            # domain: the unguarded bodies; the iterations before and after them (and every other CTA): the general form
            R = [f"            const bool fast = ({inside[0]}) && ({inside[1]});  // CTA-uniform: both tiles inside the domain",
                 f"            const int m0 = fast ? max(max({mid_range[0][0]}, {mid_range[1][0]}), 0) : nk;",
                 f"            const int m1 = fast ? max(min(min({mid_range[0][1]}, {mid_range[1][1]}), nk), m0) : nk;"]
            for lo_, hi_, fastpart in (("0", "m0", False), ("m0", "m1", True), ("m1", "nk", False)):
                R += [f"            for (int it = {lo_}; it < {hi_}; ++it) {{",
                      f"                const int kA = {kA}, kB = {kB};"]
                if has_issue[0]:
                    R.append("                cf_issueA(kA);")
                if has_issue[1]:
                    R.append("                cf_issueB(kB);")
                if has_issue[0] or has_issue[1]:
                    R.append(f"                b200_cp_commit(); b200_cp_wait<{LEAD}>();")
                fnB, fnA = ("cf_midB", "cf_midA") if fastpart else ("cf_levelB", "cf_levelA")
                R.append(f"                {fnB}(kB, parB ? nk - 1 - kB : kB);")
                R.append(f"                {fnA}(kA, parA ? nk - 1 - kA : kA);")
                R.append("            }")
            return R
        R = ["            for (int it = 0; it < nk; ++it) {", f"                const int kA = {kA}, kB = {kB}; (void)kA; (void)kB;"]
        if doA and has_issue[0]:
            R.append("                cf_issueA(kA);")
        if doB and has_issue[1]:
            R.append("                cf_issueB(kB);")
        if (doA and has_issue[0]) or (doB and has_issue[1]):
            R.append(f"                b200_cp_commit(); b200_cp_wait<{LEAD}>();")
        # the stack slot of a level: the level itself for even tiles of loop A's direction, mirrored for odd ones
        if doB:
            R.append("                cf_levelB(kB, parB ? nk - 1 - kB : kB);")
        if doA:
            R.append("                cf_levelA(kA, parA ? nk - 1 - kA : kA);")
        R.append("            }")
        return R

    S.append("    for (int r = 0; r <= cf_mine && cf_mine > 0; ++r) {")
    S.append("        const bool doA = r < cf_mine, doB = r > 0;")
    S.append("        const int parA = r & 1, parB = parA ^ 1; (void)parA; (void)parB;")
    S.append("        if (doA) { const int tile = (int)blockIdx.x + r * (int)gridDim.x;")
    S += ["            " + x for x in set_tile[0]]
    if prologue[0]:
        S.append("            " + prologue[0])
    S.append("        }")
    if prologue[1]:
        S.append("        if (doB) { " + prologue[1] + " }")
    S.append("        if (doA && doB) {")
    S += round_loop(True, True)
    S.append("        } else if (doA) {")
    S += round_loop(True, False)
    S.append("        } else {")
    S += round_loop(False, True)
    S.append("        }")
    S.append("        b200_cp_wait<0>();")
    S.append("        b200_stack_fence();")
    S.append("        if (doA) {  // the tile loop A just finished is loop B's next one")
    S.append("            const int tile = (int)blockIdx.x + r * (int)gridDim.x;")
    S += ["            " + x for x in set_tile[1]]
    S.append("        }")
    S.append("    }")
    S.append("    b200_stack_teardown(&cf_tmem_slot);")
    gen.cf_ring, gen.cf_cached = {}, {}
    src = [f"template <bool U> __device__ __forceinline__ void {kname}_impl(const Args &a)", "{"] + S + ["}"]
    for suffix, flag in (("", "false"), ("_u", "true")):
        src.append(f'extern "C" __global__ void __launch_bounds__({BX * BY}, 1) {kname}{suffix}(const Args a) '
                   f"{{ {kname}_impl<{flag}>(a); }}")
    meta = {"name": kname, "kind": "colfuse", "block": [BX, BY], "i_ext": [int(ilo), int(ihi)], "j_ext": [int(jlo), int(jhi)],
            "ring_slots": int(ring_slots), "stack_values_per_level": NP, "stack_tmem": NT, "loops": [int(x) for x in run],
            "cached": sorted(an.cached), "pushed": list(an.pushed), "pipelined": True,
            "private": sorted(nm for nm, c in an.cached.items() if not c["store"])}
    return "\n".join(src), meta
