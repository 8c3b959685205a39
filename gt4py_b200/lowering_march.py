"""Warp-march form of the fused multi-stage kernel (kernel kind ``march``).

A PARALLEL vertical loop whose horizontal executions exchange temporaries at I / J offsets (the config-4 chain,
lap-of-lap diffusion, hdiff) is lowered to ONE kernel in which every temporary lives in REGISTERS:

  * a warp owns ``32 * C`` consecutive columns (C columns per lane, interleaved: lane l holds columns C*l ..
    C*l + C - 1) of one level and marches down J over a strip of ``STRIP`` rows;
  * a temporary read at a J offset is a rolling window of rows in the lane's registers; the loop is unrolled by the
    window period, so the window slots are plain named registers and nothing is ever shifted;
  * a temporary read at an I offset comes from the neighbouring column of the same lane or, across lanes, by
    ``__shfl_up/down_sync`` -- no ``__syncthreads()``: warps are independent;
  * the stages are software-pipelined along J: stage s works ``L_s`` rows behind the front, with L_s the smallest
    lag that makes every row it reads available;
  * the read-only API inputs are copied ``LEAD`` rows ahead of the front by ``cp.async`` into a ring in shared memory
    that is private to the lane (its own columns only: no barrier, just ``cp.async.wait_group``), each element once
    per warp; the ring depth is the unroll period, so every slot is an immediate offset.  (A first version loaded
    the inputs into register windows ahead of use; ncu showed half of all stall samples on the moves of those loaded
    values -- ptxas consumed them at once -- and 43 % issue utilisation; with the ring: 76 %.)
  * the halo of the extent analysis (passes/oir_optimizations/utils.py:293-330) is recomputed per warp in I (26 of
    32 columns -- 56 of 64 with two columns per lane -- are results for a total I halo of 6) and per strip in J; a lane or row outside a stage's range
    computes values nothing valid ever reads, so only the horizontal executions that touch global memory themselves
    are guarded (the copies clamp their address instead).

The tile kernel of lowering._tile_kernel (a CTA per 64 x 16 tile, temporaries in shared memory, a barrier per stage) is
bound by shared-memory traffic and barriers on deep chains.  Counterpart in the reference: the IJ caches of the
GridTools backends (gtc/passes/oir_optimizations/caches.py, IJCacheDetection) -- there a per-block shared-memory tile,
here registers.

Measured on B200 at 512 x 512 x 128 float64 (tools/make_march_variants.py + tools/chain_bench.py,
profiles/r2_march_variants*.jsonl), ms per call, tile kernel -> this form with its first defaults (one column per lane,
4 warps per CTA, 80 registers, LEAD 2): config-4 chain 0.393 -> 0.322, lap-of-lap diffusion 0.195 -> 0.160, hdiff
0.240 -> 0.174; two-stage upwind advection stays a tile kernel (0.163, at the HBM roofline; this form: 0.19).  It is
then bound by instruction issue (7.6 warp instructions per point on the chain, of which 1.4 are float64 arithmetic),
the shallow kernels by the shuffle / shared-memory pipe.  Two columns per lane halve the shuffles; with the first ring
layout (one ring per column, two 8-byte copies and two LDS.64 per input row) they measured slower (0.395 on the chain).
Since the pair of a lane is ONE 16-byte ring element (one cp.async and one LDS.128 per input row where the pair is
whole and aligned in global memory -- every warp's first column is even, gt4py storages align the origin -- two plain
copies into the same element otherwise) two columns per lane and 2 warps per CTA are the defaults
(profiles/r2_march_pairs.jsonl, 512 x 512 x 128 / 1024 x 1024 x 80, one column -> two): chain 0.251 -> 0.248 / 0.599 ->
0.551 ms, diffusion 0.155 -> 0.139 / 0.366 -> 0.314, hdiff 0.167 -> 0.168 / 0.387 -> 0.376.

Not every unit qualifies (`emit` returns None and the tile kernel is used): every fused temporary must be written in
ONE stage and read only by later stages (or at zero offset inside its own), and the kernel must have at least
MIN_STAGES stages.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Set, Tuple

from gt4py.cartesian.gtc import common, oir
from gt4py.cartesian.gtc.passes.oir_optimizations.utils import AccessCollector

def _knob(name: str, default: int) -> int:
    """Tuning knobs, read when a kernel is emitted (experiments and tests set them; the plan records what was used)."""
    return int(os.environ.get(name, str(default)))


MAX_PERIOD = 8
MIN_STAGES = 3


def _pow2(n: int) -> int:
    p = 1
    while p < n:
        p *= 2
    return p


def _field_accesses(h):
    """[(name, di, dj, dk or None, is_write)] of a horizontal execution (static Cartesian offsets; dk None: variable)."""
    out = []
    for a in AccessCollector.apply(h).ordered_accesses():
        out.append((str(a.field), int(a.offset[0]), int(a.offset[1]), a.offset[2], bool(a.is_write)))
    return out


def emit(kname: str, unit, block_ext, fields, field_ext, scalar_names, locals_, smem_names, L
         ) -> Optional[Tuple[str, Dict[str, Any]]]:
    """(CUDA source, launch metadata) of the warp-march kernel of `unit`, or None if the unit does not qualify.
    `L` is the lowering module (ExprGen, CTYPES, helpers)."""
    C = _knob("B200_MARCH_COLUMNS", 2)          # columns per lane
    STRIP = _knob("B200_MARCH_STRIP", 64)       # rows a warp marches over
    LEAD = _knob("B200_MARCH_LEAD", 2)          # rows an input is copied into the lane's shared-memory ring ahead of use
    WARPS = _knob("B200_MARCH_WARPS", 2)        # warps per CTA (independent of each other: scheduling granularity only)
    MINB = _knob("B200_MARCH_MIN_BLOCKS", 0)    # __launch_bounds__ second argument (caps the registers per thread)
    stages = L._tile_stages(unit, smem_names)
    temps = {str(x) for x in smem_names}
    readonly = L._readonly_globals(unit, smem_names, fields)
    # ---- who writes / reads what, where -------------------------------------------------------------------------
    writer: Dict[str, int] = {}
    reads: Dict[str, List[Tuple[int, int, int]]] = {}      # temp or staged input -> [(stage, di, dj)]
    bad_input: Set[str] = set()
    for n, stage in enumerate(stages):
        for _, h in stage:
            for name, di, dj, dk, is_write in _field_accesses(h):
                if name in temps:
                    if dk is None or int(dk) != 0:
                        return None
                    if is_write:
                        if writer.setdefault(name, n) != n or di or dj:
                            return None
                    else:
                        reads.setdefault(name, []).append((n, di, dj))
                elif name in readonly:
                    f = fields[name]
                    if (not all(f["dims"]) or f["data_dims"] or dk is None or int(dk) != 0
                            or L.CTYPES[f["dtype"]] not in L.TMA_DTYPES):  # (4 / 8 bytes: what cp.async copies)
                        bad_input.add(name)
                    else:
                        reads.setdefault(name, []).append((n, di, dj))
    for name in temps:
        if name not in writer:
            return None
        for n, di, dj in reads.get(name, []):
            if n < writer[name] or (n == writer[name] and (di or dj)):
                return None
    inputs = sorted(n for n in reads if n not in temps and n not in bad_input)
    # registers that never leave one stage of one thread (no offset reads, one stage): plain locals
    regs = {t for t in temps if all(n == writer[t] and not di and not dj for n, di, dj in reads.get(t, []))}
    windows = sorted(temps - regs)
    # ---- lags: stage s works on row (front - lag[s]); inputs are at the front (lag 0) ------------------------------------
    lag = [0] * len(stages)
    for n in range(len(stages)):
        need = 0
        for name in inputs:
            for s, di, dj in reads[name]:
                if s == n:
                    need = max(need, dj)
        for name in windows:
            for s, di, dj in reads[name]:
                if s == n and writer[name] != n:
                    need = max(need, lag[writer[name]] + max(dj, 0))
        lag[n] = max(need, 0)
    depth: Dict[str, int] = {}
    for name in windows:
        depth[name] = _pow2(max([lag[s] - dj - lag[writer[name]] for s, di, dj in reads[name]] + [0]) + 1)
    U = max(list(depth.values()) + [1])
    if U > MAX_PERIOD:
        return None
    if len(stages) < MIN_STAGES and os.environ.get("B200_MARCH", "1") != "force":
        return None  # measured (module docstring): two-stage kernels are faster as shared-memory tile kernels
    # ---- geometry -------------------------------------------------------------------------------------------------------------
    exts = {id(h): block_ext[id(h)] for _, h in unit}
    s_ext = []
    for stage in stages:
        es = [exts[id(h)] for _, h in stage]
        s_ext.append((min(e[0][0] for e in es), max(e[0][1] for e in es), min(e[1][0] for e in es), max(e[1][1] for e in es)))
    in_ext = {}
    for name in inputs:  # the columns / rows of an input some stage may read
        lo_i = min(s_ext[s][0] + di for s, di, dj in reads[name])
        hi_i = max(s_ext[s][1] + di for s, di, dj in reads[name])
        lo_j = min(s_ext[s][2] + dj for s, di, dj in reads[name])
        hi_j = max(s_ext[s][3] + dj for s, di, dj in reads[name])
        fe = field_ext.get(name)
        if fe is not None:  # (mask-aware: an access inside a horizontal region does not widen the field everywhere)
            lo_i, hi_i = max(lo_i, int(fe[0][0])), min(hi_i, int(fe[0][1]))
            lo_j, hi_j = max(lo_j, int(fe[1][0])), min(hi_j, int(fe[1][1]))
        in_ext[name] = (lo_i, hi_i, lo_j, hi_j)
    ILO = min([e[0] for e in s_ext] + [e[0] for e in in_ext.values()])
    IHI = max([e[1] for e in s_ext] + [e[1] for e in in_ext.values()])
    ILOa = -((-ILO + C - 1) // C * C)                  # first column of lane 0, relative to the warp's first result
    WO = 32 * C - (IHI - ILOa)                          # result columns per warp
    # two columns per lane: the pair is ONE ring element (16 / 8 bytes: one cp.async and one LDS.128 / LDS.64 per row
    # where the pair is whole and aligned in global memory, two plain copies into the same element otherwise)
    PAIR = C == 2 and bool(_knob("B200_MARCH_PAIRS", 1))
    if PAIR:
        WO -= WO % 2                                    # (every warp's first column even: its lanes' pairs stay aligned)
    if WO < 8 * C:
        return None
    R0 = min([s_ext[n][2] + lag[n] for n in range(len(stages))] + [e[2] for e in in_ext.values()])
    NIT = max([STRIP + s_ext[n][3] - R0 + lag[n] for n in range(len(stages))] + [STRIP + e[3] - R0 for e in in_ext.values()])
    NIT = (NIT + U - 1) // U * U
    XR = max([s_ext[n][3] + lag[n] for n in range(len(stages))] + [e[3] for e in in_ext.values()])
    # per-lane input rings in shared memory: rows front - back .. front + LEAD of the lane's own columns
    back = max([lag[s_] - dj for name in inputs for s_, di, dj in reads[name]] + [0])
    RD = _pow2(LEAD + back + 1)
    T = 32 * WARPS
    # with the loop unrolled by a multiple of the ring depth every ring slot is a compile-time offset (an immediate in
    # the LDS / cp.async address) instead of three integer instructions per access
    STATIC = _knob("B200_MARCH_STATIC_SLOTS", 1) and max(U, RD) <= MAX_PERIOD
    if STATIC:
        U = max(U, RD)
        NIT = (NIT + U - 1) // U * U
    ring_off: Dict[Tuple[str, int], int] = {}
    smem_bytes = 0
    for name in inputs:
        for c in range(C):
            ring_off[(name, c)] = smem_bytes
            if not PAIR or c == C - 1:
                smem_bytes += RD * T * L.TMA_DTYPES[L.CTYPES[fields[name]["dtype"]]] * (C if PAIR else 1)

    gen = L.ExprGen(fields, scalar_names, locals_, False, regs=regs)
    gen.readonly = readonly
    ctype = {n: L.CTYPES[fields[n]["dtype"]] for n in list(temps) + inputs}

    has_while = any(isinstance(node, oir.While) for _, h in unit for node in h.walk_values())
    # horizontal regions that hang on the west / east edge of the domain (FV3-style edge and corner forms): a warp whose
    # columns stay EDGE_W / EDGE_E columns away from those edges runs a copy of the loop without them
    EDGE_W = EDGE_E = 0
    n_edge = 0
    for _, h in unit:
        for node in h.walk_values():
            if isinstance(node, oir.HorizontalRestriction):
                el = L.edge_local(node.mask.i)
                if el is not None:
                    n_edge += 1
                    if el[0] == "start":
                        EDGE_W = max(EDGE_W, el[1])
                    else:
                        EDGE_E = max(EDGE_E, el[1])
    # horizontal executions that touch global memory themselves (a write to an API field, a read of a field that is not
    # kept in register windows): they keep their guard everywhere; the others only produce register values, which a
    # lane or row outside the execution's range may compute freely -- nothing valid ever reads them
    touches_global = {}
    for _, h in unit:
        names = {nm for nm, *_ in _field_accesses(h)}
        touches_global[id(h)] = bool(names - temps - set(inputs))

    S: List[str] = []
    S.append("    const int lane = (int)(threadIdx.x & 31u);")
    # the last warp of a row would hang over the east end of the domain and take the guarded loop (twice the
    # instructions): it is moved west until its block ends at the domain's last column -- it then recomputes a few
    # columns of its neighbour and writes the same values there.  Only when no global field is read AND written by the
    # kernel (a second execution of `x = x + ...` would not be the same write).
    rw_global = False
    g_reads, g_writes = set(), set()
    for _, h in unit:
        for nm, _di, _dj, _dk, is_write in _field_accesses(h):
            if nm not in temps:
                (g_writes if is_write else g_reads).add(nm)
    rw_global = bool(g_reads & g_writes)
    S.append(f"    const int ic_raw = ((int)blockIdx.x * {WARPS} + (int)(threadIdx.x >> 5)) * {WO};")
    if rw_global:
        S.append("    const int ic0 = ic_raw;  // the warp's first result column")
    else:
        S.append(f"    const int ic0 = (ic_raw + {WO} > a.ni && a.ni >= {WO} && ic_raw < a.ni) ? a.ni - {WO} : ic_raw;"
                 "  // the warp's first result column")
    S.append(f"    const int x0 = lane * {C} + ({ILOa});                 // this lane's first column, relative to ic0")
    S.append(f"    const int tj0 = (int)blockIdx.y * {STRIP}, k = (int)blockIdx.z;")
    S.append("    if (ic_raw >= a.ni) return;  // (the whole warp: no lane is left behind at a shuffle)")
    # iterations of this strip: none once every stage is past the last row it computes (a short last strip)
    S.append(f"    const int nit = min({NIT}, a.nj + ({XR}) - (tj0 + ({R0})));")
    for name in windows:
        for q in range(depth[name]):
            S.append(f"    {ctype[name]} " + ", ".join(f"w_{name}_{q}_{c} = 0" for c in range(C)) + ";")
    if inputs:
        S.append("    extern __shared__ __align__(16) unsigned char b200_smem[];")
    for name in inputs:  # slot s of this lane's ring: rg_<field>_<column>[s * T]  (pairs: rg_<field>[s * 2 T + column])
        if PAIR:
            S.append(f"    {ctype[name]} *rg_{name} = ({ctype[name]} *)(b200_smem + {ring_off[(name, 0)]}) + 2 * (int)threadIdx.x;")
            continue
        for c in range(C):
            S.append(f"    {ctype[name]} *rg_{name}_{c} = ({ctype[name]} *)(b200_smem + {ring_off[(name, c)]}) + (int)threadIdx.x;")
    # Input rows are copied without a column guard: a lane whose column lies outside the range the stencil reads
    # (the two halo columns beyond the warp's block, the columns past the domain's end) loads the nearest column
    # inside it instead -- its value is never used (extent analysis), the address is always valid.
    for name in inputs:
        lo_i, hi_i, lo_j, hi_j = in_ext[name]
        for c in range(C):
            S.append(f"    const {ctype[name]} *p_{name}_{c};")
            S.append(f"    {{ const int i = min(max(ic0 + x0 + {c}, {lo_i}), a.ni + ({hi_i}) - 1), j = 0; "
                     f"p_{name}_{c} = g_{name} + ({gen._index(name, common.CartesianOffset.zero(), None)}); }}")
        if PAIR:  # both columns inside the field, adjacent in memory, the pair aligned in every row
            pb = 2 * L.TMA_DTYPES[ctype[name]]
            S.append(f"    const bool pr_{name} = p_{name}_1 == p_{name}_0 + 1 && ((unsigned long long)p_{name}_0 & {pb - 1}ull) == 0 "
                     f"&& (a.{name}_s[1] & 1) == 0;")

    def loop(fast: bool, inner: bool = False) -> List[str]:
        """The J march; `fast`: every column of the warp is inside the domain, and only the horizontal executions that
        touch global memory themselves are guarded."""
        def issue(row: str, slot_const: int, ind: int) -> List[str]:
            """cp.async of row `row` of every input (clamped to the rows the stencil reads: a row outside them is never
            used, the address is always valid) into ring slot (tb + slot_const) mod RD"""
            out = []
            sl = f"((tb + ({slot_const})) & {RD - 1})" if not STATIC else str(slot_const % RD)
            for name in inputs:
                lo_i, hi_i, lo_j, hi_j = in_ext[name]
                cp = "b200_cp_async8" if L.TMA_DTYPES[ctype[name]] == 8 else "b200_cp_async4"
                out.append(" " * ind + f"{{ const int jl = min(max({row}, {lo_j}), a.nj + ({hi_j}) - 1);")
                if PAIR:
                    cp2 = "b200_cp_async16" if L.TMA_DTYPES[ctype[name]] == 8 else "b200_cp_async8"
                    out.append(" " * ind + f"  if (pr_{name}) {cp2}(rg_{name} + {sl} * {2 * T}, p_{name}_0 + (idx_t)jl * (idx_t)a.{name}_s[1], true);")
                    out.append(" " * ind + f"  else {{ {cp}(rg_{name} + {sl} * {2 * T}, p_{name}_0 + (idx_t)jl * (idx_t)a.{name}_s[1], true); "
                               f"{cp}(rg_{name} + {sl} * {2 * T} + 1, p_{name}_1 + (idx_t)jl * (idx_t)a.{name}_s[1], true); }}")
                else:
                    for c in range(C):
                        out.append(" " * ind + f"  {cp}(rg_{name}_{c} + {sl} * {T}, p_{name}_{c} + (idx_t)jl * (idx_t)a.{name}_s[1], true);")
                out.append(" " * ind + "}")
            out.append(" " * ind + "b200_cp_commit();")
            return out

        B: List[str] = []
        if inputs:  # the first LEAD rows
            B.append("    { const int tb = 0;")
            for l in range(LEAD):
                B += issue(f"tj0 + ({R0}) + {l}", l, 6)
            B.append("    }")
        B += ["    #pragma unroll 1", f"    for (int tb = 0; tb < nit; tb += {U}) {{"]
        nsh = 0
        for u in range(U):
            B.append(f"        {{  // iteration tb + {u}")
            B.append(f"            const int jf = tj0 + ({R0}) + tb + {u};  // the front row")
            if inputs:
                B += issue(f"jf + {LEAD}", u + LEAD, 12)
                B.append(f"            b200_cp_wait<{LEAD}>();  // rows up to the front row have arrived")
            ring_loads: Dict[Tuple[str, int, int], str] = {}   # (input, column, row - front) -> local holding it
            top = len(B)                                        # the ring reads of the iteration go here
            for n, stage in enumerate(stages):
                ilo, ihi, jlo, jhi = s_ext[n]
                lo, hi = L._interval_cond(stage[0][0])
                B.append(f"            {{  // stage {n}: row front - {lag[n]}")
                B.append(f"                const int j = jf - {lag[n]};")
                B.append(f"                const bool row{n} = j >= tj0 + ({jlo}) && j < tj0 + {STRIP} + ({jhi}) && j < a.nj + ({jhi}) "
                         f"&& k >= {lo} && k < {hi};")
                # values of other lanes this stage reads: shuffled before any guard (every lane takes part)
                shuf: Dict[Tuple[str, int, int, int], str] = {}
                maps: List[Dict[Tuple[str, int, int], str]] = []
                for c in range(C):
                    m: Dict[Tuple[str, int, int], str] = {}
                    for name in inputs + windows:
                        for s_, di, dj in reads.get(name, []):
                            if s_ != n:
                                continue
                            x = c + di
                            delta, sc = x // C, x % C
                            if name in inputs:  # row (front - lag[n] + dj) of the lane's ring, read once per iteration
                                ro = dj - lag[n]
                                if (name, sc, ro) not in ring_loads:
                                    ring_loads[(name, sc, ro)] = f"v{u}_{name}_{sc}_{'m' if ro < 0 else 'p'}{abs(ro)}"
                                src, q = ring_loads[(name, sc, ro)], ro
                            else:
                                q = (u - lag[n] + dj) % depth[name]  # the slot of row (front - lag[n] + dj)
                                src = f"w_{name}_{q}_{sc}"
                            if delta == 0:
                                m[(name, di, dj)] = src
                            else:
                                key = (name, q, sc, delta)
                                if key not in shuf:
                                    shuf[key] = f"sh{nsh}"
                                    nsh += 1
                                    B.append(f"                const {ctype[name]} {shuf[key]} = b200_shfl({src}, {delta});")
                                m[(name, di, dj)] = shuf[key]
                    for name in windows:  # the stage's own output row
                        if writer[name] == n:
                            m[(name, 0, 0)] = f"w_{name}_{(u - lag[n]) % depth[name]}_{c}"
                    maps.append(m)
                gen.drop_edge_regions = inner
                for c in range(C):
                    gen.march = maps[c]
                    B.append(f"                {{  // column {c}")
                    B.append(f"                    const int x = x0 + {c}, i = ic0 + x;")
                    for name in sorted(regs):
                        if writer[name] == n:
                            B.append(f"                    {ctype[name]} r_{name} = 0;")
                    for _, h in stage:
                        e = exts[id(h)]
                        cond = f"row{n} && x >= ({e[0][0]}) && x < {WO} + ({e[0][1]})"
                        if not fast:
                            cond += f" && i >= ({e[0][0]}) && i < a.ni + ({e[0][1]})"
                        if e[1][0] != jlo:
                            cond += f" && j >= tj0 + ({e[1][0]})"
                        if e[1][1] != jhi:
                            cond += f" && j < tj0 + {STRIP} + ({e[1][1]}) && j < a.nj + ({e[1][1]})"
                        if fast and not touches_global[id(h)]:
                            cond = "true"
                        B.append(f"                    if ({cond}) {{")
                        for d in L._declarations(h):
                            B.append(" " * 24 + f"{L.CTYPES[d.dtype]} {d.name} = 0;")
                        B += gen.stmts(h.body, 24)
                        B.append("                    }")
                    B.append("                }")
                gen.march = None
                gen.drop_edge_regions = False
                B.append("            }")
            if PAIR:  # one 16- / 8-byte read per (input, row), then its halves
                loads, seen = [], set()
                for (name, sc, ro), var in ring_loads.items():
                    if (name, ro) not in seen:
                        seen.add((name, ro))
                        at = f"((tb + ({u + ro})) & {RD - 1}) * {2 * T}" if not STATIC else f"{((u + ro) % RD) * 2 * T}"
                        loads.append(f"            const b200_pair<{ctype[name]}> vp{u}_{name}_{'m' if ro < 0 else 'p'}{abs(ro)} = "
                                     f"*reinterpret_cast<const b200_pair<{ctype[name]}> *>(rg_{name} + {at});")
                    loads.append(f"            const {ctype[name]} {var} = vp{u}_{name}_{'m' if ro < 0 else 'p'}{abs(ro)}.{'xy'[sc]};")
            else:
                loads = [f"            const {ctype[name]} {var} = rg_{name}_{sc}["
                         + (f"((tb + ({u + ro})) & {RD - 1}) * {T}];" if not STATIC else f"{((u + ro) % RD) * T}];")
                         for (name, sc, ro), var in ring_loads.items()]
            B[top:top] = loads
            B.append("        }")
        B.append("    }")
        if inputs:
            B.append("    b200_cp_wait<0>();")
        return B

    if has_while:  # (a `while` on a value no result depends on might not terminate)
        S += loop(False)
    else:
        if n_edge:
            S.append(f"    if (ic0 + ({ILOa}) >= {EDGE_W} && ic0 + ({ILOa}) + {32 * C} <= a.ni - ({EDGE_E})) {{"
                     "  // ... and away from the west / east edge regions")
            S += ["    " + ln for ln in loop(True, True)]
            S.append("    } else")
        S.append(f"    if (ic0 + {WO} <= a.ni) {{  // every column of the warp's block is inside the domain")
        S += ["    " + ln for ln in loop(True)]
        S.append("    } else {")
        S += ["    " + ln for ln in loop(False)]
        S.append("    }")
    all_exts = list(exts.values())
    meta = {"name": kname, "kind": "march", "columns_per_lane": C, "paired_ring_elements": bool(PAIR), "strip": STRIP, "warps": WARPS, "result_columns": WO,
            "stages": len(stages), "lags": lag, "period": U, "lead": LEAD, "ring_rows": RD,
            "smem_bytes": int(smem_bytes), "input_rings": inputs,
            "register_windows": {n: depth[n] for n in windows}, "register_temporaries": sorted(regs),
            "smem_arrays": [],
            "i_ext": [int(min(e[0][0] for e in all_exts)), int(max(e[0][1] for e in all_exts))],
            "j_ext": [int(min(e[1][0] for e in all_exts)), int(max(e[1][1] for e in all_exts))]}
    return L._entry_points(kname, 32 * WARPS, S, readonly, fields, MINB), meta
