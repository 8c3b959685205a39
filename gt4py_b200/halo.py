"""IJ domain decomposition and halo refresh across the GPUs of one box.

New functionality -- ``gt4py.cartesian`` has no distributed layer (users couple it to external halo
libraries); here it is part of the backend because BASELINE.json's north star asks for it.

Model: one process per GPU; rank ``r`` owns an ``NI x NJ x NK`` block of a ``Px x Py`` block decomposition
of the global IJ plane (K is never split: vadv is a recurrence along K).  Every field lives in an array
with the same local shape / origin on every rank, its halo cells inside the allocation.  Before a stencil
call the halo cells that belong to a neighbour's compute domain are refreshed from that neighbour:

* the boxes to move are derived from the stencil's ``field_info[name].boundary`` (same metadata the
  reference uses for validation, backend/module_generator.py:80-96): for hdiff ``in_field`` a frame of width
  2 including corners, for vadv ``wcon`` one column on the +I side;
* data path: ONE kernel per refresh pulls all boxes of all registered fields straight from the neighbours'
  memory (CUDA IPC mappings, loads over NVLink); ordering between ranks by step-numbered flags in peer
  memory (csrc/halo.cu).  The process group (torch.distributed or anything with ``allgather`` / ``barrier``)
  is used only once, to exchange the IPC handles.

Open (non-periodic) boundaries: halo cells outside the global domain are left untouched.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import byref, c_void_p
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

DIRECTIONS = [(-1, -1), (0, -1), (1, -1), (-1, 0), (1, 0), (-1, 1), (0, 1), (1, 1)]  # (dx, dy)


def process_grid(world: int) -> Tuple[int, int]:
    """(Px, Py): 1x1, 2x1, 2x2, 4x2, ... -- as square as possible, Px >= Py (I is split first)."""
    if world < 1:
        raise ValueError("world size must be positive")
    py = int(np.floor(np.sqrt(world)))
    while world % py:
        py -= 1
    return world // py, py


@dataclass(frozen=True)
class Decomposition:
    world: int
    rank: int
    px_count: int
    py_count: int

    @classmethod
    def create(cls, world: int, rank: int, grid: Optional[Tuple[int, int]] = None) -> "Decomposition":
        px, py = grid if grid is not None else process_grid(world)
        if px * py != world:
            raise ValueError(f"process grid {px}x{py} does not match world size {world}")
        if not 0 <= rank < world:
            raise ValueError(f"rank {rank} outside world of {world}")
        return cls(world, rank, px, py)

    @property
    def coords(self) -> Tuple[int, int]:
        return self.rank % self.px_count, self.rank // self.px_count

    def neighbour(self, dx: int, dy: int) -> Optional[int]:
        """Rank of the block at (+dx, +dy), or None at an open boundary."""
        x, y = self.coords
        x, y = x + dx, y + dy
        if 0 <= x < self.px_count and 0 <= y < self.py_count:
            return y * self.px_count + x
        return None

    def global_offset(self, domain: Sequence[int]) -> Tuple[int, int]:
        x, y = self.coords
        return x * int(domain[0]), y * int(domain[1])


@dataclass(frozen=True)
class BoxSpec:
    """One halo box: `dst` index ranges in the local array, `src` ranges in the array of rank `src_rank`."""

    direction: Tuple[int, int]
    src_rank: int
    dst_start: Tuple[int, int, int]
    src_start: Tuple[int, int, int]
    extent: Tuple[int, int, int]


def halo_boxes(decomp: Decomposition, origin: Sequence[int], domain: Sequence[int],
               halo: Sequence[Tuple[int, int]]) -> List[BoxSpec]:
    """Boxes that refresh the halo of one field.  `halo` = ((lo_i, hi_i), (lo_j, hi_j)) widths to refresh;
    `origin` / `domain` are the local compute origin and extent (identical on every rank)."""
    (lo_i, hi_i), (lo_j, hi_j) = halo[0], halo[1]
    ni, nj, nk = (int(d) for d in domain)
    oi, oj, ok = (int(o) for o in origin)
    if lo_i > ni or hi_i > ni or lo_j > nj or hi_j > nj:
        raise ValueError("halo wider than the neighbour's compute domain")
    ranges_i = {-1: (oi - lo_i, lo_i), 0: (oi, ni), 1: (oi + ni, hi_i)}  # (start, extent) in the local array
    ranges_j = {-1: (oj - lo_j, lo_j), 0: (oj, nj), 1: (oj + nj, hi_j)}
    boxes = []
    for dx, dy in DIRECTIONS:
        (si, ei), (sj, ej) = ranges_i[dx], ranges_j[dy]
        if ei <= 0 or ej <= 0 or nk <= 0:
            continue
        src = decomp.neighbour(dx, dy)
        if src is None:
            continue
        # my cell i is the neighbour's cell i - dx * ni (same local indexing on every rank)
        boxes.append(BoxSpec((dx, dy), src, (si, sj, ok), (si - dx * ni, sj - dy * nj, ok), (ei, ej, nk)))
    return boxes


def halo_from_field_info(stencil, name: str) -> Tuple[Tuple[int, int], Tuple[int, int]]:
    """Halo widths ((lo_i, hi_i), (lo_j, hi_j)) a stencil reads around its compute domain for field `name`: the
    I / J entries of ``field_info[name].boundary`` -- the metadata the reference derives for validation
    (backend/module_generator.py:80-96) -- so the exchange is generated from the stencil, not typed in by hand."""
    fi = stencil.field_info[name]
    b = {ax: fi.boundary[n] for n, ax in enumerate("IJK")}
    return (tuple(int(x) for x in b["I"]), tuple(int(x) for x in b["J"]))


class TorchComm:
    """Adapter over torch.distributed (used only to swap IPC handles and for barriers)."""

    def __init__(self, dist):
        self.dist = dist

    def allgather(self, obj):
        out = [None] * self.dist.get_world_size()
        self.dist.all_gather_object(out, obj)
        return out

    def barrier(self):
        self.dist.barrier()


class HaloExchanger:
    """Refreshes the halos of registered device fields from the neighbour ranks.

    mode "push" (default): every rank packs its boundary faces straight into its neighbours' contiguous
    staging buffers (coalesced posted writes over NVLink), posts a step-numbered flag, and then unpacks its own
    staging buffer into its halo cells once its neighbours' flags have arrived -- ONE kernel launch per refresh
    (b200_halo_refresh mode 4), staging double buffered by step parity.  mode "push2": the same as two launches
    (push, then unpack).  mode "pull": one kernel loads the boxes out of the neighbours' fields (simple, but
    16-byte peer loads are latency bound: 0.3 ms for the 1024 x 80 faces vs 20 us pushed).

    Fields of any dtype with (I, J, K) leading axes and any number of data dimensions are supported (a box per
    data index).  ORDERING: the staging buffers are reused every second refresh, which is safe because a rank's
    refreshes execute in issue order: refreshes issued on different streams are chained with events by this class
    (`enable_multi_stream()`; `refresh` then makes the stream wait for the previous refresh of this exchanger), and
    every rank must issue the same sequence of refreshes.
    """

    ITEMS_PER_THREAD = 2
    MAX_BLOCKS_PER_BOX = 1024
    FUSED_MAX_BYTES = 512 * 1024  # mode 'push': one fused launch per refresh up to this many received bytes
    RESIDENT_BLOCKS = 444  # 3 CTAs of 256 threads per SM: co-resident even beside a persistent compute kernel

    def __init__(self, comm, decomp: Decomposition, use_flags: bool = True, mode: str = "push"):
        """`use_flags=False` skips the inter-rank flag protocol: only for single-process emulations that
        run several ranks on ONE device and order them on the host (see `refresh_local`)."""
        from . import _lib, storage

        if mode not in ("push", "push2", "pull"):
            raise ValueError("mode must be 'push', 'push2' or 'pull'")
        self.mode = mode
        self.use_flags = bool(use_flags)
        self._lib, self._storage = _lib, storage
        self.comm, self.decomp = comm, decomp
        self._groups: List[List[Tuple[object, tuple, tuple, tuple]]] = []
        self._finalized = False
        self.step = 0
        self.launches_per_refresh = 2 if mode == "push2" else 1
        self._last_stream_key = None
        self._last_event = None

    # -- registration ---------------------------------------------------------------------------------
    def add_group(self, fields: Sequence[Tuple[object, Sequence[int], Sequence[int], Sequence[Tuple[int, int]]]]) -> int:
        """Register a group of fields refreshed together: (DeviceArray, origin, domain, halo widths).
        Every rank must register the same groups (same origins, domains, halos) in the same order."""
        if self._finalized:
            raise RuntimeError("HaloExchanger: add_group after finalize")
        group = []
        for arr, origin, domain, halo in fields:
            if arr.ndim < 3 or arr.dtype.itemsize not in (1, 2, 4, 8):
                raise TypeError("halo refresh needs fields with (I, J, K) leading axes and 1/2/4/8-byte elements")
            group.append((arr, tuple(int(o) for o in origin), tuple(int(d) for d in domain),
                          tuple((int(a), int(b)) for a, b in halo)))
        self._groups.append(group)
        return len(self._groups) - 1

    def _boxes_of(self, rank: int, group) -> List[Tuple[int, BoxSpec, int]]:
        """(field index, box, staging offset in BYTES, data index) of everything `rank` receives for `group`."""
        dec = Decomposition(self.decomp.world, rank, self.decomp.px_count, self.decomp.py_count)
        out, off = [], 0
        for f, (arr, origin, domain, halo) in enumerate(group):
            item = arr.dtype.itemsize
            for b in halo_boxes(dec, origin[:3], domain, halo):
                for data_index in np.ndindex(*arr.shape[3:]):
                    out.append((f, b, off, tuple(int(x) for x in data_index)))
                    off += (b.extent[0] * b.extent[1] * b.extent[2] * item + 15) // 16 * 16
        return out

    def finalize(self):
        """Swap IPC handles with the other ranks and build the device-side box tables."""
        self._connect(self.comm.allgather(self._export()))
        self.comm.barrier()  # nobody starts before every mapping, flag block and staging buffer exists

    @staticmethod
    def finalize_local(exchangers: Sequence["HaloExchanger"]):
        """Single-process emulation (all ranks in this process, one device): connect them directly."""
        everyone = [ex._export() for ex in exchangers]
        for ex in exchangers:
            ex._connect(everyone)

    def _export(self):
        lib, L = self._lib.lib(), self._lib
        # flags: ready[8] then done[8], one slot per direction
        self._flag_buf = self._storage._Buffer(16 * 8 + 8)
        L.check(lib.b200_memset(c_void_p(self._flag_buf.ptr), 0, self._flag_buf.nbytes, None))
        # staging: what this rank receives per refresh, two parities
        def stage_bytes(g):
            entries = self._boxes_of(self.decomp.rank, g)
            if not entries:
                return 0
            f, b, off, _ = entries[-1]
            return off + (b.extent[0] * b.extent[1] * b.extent[2] * g[f][0].dtype.itemsize + 15) // 16 * 16

        self._stage_half = (max([stage_bytes(g) for g in self._groups] or [0]) + 255) // 256 * 256
        self._stage_buf = self._storage._Buffer(max(2 * self._stage_half, 256))
        L.synchronize()

        def handle_of(base_ptr):
            h = (ctypes.c_ubyte * 64)()
            L.check(lib.b200_ipc_get_mem_handle(c_void_p(base_ptr), h))
            return bytes(h)

        # (IPC handle of the allocation, its base address in the owner's address space)
        mine = {"pid": os.getpid(), "flags": (handle_of(self._flag_buf.ptr), self._flag_buf.ptr),
                "staging": (handle_of(self._stage_buf.ptr), self._stage_buf.ptr), "stage_half": self._stage_half,
                "fields": []}
        for group in self._groups:
            mine["fields"].append([((handle_of(arr._buf.ptr), arr._buf.ptr), arr.ptr - arr._buf.ptr, arr.strides)
                                   for arr, *_ in group])
        return mine

    def _connect(self, everyone):
        lib, L = self._lib.lib(), self._lib
        me = self.decomp.rank
        self._mapped: Dict[Tuple[int, bytes], int] = {}

        def mapped(rank, ref):
            handle, owner_ptr = ref
            if everyone[rank]["pid"] == os.getpid():  # same process (single-process emulation): no IPC needed
                return owner_ptr
            key = (rank, handle)
            if key not in self._mapped:
                p = c_void_p()
                hb = (ctypes.c_ubyte * 64).from_buffer_copy(handle)
                L.check(lib.b200_ipc_open_mem_handle(hb, byref(p)))
                self._mapped[key] = p.value
            return self._mapped[key]

        # synchronisation partners: every neighbour that exists in a direction used either way
        used = set()
        for group in self._groups:
            for _, b, _, _ in self._boxes_of(me, group):
                used.add(b.direction)
        for d in list(used):
            used.add((-d[0], -d[1]))
        partners = [d for d in DIRECTIONS if d in used and self.decomp.neighbour(*d) is not None]
        sync = L.HaloSync()
        sync.npeers = len(partners)
        for n, d in enumerate(partners):
            peer = self.decomp.neighbour(*d)
            peer_flags = mapped(peer, everyone[peer]["flags"])
            mine_slot = DIRECTIONS.index(d)                    # the slot the peer writes in my block
            peer_slot = DIRECTIONS.index((-d[0], -d[1]))       # the slot I write in the peer's block
            sync.post_ready[n] = peer_flags + 8 * peer_slot
            sync.post_done[n] = peer_flags + 8 * (8 + peer_slot)
            sync.wait_ready[n] = self._flag_buf.ptr + 8 * mine_slot
            sync.wait_done[n] = self._flag_buf.ptr + 8 * (8 + mine_slot)
        self._sync = sync
        self._arrive = self._storage._Buffer(8)
        L.check(lib.b200_memset(c_void_p(self._arrive.ptr), 0, 8, None))
        peers = sorted({self.decomp.neighbour(*d) for d in partners})

        def box(src, sstr, dst, dstr, extent, item=8):
            hb = L.HaloBox()
            hb.src, hb.dst = src, dst
            for d in range(3):
                hb.src_stride[d], hb.dst_stride[d], hb.extent[d] = sstr[d], dstr[d], extent[d]
            hb.elem_bytes = item
            return hb

        def at(base, strides, start, data_index):
            """Address of element (start, data_index) of an array at `base` with byte `strides`."""
            return base + sum(s * st for s, st in zip(tuple(start) + tuple(data_index), strides))

        def upload(entries, resident=False):
            """`resident`: the launch mixes pushing blocks with blocks that spin on the neighbours' flags (fused
            mode): keep the whole grid small enough to be co-resident on the GPU (148 SMs x several CTAs), so that a
            waiting block can never keep a pushing block from being scheduled."""
            n = len(entries)
            table = (L.HaloBox * max(n, 1))(*entries)
            dev = self._storage._Buffer(max(n, 1) * ctypes.sizeof(L.HaloBox))
            if n:
                L.check(lib.b200_memcpy_h2d(c_void_p(dev.ptr), ctypes.addressof(table), n * ctypes.sizeof(L.HaloBox), None))
            most = max([e.extent[0] * e.extent[1] * e.extent[2] for e in entries] or [1])
            blocks = min(self.MAX_BLOCKS_PER_BOX, max(1, -(-most // (2 * 256 * self.ITEMS_PER_THREAD))))
            if resident:
                blocks = max(1, min(blocks, self.RESIDENT_BLOCKS // max(n, 1)))
            return dev, n, blocks

        def dense(extent):
            return (1, extent[0], extent[0] * extent[1])

        # "poison" tables: every halo box this rank receives, overwritten with NaN (checks only: a refresh that does
        # not deliver a box leaves NaNs behind instead of last step's plausible values)
        self._nan = self._storage._Buffer(256)
        nan = (ctypes.c_double * 1)(float("nan"))
        L.check(lib.b200_memcpy_h2d(c_void_p(self._nan.ptr), ctypes.addressof(nan), 8, None))
        self._poison = []
        for group in self._groups:
            entries = []
            for f, b, _, di in self._boxes_of(me, group):
                arr = group[f][0]
                item = arr.dtype.itemsize
                if arr.dtype.kind != "f" or item != 8:
                    continue  # NaN poisoning is a float64 check
                entries.append(box(self._nan.ptr, (0, 0, 0), at(arr.ptr, arr.strides, b.dst_start, di),
                                   [st // item for st in arr.strides[:3]], b.extent, item))
            self._poison.append(upload(entries))
        self._tables = []
        for g, group in enumerate(self._groups):
            if self.mode == "pull":
                entries = []
                for f, b, _, di in self._boxes_of(me, group):
                    arr = group[f][0]
                    item = arr.dtype.itemsize
                    ref, offset, strides = everyone[b.src_rank]["fields"][g][f]
                    src = at(mapped(b.src_rank, ref) + offset, strides, b.src_start, di)
                    entries.append(box(src, [st // item for st in strides[:3]], at(arr.ptr, arr.strides, b.dst_start, di),
                                       [st // item for st in arr.strides[:3]], b.extent, item))
                self._tables.append({"pull": upload(entries)})
                continue
            per_parity = []
            for parity in (0, 1):
                push, unpack = [], []
                for peer in peers:  # what each neighbour receives from me -> its staging buffer
                    stage = mapped(peer, everyone[peer]["staging"]) + parity * everyone[peer]["stage_half"]
                    for f, b, off, di in self._boxes_of(peer, group):
                        if b.src_rank != me:
                            continue
                        arr = group[f][0]
                        item = arr.dtype.itemsize
                        push.append(box(at(arr.ptr, arr.strides, b.src_start, di), [st // item for st in arr.strides[:3]],
                                        stage + off, dense(b.extent), b.extent, item))
                mystage = self._stage_buf.ptr + parity * self._stage_half
                for f, b, off, di in self._boxes_of(me, group):
                    arr = group[f][0]
                    item = arr.dtype.itemsize
                    unpack.append(box(mystage + off, dense(b.extent), at(arr.ptr, arr.strides, b.dst_start, di),
                                      [st // item for st in arr.strides[:3]], b.extent, item))
                per_parity.append({"push": upload(push), "unpack": upload(unpack), "fused": upload(push + unpack, resident=True),
                                   "nsplit": len(push)})
            self._tables.append(per_parity)
        L.synchronize()
        self._finalized = True

    def poison(self, group: int, stream=None):
        """Overwrite every halo cell `refresh(group)` is going to deliver with NaN (parity checks)."""
        dev, n, blocks = self._poison[group]
        if n:
            L = self._lib
            L.check(L.lib().b200_halo_refresh(c_void_p(dev.ptr), n, blocks, None, 0, None, 0, L.stream_handle(stream)))

    def bytes_received(self, group: int) -> int:
        """Bytes this rank's halo cells of `group` receive from its neighbours per refresh."""
        total = 0
        for f, b, _, _ in self._boxes_of(self.decomp.rank, self._groups[group]):
            total += int(b.extent[0]) * int(b.extent[1]) * int(b.extent[2]) * self._groups[group][f][0].dtype.itemsize
        return total

    # -- per-step operations ---------------------------------------------------------------------------
    def _launch(self, table, mode, stream):
        dev, n, blocks = table
        L = self._lib
        L.check(L.lib().b200_halo_refresh(c_void_p(dev.ptr), n, blocks, byref(self._sync), self.step,
                                          c_void_p(self._arrive.ptr), mode, L.stream_handle(stream)))

    def refresh(self, group: int, stream=None):
        """Refresh the halos of every field of `group` (asynchronous on `stream`)."""
        if not self._finalized:
            raise RuntimeError("HaloExchanger: call finalize() first")
        if not self.use_flags:
            raise RuntimeError("HaloExchanger without flags: use HaloExchanger.refresh_local(exchangers, group)")
        self._chain(stream)
        self.step += 1
        if self.mode == "pull":
            self._launch(self._tables[group]["pull"], 1, stream)
        elif self.mode == "push2":
            t = self._tables[group][self.step & 1]
            self._launch(t["push"], 2, stream)
            self._launch(t["unpack"], 3, stream)
        elif self.bytes_received(group) > self.FUSED_MAX_BYTES:
            # large faces: two launches with full-size grids beat the co-resident single launch (measured at
            # 2048 x 2048 x 80 on 2 x B200: 52 vs 59 us for both refreshes of a step, profiles/r2_bench_n2_*.json)
            t = self._tables[group][self.step & 1]
            self._launch(t["push"], 2, stream)
            self._launch(t["unpack"], 3, stream)
        else:
            t = self._tables[group][self.step & 1]
            self._launch(t["fused"], 4 + 8 * t["nsplit"], stream)
        if self._last_event is not None:
            self._last_event.record(stream)

    def _chain(self, stream):
        """Refreshes of one exchanger must execute in issue order (staging reuse, step-numbered flags): a refresh
        issued on another stream than the previous one first waits for it."""
        L = self._lib
        key = int(L.stream_handle(stream).value or 0)
        if self._last_stream_key is not None and key != self._last_stream_key:
            if self._last_event is None:
                raise RuntimeError("HaloExchanger: refreshes on several streams need enable_multi_stream() before the "
                                   "first refresh (every rank alike)")
            L.check(L.lib().b200_stream_wait_event(L.stream_handle(stream), self._last_event.handle))
        self._last_stream_key = key

    def enable_multi_stream(self):
        """Allow refreshes on different streams (an event after every refresh chains them)."""
        if self._last_event is None:
            self._last_event = self._lib.Event()

    @staticmethod
    def refresh_local(exchangers: Sequence["HaloExchanger"], group: int, stream=None):
        """Single-process emulation: all ranks push, then all ranks unpack (host order replaces the flags)."""
        for ex in exchangers:
            ex.step += 1
        if exchangers[0].mode == "pull":
            for ex in exchangers:
                ex._launch(ex._tables[group]["pull"], 0, stream)
            return
        for ex in exchangers:
            ex._launch(ex._tables[group][ex.step & 1]["push"], 0, stream)
        for ex in exchangers:
            ex._launch(ex._tables[group][ex.step & 1]["unpack"], 0, stream)

    def wait_neighbours_done(self, stream=None):
        """pull mode only: make `stream` wait until every neighbour has finished pulling the latest refresh
        out of this rank's fields (call before overwriting a field that neighbours read).  In push mode a
        rank reads only its own fields and staging, so there is nothing to wait for."""
        L = self._lib
        if self.use_flags and self.mode == "pull":
            L.check(L.lib().b200_halo_wait_done(byref(self._sync), self.step, L.stream_handle(stream)))

    # -- bench convenience -------------------------------------------------------------------------------
    @classmethod
    def for_bench(cls, dist, rank, world, local, hsets, vsets, dom, mode="p2p"):
        ex = cls(TorchComm(dist), Decomposition.create(world, rank), mode=mode)
        for h, v in zip(hsets, vsets):  # field set s -> groups 2s (in_field) and 2s + 1 (wcon)
            ex.add_group([(h["in_field"], (2, 2, 0), dom, ((2, 2), (2, 2)))])   # field_info['in_field'].boundary
            ex.add_group([(v["wcon"], (0, 0, 0), dom, ((0, 1), (0, 0)))])       # field_info['wcon'].boundary
        ex.finalize()
        return ex


# ---- interior / boundary split: the exchange hidden behind the part of the stencil that does not need it ------------
def split_domain(domain: Sequence[int], halo: Sequence[Tuple[int, int]], has_neighbour) -> Tuple[
        Optional[Tuple[Tuple[int, int, int], Tuple[int, int, int]]], List[Tuple[Tuple[int, int, int], Tuple[int, int, int]]]]:
    """Cut a compute domain into the INTERIOR, whose points read no halo cell that a neighbour rank owns, and up to
    four BOUNDARY strips (west / east over the full J range, south / north between them), as (offset, extent) pairs
    relative to the compute origin.  `halo` = ((lo_i, hi_i), (lo_j, hi_j)) are the widths the stencil reads around a
    point (field_info.boundary); `has_neighbour(dx, dy)` says whether a rank exists on that side (at an open boundary
    the halo cells are the caller's own data, so nothing has to be waited for there).  The pieces are disjoint and
    cover the domain; a side without neighbour contributes no strip."""
    ni, nj, nk = (int(d) for d in domain)
    (lo_i, hi_i), (lo_j, hi_j) = halo
    w = lo_i if has_neighbour(-1, 0) else 0   # a point within lo_i of the west edge reads the west neighbour's cells
    e = hi_i if has_neighbour(1, 0) else 0
    s_ = lo_j if has_neighbour(0, -1) else 0
    n = hi_j if has_neighbour(0, 1) else 0
    if w + e >= ni or s_ + n >= nj:  # too thin for an interior: everything is boundary
        return None, [((0, 0, 0), (ni, nj, nk))]
    strips = []
    if w:
        strips.append(((0, 0, 0), (w, nj, nk)))
    if e:
        strips.append(((ni - e, 0, 0), (e, nj, nk)))
    if s_:
        strips.append(((w, 0, 0), (ni - w - e, s_, nk)))
    if n:
        strips.append(((w, nj - n, 0), (ni - w - e, n, nk)))
    return ((w, s_, 0), (ni - w - e, nj - s_ - n, nk)), strips


def call_on_pieces(stencil, bound: Dict[str, object], full_origin: Dict[str, Tuple[int, ...]], pieces) -> None:
    """Run `stencil` on each (offset, extent) piece of its compute domain: origins shifted by the offset."""
    for offset, extent in pieces:
        origin = {k: tuple(o + d for o, d in zip(v[:3], offset)) + tuple(v[3:]) for k, v in full_origin.items()}
        stencil(**bound, origin=origin, domain=tuple(extent))


class DistributedStencil:
    """A stencil whose halo'd input fields are refreshed from the neighbour ranks as part of the call:

        ds = DistributedStencil(stencil, exchanger, {"in_field": in_field}, origin=(2, 2, 0), domain=dom)
        ds(in_field, out_field, coeff)        # per step; every rank alike

    The halo widths come from ``stencil.field_info[name].boundary`` (backend/module_generator.py:80-96 in the
    reference), the boxes from the exchanger's decomposition.  One call = (1) the refresh of the registered fields on
    the exchanger's own stream, (2) at the same time the stencil on the INTERIOR of the domain -- the points that read
    no neighbour-owned halo cell -- on the caller's stream, (3) once the refresh has arrived, the stencil on the
    boundary strips.  This is the "exchange overlapped with interior compute on separate streams" of the north star.
    `split=False` runs refresh -> whole domain in order instead.  Fields must be registered BEFORE exchanger.finalize().
    Build the stencil with ``device_sync=False`` (the pieces are ordered by streams and events, not by host waits).
    Not for stencils with ``horizontal(region[...])`` blocks: regions are relative to the domain of each call, and
    neither the reference nor this backend has a notion of a global domain.
    """

    def __init__(self, stencil, exchanger: "HaloExchanger", halo_fields: Dict[str, object], *, origin, domain,
                 split: bool = True):
        from . import _lib

        self.stencil, self.ex, self.split = stencil, exchanger, bool(split)
        self.domain = tuple(int(d) for d in domain)
        self.origin = origin if isinstance(origin, dict) else {"_all_": tuple(int(o) for o in origin)}
        group, lo_i = [], [0, 0, 0, 0]
        for name, arr in halo_fields.items():
            h = halo_from_field_info(stencil, name)
            org = self.origin.get(name, self.origin.get("_all_"))
            group.append((arr, tuple(org)[:3], self.domain, h))
            lo_i = [max(lo_i[0], h[0][0]), max(lo_i[1], h[0][1]), max(lo_i[2], h[1][0]), max(lo_i[3], h[1][1])]
        self.halo = ((lo_i[0], lo_i[1]), (lo_i[2], lo_i[3]))
        self.group = exchanger.add_group(group)
        exchanger.enable_multi_stream()
        self.interior, self.strips = split_domain(self.domain, self.halo,
                                                  lambda dx, dy: exchanger.decomp.neighbour(dx, dy) is not None)
        self.stream_h = _lib.Stream()
        self.ev_start, self.ev_done = _lib.Event(), _lib.Event()

    def __call__(self, *args, **kwargs):
        from .stencil_object import current_stream

        st = self.stencil
        names = type(st)._arg_names_
        bound = dict(zip(names, args))
        bound.update(kwargs)
        if not hasattr(self, "_full_origin"):
            all_org = self.origin.get("_all_")
            self._full_origin = {n: tuple(self.origin.get(n, all_org)) + (0,) * len(fi.data_dims)
                                 for n, fi in st.field_info.items() if fi is not None}
        stream = current_stream()
        self.ev_start.record(stream)
        self.stream_h.wait_event(self.ev_start)          # not before the producers of the fields on the caller's stream
        self.ex.refresh(self.group, self.stream_h)
        self.ev_done.record(self.stream_h)
        do_split = self.split and self.interior is not None and self.strips
        if do_split:
            call_on_pieces(st, bound, self._full_origin, [self.interior])
        L = self.ex._lib
        L.check(L.lib().b200_stream_wait_event(L.stream_handle(stream), self.ev_done.handle))
        call_on_pieces(st, bound, self._full_origin, self.strips if do_split else [((0, 0, 0), self.domain)])
