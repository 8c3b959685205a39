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
    """Refreshes the halos of registered device fields from the neighbour ranks with one kernel launch."""

    BLOCKS_PER_BOX = 8

    def __init__(self, comm, decomp: Decomposition, use_flags: bool = True):
        """`use_flags=False` skips the inter-rank flag protocol: only for single-process emulations that
        run several ranks on ONE device and order them on the host."""
        from . import _lib, storage

        self.use_flags = bool(use_flags)
        self._lib, self._storage = _lib, storage
        self.comm, self.decomp = comm, decomp
        self._groups: List[List[Tuple[object, List[BoxSpec]]]] = []
        self._finalized = False
        self.step = 0
        self.launches_per_refresh = 1

    # -- registration ---------------------------------------------------------------------------------
    def add_group(self, fields: Sequence[Tuple[object, Sequence[int], Sequence[int], Sequence[Tuple[int, int]]]]) -> int:
        """Register a group of fields refreshed together: (DeviceArray, origin, domain, halo widths).
        Every rank must register the same groups in the same order.  Returns the group index."""
        if self._finalized:
            raise RuntimeError("HaloExchanger: add_group after finalize")
        group = []
        for arr, origin, domain, halo in fields:
            if arr.dtype != np.float64 or arr.ndim != 3:
                raise TypeError("halo refresh supports 3-D float64 fields")
            group.append((arr, halo_boxes(self.decomp, origin, domain, halo)))
        self._groups.append(group)
        return len(self._groups) - 1

    def finalize(self):
        """Swap IPC handles with the other ranks and build the device-side box tables."""
        self._connect(self.comm.allgather(self._export()))
        self.comm.barrier()  # nobody starts pulling before every mapping and flag block exists

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
        L.synchronize()

        def handle_of(base_ptr):
            h = (ctypes.c_ubyte * 64)()
            L.check(lib.b200_ipc_get_mem_handle(c_void_p(base_ptr), h))
            return bytes(h)

        # (IPC handle of the allocation, its base address in the owner's address space)
        mine = {"pid": os.getpid(), "flags": (handle_of(self._flag_buf.ptr), self._flag_buf.ptr), "fields": []}
        for group in self._groups:
            mine["fields"].append([((handle_of(arr._buf.ptr), arr._buf.ptr), arr.ptr - arr._buf.ptr, arr.strides)
                                   for arr, _ in group])
        return mine

    def _connect(self, everyone):
        lib, L = self._lib.lib(), self._lib
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
            for _, boxes in group:
                for b in boxes:
                    used.add(b.direction)
        partners = []
        for d in DIRECTIONS:
            if (d in used or (-d[0], -d[1]) in used) and self.decomp.neighbour(*d) is not None:
                partners.append(d)
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

        self._tables = []
        for g, group in enumerate(self._groups):
            entries = []
            for f, (arr, boxes) in enumerate(group):
                item = arr.dtype.itemsize
                for b in boxes:
                    handle, offset, strides = everyone[b.src_rank]["fields"][g][f]
                    src_base = mapped(b.src_rank, handle) + offset
                    hb = L.HaloBox()
                    hb.src = src_base + sum(s * st for s, st in zip(b.src_start, strides))
                    hb.dst = arr.ptr + sum(s * st for s, st in zip(b.dst_start, arr.strides))
                    for d in range(3):
                        hb.src_stride[d] = strides[d] // item
                        hb.dst_stride[d] = arr.strides[d] // item
                        hb.extent[d] = b.extent[d]
                    entries.append(hb)
            n = len(entries)
            table = (L.HaloBox * max(n, 1))(*entries)
            dev = self._storage._Buffer(max(n, 1) * ctypes.sizeof(L.HaloBox))
            if n:
                L.check(lib.b200_memcpy_h2d(c_void_p(dev.ptr), ctypes.addressof(table), n * ctypes.sizeof(L.HaloBox), None))
            self._tables.append((dev, n))
        L.synchronize()
        self._finalized = True

    # -- per-step operations ---------------------------------------------------------------------------
    def refresh(self, group: int, stream=None):
        """Refresh the halos of every field of `group` (asynchronous on `stream`)."""
        if not self._finalized:
            raise RuntimeError("HaloExchanger: call finalize() first")
        self.step += 1
        dev, n = self._tables[group]
        L = self._lib
        L.check(L.lib().b200_halo_refresh(c_void_p(dev.ptr), n, self.BLOCKS_PER_BOX, byref(self._sync),
                                          self.step, c_void_p(self._arrive.ptr), int(self.use_flags),
                                          L.stream_handle(stream)))

    def wait_neighbours_done(self, stream=None):
        """Make `stream` wait until every neighbour has finished pulling the latest refresh out of this
        rank's fields -- call before overwriting a field that neighbours read."""
        L = self._lib
        if self.use_flags:
            L.check(L.lib().b200_halo_wait_done(byref(self._sync), self.step, L.stream_handle(stream)))

    # -- bench convenience -------------------------------------------------------------------------------
    @classmethod
    def for_bench(cls, dist, rank, world, local, hsets, vsets, dom, mode="p2p"):
        if mode != "p2p":
            raise NotImplementedError("only the peer-memory pull path is implemented")
        ex = cls(TorchComm(dist), Decomposition.create(world, rank))
        for h, v in zip(hsets, vsets):
            ex.add_group([
                (h["in_field"], (2, 2, 0), dom, ((2, 2), (2, 2))),   # field_info['in_field'].boundary
                (v["wcon"], (0, 0, 0), dom, ((0, 1), (0, 0))),       # field_info['wcon'].boundary
            ])
        ex.finalize()
        return ex
