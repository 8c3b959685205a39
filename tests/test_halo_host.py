"""CPU tests of the multi-GPU host logic: process grid, neighbour map, halo boxes (from the stencils'
field_info boundaries), single-process assembly, and a world_size-2 gloo run that checks the decomposed
hdiff / vadv against the single-domain oracle."""
import os
import socket

import numpy as np
import pytest

from gt4py_b200.halo import DIRECTIONS, Decomposition, halo_boxes, process_grid
from oracle import cpu as oracle
from tests import halo_helpers as hh


def test_process_grid():
    assert [process_grid(n) for n in (1, 2, 4, 8, 6)] == [(1, 1), (2, 1), (2, 2), (4, 2), (3, 2)]
    with pytest.raises(ValueError):
        Decomposition.create(4, 0, (3, 1))


def test_neighbours_open_boundaries():
    d = Decomposition.create(8, 5)  # 4x2 grid, coords (1, 1)
    assert d.coords == (1, 1)
    assert d.neighbour(-1, 0) == 4 and d.neighbour(1, 0) == 6 and d.neighbour(0, -1) == 1
    assert d.neighbour(0, 1) is None and d.neighbour(1, 1) is None
    assert Decomposition.create(8, 0).neighbour(-1, 0) is None


def test_hdiff_boxes_include_corners():
    d = Decomposition.create(4, 0)  # 2x2, rank 0 at (0, 0): neighbours E, N, NE
    boxes = halo_boxes(d, (2, 2, 0), (16, 12, 5), ((2, 2), (2, 2)))
    got = {b.direction: b for b in boxes}
    assert set(got) == {(1, 0), (0, 1), (1, 1)}
    e = got[(1, 0)]
    assert e.src_rank == 1 and e.dst_start == (18, 2, 0) and e.src_start == (2, 2, 0) and e.extent == (2, 12, 5)
    c = got[(1, 1)]
    assert c.src_rank == 3 and c.dst_start == (18, 14, 0) and c.src_start == (2, 2, 0) and c.extent == (2, 2, 5)


def test_vadv_wcon_box_is_one_column_from_plus_i():
    boxes = halo_boxes(Decomposition.create(2, 0), (0, 0, 0), (8, 6, 4), ((0, 1), (0, 0)))
    assert len(boxes) == 1 and boxes[0].direction == (1, 0) and boxes[0].extent == (1, 6, 4)
    assert boxes[0].dst_start == (8, 0, 0) and boxes[0].src_start == (0, 0, 0)
    assert halo_boxes(Decomposition.create(2, 1), (0, 0, 0), (8, 6, 4), ((0, 1), (0, 0))) == []


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_assembly_restores_global_field(world):
    dom, hw = (10, 8, 3), 2
    G = hh.global_field(world, dom, hw, seed=world)
    decs = [Decomposition.create(world, r) for r in range(world)]
    locs = [hh.local_block(G, d, dom, hw) for d in decs]
    ref = [a.copy() for a in locs]
    for r, d in enumerate(decs):  # wipe every halo cell that a neighbour owns
        for b in halo_boxes(d, (hw, hw, 0), dom, ((hw, hw), (hw, hw))):
            locs[r][hh.box_slices(b.dst_start, b.extent)] = np.nan
    hh.apply_boxes_local(locs, world, (hw, hw, 0), dom, ((hw, hw), (hw, hw)))
    for a, b in zip(locs, ref):
        assert np.array_equal(a, b)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, dom, result):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dec = Decomposition.create(world, rank)
        ni, nj, nk = dom
        # hdiff: halo 2 everywhere
        G = hh.global_field(world, dom, 2, seed=3)
        coeff = np.full_like(G, 0.025)
        Gout = np.zeros_like(G)
        gdom = (dec.px_count * ni, dec.py_count * nj, nk)
        oracle.hdiff(G, Gout, coeff, origin=(2, 2, 0), domain=gdom)
        loc = hh.local_block(G, dec, dom, 2)
        for b in halo_boxes(dec, (2, 2, 0), dom, ((2, 2), (2, 2))):
            loc[hh.box_slices(b.dst_start, b.extent)] = -99.0  # stale halo
        hh.exchange_gloo(dist, loc, (2, 2, 0), dom, ((2, 2), (2, 2)))
        out = np.zeros_like(loc)
        oracle.hdiff(loc, out, np.full_like(loc, 0.025), origin=(2, 2, 0), domain=dom)
        ok_h = np.array_equal(out[2:-2, 2:-2], hh.local_block(Gout, dec, dom, 2)[2:-2, 2:-2])
        # vadv: wcon needs one column from +I
        rng = np.random.default_rng(5)
        gshape = (gdom[0] + 1, gdom[1], nk)
        F = {k: rng.uniform(-1, 1, gshape) for k in ("utens_stage", "u_stage", "u_pos", "utens")}
        F["wcon"] = rng.uniform(-0.1, 0.1, gshape)
        ref = F["utens_stage"].copy()
        oracle.vadv(ref, F["u_stage"], F["wcon"], F["u_pos"], F["utens"], dtr_stage=0.15, domain=gdom)
        x, y = dec.coords
        L = {k: v[x * ni: x * ni + ni + 1, y * nj: (y + 1) * nj].copy() for k, v in F.items()}
        if dec.neighbour(1, 0) is not None:
            L["wcon"][ni] = 1e9  # stale halo column
        hh.exchange_gloo(dist, L["wcon"], (0, 0, 0), dom, ((0, 1), (0, 0)))
        oracle.vadv(L["utens_stage"], L["u_stage"], L["wcon"], L["u_pos"], L["utens"], dtr_stage=0.15, domain=dom)
        ok_v = np.array_equal(L["utens_stage"][:ni], ref[x * ni: (x + 1) * ni, y * nj: (y + 1) * nj])
        result[rank] = (bool(ok_h), bool(ok_v))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_decomposed_stencils_match_global_oracle_gloo(world):
    import torch.multiprocessing as mp

    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), (12, 10, 5), result), nprocs=world, join=True)
    assert dict(result) == {r: (True, True) for r in range(world)}


def test_split_domain_pieces_are_disjoint_and_cover_the_domain():
    """Interior + boundary strips (halo.split_domain, the overlap scheme of DistributedStencil) tile the domain, the
    interior stays `halo` away from every side that has a neighbour, open sides contribute no strip."""
    from gt4py_b200.halo import Decomposition, split_domain

    for world, grid in ((1, None), (2, None), (4, None), (8, None), (6, (3, 2))):
        for rank in range(world):
            dec = Decomposition.create(world, rank, grid)
            for halo in (((2, 2), (2, 2)), ((0, 1), (0, 0)), ((3, 0), (1, 2))):
                for dom in ((20, 14, 3), (5, 4, 2), (3, 40, 1)):
                    interior, strips = split_domain(dom, halo, lambda dx, dy: dec.neighbour(dx, dy) is not None)
                    cover = np.zeros(dom[:2], int)
                    for o, e in ([interior] if interior else []) + strips:
                        assert all(x > 0 for x in e) and e[2] == dom[2]
                        cover[o[0]:o[0] + e[0], o[1]:o[1] + e[1]] += 1
                    assert (cover == 1).all(), (world, rank, halo, dom)
                    if interior is not None:
                        (oi, oj, _), (ei, ej, _) = interior
                        assert oi == (halo[0][0] if dec.neighbour(-1, 0) is not None else 0)
                        assert dom[0] - oi - ei == (halo[0][1] if dec.neighbour(1, 0) is not None else 0)
                        assert oj == (halo[1][0] if dec.neighbour(0, -1) is not None else 0)
                        assert dom[1] - oj - ej == (halo[1][1] if dec.neighbour(0, 1) is not None else 0)
                    if world == 1:
                        assert strips == [] and interior == ((0, 0, 0), dom)
