#!/usr/bin/env python
"""Multi-GPU check of the halo layer (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tools/halo_check.py

Each rank owns a block of a global field with stale halos, refreshes them over NVLink peer memory, runs
hdiff + vadv on its block and compares with the single-domain oracle; then times the refresh.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import gt4py_b200 as gb  # noqa: E402
from gt4py_b200 import _lib, stencils, storage  # noqa: E402
from gt4py_b200.halo import Decomposition, HaloExchanger, TorchComm, halo_boxes  # noqa: E402
from oracle import cpu as oracle  # noqa: E402
from tests import halo_helpers as hh  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    _lib.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dom = (96, 64, 8)
    ni, nj, nk = dom
    d = Decomposition.create(world, rank)
    G = hh.global_field(world, dom, 2, seed=21)
    gdom = (d.px_count * ni, d.py_count * nj, nk)
    Gout = np.zeros_like(G)
    oracle.hdiff(G, Gout, np.full_like(G, 0.025), origin=(2, 2, 0), domain=gdom)
    rng = np.random.default_rng(22)
    gshape = (gdom[0] + 1, gdom[1], nk)
    F = {k: rng.uniform(-1, 1, gshape) for k in ("utens_stage", "u_stage", "u_pos", "utens")}
    F["wcon"] = rng.uniform(-0.1, 0.1, gshape)
    Vref = F["utens_stage"].copy()
    oracle.vadv(Vref, F["u_stage"], F["wcon"], F["u_pos"], F["utens"], dtr_stage=0.15, domain=gdom)
    x, y = d.coords
    loc = hh.local_block(G, d, dom, 2)
    for b in halo_boxes(d, (2, 2, 0), dom, ((2, 2), (2, 2))):
        loc[hh.box_slices(b.dst_start, b.extent)] = -99.0
    L = {k: v[x * ni: x * ni + ni + 1, y * nj: (y + 1) * nj].copy() for k, v in F.items()}
    if d.neighbour(1, 0) is not None:
        L["wcon"][ni] = 1e9
    f = {"in_field": storage.from_array(loc, aligned_index=(2, 2, 0)),
         "out_field": storage.zeros(loc.shape, aligned_index=(2, 2, 0)),
         "coeff": storage.from_array(np.full_like(loc, 0.025), aligned_index=(2, 2, 0))}
    f.update({k: storage.from_array(v) for k, v in L.items()})
    ex = HaloExchanger(TorchComm(dist), d, mode=os.environ.get("HALO_MODE", "push"))
    ex.add_group([(f["in_field"], (2, 2, 0), dom, ((2, 2), (2, 2))), (f["wcon"], (0, 0, 0), dom, ((0, 1), (0, 0)))])
    ex.finalize()
    ex.refresh(0)
    stencils.horizontal_diffusion(f["in_field"], f["out_field"], f["coeff"], origin=(2, 2, 0), domain=dom)
    stencils.vertical_advection_dycore(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                                       dtr_stage=0.15, origin=(0, 0, 0), domain=dom)
    ok_h = np.array_equal(f["out_field"].asnumpy()[2:-2, 2:-2], hh.local_block(Gout, d, dom, 2)[2:-2, 2:-2])
    ok_v = np.array_equal(f["utens_stage"].asnumpy()[:ni], Vref[x * ni: (x + 1) * ni, y * nj: (y + 1) * nj])
    # timing of the refresh alone
    s = gb.Stream()
    for _ in range(5):
        ex.refresh(0, s)
    s.synchronize()
    dist.barrier()
    e0, e1 = gb.Event(), gb.Event()
    e0.record(s)
    for _ in range(50):
        ex.refresh(0, s)
    e1.record(s)
    s.synchronize()
    us = e0.elapsed_ms(e1) / 50 * 1e3
    res = [None] * world
    dist.all_gather_object(res, {"rank": rank, "hdiff_ok": bool(ok_h), "vadv_ok": bool(ok_v), "refresh_us": round(us, 1)})
    if rank == 0:
        print(json.dumps(res))
        assert all(r["hdiff_ok"] and r["vadv_ok"] for r in res)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
