"""Test-only helpers for the halo layer: a numpy transport that applies BoxSpecs between host arrays
(single process) or over a torch.distributed gloo group (world_size > 1 on CPU)."""
import numpy as np

from gt4py_b200.halo import Decomposition, halo_boxes


def box_slices(start, extent):
    return tuple(slice(s, s + e) for s, e in zip(start, extent))


def apply_boxes_local(arrays_by_rank, world, origin, domain, halo, grid=None):
    """Single process: refresh the halo of every rank's array from the other ranks' arrays."""
    snapshot = [a.copy() for a in arrays_by_rank]
    for r in range(world):
        dec = Decomposition.create(world, r, grid)
        for b in halo_boxes(dec, origin, domain, halo):
            arrays_by_rank[r][box_slices(b.dst_start, b.extent)] = snapshot[b.src_rank][box_slices(b.src_start, b.extent)]


def exchange_gloo(dist, local, origin, domain, halo, grid=None):
    """world_size > 1 on CPU: every rank sends what its neighbours pull and receives its own boxes."""
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    mine = halo_boxes(Decomposition.create(world, rank, grid), origin, domain, halo)
    sends = []
    for other in range(world):
        if other == rank:
            continue
        for b in halo_boxes(Decomposition.create(world, other, grid), origin, domain, halo):
            if b.src_rank == rank:
                sends.append((other, b))
    ops, recv_bufs = [], []
    for other, b in sends:
        t = torch.from_numpy(np.ascontiguousarray(local[box_slices(b.src_start, b.extent)]))
        ops.append(dist.P2POp(dist.isend, t, other, tag=DIR_TAG(b.direction)))
    for b in mine:
        t = torch.empty(b.extent, dtype=torch.float64)
        recv_bufs.append((b, t))
        ops.append(dist.P2POp(dist.irecv, t, b.src_rank, tag=DIR_TAG(b.direction)))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    for b, t in recv_bufs:
        local[box_slices(b.dst_start, b.extent)] = t.numpy()


def DIR_TAG(d):
    return (d[0] + 1) * 3 + (d[1] + 1)


def global_field(world, domain, halo_w, seed, grid=None):
    from gt4py_b200.halo import process_grid

    px, py = grid or process_grid(world)
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, (px * domain[0] + 2 * halo_w, py * domain[1] + 2 * halo_w, domain[2]))


def local_block(G, dec, domain, halo_w):
    x, y = dec.coords
    return G[x * domain[0]: x * domain[0] + domain[0] + 2 * halo_w,
             y * domain[1]: y * domain[1] + domain[1] + 2 * halo_w].copy()
