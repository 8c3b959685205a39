"""GPU parity (-m gpu): the CUDA path, called through the C ABI via the stencil objects, against
 (a) the golden fixtures produced by the reference numpy backend (tests/golden), and
 (b) the pinned CPU oracle on seeded inputs, incl. ragged sizes and the BASELINE config sizes.

Bar: BIT-EXACT for float64 (stricter than north_star's 1e-12 relative): the kernels keep the
reference's operator association and are built with -fmad=false.
"""
import glob
import os

import numpy as np
import pytest

import gt4py_b200 as gb
from gt4py_b200 import stencils, storage
from oracle import cpu as oracle
from oracle import inputs as gen

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _cases(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "_[0-9]*.npz")))


def _dev(a, origin):
    return storage.from_array(a, backend="b200", aligned_index=origin)


def run_hdiff(f, domain, variant, limiter=True):
    d = {k: _dev(v, (2, 2, 0)) for k, v in f.items()}
    if limiter:
        st = stencils.HorizontalDiffusion.with_variant(variant)
        st(d["in_field"], d["out_field"], d["coeff"], origin=(2, 2, 0), domain=domain)
    else:
        st = stencils.SimpleHorizontalDiffusion.with_variant(variant)
        st(d["in_field"], d["coeff"], d["out_field"], origin=(2, 2, 0), domain=domain)
    return d["out_field"].asnumpy()


def run_vadv(f, domain, variant):
    d = {k: _dev(v, (0, 0, 0)) for k, v in f.items() if k != "dtr_stage"}
    st = stencils.VerticalAdvectionDycore.with_variant(variant)
    try:
        st(d["utens_stage"], d["u_stage"], d["wcon"], d["u_pos"], d["utens"],
           dtr_stage=f["dtr_stage"], origin=(0, 0, 0), domain=domain)
    except gb.B200Error as exc:
        if exc.code == -3 and variant >= 2:  # B200_EUNSUP: this tuned kernel's preconditions fail
            pytest.skip(str(exc))
        raise
    return d["utens_stage"].asnumpy()


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("path", _cases("hdiff"), ids=os.path.basename)
def test_hdiff_golden(path, variant):
    g = np.load(path)
    dom = tuple(int(x) for x in g["shape"])
    out = run_hdiff(gen.hdiff_inputs(*dom, seed=int(g["seed"])), dom, variant)
    assert np.array_equal(out, g["out_field"])  # halo of out_field must stay untouched (zeros)


@pytest.mark.parametrize("variant", [1, 2, 4, 5])
@pytest.mark.parametrize("path", _cases("simple_hdiff"), ids=os.path.basename)
def test_simple_hdiff_golden(path, variant):
    g = np.load(path)
    dom = tuple(int(x) for x in g["shape"])
    out = run_hdiff(gen.hdiff_inputs(*dom, seed=int(g["seed"])), dom, variant, limiter=False)
    assert np.array_equal(out, g["out_field"])


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11])
@pytest.mark.parametrize("path", _cases("vadv"), ids=os.path.basename)
def test_vadv_golden(path, variant):
    g = np.load(path)
    dom = tuple(int(x) for x in g["shape"])
    out = run_vadv(gen.vadv_inputs(*dom, seed=int(g["seed"])), dom, variant)
    assert np.array_equal(out, g["utens_stage"])


@pytest.mark.parametrize("variant", [0, 1, 2, 8])
@pytest.mark.parametrize("path", _cases("tridiag"), ids=os.path.basename)
def test_tridiag_golden(path, variant):
    g = np.load(path)
    dom = tuple(int(x) for x in g["shape"])
    f = gen.tridiag_inputs(*dom, seed=int(g["seed"]))
    d = {k: _dev(v, (0, 0, 0)) for k, v in f.items()}
    st = stencils.TridiagonalSolver.with_variant(variant)
    st(d["inf"], d["diag"], d["sup"], d["rhs"], d["out"], domain=dom)
    for name in ("sup", "rhs", "out"):
        assert np.array_equal(d[name].asnumpy(), g[name]), name


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("dom", [(128, 64, 80), (61, 59, 5), (1, 1, 2), (129, 3, 9), (33, 17, 8), (40, 5, 17),
                                 (300, 2, 16), (7, 7, 81), (256, 300, 2), (70, 9, 127), (130, 66, 128), (64, 8, 131)])
def test_tridiag_vs_oracle(dom, variant):
    """ragged sizes around the prefetch depth (8) of the forward and the backward sweep"""
    f = gen.tridiag_inputs(*dom, seed=77)
    d = {k: _dev(v, (0, 0, 0)) for k, v in f.items()}
    try:
        stencils.TridiagonalSolver.with_variant(variant)(d["inf"], d["diag"], d["sup"], d["rhs"], d["out"], domain=dom)
    except gb.B200Error as exc:
        if exc.code == -3 and variant == 8:  # B200_EUNSUP: the fused kernel's preconditions fail
            pytest.skip(str(exc))
        raise
    g = {k: v.copy() for k, v in f.items()}
    oracle.tridiag(g["inf"], g["diag"], g["sup"], g["rhs"], g["out"], domain=dom)
    for name in ("sup", "rhs", "out"):
        assert np.array_equal(d[name].asnumpy(), g[name]), name


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5])
@pytest.mark.parametrize("dom", [(128, 128, 80), (61, 59, 5), (60, 8, 1), (121, 33, 2), (2, 2, 2),
                                 (300, 7, 3), (64, 100, 3), (17, 65, 2), (241, 130, 2), (180, 257, 1)])
def test_hdiff_vs_oracle(dom, variant):
    f = gen.hdiff_inputs(*dom, seed=123)
    ref = oracle.hdiff(f["in_field"], f["out_field"].copy(), f["coeff"], domain=dom)
    out = run_hdiff(f, dom, variant)
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11])
@pytest.mark.parametrize("dom", [(256, 256, 80), (33, 5, 3), (31, 7, 80), (64, 3, 17), (1, 1, 4),
                                 (100, 6, 137), (129, 2, 5), (300, 37, 7), (65, 130, 79), (40, 9, 82), (20, 4, 6),
                                 (64, 40, 128), (130, 3, 200), (260, 150, 111)])
def test_vadv_vs_oracle(dom, variant):
    f = gen.vadv_inputs(*dom, seed=99)
    ref = oracle.vadv(f["utens_stage"].copy(), f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                      dtr_stage=f["dtr_stage"], domain=dom)
    out = run_vadv(f, dom, variant)
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("dom", [(128, 128, 8), (61, 59, 5), (60, 8, 1), (121, 33, 2), (2, 2, 2), (300, 7, 3)])
def test_hdiff_float32_vs_oracle(dom, variant):
    """The pure-float32 build (float32 fields and literals) on the library kernel with T = float: bit-exact
    against the numpy float32 restatement that the reference's own output pins (tests/test_oracle.py)."""
    from oracle.hdiff_f32 import hdiff_f32

    f = {k: v.astype(np.float32) for k, v in gen.hdiff_inputs(*dom, seed=321).items()}
    ref = hdiff_f32(f["in_field"], f["out_field"].copy(), f["coeff"], domain=dom)
    d = {k: _dev(v, (2, 2, 0)) for k, v in f.items()}
    st = stencils.HorizontalDiffusionFloat32.with_variant(variant)
    st(d["in_field"], d["out_field"], d["coeff"], origin=(2, 2, 0), domain=dom)
    out = d["out_field"].asnumpy()
    assert out.dtype == np.float32 and np.array_equal(out, ref)


def test_hdiff_float32_rejects_float64_only_variants():
    dom = (32, 8, 2)
    f = {k: v.astype(np.float32) for k, v in gen.hdiff_inputs(*dom, seed=1).items()}
    d = {k: _dev(v, (2, 2, 0)) for k, v in f.items()}
    with pytest.raises(gb.B200Error) as exc:
        stencils.HorizontalDiffusionFloat32.with_variant(5)(d["in_field"], d["out_field"], d["coeff"],
                                                            origin=(2, 2, 0), domain=dom)
    assert exc.value.code == -3


def test_hdiff_subdomain_and_offset_origin():
    """origin/domain select a sub-box: points outside it must be untouched."""
    f = gen.hdiff_inputs(40, 30, 4, seed=5)
    f["out_field"][...] = -7.0
    origin, dom = (5, 4, 1), (17, 11, 2)
    d = {k: _dev(v, origin) for k, v in f.items()}
    stencils.horizontal_diffusion(d["in_field"], d["out_field"], d["coeff"], origin=origin, domain=dom)
    ref = f["out_field"].copy()
    oracle.hdiff(f["in_field"], ref, f["coeff"], origin=origin, domain=dom)
    assert np.array_equal(d["out_field"].asnumpy(), ref)


def test_hdiff_non_optimal_layout_uses_generic_kernel():
    """C-ordered (K-contiguous) device arrays: layout warning, same result."""
    f = gen.hdiff_inputs(20, 18, 6, seed=8)
    dom = (20, 18, 6)
    ref = oracle.hdiff(f["in_field"], f["out_field"].copy(), f["coeff"], domain=dom)
    d = {}
    for k, v in f.items():
        base = storage.from_array(np.ascontiguousarray(v.transpose(2, 1, 0)), backend="b200")
        d[k] = base.transpose(2, 1, 0)  # logical (I, J, K), strides K-fastest
    with pytest.warns(UserWarning, match="layout of the field"):
        stencils.horizontal_diffusion(d["in_field"], d["out_field"], d["coeff"],
                                      origin=(2, 2, 0), domain=dom)
    assert np.array_equal(d["out_field"].asnumpy(), ref)


def test_full_size_properties_hdiff():
    """BASELINE config 3 size (1024x1024x80): size-independent properties instead of a full oracle run.
    (1) constant field -> out == in exactly; (2) k-planes are independent: plane k of the full run
    equals a single-plane run; (3) a 96-row slab matches the oracle bit-for-bit."""
    ni, nj, nk = 1024, 1024, 80
    f = gen.hdiff_inputs(ni, nj, nk, seed=42)
    d = {k: _dev(v, (2, 2, 0)) for k, v in f.items()}
    stencils.horizontal_diffusion(d["in_field"], d["out_field"], d["coeff"], origin=(2, 2, 0),
                                  domain=(ni, nj, nk))
    out = d["out_field"].asnumpy()
    sl = (slice(0, ni + 4), slice(500, 600), slice(37, 39))
    sub = {k: np.ascontiguousarray(v[sl]) for k, v in f.items()}
    ref = oracle.hdiff(sub["in_field"], sub["out_field"], sub["coeff"], domain=(ni, 96, 2))
    assert np.array_equal(out[2:-2, 502:598, 37:39], ref[2:-2, 2:-2, :])
    assert not out[:2].any() and not out[:, -2:].any()
    const = storage.full((ni + 4, nj + 4, 2), 3.25, backend="b200", aligned_index=(2, 2, 0))
    o2 = storage.zeros((ni + 4, nj + 4, 2), backend="b200", aligned_index=(2, 2, 0))
    stencils.horizontal_diffusion(const, o2, d["coeff"][:, :, :2], origin=(2, 2, 0), domain=(ni, nj, 2))
    assert np.all(o2.asnumpy()[2:-2, 2:-2] == 3.25)


def test_full_size_properties_vadv():
    """1024x1024x80: columns are independent -> a strip of columns equals the oracle on that strip."""
    ni, nj, nk = 1024, 1024, 80
    f = gen.vadv_inputs(ni, nj, nk, seed=7)
    out = run_vadv(f, (ni, nj, nk), 0)
    sub = {k: (np.ascontiguousarray(v[:, 300:304]) if k != "dtr_stage" else v) for k, v in f.items()}
    ref = oracle.vadv(sub["utens_stage"], sub["u_stage"], sub["wcon"], sub["u_pos"], sub["utens"],
                      dtr_stage=sub["dtr_stage"], domain=(ni, 4, nk))
    assert np.array_equal(out[:ni, 300:304], ref[:ni])
    assert np.array_equal(out[ni], f["utens_stage"][ni])  # the i+1 halo column is not written


@pytest.mark.parametrize("mode", ["push", "pull"])
@pytest.mark.parametrize("world", [2, 4])
def test_halo_refresh_single_device_emulation(world, mode):
    """All ranks of a Px x Py decomposition emulated in this process on one device (no flags, host
    ordering): the pull kernel must restore stale halos so that the decomposed hdiff / vadv equal the
    single-domain oracle bit-for-bit."""
    from gt4py_b200.halo import Decomposition, HaloExchanger
    from tests import halo_helpers as hh

    dom = (40, 24, 6)
    ni, nj, nk = dom
    decs = [Decomposition.create(world, r) for r in range(world)]
    G = hh.global_field(world, dom, 2, seed=11)
    gdom = (decs[0].px_count * ni, decs[0].py_count * nj, nk)
    Gout = np.zeros_like(G)
    oracle.hdiff(G, Gout, np.full_like(G, 0.025), origin=(2, 2, 0), domain=gdom)
    rng = np.random.default_rng(12)
    gshape = (gdom[0] + 1, gdom[1], nk)
    F = {k: rng.uniform(-1, 1, gshape) for k in ("utens_stage", "u_stage", "u_pos", "utens")}
    F["wcon"] = rng.uniform(-0.1, 0.1, gshape)
    Vref = F["utens_stage"].copy()
    oracle.vadv(Vref, F["u_stage"], F["wcon"], F["u_pos"], F["utens"], dtr_stage=0.15, domain=gdom)

    exs, dev = [], []
    for d in decs:
        x, y = d.coords
        loc = hh.local_block(G, d, dom, 2)
        from gt4py_b200.halo import halo_boxes
        for b in halo_boxes(d, (2, 2, 0), dom, ((2, 2), (2, 2))):
            loc[hh.box_slices(b.dst_start, b.extent)] = -99.0
        L = {k: v[x * ni: x * ni + ni + 1, y * nj: (y + 1) * nj].copy() for k, v in F.items()}
        if d.neighbour(1, 0) is not None:
            L["wcon"][ni] = 1e9
        fields = {"in_field": _dev(loc, (2, 2, 0)), "out_field": _dev(np.zeros_like(loc), (2, 2, 0)),
                  "coeff": _dev(np.full_like(loc, 0.025), (2, 2, 0))}
        fields.update({k: _dev(v, (0, 0, 0)) for k, v in L.items()})
        ex = HaloExchanger(None, d, use_flags=False, mode=mode)
        ex.add_group([(fields["in_field"], (2, 2, 0), dom, ((2, 2), (2, 2))),
                      (fields["wcon"], (0, 0, 0), dom, ((0, 1), (0, 0)))])
        exs.append(ex)
        dev.append(fields)
    HaloExchanger.finalize_local(exs)
    HaloExchanger.refresh_local(exs, 0)
    HaloExchanger.refresh_local(exs, 0)  # second refresh uses the other staging parity
    gb.synchronize()
    for d, f in zip(decs, dev):
        x, y = d.coords
        stencils.horizontal_diffusion(f["in_field"], f["out_field"], f["coeff"], origin=(2, 2, 0), domain=dom)
        stencils.vertical_advection_dycore(f["utens_stage"], f["u_stage"], f["wcon"], f["u_pos"], f["utens"],
                                           dtr_stage=0.15, origin=(0, 0, 0), domain=dom)
        assert np.array_equal(f["out_field"].asnumpy()[2:-2, 2:-2], hh.local_block(Gout, d, dom, 2)[2:-2, 2:-2])
        assert np.array_equal(f["utens_stage"].asnumpy()[:ni], Vref[x * ni: (x + 1) * ni, y * nj: (y + 1) * nj])


@pytest.mark.parametrize("mode", ["push", "pull"])
def test_halo_refresh_other_dtypes_and_data_dimensions(mode):
    """float32 with a data dimension, int32 and int8 fields, asymmetric halo widths: every halo cell that has an
    owner must arrive, cells outside the global domain stay untouched (single-device emulation, 2 x 2 ranks)."""
    from gt4py_b200.halo import Decomposition, HaloExchanger, halo_boxes
    from tests import halo_helpers as hh

    world, dom, halo = 4, (12, 10, 5), ((1, 2), (2, 1))
    ni, nj, nk = dom
    decs = [Decomposition.create(world, r) for r in range(world)]
    px, py = decs[0].px_count, decs[0].py_count
    rng = np.random.default_rng(5)
    lo_i, hi_i, lo_j, hi_j = halo[0][0], halo[0][1], halo[1][0], halo[1][1]
    gshape = (px * ni + lo_i + hi_i, py * nj + lo_j + hi_j, nk)
    G = {"f32": rng.uniform(-1, 1, gshape + (2,)).astype(np.float32),
         "i32": rng.integers(-1000, 1000, gshape).astype(np.int32),
         "i8": rng.integers(-100, 100, gshape).astype(np.int8)}
    origin = (lo_i, lo_j, 0)
    exs, dev, want = [], [], []
    for d in decs:
        x, y = d.coords
        fields, expect = {}, {}
        for name, g in G.items():
            loc = g[x * ni: x * ni + ni + lo_i + hi_i, y * nj: y * nj + nj + lo_j + hi_j].copy()
            expect[name] = loc.copy()
            for b in halo_boxes(d, origin, dom, halo):  # stale values in every cell a neighbour owns
                loc[hh.box_slices(b.dst_start, b.extent)] = 77
            dims = ["I", "J", "K"] + (["0"] if loc.ndim == 4 else [])
            fields[name] = storage.from_array(loc, aligned_index=origin + (0,) * (loc.ndim - 3), dimensions=dims)
        ex = HaloExchanger(None, d, use_flags=False, mode=mode)
        ex.add_group([(fields[n], origin, dom, halo) for n in G])
        exs.append(ex)
        dev.append(fields)
        want.append(expect)
    HaloExchanger.finalize_local(exs)
    HaloExchanger.refresh_local(exs, 0)
    gb.synchronize()
    for f, w in zip(dev, want):
        for name in G:
            assert np.array_equal(f[name].asnumpy(), w[name]), name


@pytest.mark.parametrize("dom", [(70, 45, 9), (128, 64, 80)])
def test_interior_plus_boundary_strips_equal_the_whole_call(dom):
    """DistributedStencil's split (halo.split_domain + call_on_pieces): hdiff and vadv run piece by piece -- the
    interior and the four / one boundary strips of a rank with neighbours on every side -- give the same bits as one
    call over the whole domain (and as the oracle)."""
    from gt4py_b200.halo import call_on_pieces, halo_from_field_info, split_domain

    hd, va = stencils.HorizontalDiffusion(), stencils.VerticalAdvectionDycore()
    f = gen.hdiff_inputs(*dom, seed=31)
    d = {k: _dev(v, (2, 2, 0)) for k, v in f.items()}
    interior, strips = split_domain(dom, halo_from_field_info(hd, "in_field"), lambda dx, dy: True)
    assert interior is not None and len(strips) == 4
    call_on_pieces(hd, d, {k: (2, 2, 0) for k in d}, [interior] + strips)
    ref = oracle.hdiff(f["in_field"], f["out_field"].copy(), f["coeff"], domain=dom)
    assert np.array_equal(d["out_field"].asnumpy(), ref)

    f = gen.vadv_inputs(*dom, seed=32)
    dtr = f.pop("dtr_stage")
    d = {k: _dev(v, (0, 0, 0)) for k, v in f.items()}
    interior, strips = split_domain(dom, halo_from_field_info(va, "wcon"), lambda dx, dy: True)
    assert interior == ((0, 0, 0), (dom[0] - 1, dom[1], dom[2])) and strips == [((dom[0] - 1, 0, 0), (1, dom[1], dom[2]))]
    call_on_pieces(va, dict(d, dtr_stage=dtr), {k: (0, 0, 0) for k in d}, [interior] + strips)
    ref = oracle.vadv(f["utens_stage"].copy(), f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=dtr, domain=dom)
    assert np.array_equal(d["utens_stage"].asnumpy(), ref)


@pytest.mark.parametrize("extra", [[], ["--split"]], ids=["overlap", "split"])
def test_two_gpu_run_checks_halo_parity(extra):
    """With two or more devices: bench.py on 2 ranks (small blocks, few steps) -- real peers, flags and staging --
    must report halo_parity true (its model changes the exchanged values every step, oracle/model_parity.py)."""
    import json
    import subprocess
    import sys

    if gb.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(root, "bench.py"), "--gpus", "2", "--domain", "256",
           "192", "40", "--steps", "6", "--warmup", "3", "--no-e2e", *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=root, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["halo_parity"] is True and line["parity"]["ok"] is True


@pytest.mark.parametrize("chunks", [1, 3, 8])
def test_host_pipeline_matches_oracle(chunks):
    """Host-buffer call path (chunked H2D / kernel / D2H on three streams): same bits as the oracle."""
    dom = (70, 37, 13)
    f = gen.hdiff_inputs(*dom, seed=4)
    dev = {k: _dev(np.zeros_like(v), (2, 2, 0)) for k, v in f.items()}
    host = {k: storage.pinned_like(dev[k], v) for k, v in f.items()}
    pipe = gb.HostPipeline(stencils.horizontal_diffusion, dev, origin=(2, 2, 0), domain=dom, axis="K", chunks=chunks)
    h2d, d2h = pipe(host)
    ref = oracle.hdiff(f["in_field"], f["out_field"].copy(), f["coeff"], domain=dom)
    assert np.array_equal(host["out_field"], ref)
    assert h2d >= 2 * host["in_field"].nbytes and d2h >= host["out_field"].nbytes  # + row padding

    f = gen.vadv_inputs(*dom, seed=6)
    dtr = f.pop("dtr_stage")
    dev = {k: _dev(np.zeros_like(v), (0, 0, 0)) for k, v in f.items()}
    host = {k: storage.pinned_like(dev[k], v) for k, v in f.items()}
    pipe = gb.HostPipeline(stencils.vertical_advection_dycore, dev, origin=(0, 0, 0), domain=dom, axis="J", chunks=chunks)
    pipe(host, dtr_stage=dtr)
    ref = oracle.vadv(f["utens_stage"].copy(), f["u_stage"], f["wcon"], f["u_pos"], f["utens"], dtr_stage=dtr, domain=dom)
    assert np.array_equal(host["utens_stage"], ref)
    for v in host.values():
        storage.free_pinned(v)


# ---- CUDA-graph replay of a captured call sequence (gt4py_b200.Graph) -------------------------------------------
def test_graph_replay_matches_direct_calls():
    """hdiff + vadv + a generated stencil captured once, replayed on new field contents (same buffers)."""
    from gt4py_b200 import generated

    dom = (48, 40, 12)
    hf = gen.hdiff_inputs(*dom, seed=11)
    vf = gen.vadv_inputs(*dom, seed=12)
    h = {k: _dev(v, (2, 2, 0)) for k, v in hf.items()}
    v = {k: _dev(a, (0, 0, 0)) for k, a in vf.items() if k != "dtr_stage"}
    chain_dir = os.path.join(GOLDEN, "generated", "fv3_chain")
    st_chain = generated.load_plan(chain_dir)
    rng = np.random.default_rng(5)
    cf = {k: rng.uniform(-1.0, 1.5, (dom[0] + 6, dom[1] + 6, dom[2])) for k in ("phi", "u", "v")}
    c = {k: _dev(a, (3, 3, 0)) for k, a in cf.items()}
    c_out = storage.zeros(dom)
    hd, va = stencils.HorizontalDiffusion(), stencils.VerticalAdvectionDycore()

    def sequence():
        hd(h["in_field"], h["out_field"], h["coeff"], origin=(2, 2, 0), domain=dom)
        va(v["utens_stage"], v["u_stage"], v["wcon"], v["u_pos"], v["utens"], dtr_stage=vf["dtr_stage"],
           origin=(0, 0, 0), domain=dom)
        st_chain(c["phi"], c["u"], c["v"], c_out, dt=0.05, alpha=0.03,
                 origin={"_all_": (3, 3, 0), "out": (0, 0, 0)}, domain=dom)

    sequence()  # warm-up: scratch allocation, module load, descriptor caches
    gb.synchronize()
    v["utens_stage"].set(vf["utens_stage"])  # vadv updates it in place
    g = gb.Graph()
    with g:
        sequence()
    for trial in range(3):
        hf2 = gen.hdiff_inputs(*dom, seed=20 + trial)
        vf2 = gen.vadv_inputs(*dom, seed=30 + trial)
        h["in_field"].set(hf2["in_field"])
        h["coeff"].set(hf2["coeff"])
        for k in ("utens_stage", "u_stage", "wcon", "u_pos", "utens"):
            v[k].set(vf2[k])
        phi2 = rng.uniform(-1.0, 1.5, cf["phi"].shape)
        c["phi"].set(phi2)
        gb.synchronize()
        g.launch()
        g.synchronize()
        ref_h = oracle.hdiff(hf2["in_field"], hf2["out_field"].copy(), hf2["coeff"], domain=dom)
        assert np.array_equal(h["out_field"].asnumpy(), ref_h)
        ref_v = oracle.vadv(vf2["utens_stage"].copy(), vf2["u_stage"], vf2["wcon"], vf2["u_pos"], vf2["utens"],
                            dtr_stage=vf["dtr_stage"], domain=dom)
        assert np.array_equal(v["utens_stage"].asnumpy(), ref_v)
        from oracle import chain

        assert np.array_equal(c_out.asnumpy(), chain.fv3_chain(phi2, cf["u"], cf["v"], 0.05, 0.03))


# ---- L2 access-policy window of a stream (b200_stream_l2_window) --------------------------------------------------
def test_l2_window_changes_caching_not_results():
    """Two generated stencils in a row, the second reading what the first wrote, on K slabs with the stream's
    persisting-L2 window on the slab of the intermediate field (tools/l2_residency.py is the measurement): the window
    is a caching hint only -- results are bit-identical with and without it, directly and from a captured graph."""
    from gt4py_b200 import _lib, generated
    from gt4py_b200.stencil_object import use_stream

    ni, nj, nk, h, slab = 96, 70, 12, 2, 4
    st = generated.load_plan(os.path.join(GOLDEN, "generated", "horizontal_diffusion"), use_library=False)
    f = gen.hdiff_inputs(ni, nj, nk, seed=31)
    want_mid = np.zeros_like(f["out_field"])
    oracle.hdiff(f["in_field"], want_mid, f["coeff"], origin=(h, h, 0), domain=(ni, nj, nk))
    dev = {k: _dev(v, (h, h, 0)) for k, v in f.items()}
    mid = {False: storage.zeros(f["in_field"].shape, aligned_index=(h, h, 0)),
           True: storage.zeros(f["in_field"].shape, aligned_index=(h, h, 0))}
    out = {False: storage.zeros(f["in_field"].shape, aligned_index=(h, h, 0)),
           True: storage.zeros(f["in_field"].shape, aligned_index=(h, h, 0))}
    stream = _lib.Stream()

    def one_pass(window):
        level = mid[window].strides[2]
        for k0 in range(0, nk, slab):
            if window:
                granted = stream.l2_window(mid[window].ptr + k0 * level, slab * level)
                assert granted == slab * level
            st(dev["in_field"], mid[window], dev["coeff"], origin=(h, h, k0), domain=(ni, nj, slab))
            st(mid[window], out[window], dev["coeff"], origin=(2 * h, 2 * h, k0), domain=(ni - 2 * h, nj - 2 * h, slab))

    with use_stream(stream):
        one_pass(False)
        one_pass(True)
    stream.synchronize()
    assert stream.l2_window() == 0  # window off, persisting lines reset
    assert np.array_equal(mid[True].asnumpy()[h:-h, h:-h], want_mid[h:-h, h:-h])
    assert np.array_equal(mid[True].asnumpy(), mid[False].asnumpy())
    assert np.array_equal(out[True].asnumpy(), out[False].asnumpy())
    assert np.abs(out[True].asnumpy()[2 * h:-2 * h, 2 * h:-2 * h]).sum() > 0
