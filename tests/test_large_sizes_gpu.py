"""Generated kernels at sizes that cross tile / block / row-march / 32-bit-index boundaries, ON THE GPU, against the
reference's numpy backend run in the same process (the staged reference oracle/_ref) -- VERDICT r1 weak #1: every
generated-plan GPU case used to be 13 x 11 x 10.

  * one case per kernel template and entry point: `point`, `column`, `tile`, the TMA-staged `_t`, a horizontal-region
    stencil, a data-dimension stencil, at >= 300 x 200 x 40 with odd extents;
  * BASELINE configs[3]: fv3_chain at 512 x 512 x 128, checked with oracle/chain.py on levels of the full IJ plane
    (the chain is PARALLEL in K; its horizontal regions need the whole plane);
  * the 32-bit -> 64-bit index switch (`<k>_u` -> `<k>`) with a field whose element span exceeds 2^31.

Bar: bit-exact (np.array_equal).
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import ref  # noqa: E402

pytestmark = pytest.mark.gpu

GEN = os.path.join(HERE, "golden", "generated")

#: fixture name -> (domain, kernel kinds expected, entry-point suffix expected of the first kernel or None)
CASES = {
    "exec_info_advection": ((303, 201, 40), ["point"], "_u"),
    "exec_info_diffusion": ((301, 203, 41), ["march"], "_u"),           # warp-march form (lowering_march.py)
    "fv3_chain": ((300, 200, 40), ["march"], "_u"),
    "horizontal_diffusion__tile": ((322, 197, 40), ["march"], "_u"),
    "exec_info_advection__tile": ((257, 190, 33), ["tile"], "_t"),      # two stages: the tile form is the faster one
    "exec_info_diffusion__smem": ((301, 203, 41), ["tile"], "_t"),      # shared-memory tile form, TMA-staged inputs
    "fv3_chain__smem": ((300, 200, 40), ["tile"], "_t"),
    "horizontal_diffusion__smem": ((322, 197, 40), ["tile"], "_t"),
    "column_physics": ((300, 200, 64), ["colfuse"], "_u"),          # fused FORWARD + BACKWARD (lowering_colfuse.py)
    "vertical_advection_dycore": ((300, 131, 80), ["colfuse"], "_u"),  # lowered kernels (library matching off)
    "tridiagonal_solver": ((257, 66, 80), ["colfuse"], "_u"),
    "vertical_advection_dycore__kcache": ((257, 130, 65), None, "_u"),
    "horizontal_region_with_conditional": ((300, 211, 40), None, None),
    "runtime_if": ((310, 200, 40), None, None),
    "seq_exchange": ((301, 130, 48), ["column", "column"], "_u"),           # loop fission (FORWARD)
    "seq_exchange_backward": ((259, 133, 37), ["column"] * 3, "_u"),        # ... BACKWARD, three stages
    "seq_exchange__level": ((301, 130, 48), ["level", "level"], "_u"),      # per-level launches (fission off)
    "seq_exchange_backward__level": ((259, 133, 37), ["level"] * 3, "_u"),
}


EMU = os.environ.get("GT4PY_B200_EMU") == "1"  # debugging aid: the host double of the C ABI, tiny domains


def _setup():
    if EMU:
        from tests import emu_abi

        emu_abi.activate()
    if not ref.available():
        pytest.fail("the reference is not staged: run `python oracle/make_ref.py` in the build container")
    ref.activate(gpu=True)
    import gt4py_b200.backend  # noqa: F401
    import make_generated_fixtures as mk
    from gt4py import storage as gt_storage
    from gt4py.cartesian import gtscript

    return mk, gtscript, gt_storage


def _compare(gtscript, gt_storage, mk, name, fn, ext, opts, common, domain, seed):
    ref_st = gtscript.stencil("numpy", fn, externals=ext, format_source=False, rebuild=True, name=name + "_n", **common)
    dev_st = gtscript.stencil("b200", fn, externals=ext, rebuild=True, name=name + "_b", **common, **opts)
    if EMU:
        domain = tuple(max(3, d // 12) for d in domain)
    domain = (domain[0], domain[1], max(domain[2], ref_st.domain_info.min_sequential_axis_size + 3))
    fields, params, origins = mk.make_inputs(ref_st, domain, seed=seed)
    host = {k: (None if v is None else v.copy()) for k, v in fields.items()}
    dims = {n: tuple(fi.axes) + tuple(str(d) for d in range(len(fi.data_dims)))
            for n, fi in ref_st.field_info.items() if fi is not None}
    dev = {k: (None if v is None else gt_storage.from_array(v, backend="b200", aligned_index=origins[k],
                                                            dimensions=dims[k])) for k, v in fields.items()}
    ref_st(**host, **params, origin=origins, domain=domain)
    dev_st(**dev, **params, origin=origins, domain=domain)
    written = [k for k, fi in ref_st.field_info.items() if fi is not None and int(fi.access) & 2]
    assert written
    for k in written:
        got = np.asarray(dev[k])
        assert np.array_equal(got, host[k], equal_nan=got.dtype.kind == "f"), (name, k)
    return type(dev_st)._b200()


@pytest.mark.parametrize("name", sorted(CASES))
def test_generated_kernel_at_a_size_that_crosses_boundaries(name):
    mk, gtscript, gt_storage = _setup()
    domain, kinds, suffix = CASES[name]
    fn, ext, *rest = mk.definitions()[name]
    opts = dict(rest[0]) if rest else {}
    common = rest[1] if len(rest) > 1 else {}
    opts.setdefault("use_library_kernels", False)
    impl = _compare(gtscript, gt_storage, mk, name, fn, ext, opts, common, domain, seed=sum(map(ord, name)) + 1)
    plan = impl._plan_
    if kinds is not None:
        assert [k["kind"] for k in plan["kernels"]] == kinds
    if suffix is not None and not EMU:
        first = type(impl)._last_entry_points[0]
        assert first.endswith(suffix), type(impl)._last_entry_points


@pytest.mark.parametrize("name,domain", [("vertical_advection_dycore", (300, 131, 80)), ("tridiagonal_solver", (257, 66, 80)),
                                         ("column_physics", (300, 200, 64))])
def test_fused_column_kernel_with_tma_fed_rings(name, domain, monkeypatch):
    """The `<k>_t` entry point of the fused column template (lowering_colfuse.TMA_RINGS, off by default): prefetch rings
    fed by TMA tile loads issued by a producer lane in an extra row of threads, completion on an mbarrier per slot.  K offsets (vadv reads
    k - 1 .. k + 1), an I offset (wcon at i + 1), several tiles per CTA; bit-exact against the reference."""
    mk, gtscript, gt_storage = _setup()
    from gt4py_b200 import lowering_colfuse

    monkeypatch.setattr(lowering_colfuse, "TMA_RINGS", True)
    fn, ext = mk.definitions()[name][:2]
    impl = _compare(gtscript, gt_storage, mk, name + "_tma", fn, ext, {"use_library_kernels": False}, {}, domain, seed=41)
    assert "tma" in impl._plan_["kernels"][0]
    if not EMU:
        assert type(impl)._last_entry_points == ["k0_0_t"]


@pytest.mark.parametrize("nk", [3, 64, 65, 111, 200])
def test_fused_column_kernel_over_column_depths(nk):
    """Generated vadv with other externals (BET_M = 0.4: not the library's kernel) through the fused FORWARD / BACKWARD
    template: NK = 64 / 65 straddle the Tensor-Memory part of the value stack (128 values per thread, 2 per level),
    111 and 200 reach into shared memory; 3 is the minimum the stencil accepts."""
    mk, gtscript, gt_storage = _setup()
    fn, _ = mk.definitions()["vertical_advection_dycore"][:2]
    impl = _compare(gtscript, gt_storage, mk, f"vadv_bet04_nk{nk}", fn, {"BET_M": 0.4, "BET_P": 0.6}, {}, {},
                    (270, 37, nk), seed=nk)
    assert [k["kind"] for k in impl._plan_["kernels"]] == ["colfuse"] and "library" not in impl._plan_


def test_data_dimension_stencil_at_size():
    """Fields with data dimensions (test_call_interface.py:408-424 form) at 300 x 200 x 40."""
    mk, gtscript, gt_storage = _setup()
    from gt4py.cartesian.gtscript import IJK, PARALLEL, Field, computation, interval

    def vec(inp: Field[IJK, (np.float64, (3,))], w: Field[np.float64], out: Field[IJK, (np.float64, (2,))]):
        with computation(PARALLEL), interval(...):
            out[0, 0, 0][0] = inp[0, 0, 0][0] * w + inp[1, 0, 0][1]
            out[0, 0, 0][1] = inp[0, 0, 0][2] - w[0, -1, 0] * inp[0, 0, 0][1]

    impl = _compare(gtscript, gt_storage, mk, "vec_data_dims", vec, {}, {}, {}, (300, 200, 40), seed=5)
    assert len(impl._plan_["kernels"]) >= 1


def test_fv3_chain_at_baseline_config_3():
    """512 x 512 x 128 (BASELINE configs[3]) through the committed plan, checked on levels of the full plane."""
    from gt4py_b200 import generated, storage
    from oracle import chain
    from oracle import inputs as gen

    dom, h = (512, 512, 128), chain.HALO
    st = generated.load_plan(os.path.join(GEN, "fv3_chain"))
    f = {}
    for name, seed in (("phi", 91), ("u", 92), ("v", 93)):
        f[name] = storage.empty((dom[0] + 2 * h, dom[1] + 2 * h, dom[2]), aligned_index=(h, h, 0))
        storage.fill_synthetic(f[name], seed, -1.0, 1.5, global_offset=(-h, -h))
    out = storage.zeros(dom)
    st(f["phi"], f["u"], f["v"], out, dt=0.05, alpha=0.03, origin={"_all_": (h, h, 0), "out": (0, 0, 0)}, domain=dom)
    assert type(st)._last_entry_points == ["k0_0_u"] and st._plan_["kernels"][0]["kind"] == "march"
    levels = [0, 1, 63, 64, 127]
    host = {name: gen.hashed_window(-h, -h, dom[0] + 2 * h, dom[1] + 2 * h, dom[2], seed, -1.0, 1.5)[:, :, levels]
            for name, seed in (("phi", 91), ("u", 92), ("v", 93))}
    want = chain.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03)
    got = out.asnumpy()[:, :, levels]
    assert np.array_equal(got, want)
    assert 0.05 < (want == 0).mean() < 0.95


@pytest.mark.parametrize("ni", [1, 2, 55, 56, 57, 111, 112, 113, 169])
def test_warp_march_column_pairs_around_the_warp_width(ni):
    """Domains one column short of / at / past a multiple of the 56 result columns of a warp (two columns per lane):
    the warp moved west, lanes whose pair is cut by the domain's edge, a single column."""
    from gt4py_b200 import generated, storage
    from oracle import chain
    from oracle import inputs as gen

    if EMU:
        from tests import emu_abi

        emu_abi.activate()
    dom, h = (ni, 9 if EMU else 70, 2), chain.HALO
    st = generated.load_plan(os.path.join(GEN, "fv3_chain"))
    f = {}
    for name, seed in (("phi", 31), ("u", 32), ("v", 33)):
        f[name] = storage.empty((dom[0] + 2 * h, dom[1] + 2 * h, dom[2]), aligned_index=(h, h, 0))
        storage.fill_synthetic(f[name], seed, -1.0, 1.5, global_offset=(-h, -h))
    out = storage.zeros(dom)
    st(f["phi"], f["u"], f["v"], out, dt=0.05, alpha=0.03, origin={"_all_": (h, h, 0), "out": (0, 0, 0)}, domain=dom)
    host = {name: gen.hashed_window(-h, -h, dom[0] + 2 * h, dom[1] + 2 * h, dom[2], seed, -1.0, 1.5)
            for name, seed in (("phi", 31), ("u", 32), ("v", 33))}
    assert np.array_equal(out.asnumpy(), chain.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03))


@pytest.mark.parametrize("aligned_at", [(3, 3, 0), (0, 0, 0), (2, 3, 0)])
def test_warp_march_column_pairs_whatever_the_alignment(aligned_at):
    """Two columns per lane are one 16-byte ring element where the pair is aligned in global memory (origin on a
    16-byte boundary: the gt4py default, aligned_index = origin) and two 8-byte copies otherwise (origin 3 elements
    past the boundary; one field of each kind in the third case).  A domain that is not a multiple of the warp's 56
    result columns, so the last warp is moved west onto an odd or even column as well."""
    from gt4py_b200 import generated, storage
    from oracle import chain
    from oracle import inputs as gen

    if EMU:
        from tests import emu_abi

        emu_abi.activate()
    dom, h = ((130, 21, 3) if EMU else (333, 77, 5)), chain.HALO
    st = generated.load_plan(os.path.join(GEN, "fv3_chain"))
    assert st._plan_["kernels"][0]["kind"] == "march" and st._plan_["kernels"][0]["columns_per_lane"] == 2
    f = {}
    for n, (name, seed) in enumerate((("phi", 91), ("u", 92), ("v", 93))):
        at = aligned_at if aligned_at != (2, 3, 0) else ((3, 3, 0), (0, 0, 0), (2, 3, 0))[n]
        f[name] = storage.empty((dom[0] + 2 * h, dom[1] + 2 * h, dom[2]), aligned_index=at)
        storage.fill_synthetic(f[name], seed, -1.0, 1.5, global_offset=(-h, -h))
    out = storage.zeros(dom)
    st(f["phi"], f["u"], f["v"], out, dt=0.05, alpha=0.03, origin={"_all_": (h, h, 0), "out": (0, 0, 0)}, domain=dom)
    host = {name: gen.hashed_window(-h, -h, dom[0] + 2 * h, dom[1] + 2 * h, dom[2], seed, -1.0, 1.5)
            for name, seed in (("phi", 91), ("u", 92), ("v", 93))}
    want = chain.fv3_chain(host["phi"], host["u"], host["v"], 0.05, 0.03)
    assert np.array_equal(out.asnumpy(), want)


def test_index_width_switch_beyond_two_to_the_31():
    """A field whose element span exceeds 2^31 - 2^20 must take the 64-bit entry points (`<k>`, not `<k>_u`) and
    still be right: ONE input is a K-strided view of a 17.7 GB allocation (little data, huge span)."""
    from gt4py_b200 import generated, storage

    st = generated.load_plan(os.path.join(GEN, "exec_info_advection"), use_library=False)
    plan = st._plan_
    case_dom = (300, 200, 24)
    names = [f["name"] for f in plan["fields"]]
    info = st.field_info
    rng = np.random.default_rng(3)
    kstep = 1536
    small, origins = {}, {}
    for n in names:
        b = info[n].boundary
        shape = (case_dom[0] + b[0][0] + b[0][1], case_dom[1] + b[1][0] + b[1][1], case_dom[2])
        origins[n] = (b[0][0], b[1][0], 0)
        small[n] = rng.uniform(-1.0, 1.0, shape)
    dense = {n: storage.from_array(a, aligned_index=origins[n]) for n, a in small.items()}
    dense2 = {n: storage.from_array(a, aligned_index=origins[n]) for n, a in small.items()}
    first = [n for n in names if not int(info[n].access) & 2][0]  # a read-only input
    shape = small[first].shape
    whole = storage.empty((shape[0], shape[1], shape[2] * kstep), aligned_index=origins[first])
    view = whole[:, :, ::kstep]
    assert view.shape == shape and (shape[2] - 1) * (view.strides[2] // 8) > (1 << 31)
    for k in range(shape[2]):
        view[:, :, k].set(small[first][:, :, k])
    params = {p: np.float64(0.3) for p in st.parameter_info}
    st(**dense, **params, origin=origins, domain=case_dom)
    assert all(e.endswith("_u") for e in type(st)._last_entry_points)
    dense2[first] = view
    st(**dense2, **params, origin=origins, domain=case_dom)
    assert not any(e.endswith(("_u", "_t")) for e in type(st)._last_entry_points)  # 64-bit index entry points
    written = [n for n in names if int(info[n].access) & 2]
    assert written
    for n in written:
        assert np.array_equal(dense2[n].asnumpy(), dense[n].asnumpy()), n
