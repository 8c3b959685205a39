// vadv_tma.cu -- vertical advection with a TMA-fed shared-memory pipeline (variant 4, the fast path).
//
// Reference semantics: vertical_advection_dycore, stencil_definitions.py:235-313 (see vadv.cuh).
//
// Why this shape.  The Thomas solve needs (ccol, dcol) of all nk levels of a column between the
// FORWARD and the BACKWARD sweep: 16 B x nk per column.  Keeping that LIFO in shared memory (never in
// HBM) limits an SM to ~128 columns, i.e. 4 compute warps -- far too few threads to cover HBM latency
// with register prefetching (the first version of this kernel ran at 1.2 TB/s).  So memory parallelism
// is decoupled from the compute threads:
//   * a persistent CTA per SM walks over 128-column tiles (TI x TJ columns, TI along contiguous I);
//   * ONE producer thread streams the tile's input planes into an S-stage shared-memory ring with
//     3-D TMA tile loads (cp.async.bulk.tensor, completion on mbarriers): per stage KC levels of
//     u_stage(k+1), wcon(k+1) [TI+2 wide: the i+1 column], u_pos, utens, utens_stage; out-of-domain
//     parts of edge tiles are zero-filled by the TMA unit, so there is no edge code;
//   * 4 consumer warps (one column per thread) run the FORWARD sweep out of the ring, push
//     (ccol, dcol) onto the shared-memory LIFO ([level][2][128], conflict-free), then run the BACKWARD
//     sweep popping it; u_pos(k) for the backward sweep comes through a second small ring (an L2 hit:
//     the forward pass loaded it with an evict_last policy);
//   * because the CTA is persistent the producer already prefetches the next tile's first S stages
//     while the consumers are still in the backward sweep, so the load pipe never drains.
// DRAM traffic is the compulsory 48 B/point (5 fields in, 1 out).
#include "tma.cuh"
#include "vadv.cuh"

namespace b200 {

constexpr int VT_COLS = 128;       // columns per tile == consumer threads
constexpr int VT_THREADS = 160;    // 4 consumer warps + 1 producer warp
constexpr int VT_MAX_STAGES = 16;
constexpr int VT_BAR_BYTES = 1024;

template <int TI, int KC>
struct VtGeom {
    static constexpr int TJ = VT_COLS / TI;
    static constexpr int WROW = TI + 2;                 // wcon row: TI + the i+1 column, 16 B multiple
    static constexpr int WLVL = WROW * TJ;              // doubles per level (wcon)
    static constexpr int F_BYTES = KC * VT_COLS * 8;    // one field, one stage
    static constexpr int W_TX = KC * WLVL * 8;
    static constexpr int W_BYTES = (W_TX + 127) / 128 * 128;
    static constexpr int SLOT_F = 4 * F_BYTES + W_BYTES;
    static constexpr int SLOT_B = F_BYTES;
    static constexpr uint32_t TX_F = 4 * F_BYTES + W_TX;
    static constexpr uint32_t TX_B = F_BYTES;
};

template <int TI, int KC, int RCP>
__global__ void __launch_bounds__(VT_THREADS, 1)
vadv_tma_kernel(VadvArgs a, const __grid_constant__ CUtensorMap m_us, const __grid_constant__ CUtensorMap m_w,
                const __grid_constant__ CUtensorMap m_up, const __grid_constant__ CUtensorMap m_ut,
                const __grid_constant__ CUtensorMap m_uts, int S, int SB, int tiles_i, int ntiles)
{
    using G = VtGeom<TI, KC>;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *fullF = reinterpret_cast<uint64_t *>(smem);
    uint64_t *emptyF = fullF + VT_MAX_STAGES;
    uint64_t *fullB = emptyF + VT_MAX_STAGES;
    uint64_t *emptyB = fullB + VT_MAX_STAGES;
    unsigned char *ringF = smem + VT_BAR_BYTES;
    unsigned char *ringB = ringF + (size_t)S * G::SLOT_F;
    double *lifo = reinterpret_cast<double *>(ringB + (size_t)SB * G::SLOT_B);

    const int nk = a.nk;
    const int nstages = (nk + KC - 1) / KC;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&fullF[s], 1);
            mbar_init(&emptyF[s], 4);
        }
        for (int s = 0; s < SB; ++s) {
            mbar_init(&fullB[s], 1);
            mbar_init(&emptyB[s], 4);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == 4) {
        // ======================= producer: one thread issues every TMA load =======================
        if (lane != 0) return;
        tma_prefetch_desc(&m_us);
        tma_prefetch_desc(&m_w);
        tma_prefetch_desc(&m_up);
        tma_prefetch_desc(&m_ut);
        tma_prefetch_desc(&m_uts);
        int sF = 0, sB = 0;
        uint32_t phF = 0, phB = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
            for (int ks = 0; ks < nstages; ++ks) {
                mbar_wait(&emptyF[sF], phF ^ 1);
                mbar_expect_tx(&fullF[sF], G::TX_F);
                unsigned char *slot = ringF + (size_t)sF * G::SLOT_F;
                const int k0 = ks * KC;
                tma_load_3d(slot, &m_us, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                tma_load_3d(slot + G::F_BYTES, &m_up, ci, cj, k0, &fullF[sF], L2_EVICT_LAST);
                tma_load_3d(slot + 2 * G::F_BYTES, &m_ut, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                tma_load_3d(slot + 3 * G::F_BYTES, &m_uts, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                tma_load_3d(slot + 4 * G::F_BYTES, &m_w, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                if (++sF == S) { sF = 0; phF ^= 1; }
            }
            for (int kb = 0; kb < nstages; ++kb) {
                mbar_wait(&emptyB[sB], phB ^ 1);
                mbar_expect_tx(&fullB[sB], G::TX_B);
                tma_load_3d(ringB + (size_t)sB * G::SLOT_B, &m_up, ci, cj, nk - (kb + 1) * KC, &fullB[sB],
                            L2_EVICT_FIRST);
                if (++sB == SB) { sB = 0; phB ^= 1; }
            }
        }
        return;
    }

    // =========================== consumers: one column per thread ===============================
    const int t = threadIdx.x;  // 0..127
    const int ti = t % TI, tj = t / TI;
    const double dtr = a.dtr_stage;
    int sF = 0, sB = 0;
    uint32_t phF = 0, phB = 0;
    double *const lifo_t = lifo + t;

    auto first_ustage = [&](int tile) -> double {
        if (tile >= ntiles) return 0.0;
        const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
        if (i >= a.ni || j >= a.nj) return 0.0;
        return __ldg(a.u_stage.p + i + (int64_t)j * a.u_stage.sj);
    };
    // wait for the next forward stage, pull its KC levels into registers, hand the slot back
    auto fetch_stage = [&](Lvl(&l)[KC]) {
        mbar_wait(&fullF[sF], phF);
        const double *slot = reinterpret_cast<const double *>(ringF + (size_t)sF * G::SLOT_F) + t;
        const double *w = slot - t + 4 * KC * VT_COLS + tj * G::WROW + ti;
#pragma unroll
        for (int u = 0; u < KC; ++u) {
            l[u].us = slot[u * VT_COLS];
            l[u].up = slot[(KC + u) * VT_COLS];
            l[u].ut = slot[(2 * KC + u) * VT_COLS];
            l[u].uts = slot[(3 * KC + u) * VT_COLS];
            l[u].w0 = w[u * G::WLVL];
            l[u].w1 = w[u * G::WLVL + 1];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&emptyF[sF]);
        if (++sF == S) { sF = 0; phF ^= 1; }
    };
    // Thomas recurrence over the KC levels of one stage; pushes (ccol, dcol) onto the LIFO
    auto solve_stage = [&](const Pre(&p)[KC], double &ccol, double &dcol, int k0, bool guard) {
        if (RCP == 1) {
            double c = ccol, d = dcol, cs[KC], ds[KC];
            bool ok = true;
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                const double den = p[u].bcol - c * p[u].acol;
                const double num = p[u].dc - d * p[u].acol;
                ok = ok && rcp_inrange(den);
                const double divided = rcp_rn_inrange(den);
                cs[u] = c = p[u].cc * divided;
                ds[u] = d = num * divided;
            }
            if (ok) {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    if (!guard || k0 + u < nk) {
                        lifo_t[(2 * (k0 + u)) * VT_COLS] = cs[u];
                        lifo_t[(2 * (k0 + u) + 1) * VT_COLS] = ds[u];
                    }
                }
                ccol = c;
                dcol = d;
                return;
            }
        }
#pragma unroll
        for (int u = 0; u < KC; ++u) {
            rec_level(p[u], ccol, dcol);
            if (!guard || k0 + u < nk) {
                lifo_t[(2 * (k0 + u)) * VT_COLS] = ccol;
                lifo_t[(2 * (k0 + u) + 1) * VT_COLS] = dcol;
            }
        }
    };

    double us0_next = first_ustage(blockIdx.x);

    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
        const bool valid = i < a.ni && j < a.nj;

        // ---- FORWARD (software pipelined: pre_level of stage ks overlaps the recurrence of ks-1) ----
        Roll roll;
        roll.us_m1 = 0.0;
        roll.us_0 = us0_next;
        roll.wsum = 0.0;
        double ccol = 0.0, dcol = 0.0;
        Pre pend[KC];
        {
            Lvl l[KC];
            fetch_stage(l);
#pragma unroll
            for (int u = 0; u < KC; ++u) pend[u] = pre_level(l[u], roll, dtr);
        }
        for (int ks = 1; ks < nstages; ++ks) {
            Lvl l[KC];
            fetch_stage(l);
            Pre next[KC];
#pragma unroll
            for (int u = 0; u < KC; ++u) next[u] = pre_level(l[u], roll, dtr);
            solve_stage(pend, ccol, dcol, (ks - 1) * KC, false);
#pragma unroll
            for (int u = 0; u < KC; ++u) pend[u] = next[u];
        }
        solve_stage(pend, ccol, dcol, (nstages - 1) * KC, true);
        us0_next = first_ustage(tile + gridDim.x);

        // ---- BACKWARD ----  :305-312 (uniform: ccol(nk-1) == 0 and datacol starts at 0)
        double *out = a.utens_stage.p + i + (int64_t)j * a.utens_stage.sj;
        const int64_t sk = a.utens_stage.sk;
        double datacol = 0.0;
        for (int kb = 0; kb < nstages; ++kb) {
            const int khi = nk - 1 - kb * KC;  // highest level of this stage
            double cc[KC], dc[KC], up[KC];
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                const int k = khi - u >= 0 ? khi - u : 0;
                cc[u] = lifo_t[(2 * k) * VT_COLS];
                dc[u] = lifo_t[(2 * k + 1) * VT_COLS];
            }
            mbar_wait(&fullB[sB], phB);
            const double *slot = reinterpret_cast<const double *>(ringB + (size_t)sB * G::SLOT_B) + t;
#pragma unroll
            for (int u = 0; u < KC; ++u) up[u] = slot[(KC - 1 - u) * VT_COLS];
            __syncwarp();
            if (lane == 0) mbar_arrive(&emptyB[sB]);
            if (++sB == SB) { sB = 0; phB ^= 1; }
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                const int k = khi - u;
                datacol = dc[u] - cc[u] * datacol;
                if (valid && k >= 0) __stcs(out + k * sk, dtr * (datacol - up[u]));
            }
        }
    }
}

template <int TI, int KC, int RCP>
static int launch_geom(const VadvArgs &a, cudaStream_t stream, bool *handled)
{
    using G = VtGeom<TI, KC>;
    static int max_smem = -1, sms = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    static const int env_s = env_int("B200_VADV_STAGES", 0);
    static const int env_sb = env_int("B200_VADV_BSTAGES", 0);
    const int SB = env_sb > 0 ? (env_sb > VT_MAX_STAGES ? VT_MAX_STAGES : env_sb) : 4;
    const size_t lifo = (size_t)a.nk * 2 * VT_COLS * 8;
    const long avail = (long)max_smem - VT_BAR_BYTES - (long)SB * G::SLOT_B - (long)lifo;
    int S = avail > 0 ? (int)(avail / G::SLOT_F) : 0;
    const int want = env_s > 0 ? env_s : 12;
    if (S > want) S = want;
    if (S > VT_MAX_STAGES) S = VT_MAX_STAGES;
    if (S < 2) return B200_OK;  // nk too large for the shared-memory LIFO: not handled here
    const size_t smem = VT_BAR_BYTES + (size_t)S * G::SLOT_F + (size_t)SB * G::SLOT_B + lifo;

    CUtensorMap m_us, m_w, m_up, m_ut, m_uts;
    const int64_t ext[3] = {a.ni, a.nj, a.nk}, extw[3] = {a.ni + 1, a.nj, a.nk};
    const uint32_t box[3] = {TI, (uint32_t)G::TJ, KC}, boxw[3] = {TI + 2, (uint32_t)G::TJ, KC};
    int rc;
    if ((rc = encode_tmap_f64(&m_us, a.u_stage.p, ext, a.u_stage.sj, a.u_stage.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_w, a.wcon.p, extw, a.wcon.sj, a.wcon.sk, boxw))) return rc;
    if ((rc = encode_tmap_f64(&m_up, a.u_pos.p, ext, a.u_pos.sj, a.u_pos.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_ut, a.utens.p, ext, a.utens.sj, a.utens.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_uts, a.utens_stage.p, ext, a.utens_stage.sj, a.utens_stage.sk, box))) return rc;

    const int tiles_i = (a.ni + TI - 1) / TI, tiles_j = (a.nj + G::TJ - 1) / G::TJ;
    const int ntiles = tiles_i * tiles_j;
    const int grid = ntiles < sms ? ntiles : sms;
    auto kern = vadv_tma_kernel<TI, KC, RCP>;
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    note_kernel("vadv_tma_kernel");
    kern<<<grid, VT_THREADS, smem, stream>>>(a, m_us, m_w, m_up, m_ut, m_uts, S, SB, tiles_i, ntiles);
    B200_CUDA_CHECK(cudaGetLastError());
    *handled = true;
    return B200_OK;
}

// Launch the TMA kernel if the layout allows it.  *handled = false (and B200_OK) means "use another
// variant": TMA needs unit I stride, 16-byte aligned origins, even J/K strides, and the LIFO must fit.
int launch_vadv_tma(const VadvArgs &a, int variant, cudaStream_t stream, bool *handled)
{
    *handled = false;
    if (!tma_compatible(a.u_stage.p, a.u_stage.si, a.u_stage.sj, a.u_stage.sk) ||
        !tma_compatible(a.wcon.p, a.wcon.si, a.wcon.sj, a.wcon.sk) ||
        !tma_compatible(a.u_pos.p, a.u_pos.si, a.u_pos.sj, a.u_pos.sk) ||
        !tma_compatible(a.utens.p, a.utens.si, a.utens.sj, a.utens.sk) ||
        !tma_compatible(a.utens_stage.p, a.utens_stage.si, a.utens_stage.sj, a.utens_stage.sk))
        return B200_OK;
    // variants: 4 = KC 2, 5 = KC 4 (default), 6 = KC 4 with the branch-free in-range reciprocal
    if (variant == 4) return a.ni > 64 ? launch_geom<128, 2, 0>(a, stream, handled)
                           : a.ni > 32 ? launch_geom<64, 2, 0>(a, stream, handled)
                                       : launch_geom<32, 2, 0>(a, stream, handled);
    if (variant == 6) return a.ni > 64 ? launch_geom<128, 4, 1>(a, stream, handled)
                           : a.ni > 32 ? launch_geom<64, 4, 1>(a, stream, handled)
                                       : launch_geom<32, 4, 1>(a, stream, handled);
    return a.ni > 64 ? launch_geom<128, 4, 0>(a, stream, handled)
         : a.ni > 32 ? launch_geom<64, 4, 0>(a, stream, handled)
                     : launch_geom<32, 4, 0>(a, stream, handled);
}

}  // namespace b200
