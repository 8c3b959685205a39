// vadv_fused.cu -- vertical advection, the tuned fast path (variant 10 and `auto`).
//
// Reference semantics: vertical_advection_dycore, stencil_definitions.py:235-313 (see vadv.cuh).
//
// Same machine mapping as vadv_tmem.cu (persistent CTA per SM, 256 columns per tile, two single-thread
// TMA producer warps, Thomas LIFO split between Tensor Memory and shared memory) and the same fused
// schedule as its variant 9 (the BACKWARD sweep of the previous tile runs interleaved with the FORWARD
// sweep of the next one and hands its LIFO slots over stage by stage).  ncu showed variant 9 to be
// instruction-issue bound (~125 warp instructions per level and column-warp, two warps per scheduler),
// so this version is about instruction count:
//   * the reduced-operation recurrence of vadv.cuh (pre_level2 / rec_level2: 11 + 12 fp64 operations per
//     level instead of 22 + 14), with the branch-free in-range reciprocal and one range check per stage;
//   * KC = 2 levels per stage, fixed: one tcgen05.st.x8 / tcgen05.ld.x8 moves both LIFO entries;
//   * LIFO slots are addressed per stage, so the alternating stack direction costs one index computation
//     per stage instead of per level;
//   * the three schedules (first tile: forward only, steady state, last tile: backward only) are separate
//     template instances instead of predicates in the loop.
#include "vadv_tmem.cuh"

namespace b200 {

namespace {

// FK = levels per stage (2 or 4): one tcgen05.st / tcgen05.ld (.x8 / .x16) moves the LIFO entries of a stage.
// SPILL: the LIFO slots beyond the on-chip capacity (Tensor Memory + shared memory: 88 levels with the default rings)
// live in a per-CTA global scratch area that is small enough (8 KB per slot and CTA) to stay in the L2; the columns of
// a tile then cost ~16 B/level of L2 traffic for the spilled levels instead of the whole LIFO going to HBM.
template <int TI, int FK, bool SPILL>
struct Fused {
    using G = VmGeom<TI, FK>;
    static constexpr uint32_t TCOLS = 4u * FK;  // 32-bit TMEM columns per LIFO slot
    static constexpr int LIFO_SLOT = 2 * FK * VM_COLS;  // doubles per shared-memory LIFO slot

    // per-thread state of the consumer warps
    struct Ctx {
        // geometry / constants
        int t, ti, tj, lane, nk, nstages, nTs, nOn, S, SB;
        double dtr;
        int64_t sk;
        uint64_t *fullF, *emptyF, *fullB, *emptyB;
        const unsigned char *ringF, *ringB;
        double *lifo_t;     // shared-memory part of the LIFO, this thread's column
        double *spill_t;    // global-scratch part of the LIFO (slots >= nOn), this thread's column
        uint32_t tstack;    // Tensor-Memory part of the LIFO, this thread's lane / column half
        // ring cursors
        int sF, sB;
        uint32_t phF, phB;
        // forward sweep of the new tile
        Roll2 roll;
        double ccol, dcol;
        Pre2 pend[FK];
        // backward sweep of the old tile
        double datacol;
        double *op;         // utens_stage(i, j, 2 * kb) of the old tile
        bool valid_old;
    };

    // v = (ccol, dcol) of the FK levels of a stage, lowest level first
    __device__ static __forceinline__ void push(Ctx &c, int q, const double (&v)[2 * FK])
    {
        if (q < c.nTs) {
            tmem_push_n<FK>(c.tstack + TCOLS * (uint32_t)q, v);
        } else if (!SPILL || q < c.nOn) {
            double *s = c.lifo_t + (size_t)(q - c.nTs) * LIFO_SLOT;
#pragma unroll
            for (int u = 0; u < 2 * FK; ++u) s[u * VM_COLS] = v[u];
        } else {
            double *g = c.spill_t + (size_t)(q - c.nOn) * LIFO_SLOT;
#pragma unroll
            for (int u = 0; u < 2 * FK; ++u) __stcg(g + u * VM_COLS, v[u]);
        }
    }

    // Thomas recurrence over the pending stage; result pushed into LIFO slot q
    __device__ static __forceinline__ void solve_push(Ctx &c, int q)
    {
        double cc = c.ccol, dc = c.dcol;
        double v[2 * FK];
        bool ok = true;
#pragma unroll
        for (int u = 0; u < FK; ++u) {
            rec_level2(c.pend[u], cc, dc, ok);
            v[2 * u] = cc;
            v[2 * u + 1] = dc;
        }
        if (!ok) {  // a divisor outside 2^-999 .. 2^999 (or not finite): IEEE division, same rounding
            cc = c.ccol;
            dc = c.dcol;
#pragma unroll
            for (int u = 0; u < FK; ++u) {
                rec_level2_exact(c.pend[u], cc, dc);
                v[2 * u] = cc;
                v[2 * u + 1] = dc;
            }
        }
        c.ccol = cc;
        c.dcol = dc;
        push(c, q, v);
    }

    template <bool FWD, bool BWD>
    __device__ static __forceinline__ void sweep(Ctx &c, const int dir)
    {
        const int t = c.t;
        const int ntop = c.nk - (c.nstages - 1) * FK;  // levels that exist in the top stage (1 .. FK)
        int qprev = 0;
#pragma unroll 1
        for (int ks = 0; ks < c.nstages; ++ks) {
            const int q = dir ? c.nstages - 1 - ks : ks;  // LIFO slot popped now, refilled next iteration
            const bool tm = q < c.nTs;
            RawN<FK> raw;
            double v[2 * FK];  // (ccol, dcol) of the popped stage
            if (BWD) {
                if (tm) {
                    tmem_pop_n_issue<FK>(c.tstack + TCOLS * (uint32_t)q, raw);
                } else if (!SPILL || q < c.nOn) {
                    const double *s = c.lifo_t + (size_t)(q - c.nTs) * LIFO_SLOT;
#pragma unroll
                    for (int u = 0; u < 2 * FK; ++u) v[u] = s[u * VM_COLS];
                } else {  // written by this very thread one tile earlier: L2
                    const double *g = c.spill_t + (size_t)(q - c.nOn) * LIFO_SLOT;
#pragma unroll
                    for (int u = 0; u < 2 * FK; ++u) v[u] = __ldcg(g + u * VM_COLS);
                }
            }
            if (FWD) {
                mbar_wait(&c.fullF[c.sF], c.phF);
                const double *slot = reinterpret_cast<const double *>(c.ringF + (size_t)c.sF * G::SLOT_F) + t;
                const double *w = slot - t + 4 * FK * VM_COLS + c.tj * G::WROW + c.ti;
                Lvl l[FK];
#pragma unroll
                for (int u = 0; u < FK; ++u) {
                    l[u].us = slot[u * VM_COLS];
                    l[u].up = slot[(FK + u) * VM_COLS];
                    l[u].ut = slot[(2 * FK + u) * VM_COLS];
                    l[u].uts = slot[(3 * FK + u) * VM_COLS];
                    l[u].w0 = w[u * G::WLVL];
                    l[u].w1 = w[u * G::WLVL + 1];
                }
                __syncwarp();
                if (c.lane == 0) mbar_arrive(&c.emptyF[c.sF]);
                if (++c.sF == c.S) { c.sF = 0; c.phF ^= 1; }
                if (ks > 0) solve_push(c, qprev);
#pragma unroll
                for (int u = 0; u < FK; ++u) c.pend[u] = pre_level2(l[u], c.roll, c.dtr);
            }
            if (BWD) {
                mbar_wait(&c.fullB[c.sB], c.phB);
                const double *slot = reinterpret_cast<const double *>(c.ringB + (size_t)c.sB * G::SLOT_B) + t;
                double up[FK];
#pragma unroll
                for (int u = 0; u < FK; ++u) up[u] = slot[u * VM_COLS];
                __syncwarp();
                if (c.lane == 0) mbar_arrive(&c.emptyB[c.sB]);
                if (++c.sB == c.SB) { c.sB = 0; c.phB ^= 1; }
                if (tm) {
                    tmem_pop_n_wait<FK>(raw);
#pragma unroll
                    for (int u = 0; u < 2 * FK; ++u)
                        v[u] = __hiloint2double((int)raw.r[2 * u + 1], (int)raw.r[2 * u]);
                }
                // levels FK*kb + FK-1 down to FK*kb of the old tile (kb = nstages-1-ks)  :305-312
#pragma unroll
                for (int u = FK - 1; u >= 0; --u) {
                    if (u == 0 || ks > 0 || u < ntop) {  // the top stage may be partly beyond nk
                        c.datacol = v[2 * u + 1] - v[2 * u] * c.datacol;
                        if (c.valid_old) __stcs(c.op + u * c.sk, c.dtr * (c.datacol - up[u]));
                    }
                }
                c.op -= FK * c.sk;
            }
            qprev = q;
        }
        if (FWD) solve_push(c, qprev);
    }
};

}  // namespace

template <int TI, int FK, bool SPILL>
__global__ void __launch_bounds__(VM_THREADS, 1)
vadv_fused2_kernel(VadvArgs a, const __grid_constant__ CUtensorMap m_us, const __grid_constant__ CUtensorMap m_w,
                   const __grid_constant__ CUtensorMap m_up, const __grid_constant__ CUtensorMap m_ut,
                   const __grid_constant__ CUtensorMap m_uts, int S, int SB, int nTs, int tiles_i, int ntiles,
                   uint64_t pol_stream, uint64_t pol_keep, uint64_t pol_bwd, int nOn, double *spill)
{
    using F = Fused<TI, FK, SPILL>;
    using G = typename F::G;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *fullF = reinterpret_cast<uint64_t *>(smem);
    uint64_t *emptyF = fullF + VM_MAX_STAGES;
    uint64_t *fullB = emptyF + VM_MAX_STAGES;
    uint64_t *emptyB = fullB + VM_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(emptyB + VM_MAX_STAGES);
    unsigned char *ringF = smem + VM_BAR_BYTES;
    unsigned char *ringB = ringF + (size_t)S * G::SLOT_F;
    double *lifo = reinterpret_cast<double *>(ringB + (size_t)SB * G::SLOT_B);  // stage slots [nTs, nstages)

    const int nk = a.nk;
    const int nstages = (nk + FK - 1) / FK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&fullF[s], 1);
            mbar_init(&emptyF[s], VM_CWARPS);
        }
        for (int s = 0; s < SB; ++s) {
            mbar_init(&fullB[s], 1);
            mbar_init(&emptyB[s], VM_CWARPS);
        }
        mbar_fence_init();
    }
    if (warp == VM_CWARPS) tmem_alloc(tmem_slot, VM_TMEM_COLS);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == VM_CWARPS) {
        // ================= forward-ring producer: one thread issues the TMA loads =====================
        if (lane == 0) {
            tma_prefetch_desc(&m_us);
            tma_prefetch_desc(&m_w);
            tma_prefetch_desc(&m_up);
            tma_prefetch_desc(&m_ut);
            tma_prefetch_desc(&m_uts);
            int sF = 0;
            uint32_t phF = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
                for (int ks = 0; ks < nstages; ++ks) {
                    mbar_wait(&emptyF[sF], phF ^ 1);
                    mbar_expect_tx(&fullF[sF], G::TX_F);
                    unsigned char *slot = ringF + (size_t)sF * G::SLOT_F;
                    const int k0 = ks * FK;
                    tma_load_3d(slot, &m_us, ci, cj, k0 + 1, &fullF[sF], pol_stream);
                    tma_load_3d(slot + G::F_BYTES, &m_up, ci, cj, k0, &fullF[sF], pol_keep);
                    tma_load_3d(slot + 2 * G::F_BYTES, &m_ut, ci, cj, k0, &fullF[sF], pol_stream);
                    tma_load_3d(slot + 3 * G::F_BYTES, &m_uts, ci, cj, k0, &fullF[sF], pol_stream);
                    tma_load_3d(slot + 4 * G::F_BYTES, &m_w, ci, cj, k0 + 1, &fullF[sF], pol_stream);
                    if (++sF == S) { sF = 0; phF ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == VM_CWARPS + 1) {
        // ================= backward-ring producer: u_pos stages, top stage first ========================
        if (lane == 0) {
            int sB = 0;
            uint32_t phB = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
                for (int ks = nstages - 1; ks >= 0; --ks) {
                    mbar_wait(&emptyB[sB], phB ^ 1);
                    mbar_expect_tx(&fullB[sB], G::TX_B);
                    tma_load_3d(ringB + (size_t)sB * G::SLOT_B, &m_up, ci, cj, ks * FK, &fullB[sB], pol_bwd);
                    if (++sB == SB) { sB = 0; phB ^= 1; }
                }
            }
        }
        __syncwarp();
    } else {
        // =========================== consumers: one column per thread ===============================
        typename F::Ctx c;
        c.t = threadIdx.x;
        c.ti = c.t % TI;
        c.tj = c.t / TI;
        c.lane = lane;
        c.nk = nk;
        c.nstages = nstages;
        c.nTs = nTs;
        c.nOn = nOn;
        c.spill_t = SPILL ? spill + (size_t)blockIdx.x * (size_t)(nstages - nOn) * F::LIFO_SLOT + c.t : nullptr;
        c.S = S;
        c.SB = SB;
        c.dtr = a.dtr_stage;
        c.sk = a.utens_stage.sk;
        c.fullF = fullF;
        c.emptyF = emptyF;
        c.fullB = fullB;
        c.emptyB = emptyB;
        c.ringF = ringF;
        c.ringB = ringB;
        c.lifo_t = lifo + c.t;
        c.tstack = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 256u;
        c.sF = c.sB = 0;
        c.phF = c.phB = 0;
        c.op = nullptr;
        c.valid_old = false;
        c.datacol = 0.0;

        auto first_ustage = [&](int tile) -> double {
            if (tile >= ntiles) return 0.0;
            const int i = (tile % tiles_i) * TI + c.ti, j = (tile / tiles_i) * G::TJ + c.tj;
            if (i >= a.ni || j >= a.nj) return 0.0;
            return __ldg(a.u_stage.p + i + (int64_t)j * a.u_stage.sj);
        };
        const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
        double us0_next = first_ustage(blockIdx.x);

        for (int n = 0; n <= my_tiles && my_tiles > 0; ++n) {
            const int tile = blockIdx.x + n * gridDim.x;
            const int dir = n & 1;  // stack direction of the tile in its forward sweep
            c.roll.us_0 = us0_next;
            c.roll.ccp = 0.0;
            c.roll.qp = 0.0;
            c.ccol = c.dcol = 0.0;
            c.datacol = 0.0;
            if (n < my_tiles) us0_next = first_ustage(tile + gridDim.x);
            if (n == 0)
                F::template sweep<true, false>(c, dir);
            else if (n < my_tiles)
                F::template sweep<true, true>(c, dir);
            else
                F::template sweep<false, true>(c, dir);
            if (n < my_tiles) {  // this tile is the "old" one of the next round
                const int i = (tile % tiles_i) * TI + c.ti, j = (tile / tiles_i) * G::TJ + c.tj;
                c.valid_old = i < a.ni && j < a.nj;
                c.op = a.utens_stage.p + i + (int64_t)j * a.utens_stage.sj + (int64_t)(nstages - 1) * FK * c.sk;
            }
            tmem_push_wait();
        }
    }

    // every warp arrives here; TMEM is released only after all consumers finished their last pop
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == VM_CWARPS) tmem_dealloc(tmem_base, VM_TMEM_COLS);
}

template <int TI, int FK>
static int launch_fused2(const VadvArgs &a, cudaStream_t stream, bool *handled)
{
    using G = VmGeom<TI, FK>;
    static int max_smem = -1, sms = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    const int env_s = env_int("B200_VADV_STAGES", 0);   // tuning hooks, read at every launch
    const int env_sb = env_int("B200_VADV_BSTAGES", 0);
    const int nstages = (a.nk + FK - 1) / FK;
    const int tmem_slots = VM_TMEM_LEVELS / FK;  // stage slots of 4 * FK columns per thread
    const int nTs = nstages < tmem_slots ? nstages : tmem_slots;
    const size_t slot_bytes = (size_t)2 * FK * VM_COLS * 8;
    // ring depths, measured on B200 (tools/vadv_knob_sweep.py, profiles/r2_vadv_ring_depth.jsonl): a forward ring of 7
    // stages beats 5 by 2-3 % at 2048 x 2048 x 80 (2.72 -> 2.66 ms) and by 0.6 % at 1024 x 1024 x 80; the backward ring
    // (u_pos only) gives up two of its 6 stages when that is what lets the seventh forward stage fit (nk = 80: the
    // LIFO's shared-memory part takes 64 KB)
    const int want_s = env_s > 0 ? env_s : (FK == 2 ? 7 : 3);
    const int want_sb = env_sb > 0 ? env_sb : (FK == 2 ? 6 : 3);
    // 1. whole LIFO on chip, rings shrunk if need be (down to 2 stages)
    size_t lifo = (size_t)(nstages - nTs) * slot_bytes;
    int nOn = nstages;
    int SB = want_sb > VM_MAX_STAGES ? VM_MAX_STAGES : want_sb;
    long avail = (long)max_smem - VM_BAR_BYTES - (long)lifo;
    while (SB > 2 && (long)SB * G::SLOT_B > avail / 4) --SB;
    if (env_sb <= 0 && FK == 2)  // the forward ring first: shrink the backward ring to 4 if that buys a forward stage
        while (SB > 4 && (avail - (long)SB * G::SLOT_B) / (long)G::SLOT_F < want_s) --SB;
    avail -= (long)SB * G::SLOT_B;
    int S = avail > 0 ? (int)(avail / G::SLOT_F) : 0;
    if (S > want_s) S = want_s;
    if (S > VM_MAX_STAGES) S = VM_MAX_STAGES;
    if (S < 2) {
        // 2. nk too large for that (> ~108): full-depth rings, as many LIFO slots in shared memory as then fit,
        //    the rest in the per-CTA global scratch (SPILL)
        // (here the old ring depth of 5 stays the default: every ring stage costs 2.5 LIFO slots that would spill --
        // 512 x 512 x 128: 0.372 ms with 5 stages, 0.398 with 7)
        const int want_s2 = env_s > 0 ? env_s : (FK == 2 ? 5 : 3);
        S = want_s2 > VM_MAX_STAGES ? VM_MAX_STAGES : want_s2;
        SB = want_sb > VM_MAX_STAGES ? VM_MAX_STAGES : want_sb;
        avail = (long)max_smem - VM_BAR_BYTES - (long)S * G::SLOT_F - (long)SB * G::SLOT_B;
        if (avail < 0) return B200_OK;
        const int nSs = (int)(avail / (long)slot_bytes);
        nOn = nTs + nSs;
        if (nOn >= nstages) nOn = nstages;  // cannot happen: case 1 would have fitted
        lifo = (size_t)(nOn - nTs) * slot_bytes;
    }
    const size_t smem = VM_BAR_BYTES + (size_t)S * G::SLOT_F + (size_t)SB * G::SLOT_B + lifo;

    CUtensorMap m_us, m_w, m_up, m_ut, m_uts;
    const int64_t ext[3] = {a.ni, a.nj, a.nk}, extw[3] = {a.ni + 1, a.nj, a.nk};
    const uint32_t box[3] = {TI, (uint32_t)G::TJ, FK}, boxw[3] = {TI + 2, (uint32_t)G::TJ, FK};
    int rc;
    if ((rc = encode_tmap_f64(&m_us, a.u_stage.p, ext, a.u_stage.sj, a.u_stage.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_w, a.wcon.p, extw, a.wcon.sj, a.wcon.sk, boxw))) return rc;
    if ((rc = encode_tmap_f64(&m_up, a.u_pos.p, ext, a.u_pos.sj, a.u_pos.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_ut, a.utens.p, ext, a.utens.sj, a.utens.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_uts, a.utens_stage.p, ext, a.utens_stage.sj, a.utens_stage.sk, box))) return rc;

    const int tiles_i = (a.ni + TI - 1) / TI, tiles_j = (a.nj + G::TJ - 1) / G::TJ;
    const int ntiles = tiles_i * tiles_j;
    const int grid = ntiles < sms ? ntiles : sms;
    // L2 policies of the streaming loads, the forward u_pos load (re-read by the BACKWARD sweep one tile later)
    // and that re-read; tuning hooks: 0 = normal, 1 = evict_first, 2 = evict_last
    static const uint64_t pols[3] = {L2_EVICT_NORMAL, L2_EVICT_FIRST, L2_EVICT_LAST};
    const uint64_t pol_stream = pols[env_int("B200_VADV_POL_STREAM", 1) % 3];
    const uint64_t pol_keep = pols[env_int("B200_VADV_POL_KEEP", 2) % 3];
    const uint64_t pol_bwd = pols[env_int("B200_VADV_POL_BWD", 1) % 3];
    if (nOn < nstages) {
        // spill scratch: grown on demand, per process (launches that need it are ordered by the caller's stream;
        // concurrent spilling launches on different streams are not supported -- same rule as vadv.cu's scratch)
        static double *scratch = nullptr;
        static size_t scratch_bytes = 0;
        const size_t need = (size_t)grid * (size_t)(nstages - nOn) * slot_bytes;
        if (need > scratch_bytes) {
            if (scratch) {
                B200_CUDA_CHECK(cudaDeviceSynchronize());
                B200_CUDA_CHECK(cudaFree(scratch));
            }
            scratch = nullptr;
            scratch_bytes = 0;
            B200_CUDA_CHECK(cudaMalloc(&scratch, need));
            scratch_bytes = need;
        }
        auto kern = vadv_fused2_kernel<TI, FK, true>;
        B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        note_kernel("vadv_fused2_kernel");
        kern<<<grid, VM_THREADS, smem, stream>>>(a, m_us, m_w, m_up, m_ut, m_uts, S, SB, nTs, tiles_i, ntiles,
                                                 pol_stream, pol_keep, pol_bwd, nOn, scratch);
    } else {
        auto kern = vadv_fused2_kernel<TI, FK, false>;
        B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        note_kernel("vadv_fused2_kernel");
        kern<<<grid, VM_THREADS, smem, stream>>>(a, m_us, m_w, m_up, m_ut, m_uts, S, SB, nTs, tiles_i, ntiles,
                                                 pol_stream, pol_keep, pol_bwd, nOn, nullptr);
    }
    B200_CUDA_CHECK(cudaGetLastError());
    *handled = true;
    return B200_OK;
}

// Launch the tuned fused kernel if the layout allows it (same preconditions as vadv_tma.cu); *handled =
// false with B200_OK means "use another variant".
int launch_vadv_fused(const VadvArgs &a, int levels_per_stage, cudaStream_t stream, bool *handled)
{
    *handled = false;
    if (!tma_compatible(a.u_stage.p, a.u_stage.si, a.u_stage.sj, a.u_stage.sk) ||
        !tma_compatible(a.wcon.p, a.wcon.si, a.wcon.sj, a.wcon.sk) ||
        !tma_compatible(a.u_pos.p, a.u_pos.si, a.u_pos.sj, a.u_pos.sk) ||
        !tma_compatible(a.utens.p, a.utens.si, a.utens.sj, a.utens.sk) ||
        !tma_compatible(a.utens_stage.p, a.utens_stage.si, a.utens_stage.sj, a.utens_stage.sk))
        return B200_OK;
    if (levels_per_stage == 4)
        return a.ni > 64 ? launch_fused2<128, 4>(a, stream, handled)
             : a.ni > 32 ? launch_fused2<64, 4>(a, stream, handled)
                         : launch_fused2<32, 4>(a, stream, handled);
    return a.ni > 64 ? launch_fused2<128, 2>(a, stream, handled)
         : a.ni > 32 ? launch_fused2<64, 2>(a, stream, handled)
                     : launch_fused2<32, 2>(a, stream, handled);
}

}  // namespace b200
