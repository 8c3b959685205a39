// vadv_tmem.cu -- vertical advection, 256 columns per SM with the Thomas LIFO in TENSOR MEMORY
// (variant 7, the fast path on B200).
//
// Reference semantics: vertical_advection_dycore, stencil_definitions.py:235-313 (see vadv.cuh).
//
// The bottleneck of this stencil on B200 is not HBM but on-chip state: between the FORWARD and the
// BACKWARD sweep every column must park (ccol, dcol) of all nk levels, 16 B x nk.  With that LIFO in
// shared memory alone an SM holds 128 columns = ONE warp per scheduler, and a single warp cannot hide
// the ~90-cycle dependent chain of the Thomas recurrence (fp64 multiply, subtract, reciprocal, multiply;
// measured DFMA latency 8.8 cycles, 1/x 75 cycles, tools/ubench_fp64.cu): vadv_tma.cu stalls at
// ~4 TB/s.  Blackwell's 256 KB of Tensor Memory per SM sits idle in a stencil code, and tcgen05.st /
// tcgen05.ld in the 32x32b shape give every thread of a warp a private, run-time indexed column range
// of its own TMEM lane -- exactly a per-thread stack.  So:
//   * the LIFO of levels [0, 64) lives in TMEM (4 x 32-bit columns per level; the two consumer warps
//     that share a TMEM lane quarter get 256 columns each), levels [64, nk) in shared memory;
//   * that frees enough shared memory for 256 columns per SM = two consumer warps per scheduler whose
//     recurrence chains interleave, plus deep TMA rings;
//   * the tile's input planes arrive through 3-D TMA tile loads issued by two single-thread producer
//     warps: one feeds the FORWARD ring (u_stage, wcon [TI+2 wide for the i+1 column], u_pos, utens,
//     utens_stage; KC levels per stage), the other the BACKWARD ring (u_pos again, an L2 hit thanks to
//     the evict_last hint of the forward load).  The CTA is persistent, so both producers run ahead
//     into the next tile while the consumers are still busy; edge tiles are zero-filled by the TMA unit.
// DRAM traffic is the compulsory 48 B/point (5 fields in, 1 out).  Arithmetic is the branch-free
// formulation of vadv.cuh, bit-identical to the reference.
#include "vadv_tmem.cuh"

namespace b200 {

template <int TI, int KC>
__global__ void __launch_bounds__(VM_THREADS, 1)
vadv_tmem_kernel(VadvArgs a, const __grid_constant__ CUtensorMap m_us, const __grid_constant__ CUtensorMap m_w,
                 const __grid_constant__ CUtensorMap m_up, const __grid_constant__ CUtensorMap m_ut,
                 const __grid_constant__ CUtensorMap m_uts, int S, int SB, int nT, int tiles_i, int ntiles)
{
    using G = VmGeom<TI, KC>;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *fullF = reinterpret_cast<uint64_t *>(smem);
    uint64_t *emptyF = fullF + VM_MAX_STAGES;
    uint64_t *fullB = emptyF + VM_MAX_STAGES;
    uint64_t *emptyB = fullB + VM_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(emptyB + VM_MAX_STAGES);
    unsigned char *ringF = smem + VM_BAR_BYTES;
    unsigned char *ringB = ringF + (size_t)S * G::SLOT_F;
    double *lifo = reinterpret_cast<double *>(ringB + (size_t)SB * G::SLOT_B);  // levels [nT, nk)

    const int nk = a.nk;
    const int nstages = (nk + KC - 1) / KC;
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
                    const int k0 = ks * KC;
                    tma_load_3d(slot, &m_us, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + G::F_BYTES, &m_up, ci, cj, k0, &fullF[sF], L2_EVICT_LAST);
                    tma_load_3d(slot + 2 * G::F_BYTES, &m_ut, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 3 * G::F_BYTES, &m_uts, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 4 * G::F_BYTES, &m_w, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                    if (++sF == S) { sF = 0; phF ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == VM_CWARPS + 1) {
        // ================= backward-ring producer: u_pos stages, top level first =======================
        if (lane == 0) {
            int sB = 0;
            uint32_t phB = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
                for (int ks = nstages - 1; ks >= 0; --ks) {
                    mbar_wait(&emptyB[sB], phB ^ 1);
                    mbar_expect_tx(&fullB[sB], G::TX_B);
                    tma_load_3d(ringB + (size_t)sB * G::SLOT_B, &m_up, ci, cj, ks * KC, &fullB[sB], L2_EVICT_FIRST);
                    if (++sB == SB) { sB = 0; phB ^= 1; }
                }
            }
        }
        __syncwarp();
    } else {
        // =========================== consumers: one column per thread ===============================
        const int t = threadIdx.x;  // 0..255
        const int ti = t % TI, tj = t / TI;
        const double dtr = a.dtr_stage;
        int sF = 0, sB = 0;
        uint32_t phF = 0, phB = 0;
        double *const lifo_t = lifo + t;
        // this thread's TMEM stack: lane quarter of the warp, column half of the warp group
        const uint32_t tstack = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 256u;

        auto first_ustage = [&](int tile) -> double {
            if (tile >= ntiles) return 0.0;
            const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
            if (i >= a.ni || j >= a.nj) return 0.0;
            return __ldg(a.u_stage.p + i + (int64_t)j * a.u_stage.sj);
        };
        auto fetch_stage = [&](Lvl(&l)[KC]) {
            mbar_wait(&fullF[sF], phF);
            const double *slot = reinterpret_cast<const double *>(ringF + (size_t)sF * G::SLOT_F) + t;
            const double *w = slot - t + 4 * KC * VM_COLS + tj * G::WROW + ti;
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                l[u].us = slot[u * VM_COLS];
                l[u].up = slot[(KC + u) * VM_COLS];
                l[u].ut = slot[(2 * KC + u) * VM_COLS];
                l[u].uts = slot[(3 * KC + u) * VM_COLS];
                l[u].w0 = w[u * G::WLVL];
                l[u].w1 = w[u * G::WLVL + 1];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&emptyF[sF]);
            if (++sF == S) { sF = 0; phF ^= 1; }
        };
        // Thomas recurrence over one stage; pushes (ccol, dcol) of its levels onto the LIFO
        auto solve_stage = [&](const Pre(&p)[KC], double &ccol, double &dcol, int k0) {
            double cs[KC], ds[KC];
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                rec_level(p[u], ccol, dcol);
                cs[u] = ccol;
                ds[u] = dcol;
            }
            if (k0 < nT) {  // warp-uniform
#pragma unroll
                for (int u = 0; u < KC; ++u) tmem_push(tstack + 4u * (uint32_t)(k0 + u), cs[u], ds[u]);
            } else {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    if (k0 + u < nk) {
                        lifo_t[(2 * (k0 + u - nT)) * VM_COLS] = cs[u];
                        lifo_t[(2 * (k0 + u - nT) + 1) * VM_COLS] = ds[u];
                    }
                }
            }
        };
        // start popping the LIFO entries of stage ks (TMEM loads complete at pop_finish)
        auto pop_begin = [&](int ks, Raw4(&raw)[KC], double(&cc)[KC], double(&dc)[KC]) {
            const int k0 = ks * KC;
            if (k0 < nT) {
#pragma unroll
                for (int u = 0; u < KC; ++u) tmem_pop_issue(tstack + 4u * (uint32_t)(k0 + u), raw[u]);
            } else {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    const int k = k0 + u < nk ? k0 + u : nk - 1;
                    cc[u] = lifo_t[(2 * (k - nT)) * VM_COLS];
                    dc[u] = lifo_t[(2 * (k - nT) + 1) * VM_COLS];
                }
            }
        };
        auto pop_finish = [&](int ks, Raw4(&raw)[KC], double(&cc)[KC], double(&dc)[KC]) {
            if (ks * KC < nT) {
                tmem_pop_wait<KC>(raw);
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    cc[u] = __hiloint2double((int)raw[u].r[1], (int)raw[u].r[0]);
                    dc[u] = __hiloint2double((int)raw[u].r[3], (int)raw[u].r[2]);
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
                solve_stage(pend, ccol, dcol, (ks - 1) * KC);
#pragma unroll
                for (int u = 0; u < KC; ++u) pend[u] = next[u];
            }
            solve_stage(pend, ccol, dcol, (nstages - 1) * KC);
            us0_next = first_ustage(tile + gridDim.x);
            tmem_push_wait();

            // ---- BACKWARD ----  :305-312 (uniform: ccol(nk-1) == 0 and datacol starts at 0)
            double *outp = a.utens_stage.p + i + (int64_t)j * a.utens_stage.sj +
                           (int64_t)(nstages - 1) * KC * a.utens_stage.sk;
            const int64_t sk = a.utens_stage.sk;
            double datacol = 0.0;
            Raw4 raw[KC];
            double cc[KC], dc[KC];
            pop_begin(nstages - 1, raw, cc, dc);
            pop_finish(nstages - 1, raw, cc, dc);
            for (int ks = nstages - 1; ks >= 0; --ks) {
                Raw4 nraw[KC];
                double ncc[KC], ndc[KC];
                if (ks > 0) pop_begin(ks - 1, nraw, ncc, ndc);
                mbar_wait(&fullB[sB], phB);
                const double *slot = reinterpret_cast<const double *>(ringB + (size_t)sB * G::SLOT_B) + t;
                double up[KC];
#pragma unroll
                for (int u = 0; u < KC; ++u) up[u] = slot[u * VM_COLS];
                __syncwarp();
                if (lane == 0) mbar_arrive(&emptyB[sB]);
                if (++sB == SB) { sB = 0; phB ^= 1; }
#pragma unroll
                for (int u = KC - 1; u >= 0; --u) {
                    const bool live = ks * KC + u < nk;  // pad levels exist only in the top stage
                    const double d = dc[u] - cc[u] * datacol;
                    datacol = live ? d : datacol;
                    if (valid && live) __stcs(outp + u * sk, dtr * (datacol - up[u]));
                }
                outp -= KC * sk;
                if (ks > 0) {
                    pop_finish(ks - 1, nraw, ncc, ndc);
#pragma unroll
                    for (int u = 0; u < KC; ++u) {
                        cc[u] = ncc[u];
                        dc[u] = ndc[u];
                    }
                }
            }
        }
    }

    // every warp arrives here; TMEM is released only after all consumers finished their last pop
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == VM_CWARPS) tmem_dealloc(tmem_base, VM_TMEM_COLS);
}


// ---------------------------------------------------------------------------------------------------
// Fused sweeps (variant 9).  Same resources as vadv_tmem_kernel, different schedule: every consumer
// thread runs the BACKWARD sweep of its previous tile interleaved, stage by stage, with the FORWARD
// sweep of its next tile.  Backward stage (nstages-1-ks) of the old tile pops exactly the LIFO slots
// that forward stage ks of the new tile then fills (the stack direction alternates from tile to tile),
// so one LIFO serves both.  Effects: (1) HBM demand is flat -- every iteration drains one forward-ring
// stage -- instead of bursting during forward sweeps and idling during backward sweeps; (2) each
// thread carries two independent dependency chains, which fills the issue slots the Thomas recurrence
// leaves empty.
template <int TI, int KC>
__global__ void __launch_bounds__(VM_THREADS, 1)
vadv_fused_kernel(VadvArgs a, const __grid_constant__ CUtensorMap m_us, const __grid_constant__ CUtensorMap m_w,
                  const __grid_constant__ CUtensorMap m_up, const __grid_constant__ CUtensorMap m_ut,
                  const __grid_constant__ CUtensorMap m_uts, int S, int SB, int nT, int tiles_i, int ntiles)
{
    using G = VmGeom<TI, KC>;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *fullF = reinterpret_cast<uint64_t *>(smem);
    uint64_t *emptyF = fullF + VM_MAX_STAGES;
    uint64_t *fullB = emptyF + VM_MAX_STAGES;
    uint64_t *emptyB = fullB + VM_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(emptyB + VM_MAX_STAGES);
    unsigned char *ringF = smem + VM_BAR_BYTES;
    unsigned char *ringB = ringF + (size_t)S * G::SLOT_F;
    double *lifo = reinterpret_cast<double *>(ringB + (size_t)SB * G::SLOT_B);  // physical slots [nT, L)

    const int nk = a.nk;
    const int nstages = (nk + KC - 1) / KC;
    const int L = nstages * KC;  // physical LIFO slots (levels padded to whole stages)
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
                    const int k0 = ks * KC;
                    tma_load_3d(slot, &m_us, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + G::F_BYTES, &m_up, ci, cj, k0, &fullF[sF], L2_EVICT_LAST);
                    tma_load_3d(slot + 2 * G::F_BYTES, &m_ut, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 3 * G::F_BYTES, &m_uts, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 4 * G::F_BYTES, &m_w, ci, cj, k0 + 1, &fullF[sF], L2_EVICT_FIRST);
                    if (++sF == S) { sF = 0; phF ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == VM_CWARPS + 1) {
        if (lane == 0) {
            int sB = 0;
            uint32_t phB = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
                for (int ks = nstages - 1; ks >= 0; --ks) {
                    mbar_wait(&emptyB[sB], phB ^ 1);
                    mbar_expect_tx(&fullB[sB], G::TX_B);
                    tma_load_3d(ringB + (size_t)sB * G::SLOT_B, &m_up, ci, cj, ks * KC, &fullB[sB], L2_EVICT_FIRST);
                    if (++sB == SB) { sB = 0; phB ^= 1; }
                }
            }
        }
        __syncwarp();
    } else {
        const int t = threadIdx.x;  // 0..255
        const int ti = t % TI, tj = t / TI;
        const double dtr = a.dtr_stage;
        const int64_t sk = a.utens_stage.sk;
        int sF = 0, sB = 0;
        uint32_t phF = 0, phB = 0;
        double *const lifo_t = lifo + t;
        const uint32_t tstack = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 256u;

        auto first_ustage = [&](int tile) -> double {
            if (tile >= ntiles) return 0.0;
            const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
            if (i >= a.ni || j >= a.nj) return 0.0;
            return __ldg(a.u_stage.p + i + (int64_t)j * a.u_stage.sj);
        };
        auto fetch_stage = [&](Lvl(&l)[KC]) {
            mbar_wait(&fullF[sF], phF);
            const double *slot = reinterpret_cast<const double *>(ringF + (size_t)sF * G::SLOT_F) + t;
            const double *w = slot - t + 4 * KC * VM_COLS + tj * G::WROW + ti;
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                l[u].us = slot[u * VM_COLS];
                l[u].up = slot[(KC + u) * VM_COLS];
                l[u].ut = slot[(2 * KC + u) * VM_COLS];
                l[u].uts = slot[(3 * KC + u) * VM_COLS];
                l[u].w0 = w[u * G::WLVL];
                l[u].w1 = w[u * G::WLVL + 1];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&emptyF[sF]);
            if (++sF == S) { sF = 0; phF ^= 1; }
        };
        // Physical slots of logical stage `ks` for a tile whose stack grows up (dir = 0) or down (dir = 1):
        // level ks*KC + u sits in slot  dir ? L-1-(ks*KC+u) : ks*KC+u.  Stage ranges stay KC-aligned, so
        // "TMEM or shared memory" is a per-stage (warp-uniform) decision.
        auto push_stage = [&](int ks, int dir, const double(&cs)[KC], const double(&ds)[KC]) {
            const int base = dir ? L - KC * (ks + 1) : KC * ks;  // lowest physical slot of the stage
            if (base < nT) {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    const int p = dir ? base + (KC - 1 - u) : base + u;
                    tmem_push(tstack + 4u * (uint32_t)p, cs[u], ds[u]);
                }
            } else {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    const int p = (dir ? base + (KC - 1 - u) : base + u) - nT;
                    lifo_t[(2 * p) * VM_COLS] = cs[u];
                    lifo_t[(2 * p + 1) * VM_COLS] = ds[u];
                }
            }
        };
        auto pop_begin = [&](int ks, int dir, Raw4(&raw)[KC], double(&cc)[KC], double(&dc)[KC]) {
            const int base = dir ? L - KC * (ks + 1) : KC * ks;
            if (base < nT) {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    const int p = dir ? base + (KC - 1 - u) : base + u;
                    tmem_pop_issue(tstack + 4u * (uint32_t)p, raw[u]);
                }
            } else {
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    const int p = (dir ? base + (KC - 1 - u) : base + u) - nT;
                    cc[u] = lifo_t[(2 * p) * VM_COLS];
                    dc[u] = lifo_t[(2 * p + 1) * VM_COLS];
                }
            }
        };
        auto pop_finish = [&](int ks, int dir, Raw4(&raw)[KC], double(&cc)[KC], double(&dc)[KC]) {
            const int base = dir ? L - KC * (ks + 1) : KC * ks;
            if (base < nT) {
                tmem_pop_wait<KC>(raw);
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    cc[u] = __hiloint2double((int)raw[u].r[1], (int)raw[u].r[0]);
                    dc[u] = __hiloint2double((int)raw[u].r[3], (int)raw[u].r[2]);
                }
            }
        };
        auto solve = [&](const Pre(&p)[KC], double &ccol, double &dcol, double(&cs)[KC], double(&ds)[KC]) {
#pragma unroll
            for (int u = 0; u < KC; ++u) {
                rec_level(p[u], ccol, dcol);
                cs[u] = ccol;
                ds[u] = dcol;
            }
        };

        const int my_tiles = blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
        double us0_next = first_ustage(blockIdx.x);
        double *outp = nullptr;  // utens_stage column of the tile in its backward sweep
        bool valid_old = false;

        for (int n = 0; n <= my_tiles; ++n) {
            const bool has_fwd = n < my_tiles, has_bwd = n > 0;  // CTA-uniform
            const int tile = blockIdx.x + n * gridDim.x;
            const int dir = n & 1;  // stack direction of the tile going forward; the old tile has dir ^ 1
            Roll roll;
            roll.us_m1 = 0.0;
            roll.us_0 = us0_next;
            roll.wsum = 0.0;
            double ccol = 0.0, dcol = 0.0, datacol = 0.0;
            Pre pend[KC];
            if (has_fwd) us0_next = first_ustage(tile + gridDim.x);
            double *op = outp + (int64_t)(nstages - 1) * KC * sk;

            for (int ks = 0; ks < nstages; ++ks) {
                const int kb = nstages - 1 - ks;  // backward stage of the old tile
                Raw4 raw[KC];
                double cc[KC], dc[KC];
                if (has_bwd) pop_begin(kb, dir ^ 1, raw, cc, dc);
                Pre next[KC];
                if (has_fwd) {
                    Lvl l[KC];
                    fetch_stage(l);
#pragma unroll
                    for (int u = 0; u < KC; ++u) next[u] = pre_level(l[u], roll, dtr);
                }
                if (has_bwd) {
                    mbar_wait(&fullB[sB], phB);
                    const double *slot = reinterpret_cast<const double *>(ringB + (size_t)sB * G::SLOT_B) + t;
                    double up[KC];
#pragma unroll
                    for (int u = 0; u < KC; ++u) up[u] = slot[u * VM_COLS];
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&emptyB[sB]);
                    if (++sB == SB) { sB = 0; phB ^= 1; }
                    pop_finish(kb, dir ^ 1, raw, cc, dc);
#pragma unroll
                    for (int u = KC - 1; u >= 0; --u) {
                        const bool live = kb * KC + u < nk;  // pad levels exist only in the top stage
                        const double d = dc[u] - cc[u] * datacol;
                        datacol = live ? d : datacol;
                        if (valid_old && live) __stcs(op + u * sk, dtr * (datacol - up[u]));
                    }
                    op -= KC * sk;
                }
                if (has_fwd) {
                    if (ks > 0) {
                        double cs[KC], ds[KC];
                        solve(pend, ccol, dcol, cs, ds);
                        push_stage(ks - 1, dir, cs, ds);
                    }
#pragma unroll
                    for (int u = 0; u < KC; ++u) pend[u] = next[u];
                }
            }
            if (has_fwd) {
                double cs[KC], ds[KC];
                solve(pend, ccol, dcol, cs, ds);
                push_stage(nstages - 1, dir, cs, ds);
                const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
                valid_old = i < a.ni && j < a.nj;
                outp = a.utens_stage.p + i + (int64_t)j * a.utens_stage.sj;
            }
            tmem_push_wait();
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == VM_CWARPS) tmem_dealloc(tmem_base, VM_TMEM_COLS);
}

template <int TI, int KC, bool FUSED>
static int launch_geom(const VadvArgs &a, cudaStream_t stream, bool *handled)
{
    using G = VmGeom<TI, KC>;
    static int max_smem = -1, sms = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    static const int env_s = env_int("B200_VADV_STAGES", 0);
    static const int env_sb = env_int("B200_VADV_BSTAGES", 0);
    const int L = (a.nk + KC - 1) / KC * KC;  // LIFO slots (levels padded to whole stages)
    const int nT = L < VM_TMEM_LEVELS ? L : VM_TMEM_LEVELS;
    const size_t lifo = (size_t)(L - nT) * 2 * VM_COLS * 8;
    int SB = env_sb > 0 ? env_sb : 8;
    if (SB > VM_MAX_STAGES) SB = VM_MAX_STAGES;
    long avail = (long)max_smem - VM_BAR_BYTES - (long)lifo;
    // give the backward ring at most a quarter of what is left
    while (SB > 2 && (long)SB * G::SLOT_B > avail / 4) --SB;
    avail -= (long)SB * G::SLOT_B;
    int S = avail > 0 ? (int)(avail / G::SLOT_F) : 0;
    const int want = env_s > 0 ? env_s : 6;
    if (S > want) S = want;
    if (S > VM_MAX_STAGES) S = VM_MAX_STAGES;
    if (S < 2) return B200_OK;  // LIFO does not fit on chip: not handled here
    const size_t smem = VM_BAR_BYTES + (size_t)S * G::SLOT_F + (size_t)SB * G::SLOT_B + lifo;

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
    auto kern = FUSED ? vadv_fused_kernel<TI, KC> : vadv_tmem_kernel<TI, KC>;
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    note_kernel("vadv_tmem_kernel");
    kern<<<grid, VM_THREADS, smem, stream>>>(a, m_us, m_w, m_up, m_ut, m_uts, S, SB, nT, tiles_i, ntiles);
    B200_CUDA_CHECK(cudaGetLastError());
    *handled = true;
    return B200_OK;
}

// Launch the TMEM kernel if the layout allows it (same preconditions as vadv_tma.cu); *handled = false
// with B200_OK means "use another variant".  variant 7: KC = 2, variant 8: KC = 4, variant 9 (and auto):
// fused forward/backward sweeps.
int launch_vadv_tmem(const VadvArgs &a, int variant, cudaStream_t stream, bool *handled)
{
    *handled = false;
    if (!tma_compatible(a.u_stage.p, a.u_stage.si, a.u_stage.sj, a.u_stage.sk) ||
        !tma_compatible(a.wcon.p, a.wcon.si, a.wcon.sj, a.wcon.sk) ||
        !tma_compatible(a.u_pos.p, a.u_pos.si, a.u_pos.sj, a.u_pos.sk) ||
        !tma_compatible(a.utens.p, a.utens.si, a.utens.sj, a.utens.sk) ||
        !tma_compatible(a.utens_stage.p, a.utens_stage.si, a.utens_stage.sj, a.utens_stage.sk))
        return B200_OK;
    if (variant == 8)
        return a.ni > 64 ? launch_geom<128, 4, false>(a, stream, handled)
             : a.ni > 32 ? launch_geom<64, 4, false>(a, stream, handled)
                         : launch_geom<32, 4, false>(a, stream, handled);
    if (variant == 7)
        return a.ni > 64 ? launch_geom<128, 2, false>(a, stream, handled)
             : a.ni > 32 ? launch_geom<64, 2, false>(a, stream, handled)
                         : launch_geom<32, 2, false>(a, stream, handled);
    return a.ni > 64 ? launch_geom<128, 2, true>(a, stream, handled)
         : a.ni > 32 ? launch_geom<64, 2, true>(a, stream, handled)
                     : launch_geom<32, 2, true>(a, stream, handled);
}

}  // namespace b200
