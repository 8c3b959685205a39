// tridiag_fused.cu -- generic Thomas solver on the machine mapping of vadv_fused.cu (variant 8 and `auto`).
//
// Reference semantics: tridiagonal_solver, tests/cartesian_tests/integration_tests/multi_feature_tests/
// stencil_definitions.py:219-232 of the reference:
//   FORWARD  [0,1):    sup = sup / diag;                         rhs = rhs / diag
//            [1,None): sup = sup / (diag - sup[k-1] * inf);      rhs = (rhs - inf * rhs[k-1]) / (diag - sup[k-1] * inf)
//   BACKWARD [-1,None): out = rhs;    [0,-1): out = rhs - sup * out[k+1]
// sup and rhs are API fields: the normalised values are results and are written back, but the BACKWARD sweep does
// not re-read them from HBM -- each column parks its (sup, rhs) of all NK levels on chip, which is the same
// 16 B x NK per column LIFO as vadv's (ccol, dcol).  So: a persistent CTA per SM walks 256-column tiles; one
// producer thread feeds a ring of TMA tile loads (inf, diag, sup, rhs; 2 levels per stage; out-of-range columns
// and the odd top level are zero-filled by the TMA unit and never stored); eight consumer warps own one column per
// thread, keep levels [0, 64) of the LIFO in Tensor Memory (tcgen05.st / tcgen05.ld as a per-thread stack) and the
// rest in shared memory, and interleave the BACKWARD sweep of their previous tile with the FORWARD sweep of the
// next one (the stack direction alternates per tile, so a popped slot is refilled in the same iteration).
// Divisions are IEEE (-prec-div=true), operations in the reference's order, no FMA contraction: bit-identical.
// Algorithmic traffic 56 B/point (4 reads + 3 writes) is also the DRAM traffic.
#include "vadv_tmem.cuh"

namespace b200 {

struct TriFArgs {
    View3<const double> inf, diag;
    View3<double> sup, rhs, out;
    int ni, nj, nk;
};

namespace {

constexpr int TK = 2;  // levels per stage

template <int TI>
struct TriGeom {
    static constexpr int TJ = VM_COLS / TI;
    static constexpr int F_BYTES = TK * VM_COLS * 8;
    static constexpr int SLOT_F = 4 * F_BYTES;
    static constexpr uint32_t TX_F = 4 * F_BYTES;
};

template <int TI>
struct TriFused {
    using G = TriGeom<TI>;

    struct Ctx {
        int t, lane, nk, nstages, nTs, S;
        uint64_t *fullF, *emptyF;
        const unsigned char *ringF;
        double *lifo_t;   // shared-memory part of the LIFO, this thread's column
        uint32_t tstack;  // Tensor-Memory part
        int sF;
        uint32_t phF;
        // forward sweep of the new tile
        double sup_m1, rhs_m1;
        double *sp, *rp;  // sup(i, j, 2 * ks), rhs(i, j, 2 * ks) of the new tile
        int64_t ssk, rsk;
        bool valid_new;
        // backward sweep of the old tile
        double outp;
        double *op;       // out(i, j, 2 * kb) of the old tile
        int64_t osk;
        bool valid_old;
    };

    __device__ static __forceinline__ void push(Ctx &c, int q, double s0, double r0, double s1, double r1)
    {
        if (q < c.nTs) {
            tmem_push2(c.tstack + 8u * (uint32_t)q, s0, r0, s1, r1);
        } else {
            double *s = c.lifo_t + (size_t)(q - c.nTs) * 4 * VM_COLS;
            s[0] = s0;
            s[VM_COLS] = r0;
            s[2 * VM_COLS] = s1;
            s[3 * VM_COLS] = r1;
        }
    }

    template <bool FWD, bool BWD>
    __device__ static __forceinline__ void sweep(Ctx &c, const int dir)
    {
        const int t = c.t;
        const bool pad_top = (c.nk & 1) != 0;  // level 2 * nstages - 1 does not exist
#pragma unroll 1
        for (int ks = 0; ks < c.nstages; ++ks) {
            const int q = dir ? c.nstages - 1 - ks : ks;  // LIFO slot popped now and refilled below
            const bool tm = q < c.nTs;
            Raw8 raw;
            double s0, r0, s1, r1;
            if (BWD) {
                if (tm) {
                    tmem_pop2_issue(c.tstack + 8u * (uint32_t)q, raw);
                } else {
                    const double *s = c.lifo_t + (size_t)(q - c.nTs) * 4 * VM_COLS;
                    s0 = s[0];
                    r0 = s[VM_COLS];
                    s1 = s[2 * VM_COLS];
                    r1 = s[3 * VM_COLS];
                }
            }
            double fs0, fr0, fs1, fr1;
            if (FWD) {
                mbar_wait(&c.fullF[c.sF], c.phF);
                const double *slot = reinterpret_cast<const double *>(c.ringF + (size_t)c.sF * G::SLOT_F) + t;
                const double inf0 = slot[0], inf1 = slot[VM_COLS];
                const double dia0 = slot[2 * VM_COLS], dia1 = slot[3 * VM_COLS];
                const double sup0 = slot[4 * VM_COLS], sup1 = slot[5 * VM_COLS];
                const double rhs0 = slot[6 * VM_COLS], rhs1 = slot[7 * VM_COLS];
                __syncwarp();
                if (c.lane == 0) mbar_arrive(&c.emptyF[c.sF]);
                if (++c.sF == c.S) { c.sF = 0; c.phF ^= 1; }
                if (ks == 0) {  // interval(0, 1)  :222-224
                    fs0 = sup0 / dia0;
                    fr0 = rhs0 / dia0;
                } else {        // interval(1, None)  :225-227
                    fs0 = sup0 / (dia0 - c.sup_m1 * inf0);
                    fr0 = (rhs0 - inf0 * c.rhs_m1) / (dia0 - c.sup_m1 * inf0);
                }
                fs1 = sup1 / (dia1 - fs0 * inf1);
                fr1 = (rhs1 - inf1 * fr0) / (dia1 - fs0 * inf1);
                c.sup_m1 = fs1;
                c.rhs_m1 = fr1;
                if (c.valid_new) {
                    __stcs(c.sp, fs0);
                    __stcs(c.rp, fr0);
                    if (!(pad_top && ks == c.nstages - 1)) {
                        __stcs(c.sp + c.ssk, fs1);
                        __stcs(c.rp + c.rsk, fr1);
                    }
                }
                c.sp += 2 * c.ssk;
                c.rp += 2 * c.rsk;
            }
            if (BWD) {
                if (tm) {
                    tmem_pop2_wait(raw);
                    s0 = __hiloint2double((int)raw.r[1], (int)raw.r[0]);
                    r0 = __hiloint2double((int)raw.r[3], (int)raw.r[2]);
                    s1 = __hiloint2double((int)raw.r[5], (int)raw.r[4]);
                    r1 = __hiloint2double((int)raw.r[7], (int)raw.r[6]);
                }
                // levels 2*kb+1 then 2*kb of the old tile (kb = nstages-1-ks)  :228-232
                if (ks == 0) {
                    if (pad_top) {
                        c.outp = r0;  // level nk-1 = 2*kb: out = rhs
                    } else {
                        const double o1 = r1;  // level nk-1 = 2*kb+1
                        if (c.valid_old) __stcs(c.op + c.osk, o1);
                        c.outp = r0 - s0 * o1;
                    }
                } else {
                    const double o1 = r1 - s1 * c.outp;
                    if (c.valid_old) __stcs(c.op + c.osk, o1);
                    c.outp = r0 - s0 * o1;
                }
                if (c.valid_old) __stcs(c.op, c.outp);
                c.op -= 2 * c.osk;
            }
            if (FWD) push(c, q, fs0, fr0, fs1, fr1);
        }
    }
};

}  // namespace

template <int TI>
__global__ void __launch_bounds__(VM_COLS + 32, 1)
tridiag_fused_kernel(TriFArgs a, const __grid_constant__ CUtensorMap m_inf, const __grid_constant__ CUtensorMap m_diag,
                     const __grid_constant__ CUtensorMap m_sup, const __grid_constant__ CUtensorMap m_rhs, int S, int nTs,
                     int tiles_i, int ntiles)
{
    using F = TriFused<TI>;
    using G = typename F::G;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *fullF = reinterpret_cast<uint64_t *>(smem);
    uint64_t *emptyF = fullF + VM_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(emptyF + VM_MAX_STAGES);
    unsigned char *ringF = smem + VM_BAR_BYTES;
    double *lifo = reinterpret_cast<double *>(ringF + (size_t)S * G::SLOT_F);  // stage slots [nTs, nstages)

    const int nk = a.nk;
    const int nstages = (nk + TK - 1) / TK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&fullF[s], 1);
            mbar_init(&emptyF[s], VM_CWARPS);
        }
        mbar_fence_init();
    }
    if (warp == VM_CWARPS) tmem_alloc(tmem_slot, VM_TMEM_COLS);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == VM_CWARPS) {
        // ================= producer: one thread issues the TMA loads ==================================
        if (lane == 0) {
            tma_prefetch_desc(&m_inf);
            tma_prefetch_desc(&m_diag);
            tma_prefetch_desc(&m_sup);
            tma_prefetch_desc(&m_rhs);
            int sF = 0;
            uint32_t phF = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const int ci = (tile % tiles_i) * TI, cj = (tile / tiles_i) * G::TJ;
                for (int ks = 0; ks < nstages; ++ks) {
                    mbar_wait(&emptyF[sF], phF ^ 1);
                    mbar_expect_tx(&fullF[sF], G::TX_F);
                    unsigned char *slot = ringF + (size_t)sF * G::SLOT_F;
                    const int k0 = ks * TK;
                    tma_load_3d(slot, &m_inf, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + G::F_BYTES, &m_diag, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 2 * G::F_BYTES, &m_sup, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    tma_load_3d(slot + 3 * G::F_BYTES, &m_rhs, ci, cj, k0, &fullF[sF], L2_EVICT_FIRST);
                    if (++sF == S) { sF = 0; phF ^= 1; }
                }
            }
        }
        __syncwarp();
    } else {
        // =========================== consumers: one column per thread ===============================
        typename F::Ctx c;
        c.t = threadIdx.x;
        const int ti = c.t % TI, tj = c.t / TI;
        c.lane = lane;
        c.nk = nk;
        c.nstages = nstages;
        c.nTs = nTs;
        c.S = S;
        c.fullF = fullF;
        c.emptyF = emptyF;
        c.ringF = ringF;
        c.lifo_t = lifo + c.t;
        c.tstack = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(warp >> 2) * 256u;
        c.sF = 0;
        c.phF = 0;
        c.ssk = a.sup.sk;
        c.rsk = a.rhs.sk;
        c.osk = a.out.sk;
        c.op = nullptr;
        c.valid_old = false;
        c.outp = 0.0;

        const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
        for (int n = 0; n <= my_tiles && my_tiles > 0; ++n) {
            const int tile = blockIdx.x + n * gridDim.x;
            const int dir = n & 1;  // stack direction of the tile in its forward sweep
            const int i = (tile % tiles_i) * TI + ti, j = (tile / tiles_i) * G::TJ + tj;
            c.sup_m1 = c.rhs_m1 = 0.0;
            c.valid_new = n < my_tiles && i < a.ni && j < a.nj;
            c.sp = a.sup.p + (int64_t)i * a.sup.si + (int64_t)j * a.sup.sj;
            c.rp = a.rhs.p + (int64_t)i * a.rhs.si + (int64_t)j * a.rhs.sj;
            if (n == 0)
                F::template sweep<true, false>(c, dir);
            else if (n < my_tiles)
                F::template sweep<true, true>(c, dir);
            else
                F::template sweep<false, true>(c, dir);
            if (n < my_tiles) {  // this tile is the "old" one of the next round
                c.valid_old = i < a.ni && j < a.nj;
                c.op = a.out.p + (int64_t)i * a.out.si + (int64_t)j * a.out.sj + (int64_t)(nstages - 1) * TK * c.osk;
            }
            tmem_push_wait();
        }
    }

    // every warp arrives here; TMEM is released only after all consumers finished their last pop
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == VM_CWARPS) tmem_dealloc(tmem_base, VM_TMEM_COLS);
}

template <int TI>
static int launch_tri(const TriFArgs &a, cudaStream_t stream, bool *handled)
{
    using G = TriGeom<TI>;
    static int max_smem = -1, sms = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    static const int env_s = env_int("B200_TRIDIAG_STAGES", 0);
    const int nstages = (a.nk + TK - 1) / TK;
    const int tmem_slots = VM_TMEM_LEVELS / TK;  // 32 stage slots of 8 columns per thread
    const int nTs = nstages < tmem_slots ? nstages : tmem_slots;
    const size_t lifo = (size_t)(nstages - nTs) * 2 * TK * VM_COLS * 8;
    const long avail = (long)max_smem - VM_BAR_BYTES - (long)lifo;
    int S = avail > 0 ? (int)(avail / G::SLOT_F) : 0;
    const int want = env_s > 0 ? env_s : 4;  // measured at 1024x1024x80: S = 4 0.73 ms, 3 0.77, 5 0.76, 8 0.83 (deeper rings lose DRAM locality)
    if (S > want) S = want;
    if (S > VM_MAX_STAGES) S = VM_MAX_STAGES;
    if (S < 2) return B200_OK;  // LIFO does not fit on chip: not handled here
    const size_t smem = VM_BAR_BYTES + (size_t)S * G::SLOT_F + lifo;

    CUtensorMap m_inf, m_diag, m_sup, m_rhs;
    const int64_t ext[3] = {a.ni, a.nj, a.nk};
    const uint32_t box[3] = {TI, (uint32_t)G::TJ, TK};
    int rc;
    if ((rc = encode_tmap_f64(&m_inf, a.inf.p, ext, a.inf.sj, a.inf.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_diag, a.diag.p, ext, a.diag.sj, a.diag.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_sup, a.sup.p, ext, a.sup.sj, a.sup.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_rhs, a.rhs.p, ext, a.rhs.sj, a.rhs.sk, box))) return rc;

    const int tiles_i = (a.ni + TI - 1) / TI, tiles_j = (a.nj + G::TJ - 1) / G::TJ;
    const int ntiles = tiles_i * tiles_j;
    const int grid = ntiles < sms ? ntiles : sms;
    auto kern = tridiag_fused_kernel<TI>;
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    note_kernel("tridiag_fused_kernel");
    kern<<<grid, VM_COLS + 32, smem, stream>>>(a, m_inf, m_diag, m_sup, m_rhs, S, nTs, tiles_i, ntiles);
    B200_CUDA_CHECK(cudaGetLastError());
    *handled = true;
    return B200_OK;
}

// Launch the fused kernel if the layout allows it; *handled = false with B200_OK means "use another variant".
int launch_tridiag_fused(const View3<const double> &inf, const View3<const double> &diag, const View3<double> &sup,
                         const View3<double> &rhs, const View3<double> &out, int ni, int nj, int nk,
                         cudaStream_t stream, bool *handled)
{
    *handled = false;
    if (nk < 2) return B200_OK;
    if (!tma_compatible(inf.p, inf.si, inf.sj, inf.sk) || !tma_compatible(diag.p, diag.si, diag.sj, diag.sk) ||
        !tma_compatible(sup.p, sup.si, sup.sj, sup.sk) || !tma_compatible(rhs.p, rhs.si, rhs.sj, rhs.sk))
        return B200_OK;
    TriFArgs a{inf, diag, sup, rhs, out, ni, nj, nk};
    return ni > 64 ? launch_tri<128>(a, stream, handled)
         : ni > 32 ? launch_tri<64>(a, stream, handled)
                   : launch_tri<32>(a, stream, handled);
}

}  // namespace b200
