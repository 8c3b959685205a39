// hdiff_tma.cu -- horizontal diffusion with TMA-fed shared-memory rings (variant 5).
//
// Reference semantics: horizontal_diffusion, tests/cartesian_tests/integration_tests/multi_feature_tests/
// stencil_definitions.py:316-328 (limiter) and :206-216 (no limiter); arithmetic, operator association and
// rolling-window schedule are those of hdiff_jmarch (hdiff.cu), so the results are bit-identical to it
// and to the reference numpy backend.
//
// Machine mapping.  A CTA owns a strip of 60*NW output columns of one K-plane and marches over a chunk of
// CJ rows.  One elected producer thread streams the strip through a ring of NS stages with the TMA unit: per
// stage two 3-D tile loads of RS rows, in_field (strip + 2 halo columns each side, starting 2 rows above the
// chunk) and coeff (same box, shifted by 4 rows so that row r of a stage is the coefficient of the output row
// produced when row r of in_field arrives).  Out-of-range rows / columns are zero-filled by the TMA unit:
// no edge code on the load side.  NW consumer warps (64 loaded / 60 produced columns each, two columns per
// lane) read the rows from shared memory with one LDS.128 per field and row, keep the 3-row window of
// in_field and 2 rows of the Laplacian in registers, exchange the I+-1 neighbours with warp shuffles and
// store the result with 128-bit streaming stores.  Compared with hdiff_jmarch the consumers issue no global
// loads, no address arithmetic and no load predicates, and the number of bytes in flight per SM is set by the
// ring (NS * RS rows * 2 fields per CTA), not by registers.
#include "hdiff.cuh"
#include "tma.cuh"

namespace b200 {

namespace {

constexpr int HT_MAX_STAGES = 16;
constexpr int HT_BAR_BYTES = 256;  // full[16] + empty[16] mbarriers

template <bool LIMITER>
__device__ __forceinline__ double lim(double res, double d)
{
    if (LIMITER) return (res * d > 0.0) ? 0.0 : res;
    return res;
}

template <int NW, int RS>
struct HtGeom {
    static constexpr int OUT_COLS = 60 * NW;       // columns produced by a CTA
    static constexpr int BW = OUT_COLS + 4;        // columns loaded (box width)
    static constexpr int ROWB = BW * 8;            // bytes per box row in shared memory
    static constexpr int SUB = (RS * ROWB + 127) / 128 * 128;  // one field's rows of a stage
    static constexpr int SLOT = 2 * SUB;           // in_field rows, then coeff rows
    static constexpr uint32_t TX = 2u * RS * ROWB;
};

template <bool LIMITER, int NW, int RS>
__global__ void __launch_bounds__(32 * (NW + 1))
hdiff_tma_kernel(const __grid_constant__ CUtensorMap m_in, const __grid_constant__ CUtensorMap m_co,
                 View3<double> out, int ni, int nj, int cj, int NS)
{
    using G = HtGeom<NW, RS>;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *empty = full + HT_MAX_STAGES;
    unsigned char *ring = smem + HT_BAR_BYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i0 = blockIdx.x * G::OUT_COLS;
    const int j0 = blockIdx.y * cj;
    const int k = blockIdx.z;
    const int jend = min(j0 + cj, nj);        // exclusive end of the output rows of this chunk
    const int nsteps = jend - j0 + 4;         // rows of in_field to stream: [j0 - 2, jend + 2)
    const int nstages = (nsteps + RS - 1) / RS;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == NW) {
        // ============================ producer: one thread issues the TMA loads =========================
        if (lane == 0) {
            tma_prefetch_desc(&m_in);
            tma_prefetch_desc(&m_co);
            int slot = 0;
            uint32_t ph = 0;
            for (int s = 0; s < nstages; ++s) {
                mbar_wait(&empty[slot], ph ^ 1);
                mbar_expect_tx(&full[slot], G::TX);
                unsigned char *dst = ring + (size_t)slot * G::SLOT;
                // m_in is based at the halo corner (-2, -2): coordinate (i0, j0 + s*RS) is real (i0-2, j0-2+s*RS)
                tma_load_3d(dst, &m_in, i0, j0 + s * RS, k, &full[slot], L2_EVICT_NORMAL);
                tma_load_3d(dst + G::SUB, &m_co, i0 - 2, j0 - 4 + s * RS, k, &full[slot], L2_EVICT_FIRST);
                if (++slot == NS) { slot = 0; ph ^= 1; }
            }
        }
        return;
    }

    // ================================ consumers: two columns per lane ===================================
    const int cb = 60 * warp + 2 * lane;  // box column of the lane's first column
    const int col = i0 - 2 + cb;          // its real column
    const bool store_x = lane >= 1 && lane <= 30 && col < ni;
    const bool store_y = lane >= 1 && lane <= 30 && col + 1 < ni;

    double2 in0 = make_double2(0.0, 0.0), in1 = in0, lap0 = in0, flym1 = in0;
    double r0 = 0.0, lapr0 = 0.0;
    int jr = j0 - 4;  // output row of the current step (valid from j0 on)
    double *prow = out.p + (int64_t)k * out.sk + (int64_t)jr * out.sj + col;
    const int64_t osj = out.sj;

    int slot = 0;
    uint32_t ph = 0;
#pragma unroll 1
    for (int s = 0; s < nstages; ++s) {
        mbar_wait(&full[slot], ph);
        const unsigned char *base = ring + (size_t)slot * G::SLOT + cb * 8;
#pragma unroll
        for (int r = 0; r < RS; ++r) {
            const double2 in2 = *reinterpret_cast<const double2 *>(base + r * G::ROWB);
            const double2 co = *reinterpret_cast<const double2 *>(base + G::SUB + r * G::ROWB);
            // Laplacian of row jr+1 on the own two columns
            const double l1 = __shfl_up_sync(0xffffffffu, in1.y, 1);    // in(a-1, jr+1)
            const double r1 = __shfl_down_sync(0xffffffffu, in1.x, 1);  // in(b+1, jr+1)
            double2 lap1;
            lap1.x = 4.0 * in1.x - (((in1.y + l1) + in2.x) + in0.x);
            lap1.y = 4.0 * in1.y - (((r1 + in1.x) + in2.y) + in0.y);
            const double lapr1 = __shfl_down_sync(0xffffffffu, lap1.x, 1);  // lap(b+1, jr+1)
            // fluxes of row jr
            double2 fly;
            fly.x = lim<LIMITER>(lap1.x - lap0.x, in1.x - in0.x);
            fly.y = lim<LIMITER>(lap1.y - lap0.y, in1.y - in0.y);
            const double fa = lim<LIMITER>(lap0.y - lap0.x, in0.y - in0.x);  // flx(a, jr)
            const double fb = lim<LIMITER>(lapr0 - lap0.y, r0 - in0.y);      // flx(b, jr)
            const double fl = __shfl_up_sync(0xffffffffu, fb, 1);            // flx(a-1, jr)
            if (jr >= j0 && jr < jend) {
                double2 o;
                o.x = in0.x - co.x * (((fa - fl) + fly.x) - flym1.x);
                o.y = in0.y - co.y * (((fb - fa) + fly.y) - flym1.y);
                if (store_y) __stcs(reinterpret_cast<double2 *>(prow), o);
                else if (store_x) __stcs(prow, o.x);
            }
            in0 = in1; in1 = in2; lap0 = lap1; flym1 = fly; r0 = r1; lapr0 = lapr1;
            ++jr;
            prow += osj;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        if (++slot == NS) { slot = 0; ph ^= 1; }
    }
}

template <bool LIMITER, int NW, int RS>
int launch_cfg(const View3<const double> &in, const View3<double> &out, const View3<const double> &co, int ni,
               int nj, int nk, int cj, int ns_want, cudaStream_t stream)
{
    using G = HtGeom<NW, RS>;
    static int max_smem = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    int NS = ns_want;
    if (NS > HT_MAX_STAGES) NS = HT_MAX_STAGES;
    while (NS > 2 && HT_BAR_BYTES + NS * G::SLOT > max_smem) --NS;
    const size_t smem = HT_BAR_BYTES + (size_t)NS * G::SLOT;

    CUtensorMap m_in, m_co;
    const int64_t ext_in[3] = {ni + 4, nj + 4, nk}, ext_co[3] = {ni, nj, nk};
    const uint32_t box[3] = {(uint32_t)G::BW, (uint32_t)RS, 1};
    int rc;
    if ((rc = encode_tmap_f64(&m_in, in.p - 2 - 2 * in.sj, ext_in, in.sj, in.sk, box))) return rc;
    if ((rc = encode_tmap_f64(&m_co, co.p, ext_co, co.sj, co.sk, box))) return rc;

    auto kern = hdiff_tma_kernel<LIMITER, NW, RS>;
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((ni + G::OUT_COLS - 1) / G::OUT_COLS, (nj + cj - 1) / cj, nk);
    note_kernel("hdiff_tma_kernel");
    kern<<<grid, 32 * (NW + 1), smem, stream>>>(m_in, m_co, out, ni, nj, cj, NS);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}

template <bool LIMITER, int RS>
int launch_nw(int nw, const View3<const double> &in, const View3<double> &out, const View3<const double> &co,
              int ni, int nj, int nk, int cj, int ns, cudaStream_t stream)
{
    switch (nw) {
    case 1: return launch_cfg<LIMITER, 1, RS>(in, out, co, ni, nj, nk, cj, ns, stream);
    case 2: return launch_cfg<LIMITER, 2, RS>(in, out, co, ni, nj, nk, cj, ns, stream);
    case 3: return launch_cfg<LIMITER, 3, RS>(in, out, co, ni, nj, nk, cj, ns, stream);
    default: return launch_cfg<LIMITER, 4, RS>(in, out, co, ni, nj, nk, cj, ns, stream);
    }
}

}  // namespace

// Launch the TMA-ring kernel if the layout allows it; *handled = false with B200_OK means "use another
// variant".  Tuning hooks (read at every launch, for tools/hdiff_sweep.py): B200_HDIFF_NW (consumer warps per
// CTA, 1..4), B200_HDIFF_RS (rows per stage, 4 or 8), B200_HDIFF_NS (ring depth), B200_HDIFF_CJ (rows per
// chunk).
int launch_hdiff_tma(bool limiter, const View3<const double> &in, const View3<double> &out,
                     const View3<const double> &co, int ni, int nj, int nk, cudaStream_t stream, bool *handled)
{
    *handled = false;
    if (!tma_compatible(in.p, in.si, in.sj, in.sk) || !tma_compatible(co.p, co.si, co.sj, co.sk) ||
        !(out.si == 1 && aligned16(out.p) && out.sj % 2 == 0 && out.sk % 2 == 0))
        return B200_OK;
    int nw = env_int("B200_HDIFF_NW", 0);
    if (nw < 1 || nw > 4) {  // fewest padded columns, wider strips preferred
        int best = 1 << 30;
        nw = 1;
        for (int w = 4; w >= 1; --w) {
            const int padded = (ni + 60 * w - 1) / (60 * w) * (60 * w);
            if (padded < best) { best = padded; nw = w; }
        }
    }
    const int rs = env_int("B200_HDIFF_RS", 4);
    int ns = env_int("B200_HDIFF_NS", 4);
    if (ns < 2) ns = 2;
    int cj = env_int("B200_HDIFF_CJ", 0);
    if (cj <= 0) cj = 32;  // measured optimum at 512^2 .. 2048^2 (tools/kbench.py --sweep): many short marches
    if (cj > nj) cj = nj;
    int rc;
    if (rs == 2)
        rc = limiter ? launch_nw<true, 2>(nw, in, out, co, ni, nj, nk, cj, ns, stream)
                     : launch_nw<false, 2>(nw, in, out, co, ni, nj, nk, cj, ns, stream);
    else if (rs == 8)
        rc = limiter ? launch_nw<true, 8>(nw, in, out, co, ni, nj, nk, cj, ns, stream)
                     : launch_nw<false, 8>(nw, in, out, co, ni, nj, nk, cj, ns, stream);
    else
        rc = limiter ? launch_nw<true, 4>(nw, in, out, co, ni, nj, nk, cj, ns, stream)
                     : launch_nw<false, 4>(nw, in, out, co, ni, nj, nk, cj, ns, stream);
    if (rc == B200_OK) *handled = true;
    return rc;
}

}  // namespace b200
