// vadv.cu -- COSMO vertical advection (tridiagonal set-up + Thomas solve), hand-written for sm_100a.
//
// Reference semantics: vertical_advection_dycore, tests/cartesian_tests/integration_tests/
// multi_feature_tests/stencil_definitions.py:235-313 of the reference (externals BET_M=BET_P=0.5).
// Operator association follows the definition exactly; compiled with -fmad=false and IEEE division,
// so results are bit-identical to the reference numpy backend.
//
// Kernels
//   vadv_column<PF> (variant 2): one thread per (i, j) column, warps along the contiguous I axis.
//       FORWARD sweep keeps u_stage[k-1..k+1] and the wcon pair sums in registers (rolling buffer),
//       prefetches PF levels ahead, and pushes the Thomas coefficients ccol/dcol of every level
//       onto a per-thread LIFO in SHARED MEMORY ([level][thread], conflict-free); the BACKWARD sweep
//       pops them.  ccol/dcol/datacol therefore never touch HBM: DRAM traffic is the 5 input
//       fields + 1 output (48 B/point) plus the backward re-read of u_pos (normally an L2 hit).
//   vadv_column_gmem (variant 1): same arithmetic, LIFO in a global scratch buffer; works for any
//       nk and any strides (safe path, also the "GridTools-style" 80 B/point baseline).
#include "vadv.cuh"

namespace b200 {

struct Ptrs {
    const double *__restrict__ us, *__restrict__ w0, *__restrict__ w1, *__restrict__ up,
        *__restrict__ ut;
    double *__restrict__ uts;
    int64_t sk_us, sk_w, sk_up, sk_ut, sk_uts;
};

__device__ __forceinline__ Lvl load_level(const Ptrs &p, int k, int nk)
{
    Lvl l;
    if (k < nk) {
        l.up = __ldg(p.up + k * p.sk_up);
        l.ut = __ldcs(p.ut + k * p.sk_ut);
        l.uts = __ldcs(p.uts + k * p.sk_uts);
    } else {
        l.up = l.ut = l.uts = 0.0;
    }
    if (k + 1 < nk) {
        l.w0 = __ldg(p.w0 + (k + 1) * p.sk_w);
        l.w1 = __ldg(p.w1 + (k + 1) * p.sk_w);
        l.us = __ldcs(p.us + (k + 1) * p.sk_us);
    } else {
        l.w0 = l.w1 = l.us = 0.0;
    }
    return l;
}

__device__ __forceinline__ Ptrs make_ptrs(const VadvArgs &a, int i, int j)
{
    Ptrs p;
    p.us = a.u_stage.p + i * a.u_stage.si + j * a.u_stage.sj;
    p.w0 = a.wcon.p + i * a.wcon.si + j * a.wcon.sj;
    p.w1 = p.w0 + a.wcon.si;
    p.up = a.u_pos.p + i * a.u_pos.si + j * a.u_pos.sj;
    p.ut = a.utens.p + i * a.utens.si + j * a.utens.sj;
    p.uts = a.utens_stage.p + i * a.utens_stage.si + j * a.utens_stage.sj;
    p.sk_us = a.u_stage.sk;
    p.sk_w = a.wcon.sk;
    p.sk_up = a.u_pos.sk;
    p.sk_ut = a.utens.sk;
    p.sk_uts = a.utens_stage.sk;
    return p;
}

// LIFO accessors: STASH(k) -> reference to the (ccol, dcol) slot of level k for this thread.
template <int PF, typename Stash>
__device__ __forceinline__ void vadv_sweeps(const VadvArgs &a, const Ptrs &p, Stash stash)
{
    const int nk = a.nk;
    const double dtr = a.dtr_stage;
    // ---- FORWARD ----
    Fwd s;
    s.us_m1 = 0.0;
    s.us_0 = __ldcs(p.us);
    s.wsum = 0.0;
    s.ccol_m1 = s.dcol_m1 = 0.0;
    Lvl q[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) q[u] = load_level(p, u, nk);
    for (int kb = 0; kb < nk; kb += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int k = kb + u;
            if (k < nk) {
                Lvl l = q[u];
                q[u] = load_level(p, k + PF, nk);
                double cc, dc;
                forward_level(l, s, dtr, k, nk, cc, dc);
                stash(k, cc, dc, /*store=*/true);
            }
        }
    }
    // ---- BACKWARD ----  :305-312
    double upq[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) upq[u] = (nk - 1 - u >= 0) ? __ldg(p.up + (nk - 1 - u) * p.sk_up) : 0.0;
    double datacol = 0.0;
    for (int kb = nk - 1; kb >= 0; kb -= PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int k = kb - u;
            if (k >= 0) {
                double up = upq[u];
                upq[u] = (k - PF >= 0) ? __ldg(p.up + (k - PF) * p.sk_up) : 0.0;
                double cc, dc;
                stash(k, cc, dc, /*store=*/false);
                datacol = (k == nk - 1) ? dc : dc - cc * datacol;
                __stcs(p.uts + k * p.sk_uts, dtr * (datacol - up));
            }
        }
    }
}

template <int PF>
__global__ void __launch_bounds__(160, 1) vadv_column(VadvArgs a, int tpb)
{
    extern __shared__ double smem[];  // [nk][2][tpb]
    const int t = threadIdx.x;
    const int i = blockIdx.x * 32 + (t & 31);
    const int j = blockIdx.y * (tpb >> 5) + (t >> 5);
    if (i >= a.ni || j >= a.nj) return;
    Ptrs p = make_ptrs(a, i, j);
    double *base = smem + t;
    const int lv = 2 * tpb;
    vadv_sweeps<PF>(a, p, [=](int k, double &cc, double &dc, bool store) {
        double *slot = base + k * lv;
        if (store) {
            slot[0] = cc;
            slot[tpb] = dc;
        } else {
            cc = slot[0];
            dc = slot[tpb];
        }
    });
}

__global__ void __launch_bounds__(128) vadv_column_gmem(VadvArgs a, double *scratch)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= a.ni || j >= a.nj) return;
    Ptrs p = make_ptrs(a, i, j);
    const int64_t cols = (int64_t)a.ni * a.nj;
    double *base = scratch + (int64_t)j * a.ni + i;
    vadv_sweeps<2>(a, p, [=](int k, double &cc, double &dc, bool store) {
        double *slot = base + (int64_t)k * 2 * cols;
        if (store) {
            slot[0] = cc;
            slot[cols] = dc;
        } else {
            cc = slot[0];
            dc = slot[cols];
        }
    });
}

// scratch for the global-memory variant, grown on demand (per process; guarded by the caller's
// stream ordering: launches on different streams that need scratch are not supported concurrently)
static double *g_scratch = nullptr;
static size_t g_scratch_bytes = 0;

// ---- b200_vadv_noalloc: the same kernel with its LIFO in a fixed device array (no allocation, ever) -------------
// File-scope scratch: B200_VADV_NOALLOC_MAX_POINTS x (ccol, dcol) = 32 MiB, held by every GPU that loads the
// library until the process exits.  In HOST code the name `g_noalloc_scratch` is only a stand-in for the symbol, not
// a device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
// Nothing in it has to start at zero: the FORWARD sweep stores slot (k, column) before the BACKWARD sweep of the
// same thread reads it, so the previous call's contents are never observed.
static __device__ __align__(256) double g_noalloc_scratch[2 * B200_VADV_NOALLOC_MAX_POINTS];
constexpr int NOALLOC_MAX_DEVICES = 64;
static double *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};  // device address of the array, per device ordinal

static int noalloc_scratch(double **out)
{
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "vadv_noalloc: device ordinal %d out of range", dev);
    double *p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {  // set-up, once per device (a race between two threads resolves to the same address)
        void *sym = nullptr;
        B200_CUDA_CHECK(cudaGetSymbolAddress(&sym, g_noalloc_scratch));
        B200_REQUIRE(sym != nullptr, "vadv_noalloc: cudaGetSymbolAddress returned a null address");
        p = static_cast<double *>(sym);
        __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return B200_OK;
}

}  // namespace b200

extern "C" int b200_vadv_noalloc(const b200_field *utens_stage, const b200_field *u_stage, const b200_field *wcon,
                                 const b200_field *u_pos, const b200_field *utens, double dtr_stage,
                                 const int64_t domain[3], void *stream_)
{
    using namespace b200;
    cudaStream_t stream = (cudaStream_t)stream_;
    B200_REQUIRE(domain, "domain is null");
    B200_REQUIRE(domain[0] >= 0 && domain[1] >= 0 && domain[2] >= 0, "negative domain");
    B200_REQUIRE(domain[0] < (1 << 30) && domain[1] < (1 << 30) && domain[2] < (1 << 30), "domain too large");
    if (domain[0] == 0 || domain[1] == 0 || domain[2] == 0) return B200_OK;
    B200_REQUIRE(domain[2] >= 3, "vadv needs at least 3 vertical levels (got %lld)", (long long)domain[2]);
    const long long points = (long long)domain[0] * domain[1] * domain[2];
    B200_REQUIRE(points <= (long long)B200_VADV_NOALLOC_MAX_POINTS,
                 "vadv_noalloc: domain %lld x %lld x %lld = %lld grid points exceeds the fixed scratch of %lld points "
                 "(B200_VADV_NOALLOC_MAX_POINTS); use b200_vadv for larger domains",
                 (long long)domain[0], (long long)domain[1], (long long)domain[2], points,
                 (long long)B200_VADV_NOALLOC_MAX_POINTS);
    const int h0[3] = {0, 0, 0}, hw[3] = {1, 0, 0};
    VadvArgs a;
    int rc;
    if ((rc = make_view<double>(utens_stage, "utens_stage", B200_FLOAT64, domain, h0, h0, &a.utens_stage))) return rc;
    if ((rc = make_view<const double>(u_stage, "u_stage", B200_FLOAT64, domain, h0, h0, &a.u_stage))) return rc;
    if ((rc = make_view<const double>(wcon, "wcon", B200_FLOAT64, domain, h0, hw, &a.wcon))) return rc;
    if ((rc = make_view<const double>(u_pos, "u_pos", B200_FLOAT64, domain, h0, h0, &a.u_pos))) return rc;
    if ((rc = make_view<const double>(utens, "utens", B200_FLOAT64, domain, h0, h0, &a.utens))) return rc;
    a.dtr_stage = dtr_stage;
    a.ni = (int)domain[0];
    a.nj = (int)domain[1];
    a.nk = (int)domain[2];
    double *scratch = nullptr;
    if ((rc = noalloc_scratch(&scratch))) return rc;
    dim3 grid((a.ni + 127) / 128, a.nj, 1);
    note_kernel("vadv_column_gmem");
    vadv_column_gmem<<<grid, 128, 0, stream>>>(a, scratch);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_vadv(const b200_field *utens_stage, const b200_field *u_stage,
                         const b200_field *wcon, const b200_field *u_pos, const b200_field *utens,
                         double dtr_stage, const int64_t domain[3], int variant, void *stream_)
{
    using namespace b200;
    cudaStream_t stream = (cudaStream_t)stream_;
    B200_REQUIRE(domain, "domain is null");
    B200_REQUIRE(domain[0] >= 0 && domain[1] >= 0 && domain[2] >= 0, "negative domain");
    B200_REQUIRE(domain[0] < (1 << 30) && domain[1] < (1 << 30), "domain too large");
    if (domain[0] == 0 || domain[1] == 0 || domain[2] == 0) return B200_OK;
    B200_REQUIRE(domain[2] >= 3, "vadv needs at least 3 vertical levels (got %lld)",
                 (long long)domain[2]);
    const int h0[3] = {0, 0, 0}, hw[3] = {1, 0, 0};
    VadvArgs a;
    int rc;
    if ((rc = make_view<double>(utens_stage, "utens_stage", B200_FLOAT64, domain, h0, h0, &a.utens_stage))) return rc;
    if ((rc = make_view<const double>(u_stage, "u_stage", B200_FLOAT64, domain, h0, h0, &a.u_stage))) return rc;
    if ((rc = make_view<const double>(wcon, "wcon", B200_FLOAT64, domain, h0, hw, &a.wcon))) return rc;
    if ((rc = make_view<const double>(u_pos, "u_pos", B200_FLOAT64, domain, h0, h0, &a.u_pos))) return rc;
    if ((rc = make_view<const double>(utens, "utens", B200_FLOAT64, domain, h0, h0, &a.utens))) return rc;
    a.dtr_stage = dtr_stage;
    a.ni = (int)domain[0];
    a.nj = (int)domain[1];
    a.nk = (int)domain[2];

    static int max_smem = -1;
    if (max_smem < 0) {
        int dev = 0;
        B200_CUDA_CHECK(cudaGetDevice(&dev));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    }
    if (variant == 0 || variant == 10 || variant == 11) {  // tuned fused-sweep kernel (vadv_fused.cu)
        bool handled = false;  // variant 10: 2 levels per ring stage, 11: 4 levels per stage
        if ((rc = launch_vadv_fused(a, variant == 11 ? 4 : 2, stream, &handled))) return rc;
        if (handled) return B200_OK;
        if (variant != 0) {
            set_error("vadv variant %d needs unit I stride, 16-byte aligned origins, even J/K strides and nk "
                      "small enough for the on-chip LIFO", variant);
            return B200_EUNSUP;
        }
    }
    if (variant >= 7 && variant <= 9) {  // TMEM-LIFO kernel, 256 columns / SM (vadv_tmem.cu)
        bool handled = false;
        if ((rc = launch_vadv_tmem(a, variant, stream, &handled))) return rc;
        if (handled) return B200_OK;
        if (variant != 0) {
            set_error("vadv variant %d (TMEM) needs unit I stride, 16-byte aligned origins, even J/K strides "
                      "and nk small enough for the on-chip LIFO", variant);
            return B200_EUNSUP;
        }
    }
    if (variant == 0 || (variant >= 4 && variant <= 6)) {  // TMA-fed shared-memory LIFO (vadv_tma.cu)
        bool handled = false;
        if ((rc = launch_vadv_tma(a, variant, stream, &handled))) return rc;
        if (handled) return B200_OK;
        if (variant != 0) {
            set_error("vadv variant %d (TMA) needs unit I stride, 16-byte aligned origins, even J/K strides "
                      "and nk small enough for the shared-memory LIFO", variant);
            return B200_EUNSUP;
        }
    }
    // threads per block for the smem-LIFO kernel: as many warps as the LIFO allows, at most 5
    int warps = (int)(max_smem / ((size_t)a.nk * 16 * 32));
    if (warps > 5) warps = 5;
    if (variant == 0) variant = 1;  // generic fallback: global-memory LIFO, any strides / nk

    if (variant == 2 || variant == 3) {
        if (warps < 1) {
            set_error("vadv variant %d: nk=%d does not fit the shared-memory LIFO", variant, a.nk);
            return B200_EUNSUP;
        }
        static const int env_w = env_int("B200_VADV_WARPS", 0);
        if (env_w > 0 && env_w < warps) warps = env_w;
        const int tpb = warps * 32;
        const size_t smem = (size_t)a.nk * 16 * tpb;
        dim3 grid((a.ni + 31) / 32, (a.nj + warps - 1) / warps, 1);
        if (variant == 2) {
            B200_CUDA_CHECK(cudaFuncSetAttribute(vadv_column<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            note_kernel("vadv_column");
            vadv_column<4><<<grid, tpb, smem, stream>>>(a, tpb);
        } else {
            B200_CUDA_CHECK(cudaFuncSetAttribute(vadv_column<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            note_kernel("vadv_column");
            vadv_column<8><<<grid, tpb, smem, stream>>>(a, tpb);
        }
    } else if (variant == 1) {
        size_t need = (size_t)a.ni * a.nj * a.nk * 16;
        if (need > g_scratch_bytes) {
            if (g_scratch) B200_CUDA_CHECK(cudaFree(g_scratch));
            g_scratch = nullptr;
            g_scratch_bytes = 0;
            B200_CUDA_CHECK(cudaMalloc(&g_scratch, need));
            g_scratch_bytes = need;
        }
        dim3 grid((a.ni + 127) / 128, a.nj, 1);
        note_kernel("vadv_column_gmem");
        vadv_column_gmem<<<grid, 128, 0, stream>>>(a, g_scratch);
    } else {
        set_error("vadv: unknown variant %d", variant);
        return B200_EUNSUP;
    }
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}
