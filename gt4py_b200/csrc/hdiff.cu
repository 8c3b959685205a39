// hdiff.cu -- horizontal diffusion (lap-of-lap with optional flux limiter), hand-written for sm_100a.
//
// Reference semantics: tests/cartesian_tests/integration_tests/multi_feature_tests/
// stencil_definitions.py:316-328 (limiter) and :206-216 (no limiter) of the reference.  Operator
// association is kept exactly as written there and the file is compiled with -fmad=false, so the
// results are bit-identical to the reference numpy backend (which evaluates node by node).
//
// Kernels
//   hdiff_point   : one thread per grid point, any strides/alignment (variant 1, the safe path).
//   hdiff_jmarch  : the HBM-bound path (variant 2).  A warp owns 64 consecutive I columns (two per
//                   lane, 128-bit loads), marches along J keeping a 3-row rolling window of
//                   in_field and 2 rows of the Laplacian in registers, and exchanges the I+-1
//                   neighbours with warp shuffles.  Every in_field element is loaded once per
//                   J-strip; the 4 overlap columns/rows between neighbouring warps/strips are L1/L2
//                   hits, so DRAM traffic stays at ~24 B/point (8 in + 8 coeff + 8 out).
#include "hdiff.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// generic point kernel
// ------------------------------------------------------------------------------------------------
// T = double, or float for the pure-float32 build of the same definition (float32 fields and
// `literal_float_precision=32`: every literal and every temporary is float32, gtscript.py:242-290).
template <typename T>
__device__ __forceinline__ T lap_pt(const T *__restrict__ in, int64_t si, int64_t sj)
{
    return T(4.0) * in[0] - (((in[si] + in[-si]) + in[sj]) + in[-sj]);
}

template <bool LIMITER, typename T>
__device__ __forceinline__ T limit(T res, T d)
{
    if (LIMITER) return (res * d > T(0.0)) ? T(0.0) : res;
    return res;
}

template <typename T> struct Vec2;
template <> struct Vec2<double> { typedef double2 type; };
template <> struct Vec2<float> { typedef float2 type; };

template <bool LIMITER, typename T>
__global__ void __launch_bounds__(256)
hdiff_point(View3<const T> in, View3<T> out, View3<const T> coeff, int ni, int nj, int nk)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int j = blockIdx.y * blockDim.y + threadIdx.y;
    int k = blockIdx.z;
    if (i >= ni || j >= nj || k >= nk) return;
    const T *p = in.p + i * in.si + j * in.sj + k * in.sk;
    const int64_t si = in.si, sj = in.sj;
    T l0 = lap_pt(p, si, sj);
    T lxp = lap_pt(p + si, si, sj), lxm = lap_pt(p - si, si, sj);
    T lyp = lap_pt(p + sj, si, sj), lym = lap_pt(p - sj, si, sj);
    T flx = limit<LIMITER, T>(lxp - l0, p[si] - p[0]);
    T flxm = limit<LIMITER, T>(l0 - lxm, p[0] - p[-si]);
    T fly = limit<LIMITER, T>(lyp - l0, p[sj] - p[0]);
    T flym = limit<LIMITER, T>(l0 - lym, p[0] - p[-sj]);
    T c = coeff.p[i * coeff.si + j * coeff.sj + k * coeff.sk];
    out.p[i * out.si + j * out.sj + k * out.sk] = p[0] - c * (((flx - flxm) + fly) - flym);
}

// ------------------------------------------------------------------------------------------------
// J-marching register kernel
// ------------------------------------------------------------------------------------------------
constexpr int JM_COLS_PER_WARP = 60;  // 64 loaded, 60 produced (lanes 1..30)

// Load columns (col, col+1) of one row; columns outside [-2, ni+2) read as 0 (never used for a
// stored result).  `row` points at column 0 of the row.
template <typename T>
__device__ __forceinline__ typename Vec2<T>::type load_pair(const T *__restrict__ row, int col, int ni)
{
    typedef typename Vec2<T>::type V;
    V v;
    v.x = v.y = T(0.0);
    if (col >= -2) {
        if (col + 1 < ni + 2) {
            v = __ldg(reinterpret_cast<const V *>(row + col));
        } else if (col < ni + 2) {
            v.x = __ldg(row + col);
        }
    }
    return v;
}

template <bool LIMITER, int PF, typename T>
__global__ void __launch_bounds__(256)
hdiff_jmarch(View3<const T> in, View3<T> out, View3<const T> coeff, int ni, int nj, int nk, int tj)
{
    typedef typename Vec2<T>::type V;
    const V zero2 = {T(0.0), T(0.0)};
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const int col = (blockIdx.x * wpb + warp) * JM_COLS_PER_WARP - 2 + 2 * lane;
    const int j0 = blockIdx.y * tj;
    const int k = blockIdx.z;
    if (col - 2 * lane + 2 >= ni) return;  // first produced column >= ni: whole warp idle (uniform)
    const int jend = min(j0 + tj, nj);  // exclusive end of output rows
    const int jmax = nj + 2;            // exclusive end of readable rows

    const T *__restrict__ pin = in.p + (int64_t)k * in.sk;
    const T *__restrict__ pco = coeff.p + (int64_t)k * coeff.sk;
    T *__restrict__ pout = out.p + (int64_t)k * out.sk;
    const int64_t sj = in.sj;

    const bool store_x = lane >= 1 && lane <= 30 && col < ni;
    const bool store_y = lane >= 1 && lane <= 30 && col + 1 < ni;

    auto load_in = [&](int r) -> V {
        return (r < jmax) ? load_pair<T>(pin + (int64_t)r * sj, col, ni) : zero2;
    };
    auto load_co = [&](int r) -> V {
        V v = zero2;
        if (r >= j0 && r < jend) {
            const T *row = pco + (int64_t)r * coeff.sj;
            if (store_y) v = __ldcs(reinterpret_cast<const V *>(row + col));
            else if (store_x) v.x = __ldcs(row + col);
        }
        return v;
    };

    // rolling window
    V in0 = load_in(j0 - 2);
    V in1 = load_in(j0 - 1);
    V nin[PF], nco[PF];  // prefetch queues: rows jr+2 .. jr+1+PF of in, rows jr .. of coeff
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        nin[u] = load_in(j0 + u);
        nco[u] = load_co(j0 - 2 + u);
    }
    V lap0 = zero2, flym1 = zero2;
    T r0 = T(0.0), lapr0 = T(0.0);  // in(b+1, row jr) and lap(b+1, row jr)

    for (int jb = j0 - 2; jb < jend; jb += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int jr = jb + u;  // the row whose output is produced in this step
            V in2 = nin[u];
            V co = nco[u];
            nin[u] = load_in(jr + 2 + PF);
            nco[u] = load_co(jr + PF);
            // Laplacian of row jr+1 on the own two columns
            T l1 = __shfl_up_sync(0xffffffffu, in1.y, 1);    // in(a-1, jr+1)
            T r1 = __shfl_down_sync(0xffffffffu, in1.x, 1);  // in(b+1, jr+1)
            V lap1;
            lap1.x = T(4.0) * in1.x - (((in1.y + l1) + in2.x) + in0.x);
            lap1.y = T(4.0) * in1.y - (((r1 + in1.x) + in2.y) + in0.y);
            T lapr1 = __shfl_down_sync(0xffffffffu, lap1.x, 1);  // lap(b+1, jr+1)
            // fluxes of row jr
            V fly;
            fly.x = limit<LIMITER, T>(lap1.x - lap0.x, in1.x - in0.x);
            fly.y = limit<LIMITER, T>(lap1.y - lap0.y, in1.y - in0.y);
            T fa = limit<LIMITER, T>(lap0.y - lap0.x, in0.y - in0.x);  // flx(a, jr)
            T fb = limit<LIMITER, T>(lapr0 - lap0.y, r0 - in0.y);      // flx(b, jr)
            T fl = __shfl_up_sync(0xffffffffu, fb, 1);                 // flx(a-1, jr)
            if (jr >= j0 && jr < jend) {
                V o;
                o.x = in0.x - co.x * (((fa - fl) + fly.x) - flym1.x);
                o.y = in0.y - co.y * (((fb - fa) + fly.y) - flym1.y);
                T *row = pout + (int64_t)jr * out.sj;
                if (store_y) __stcs(reinterpret_cast<V *>(row + col), o);
                else if (store_x) __stcs(row + col, o.x);
            }
            // roll
            in0 = in1; in1 = in2; lap0 = lap1; flym1 = fly; r0 = r1; lapr0 = lapr1;
        }
    }
}

// One J-strip of one K-plane.  REV = false marches towards +J, REV = true towards -J: the marching row r of a
// reversed strip is the real row (j0 + jend - 1 - r).  Neighbouring strips march in OPPOSITE directions, so
// the 4 rows two strips share are read by both at the same time (both at the start or both at the end of
// their march) and the second read is an L2 hit; with a common direction the reads are a whole march
// (~30 us) apart and the line has long been evicted by the streaming traffic (ncu: +12.5 % in_field bytes).
// The reversed strip evaluates the same expression tree: under j -> -j the limited fluxes change sign
// exactly (the limiter test res * d is invariant), x + (-a) == x - a and x - (-b) == x + b in IEEE
// arithmetic, so ((X + fly[j]) - fly[j-1]) is computed as ((X - prev) + cur).
template <bool LIMITER, int PF, bool REV>
__device__ __forceinline__ void hdiff_strip(const View3<const double> &in, const View3<double> &out,
                                            const View3<const double> &coeff, int ni, int nj, int tj)
{
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const int col = (blockIdx.x * wpb + warp) * JM_COLS_PER_WARP - 2 + 2 * lane;
    const int j0 = blockIdx.y * tj;
    const int k = blockIdx.z;
    if (col - 2 * lane + 2 >= ni) return;  // first produced column >= ni: whole warp idle (uniform)
    const int jend = min(j0 + tj, nj);  // exclusive end of output rows
    const int mirror = j0 + jend - 1;   // REV: real row = mirror - marching row

    const double *__restrict__ pin = in.p + (int64_t)k * in.sk;
    const double *__restrict__ pco = coeff.p + (int64_t)k * coeff.sk;
    double *__restrict__ pout = out.p + (int64_t)k * out.sk;
    const int64_t sj = in.sj;

    const bool store_x = lane >= 1 && lane <= 30 && col < ni;
    const bool store_y = lane >= 1 && lane <= 30 && col + 1 < ni;

    auto load_in = [&](int r) -> double2 {
        // forward: r >= j0 - 2 >= -2 always; reverse: real <= jend + 1 <= nj + 1 always -> one compare each
        const int real = REV ? mirror - r : r;
        const bool ok = REV ? real >= -2 : real < nj + 2;
        return ok ? load_pair<double>(pin + (int64_t)real * sj, col, ni) : make_double2(0.0, 0.0);
    };
    auto load_co = [&](int r) -> double2 {
        double2 v = make_double2(0.0, 0.0);
        if (r >= j0 && r < jend) {
            const double *row = pco + (int64_t)(REV ? mirror - r : r) * coeff.sj;
            if (store_y) v = __ldcs(reinterpret_cast<const double2 *>(row + col));
            else if (store_x) v.x = __ldcs(row + col);
        }
        return v;
    };

    // rolling window (marching rows)
    double2 in0 = load_in(j0 - 2);
    double2 in1 = load_in(j0 - 1);
    double2 nin[PF], nco[PF];  // prefetch queues: rows jr+2 .. jr+1+PF of in, rows jr .. of coeff
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        nin[u] = load_in(j0 + u);
        nco[u] = load_co(j0 - 2 + u);
    }
    double2 lap0 = make_double2(0.0, 0.0), flym1 = make_double2(0.0, 0.0);
    double r0 = 0.0, lapr0 = 0.0;  // in(b+1, row jr) and lap(b+1, row jr)

    for (int jb = j0 - 2; jb < jend; jb += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int jr = jb + u;  // the marching row whose output is produced in this step
            double2 in2 = nin[u];
            double2 co = nco[u];
            nin[u] = load_in(jr + 2 + PF);
            nco[u] = load_co(jr + PF);
            // Laplacian of row jr+1 on the own two columns:  4*in - (((in[+1,0] + in[-1,0]) + in[0,+1]) + in[0,-1])
            double l1 = __shfl_up_sync(0xffffffffu, in1.y, 1);    // in(a-1, jr+1)
            double r1 = __shfl_down_sync(0xffffffffu, in1.x, 1);  // in(b+1, jr+1)
            const double2 jp = REV ? in0 : in2;  // real row +1
            const double2 jm = REV ? in2 : in0;  // real row -1
            double2 lap1;
            lap1.x = 4.0 * in1.x - (((in1.y + l1) + jp.x) + jm.x);
            lap1.y = 4.0 * in1.y - (((r1 + in1.x) + jp.y) + jm.y);
            double lapr1 = __shfl_down_sync(0xffffffffu, lap1.x, 1);  // lap(b+1, jr+1)
            // fluxes of row jr
            double2 fly;
            fly.x = limit<LIMITER, double>(lap1.x - lap0.x, in1.x - in0.x);
            fly.y = limit<LIMITER, double>(lap1.y - lap0.y, in1.y - in0.y);
            double fa = limit<LIMITER, double>(lap0.y - lap0.x, in0.y - in0.x);  // flx(a, jr)
            double fb = limit<LIMITER, double>(lapr0 - lap0.y, r0 - in0.y);      // flx(b, jr)
            double fl = __shfl_up_sync(0xffffffffu, fb, 1);              // flx(a-1, jr)
            if (jr >= j0 && jr < jend) {
                double2 o;
                if (REV) {  // fly == -fly_real[j-1], flym1 == -fly_real[j]
                    o.x = in0.x - co.x * (((fa - fl) - flym1.x) + fly.x);
                    o.y = in0.y - co.y * (((fb - fa) - flym1.y) + fly.y);
                } else {
                    o.x = in0.x - co.x * (((fa - fl) + fly.x) - flym1.x);
                    o.y = in0.y - co.y * (((fb - fa) + fly.y) - flym1.y);
                }
                double *row = pout + (int64_t)(REV ? mirror - jr : jr) * out.sj;
                if (store_y) __stcs(reinterpret_cast<double2 *>(row + col), o);
                else if (store_x) __stcs(row + col, o.x);
            }
            // roll
            in0 = in1; in1 = in2; lap0 = lap1; flym1 = fly; r0 = r1; lapr0 = lapr1;
        }
    }
}

template <bool LIMITER, int PF, int MINB>
__global__ void __launch_bounds__(256, MINB)
hdiff_jmarch_alt(View3<const double> in, View3<double> out, View3<const double> coeff, int ni, int nj,
                 int nk, int tj)
{
    if (blockIdx.y & 1)
        hdiff_strip<LIMITER, PF, true>(in, out, coeff, ni, nj, tj);
    else
        hdiff_strip<LIMITER, PF, false>(in, out, coeff, ni, nj, tj);
}

// ------------------------------------------------------------------------------------------------
// host dispatch
// ------------------------------------------------------------------------------------------------
template <bool LIMITER, typename T>
static int launch_hdiff(View3<const T> in, View3<T> out, View3<const T> co, const int64_t domain[3], int variant,
                        cudaStream_t stream)
{
    constexpr bool F64 = sizeof(T) == 8;
    const int ni = (int)domain[0], nj = (int)domain[1], nk = (int)domain[2];
    // two-element vector accesses: 16-byte (float64) / 8-byte (float32) aligned origins, even J/K strides
    auto aligned2 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & (2 * sizeof(T) - 1)) == 0; };
    const bool fast_ok = in.si == 1 && out.si == 1 && co.si == 1 && aligned2(in.p) && aligned2(out.p) &&
                         aligned2(co.p) && (in.sj % 2 == 0) && (in.sk % 2 == 0) && (out.sj % 2 == 0) &&
                         (out.sk % 2 == 0) && (co.sj % 2 == 0) && (co.sk % 2 == 0);
    if (variant == 0) {  // auto: TMA-ring kernel (float64), else the register J-march, else the point kernel
        if constexpr (F64) {
            // The TMA ring wins where the fields stream from HBM (512x512x128: 0.134 vs 0.150 ms, 1024x1024x80:
            // 0.312 vs 0.342); when the three fields fit the 126 MB L2 and calls run back to back (CUDA-graph
            // replay at 256x256x80: 29.1 vs 25.2 us) the register march with direct loads is faster.
            const int64_t bytes = (int64_t)ni * nj * nk * 24;
            static const int64_t tma_min = (int64_t)env_int("B200_HDIFF_TMA_MIN_MB", 192) << 20;
            if (fast_ok && bytes >= tma_min) {
                bool handled = false;
                int rc = launch_hdiff_tma(LIMITER, in, out, co, ni, nj, nk, stream, &handled);
                if (rc != B200_OK || handled) return rc;
            }
        }
        variant = fast_ok ? 2 : 1;
    }
    if (!F64 && variant >= 4) {
        set_error("hdiff variant %d is float64-only (float32 fields: variants 1-3)", variant);
        return B200_EUNSUP;
    }
    if (variant == 1) {
        dim3 block(64, 4, 1);
        dim3 grid((ni + 63) / 64, (nj + 3) / 4, nk);
        note_kernel("hdiff_point");
        hdiff_point<LIMITER, T><<<grid, block, 0, stream>>>(in, out, co, ni, nj, nk);
    } else if (variant >= 2 && variant <= 4) {
        if (!fast_ok) {
            set_error("hdiff variant %d needs unit I stride, origins aligned to two elements and even J/K strides",
                      variant);
            return B200_EUNSUP;
        }
        const int nwarps = (ni + JM_COLS_PER_WARP - 1) / JM_COLS_PER_WARP;
        int wpb = nwarps < 8 ? nwarps : 8;
        if (nwarps > 8) {  // fewest padded warps, prefer wider blocks
            int best = 1 << 30;
            for (int w = 8; w >= 4; --w) {
                int padded = (nwarps + w - 1) / w * w;
                if (padded < best) { best = padded; wpb = w; }
            }
        }
        static const int env_tj = env_int("B200_HDIFF_TJ", 0);  // tuning hook (read once)
        const int tj = env_tj > 0 ? env_tj : (nj >= 256 ? 32 : (nj >= 64 ? 16 : 8));
        dim3 block(32 * wpb, 1, 1);
        dim3 grid((nwarps + wpb - 1) / wpb, (nj + tj - 1) / tj, nk);
        if (variant == 2)
            hdiff_jmarch<LIMITER, 2, T><<<grid, block, 0, stream>>>(in, out, co, ni, nj, nk, tj), note_kernel("hdiff_jmarch");
        else if (variant == 3)
            hdiff_jmarch<LIMITER, 4, T><<<grid, block, 0, stream>>>(in, out, co, ni, nj, nk, tj), note_kernel("hdiff_jmarch");
        else if constexpr (F64)  // variant 4: alternating march directions (less DRAM traffic, not faster: DESIGN.md 3.1)
            hdiff_jmarch_alt<LIMITER, 2, 4><<<grid, block, 0, stream>>>(in, out, co, ni, nj, nk, tj), note_kernel("hdiff_jmarch_alt");
    } else if (variant == 5) {
        if constexpr (F64) {
            bool handled = false;
            int rc = launch_hdiff_tma(LIMITER, in, out, co, ni, nj, nk, stream, &handled);
            if (rc != B200_OK) return rc;
            if (!handled) {
                set_error("hdiff variant 5 needs unit I stride, 16-byte aligned origins and even J/K strides");
                return B200_EUNSUP;
            }
        }
        return B200_OK;
    } else {
        set_error("hdiff: unknown variant %d", variant);
        return B200_EUNSUP;
    }
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}

template <typename T>
static int hdiff_entry(const b200_field *in_field, const b200_field *out_field, const b200_field *coeff,
                       const int64_t domain[3], int limiter, int variant, void *stream, int dtype)
{
    const int h2[3] = {2, 2, 0}, h0[3] = {0, 0, 0};
    View3<const T> in, co;
    View3<T> out;
    int rc;
    if ((rc = make_view<const T>(in_field, "in_field", dtype, domain, h2, h2, &in))) return rc;
    if ((rc = make_view<T>(out_field, "out_field", dtype, domain, h0, h0, &out))) return rc;
    if ((rc = make_view<const T>(coeff, "coeff", dtype, domain, h0, h0, &co))) return rc;
    if (domain[0] == 0 || domain[1] == 0 || domain[2] == 0) return B200_OK;
    return limiter ? launch_hdiff<true, T>(in, out, co, domain, variant, (cudaStream_t)stream)
                   : launch_hdiff<false, T>(in, out, co, domain, variant, (cudaStream_t)stream);
}

}  // namespace b200

extern "C" int b200_hdiff(const b200_field *in_field, const b200_field *out_field,
                          const b200_field *coeff, const int64_t domain[3], int limiter,
                          int variant, void *stream)
{
    using namespace b200;
    B200_REQUIRE(domain, "domain is null");
    B200_REQUIRE(domain[0] >= 0 && domain[1] >= 0 && domain[2] >= 0, "negative domain");
    B200_REQUIRE(domain[0] < (1 << 30) && domain[1] < (1 << 30) && domain[2] < 65536,
                 "domain too large");
    // float64, or all three fields float32 (the pure-float32 build of the definition: float32 literals and
    // temporaries); mixed dtypes are a different computation and run through the generated kernels
    if (in_field && in_field->dtype == B200_FLOAT32)
        return hdiff_entry<float>(in_field, out_field, coeff, domain, limiter, variant, stream, B200_FLOAT32);
    return hdiff_entry<double>(in_field, out_field, coeff, domain, limiter, variant, stream, B200_FLOAT64);
}
