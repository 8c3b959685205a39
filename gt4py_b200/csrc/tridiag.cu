// tridiag.cu -- generic Thomas solver, one thread per (i, j) column, warps along I.
//
// Reference semantics: tridiagonal_solver, tests/cartesian_tests/integration_tests/
// multi_feature_tests/stencil_definitions.py:219-232 of the reference.  sup and rhs are API fields
// that the stencil overwrites (KCache fill+flush in the reference's OIR), so the forward sweep
// stores them; the backward sweep re-reads them (L2-resident for the in-flight columns).
#include "common.cuh"

namespace b200 {

struct TriArgs {
    View3<const double> inf, diag;
    View3<double> sup, rhs, out;
    int ni, nj, nk;
};

template <int PF>
__global__ void __launch_bounds__(128) tridiag_column(TriArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= a.ni || j >= a.nj) return;
    const double *__restrict__ inf = a.inf.p + i * a.inf.si + j * a.inf.sj;
    const double *__restrict__ diag = a.diag.p + i * a.diag.si + j * a.diag.sj;
    double *__restrict__ sup = a.sup.p + i * a.sup.si + j * a.sup.sj;
    double *__restrict__ rhs = a.rhs.p + i * a.rhs.si + j * a.rhs.sj;
    double *__restrict__ out = a.out.p + i * a.out.si + j * a.out.sj;
    const int nk = a.nk;

    struct L { double inf, diag, sup, rhs; };
    auto load = [&](int k) -> L {
        L l = {0.0, 1.0, 0.0, 0.0};
        if (k < nk) {
            l.inf = __ldcs(inf + k * a.inf.sk);
            l.diag = __ldcs(diag + k * a.diag.sk);
            l.sup = sup[k * a.sup.sk];
            l.rhs = rhs[k * a.rhs.sk];
        }
        return l;
    };
    L q[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) q[u] = load(u);
    double sup_m1 = 0.0, rhs_m1 = 0.0;
    for (int kb = 0; kb < nk; kb += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int k = kb + u;
            if (k < nk) {
                L l = q[u];
                q[u] = load(k + PF);
                double s, r;
                if (k == 0) {  // interval(0, 1)  :222-224
                    s = l.sup / l.diag;
                    r = l.rhs / l.diag;
                } else {  // interval(1, None)  :225-227
                    s = l.sup / (l.diag - sup_m1 * l.inf);
                    r = (l.rhs - l.inf * rhs_m1) / (l.diag - sup_m1 * l.inf);
                }
                sup[k * a.sup.sk] = s;
                rhs[k * a.rhs.sk] = r;
                sup_m1 = s;
                rhs_m1 = r;
            }
        }
    }
    // BACKWARD  :228-232  (level nk-1 values are still in registers)
    double o = rhs_m1;
    __stcs(out + (int64_t)(nk - 1) * a.out.sk, o);
    for (int k = nk - 2; k >= 0; --k) {
        o = rhs[k * a.rhs.sk] - sup[k * a.sup.sk] * o;
        __stcs(out + k * a.out.sk, o);
    }
}

// Variants 2..: the same column walk with (a) a register budget that lets more columns be resident per SM
// (the kernel is bound by load latency: ncu of variant 1 shows 19 of 64 warps active at 84 registers, 51 % of DRAM
// throughput, long-scoreboard stalls), (b) a prefetch queue in the BACKWARD sweep too, whose loads of the parked
// (sup, rhs) otherwise each wait out an L2 / DRAM round trip.
template <int PF, int BPF, int MINB>
__global__ void __launch_bounds__(128, MINB) tridiag_column_q(TriArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= a.ni || j >= a.nj) return;
    const double *__restrict__ inf = a.inf.p + i * a.inf.si + j * a.inf.sj;
    const double *__restrict__ diag = a.diag.p + i * a.diag.si + j * a.diag.sj;
    double *__restrict__ sup = a.sup.p + i * a.sup.si + j * a.sup.sj;
    double *__restrict__ rhs = a.rhs.p + i * a.rhs.si + j * a.rhs.sj;
    double *__restrict__ out = a.out.p + i * a.out.si + j * a.out.sj;
    const int nk = a.nk;

    struct L { double inf, diag, sup, rhs; };
    auto load = [&](int k) -> L {
        L l = {0.0, 1.0, 0.0, 0.0};
        if (k < nk) {
            l.inf = __ldcs(inf + k * a.inf.sk);
            l.diag = __ldcs(diag + k * a.diag.sk);
            l.sup = __ldcs(sup + k * a.sup.sk);
            l.rhs = __ldcs(rhs + k * a.rhs.sk);
        }
        return l;
    };
    L q[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) q[u] = load(u);
    double sup_m1 = 0.0, rhs_m1 = 0.0;
    for (int kb = 0; kb < nk; kb += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int k = kb + u;
            if (k < nk) {
                L l = q[u];
                q[u] = load(k + PF);
                double s, r;
                if (k == 0) {
                    s = l.sup / l.diag;
                    r = l.rhs / l.diag;
                } else {
                    s = l.sup / (l.diag - sup_m1 * l.inf);
                    r = (l.rhs - l.inf * rhs_m1) / (l.diag - sup_m1 * l.inf);
                }
                sup[k * a.sup.sk] = s;
                rhs[k * a.rhs.sk] = r;
                sup_m1 = s;
                rhs_m1 = r;
            }
        }
    }
    double o = rhs_m1;
    __stcs(out + (int64_t)(nk - 1) * a.out.sk, o);
    double qs[BPF], qr[BPF];
    int top = nk - 2;
#pragma unroll
    for (int u = 0; u < BPF; ++u) {
        const int k = top - u;
        qs[u] = k >= 0 ? sup[k * a.sup.sk] : 0.0;
        qr[u] = k >= 0 ? rhs[k * a.rhs.sk] : 0.0;
    }
    for (; top >= 0; top -= BPF) {
#pragma unroll
        for (int u = 0; u < BPF; ++u) {
            const int k = top - u;
            if (k >= 0) {
                const double s = qs[u], r = qr[u];
                const int kn = k - BPF;
                qs[u] = kn >= 0 ? sup[kn * a.sup.sk] : 0.0;
                qr[u] = kn >= 0 ? rhs[kn * a.rhs.sk] : 0.0;
                o = r - s * o;
                __stcs(out + k * a.out.sk, o);
            }
        }
    }
}

// tridiag_fused.cu: TMA ring + Tensor-Memory LIFO, fused sweeps
int launch_tridiag_fused(const View3<const double> &inf, const View3<const double> &diag, const View3<double> &sup,
                         const View3<double> &rhs, const View3<double> &out, int ni, int nj, int nk,
                         cudaStream_t stream, bool *handled);

}  // namespace b200

extern "C" int b200_tridiag(const b200_field *inf, const b200_field *diag, const b200_field *sup,
                            const b200_field *rhs, const b200_field *out, const int64_t domain[3],
                            int variant, void *stream)
{
    using namespace b200;
    B200_REQUIRE(domain, "domain is null");
    B200_REQUIRE(domain[0] >= 0 && domain[1] >= 0 && domain[2] >= 0, "negative domain");
    B200_REQUIRE(domain[0] < (1 << 30) && domain[1] < 65536, "domain too large");
    if (domain[0] == 0 || domain[1] == 0 || domain[2] == 0) return B200_OK;
    const int h0[3] = {0, 0, 0};
    TriArgs a;
    int rc;
    if ((rc = make_view<const double>(inf, "inf", B200_FLOAT64, domain, h0, h0, &a.inf))) return rc;
    if ((rc = make_view<const double>(diag, "diag", B200_FLOAT64, domain, h0, h0, &a.diag))) return rc;
    if ((rc = make_view<double>(sup, "sup", B200_FLOAT64, domain, h0, h0, &a.sup))) return rc;
    if ((rc = make_view<double>(rhs, "rhs", B200_FLOAT64, domain, h0, h0, &a.rhs))) return rc;
    if ((rc = make_view<double>(out, "out", B200_FLOAT64, domain, h0, h0, &a.out))) return rc;
    a.ni = (int)domain[0];
    a.nj = (int)domain[1];
    a.nk = (int)domain[2];
    if (variant < 0 || variant > 8) {
        set_error("tridiag: unknown variant %d", variant);
        return B200_EUNSUP;
    }
    dim3 grid((a.ni + 127) / 128, a.nj, 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (variant == 0 || variant == 8) {  // fused TMEM kernel when the layout allows it (tridiag_fused.cu)
        bool handled = false;
        if ((rc = launch_tridiag_fused(a.inf, a.diag, a.sup, a.rhs, a.out, a.ni, a.nj, a.nk, st, &handled))) return rc;
        if (handled) return B200_OK;
        if (variant == 8) {
            set_error("tridiag variant 8 (fused TMEM kernel) needs TMA-compatible fields and nk >= 2");
            return B200_EUNSUP;
        }
    }
    switch (variant) {
    case 1: note_kernel("tridiag_column"); tridiag_column<4><<<grid, 128, 0, st>>>(a); break;
    case 2: note_kernel("tridiag_column_q"); tridiag_column_q<4, 4, 8><<<grid, 128, 0, st>>>(a); break;
    case 3: note_kernel("tridiag_column_q"); tridiag_column_q<6, 4, 6><<<grid, 128, 0, st>>>(a); break;
    case 4: note_kernel("tridiag_column_q"); tridiag_column_q<8, 8, 4><<<grid, 128, 0, st>>>(a); break;
    case 5: note_kernel("tridiag_column_q"); tridiag_column_q<4, 8, 8><<<grid, 128, 0, st>>>(a); break;
    case 6: note_kernel("tridiag_column_q"); tridiag_column_q<3, 4, 10><<<grid, 128, 0, st>>>(a); break;
    case 7: note_kernel("tridiag_column_q"); tridiag_column_q<2, 4, 12><<<grid, 128, 0, st>>>(a); break;
    default: note_kernel("tridiag_column_q"); tridiag_column_q<3, 4, 10><<<grid, 128, 0, st>>>(a); break;  // variant 0 = 6, the fastest measured
    }
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}
