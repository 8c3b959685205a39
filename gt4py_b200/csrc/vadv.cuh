// vadv.cuh -- arithmetic of vertical_advection_dycore shared by all vadv kernels (vadv.cu, vadv_tma.cu).
// Reference: tests/cartesian_tests/integration_tests/multi_feature_tests/stencil_definitions.py:235-313.
#pragma once
#include "common.cuh"

namespace b200 {

struct VadvArgs {
    View3<double> utens_stage;
    View3<const double> u_stage, wcon, u_pos, utens;
    double dtr_stage;
    int ni, nj, nk;
};

constexpr double BET_M = 0.5, BET_P = 0.5;

// One level's worth of inputs (what a forward step loads).
struct Lvl {
    double w0, w1;  // wcon(i, k+1), wcon(i+1, k+1)
    double us;      // u_stage(k+1)
    double up, ut, uts;  // u_pos(k), utens(k), utens_stage(k)
};

// Rolling state carried through the forward sweep.
struct Fwd {
    double us_m1, us_0;  // u_stage(k-1), u_stage(k)
    double wsum;         // wcon(i+1,k) + wcon(i,k)
    double ccol_m1, dcol_m1;
};

// Forward step of level k (section chosen by k).  Returns (ccol, dcol) of the level.
__device__ __forceinline__ void forward_level(const Lvl &l, Fwd &s, double dtr, int k, int nk,
                                              double &ccol, double &dcol)
{
    const double wsum_p1 = l.w1 + l.w0;  // wcon[1,0,1] + wcon[0,0,1]
    const double us_p1 = l.us;
    if (k == 0) {  // interval(0, 1)  :248-264
        double gcv = 0.25 * wsum_p1;
        double cs = gcv * BET_M;
        double cc = gcv * BET_P;
        double bcol = dtr - cc;
        double correction = -cs * (us_p1 - s.us_0);
        double dc = ((dtr * l.up + l.ut) + l.uts) + correction;
        double divided = 1.0 / bcol;
        ccol = cc * divided;
        dcol = dc * divided;
    } else if (k < nk - 1) {  // interval(1, -1)  :266-287
        double gav = -0.25 * s.wsum;
        double gcv = 0.25 * wsum_p1;
        double as_ = gav * BET_M;
        double cs = gcv * BET_M;
        double acol = gav * BET_P;
        double cc = gcv * BET_P;
        double bcol = (dtr - acol) - cc;
        double correction = -as_ * (s.us_m1 - s.us_0) - cs * (us_p1 - s.us_0);
        double dc = ((dtr * l.up + l.ut) + l.uts) + correction;
        double divided = 1.0 / (bcol - s.ccol_m1 * acol);
        ccol = cc * divided;
        dcol = (dc - s.dcol_m1 * acol) * divided;
    } else {  // interval(-1, None)  :289-303
        double gav = -0.25 * s.wsum;
        double as_ = gav * BET_M;
        double acol = gav * BET_P;
        double bcol = dtr - acol;
        double correction = -as_ * (s.us_m1 - s.us_0);
        double dc = ((dtr * l.up + l.ut) + l.uts) + correction;
        double divided = 1.0 / (bcol - s.ccol_m1 * acol);
        ccol = s.ccol_m1;  // ccol is not assigned on the last level (unused by the backward sweep)
        dcol = (dc - s.dcol_m1 * acol) * divided;
    }
    s.us_m1 = s.us_0;
    s.us_0 = us_p1;
    s.wsum = wsum_p1;
    s.ccol_m1 = ccol;
    s.dcol_m1 = dcol;
}


// ---- branch-free formulation used by the pipelined kernel (vadv_tma.cu) ----------------------------
// The three FORWARD sections (:248-264, :266-287, :289-303) are the middle-section formula with the
// terms of the missing neighbour level equal to zero, and multiplying / subtracting an exact zero is
// exact in IEEE arithmetic.  So with the rolling state initialised to wsum(k=0) := 0, us(k=-1) := 0,
// ccol(-1) = dcol(-1) := 0 and with wcon / u_stage of level nk read as 0 (the TMA unit zero-fills the
// out-of-domain level) one formula serves every level bit-for-bit (only the sign of an exact zero can
// differ, which no comparison or later result observes).  The level is split into the part that does
// not depend on the Thomas recurrence (`pre_level`, free to run ahead) and the recurrence itself
// (`rec_level`: multiply, subtract, reciprocal, multiply).
struct Pre {
    double acol, cc, bcol, dc;
};

struct Roll {
    double us_m1, us_0;  // u_stage(k-1), u_stage(k)
    double wsum;         // wcon(i+1,k) + wcon(i,k)
};

__device__ __forceinline__ Pre pre_level(const Lvl &l, Roll &r, double dtr)
{
    const double wsum_p1 = l.w1 + l.w0;
    const double gav = -0.25 * r.wsum;
    const double gcv = 0.25 * wsum_p1;
    const double as_ = gav * BET_M;
    const double cs = gcv * BET_M;
    Pre p;
    p.acol = gav * BET_P;
    p.cc = gcv * BET_P;
    p.bcol = (dtr - p.acol) - p.cc;
    const double correction = -as_ * (r.us_m1 - r.us_0) - cs * (l.us - r.us_0);
    p.dc = ((dtr * l.up + l.ut) + l.uts) + correction;
    r.us_m1 = r.us_0;
    r.us_0 = l.us;
    r.wsum = wsum_p1;
    return p;
}

// Correctly rounded reciprocal for 2^-1000 <= |x| <= 2^1000: the Newton sequence nvcc emits inline for
// 1.0 / x (MUFU.RCP64H seed, two fused refinement steps), without its per-call range branch.
__device__ __forceinline__ double rcp_rn_inrange(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    e = fma(e, e, e);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

__device__ __forceinline__ bool rcp_inrange(double x)
{
    const int e = (__double2hiint(x) >> 20) & 0x7ff;
    return e >= 24 && e <= 2022;  // biased exponent: 2^-999 .. 2^999
}

__device__ __forceinline__ void rec_level(const Pre &p, double &ccol_m1, double &dcol_m1)
{
    const double den = p.bcol - ccol_m1 * p.acol;
    const double num = p.dc - dcol_m1 * p.acol;
    const double divided = 1.0 / den;
    ccol_m1 = p.cc * divided;
    dcol_m1 = num * divided;
}

// ---- reduced-operation form of pre_level / rec_level (vadv_fused kernel) ---------------------------------
// With BET_M = BET_P = 0.5 (the stencil's fixed externals) and power-of-two scalings being exact:
//     cc(k)   = (0.25 * wsum(k+1)) * 0.5           == 0.125 * wsum(k+1)
//     acol(k) = (-0.25 * wsum(k)) * 0.5            == -cc(k-1)            (as_ == acol, cs == cc)
//     -as_ * (us(k-1) - us(k))                     == -(cc(k-1) * (us(k) - us(k-1)))  == -q(k-1)
//      cs  * (us(k+1) - us(k))                     ==   cc(k) * d(k)                  ==  q(k)
// so every level needs one new product q(k) = cc(k) * d(k), d(k) = us(k+1) - us(k), and reuses q(k-1).
// All identities are exact in IEEE-754 binary64 (sign-symmetric rounding; scaling by 2^-n is exact unless
// the result is subnormal, i.e. |wcon| < 2^-1019), so results stay bit-identical to the reference.
struct Pre2 {
    double ccp;   // cc(k-1) == -acol(k)
    double cc;    // cc(k)
    double bcol, dc;
};

struct Roll2 {
    double us_0;   // u_stage(k)
    double ccp;    // cc(k-1)
    double qp;     // q(k-1) = cc(k-1) * (us(k) - us(k-1))
};

__device__ __forceinline__ Pre2 pre_level2(const Lvl &l, Roll2 &r, double dtr)
{
    Pre2 p;
    p.ccp = r.ccp;
    p.cc = 0.125 * (l.w1 + l.w0);
    p.bcol = (dtr + r.ccp) - p.cc;                       // (dtr - acol) - ccol
    const double q = p.cc * (l.us - r.us_0);
    const double correction = -r.qp - q;                 // -as*(us[-1]-us) - cs*(us[+1]-us)
    p.dc = ((dtr * l.up + l.ut) + l.uts) + correction;
    r.us_0 = l.us;
    r.ccp = p.cc;
    r.qp = q;
    return p;
}

// One step of the Thomas recurrence with the in-range reciprocal; `ok` is cleared if the divisor left
// the range in which rcp_rn_inrange is correctly rounded (the caller then redoes the stage with `/`).
__device__ __forceinline__ void rec_level2(const Pre2 &p, double &ccol, double &dcol, bool &ok)
{
    const double den = p.bcol + ccol * p.ccp;            // bcol - ccol[k-1] * acol
    const double num = p.dc + dcol * p.ccp;              // dcol - dcol[k-1] * acol
    ok = ok && rcp_inrange(den);
    const double divided = rcp_rn_inrange(den);
    ccol = p.cc * divided;
    dcol = num * divided;
}

__device__ __forceinline__ void rec_level2_exact(const Pre2 &p, double &ccol, double &dcol)
{
    const double den = p.bcol + ccol * p.ccp;
    const double num = p.dc + dcol * p.ccp;
    const double divided = 1.0 / den;
    ccol = p.cc * divided;
    dcol = num * divided;
}

// host launchers (vadv_tma.cu)
int launch_vadv_tma(const VadvArgs &a, int variant, cudaStream_t stream, bool *handled);
// (vadv_fused.cu)
int launch_vadv_fused(const VadvArgs &a, int levels_per_stage, cudaStream_t stream, bool *handled);
// (vadv_tmem.cu)
int launch_vadv_tmem(const VadvArgs &a, int variant, cudaStream_t stream, bool *handled);

}  // namespace b200
