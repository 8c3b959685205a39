// vadv_tmem.cuh -- pieces shared by the Tensor-Memory vadv kernels (vadv_tmem.cu, vadv_fused.cu):
// tile geometry and the tcgen05.st / tcgen05.ld wrappers that turn TMEM into a per-thread stack.
#pragma once
#include "tma.cuh"
#include "vadv.cuh"

namespace b200 {

constexpr int VM_COLS = 256;                  // columns per tile == consumer threads
constexpr int VM_CWARPS = VM_COLS / 32;       // 8 consumer warps
constexpr int VM_THREADS = VM_COLS + 64;      // + forward-ring producer warp + backward-ring producer warp
constexpr int VM_MAX_STAGES = 16;
constexpr int VM_BAR_BYTES = 1024;
constexpr int VM_TMEM_LEVELS = 64;            // levels of LIFO per thread in TMEM (256 columns x 32 bit)
constexpr int VM_TMEM_COLS = 512;

template <int TI, int KC>
struct VmGeom {
    static constexpr int TJ = VM_COLS / TI;
    static constexpr int WROW = TI + 2;
    static constexpr int WLVL = WROW * TJ;
    static constexpr int F_BYTES = KC * VM_COLS * 8;
    static constexpr int W_TX = KC * WLVL * 8;
    static constexpr int W_BYTES = (W_TX + 127) / 128 * 128;
    static constexpr int SLOT_F = 4 * F_BYTES + W_BYTES;
    static constexpr int SLOT_B = F_BYTES;
    static constexpr uint32_t TX_F = 4 * F_BYTES + W_TX;
    static constexpr uint32_t TX_B = F_BYTES;
};

// ---- Tensor Memory as a per-thread stack -------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t ncols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_push(uint32_t taddr, double a, double b)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr),
                 "r"(__double2loint(a)), "r"(__double2hiint(a)), "r"(__double2loint(b)), "r"(__double2hiint(b))
                 : "memory");
}
struct Raw4 {
    uint32_t r[4];
};
__device__ __forceinline__ void tmem_pop_issue(uint32_t taddr, Raw4 &v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3])
                 : "r"(taddr)
                 : "memory");
}
// The loaded registers are only valid after wait::ld; threading them through the asm keeps the compiler
// from scheduling any use before it.
template <int N>
__device__ __forceinline__ void tmem_pop_wait(Raw4 (&v)[N])
{
    static_assert(N == 1 || N == 2 || N == 4, "unsupported");
    if (N == 1)
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(v[0].r[0]), "+r"(v[0].r[1]), "+r"(v[0].r[2]), "+r"(v[0].r[3])::"memory");
    if (N == 2)
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(v[0].r[0]), "+r"(v[0].r[1]), "+r"(v[0].r[2]), "+r"(v[0].r[3]), "+r"(v[N > 1 ? 1 : 0].r[0]),
                       "+r"(v[N > 1 ? 1 : 0].r[1]), "+r"(v[N > 1 ? 1 : 0].r[2]), "+r"(v[N > 1 ? 1 : 0].r[3])::"memory");
    if (N == 4)
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(v[0].r[0]), "+r"(v[0].r[1]), "+r"(v[0].r[2]), "+r"(v[0].r[3]), "+r"(v[N > 1 ? 1 : 0].r[0]),
                       "+r"(v[N > 1 ? 1 : 0].r[1]), "+r"(v[N > 1 ? 1 : 0].r[2]), "+r"(v[N > 1 ? 1 : 0].r[3]),
                       "+r"(v[N > 2 ? 2 : 0].r[0]), "+r"(v[N > 2 ? 2 : 0].r[1]), "+r"(v[N > 2 ? 2 : 0].r[2]),
                       "+r"(v[N > 2 ? 2 : 0].r[3]), "+r"(v[N > 3 ? 3 : 0].r[0]), "+r"(v[N > 3 ? 3 : 0].r[1]),
                       "+r"(v[N > 3 ? 3 : 0].r[2]), "+r"(v[N > 3 ? 3 : 0].r[3])::"memory");
}
__device__ __forceinline__ void tmem_push_wait()
{
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}


// x8 forms: two LIFO entries (ccol, dcol of two consecutive slots) per instruction
__device__ __forceinline__ void tmem_push2(uint32_t taddr, double a0, double b0, double a1, double b1)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
                 "r"(__double2loint(a0)), "r"(__double2hiint(a0)), "r"(__double2loint(b0)), "r"(__double2hiint(b0)),
                 "r"(__double2loint(a1)), "r"(__double2hiint(a1)), "r"(__double2loint(b1)), "r"(__double2hiint(b1))
                 : "memory");
}
struct Raw8 {
    uint32_t r[8];
};
__device__ __forceinline__ void tmem_pop2_issue(uint32_t taddr, Raw8 &v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]),
                   "=r"(v.r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_pop2_wait(Raw8 &v)
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v.r[0]), "+r"(v.r[1]), "+r"(v.r[2]), "+r"(v.r[3]), "+r"(v.r[4]), "+r"(v.r[5]), "+r"(v.r[6]),
                   "+r"(v.r[7])::"memory");
}

// Generic forms over L levels per LIFO slot (L = 2: .x8, L = 4: .x16); v = (ccol, dcol) per level.
template <int L>
struct RawN {
    uint32_t r[4 * L];
};
template <int L>
__device__ __forceinline__ void tmem_push_n(uint32_t taddr, const double (&v)[2 * L])
{
    static_assert(L == 2 || L == 4, "levels per stage must be 2 or 4");
    if constexpr (L == 2) {
        tmem_push2(taddr, v[0], v[1], v[2], v[3]);
    } else {
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, "
            "%14, %15, %16};" ::"r"(taddr),
            "r"(__double2loint(v[0])), "r"(__double2hiint(v[0])), "r"(__double2loint(v[1])), "r"(__double2hiint(v[1])),
            "r"(__double2loint(v[2])), "r"(__double2hiint(v[2])), "r"(__double2loint(v[3])), "r"(__double2hiint(v[3])),
            "r"(__double2loint(v[4])), "r"(__double2hiint(v[4])), "r"(__double2loint(v[5])), "r"(__double2hiint(v[5])),
            "r"(__double2loint(v[6])), "r"(__double2hiint(v[6])), "r"(__double2loint(v[7])), "r"(__double2hiint(v[7]))
            : "memory");
    }
}
template <int L>
__device__ __forceinline__ void tmem_pop_n_issue(uint32_t taddr, RawN<L> &v)
{
    if constexpr (L == 2) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]),
                       "=r"(v.r[6]), "=r"(v.r[7])
                     : "r"(taddr)
                     : "memory");
    } else {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
            "%15}, [%16];"
            : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]),
              "=r"(v.r[7]), "=r"(v.r[8]), "=r"(v.r[9]), "=r"(v.r[10]), "=r"(v.r[11]), "=r"(v.r[12]), "=r"(v.r[13]),
              "=r"(v.r[14]), "=r"(v.r[15])
            : "r"(taddr)
            : "memory");
    }
}
template <int L>
__device__ __forceinline__ void tmem_pop_n_wait(RawN<L> &v)
{
    if constexpr (L == 2) {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(v.r[0]), "+r"(v.r[1]), "+r"(v.r[2]), "+r"(v.r[3]), "+r"(v.r[4]), "+r"(v.r[5]),
                       "+r"(v.r[6]), "+r"(v.r[7])::"memory");
    } else {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(v.r[0]), "+r"(v.r[1]), "+r"(v.r[2]), "+r"(v.r[3]), "+r"(v.r[4]), "+r"(v.r[5]),
                       "+r"(v.r[6]), "+r"(v.r[7]), "+r"(v.r[8]), "+r"(v.r[9]), "+r"(v.r[10]), "+r"(v.r[11]),
                       "+r"(v.r[12]), "+r"(v.r[13]), "+r"(v.r[14]), "+r"(v.r[15])::"memory");
    }
}

}  // namespace b200
