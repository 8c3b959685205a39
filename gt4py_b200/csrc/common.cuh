// common.cuh -- shared helpers of libgt4py_b200 (error plumbing, field views).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../../include/gt4py_b200.h"

namespace b200 {

void set_error(const char *fmt, ...);
// name of the kernel a library entry point (b200_hdiff / b200_vadv / b200_tridiag) is about to launch on this
// thread: b200_last_kernel() reports it, so that measurements name what really ran
void note_kernel(const char *name);

#define B200_CUDA_CHECK(expr)                                                              \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),        \
                            __FILE__, __LINE__);                                           \
            return B200_ECUDA;                                                             \
        }                                                                                  \
    } while (0)

#define B200_REQUIRE(cond, ...)                \
    do {                                       \
        if (!(cond)) {                         \
            b200::set_error(__VA_ARGS__);      \
            return B200_EINVAL;                \
        }                                      \
    } while (0)

// A 3-D view whose pointer is already shifted to the origin of the compute domain.
template <typename T>
struct View3 {
    T *p;
    int64_t si, sj, sk;
};

// Validate a 3-D field against a compute domain with the given halo and build the origin view.
// lo/hi are the halo widths the stencil reads/writes around the domain: {i, j, k}.
template <typename T>
int make_view(const b200_field *f, const char *name, int dtype, const int64_t domain[3],
              const int lo[3], const int hi[3], View3<T> *v)
{
    B200_REQUIRE(f != nullptr && f->ptr != nullptr, "field '%s' is null", name);
    B200_REQUIRE(f->ndim == 3, "field '%s': expected 3 dimensions, got %d", name, f->ndim);
    B200_REQUIRE(f->dtype == dtype, "field '%s': dtype code %d, expected %d", name, f->dtype, dtype);
    for (int d = 0; d < 3; ++d) {
        B200_REQUIRE(f->origin[d] - lo[d] >= 0, "field '%s': origin[%d]=%d smaller than halo %d",
                     name, d, f->origin[d], lo[d]);
        B200_REQUIRE(f->origin[d] + domain[d] + hi[d] <= f->shape[d],
                     "field '%s': shape[%d]=%lld too small for origin %d + domain %lld + halo %d",
                     name, d, (long long)f->shape[d], f->origin[d], (long long)domain[d], hi[d]);
    }
    v->si = f->stride[0];
    v->sj = f->stride[1];
    v->sk = f->stride[2];
    v->p = static_cast<T *>(f->ptr) + f->origin[0] * v->si + f->origin[1] * v->sj +
           f->origin[2] * v->sk;
    return B200_OK;
}

inline int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace b200
