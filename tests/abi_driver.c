/* abi_driver.c -- TEST: a host WITHOUT Python runs a generated stencil (the config-4 chain, fv3_chain) through the
 * plan-level C ABI of include/gt4py_b200.h: b200_plan_create / b200_plan_launch / b200_plan_destroy.
 *
 *   abi_driver <libgt4py_b200.so | host double> <plan.desc> <cubin | image file> <ni> <nj> <nk> <out.bin>
 *
 * Inputs are the synthetic fields of b200_fill_synthetic (seeds 91 / 92 / 93, range [-1, 1.5), halo 3), the result
 * (ni x nj x nk doubles, I fastest) is written to <out.bin>; tests/test_abi.py compares it with oracle/chain.py.
 * The library is opened with dlopen so that the same binary drives the CUDA library and its host double. */
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/gt4py_b200.h"

#define LOAD(name) *(void **)(&name##_p) = dlsym(lib, #name); if (!name##_p) { fprintf(stderr, "missing symbol %s\n", #name); return 2; }
#define CHECK(call) do { int rc_ = (call); if (rc_) { fprintf(stderr, "%s -> %d: %s\n", #call, rc_, b200_last_error_p()); return 3; } } while (0)

static char *slurp(const char *path, size_t *n)
{
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    *n = (size_t)ftell(f);
    fseek(f, 0, SEEK_SET);
    char *buf = (char *)malloc(*n + 1);
    if (fread(buf, 1, *n, f) != *n) { fclose(f); return NULL; }
    buf[*n] = 0;
    fclose(f);
    return buf;
}

int main(int argc, char **argv)
{
    if (argc != 8) { fprintf(stderr, "usage: %s lib plan.desc image ni nj nk out.bin\n", argv[0]); return 1; }
    void *lib = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!lib) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    int (*b200_plan_create_p)(const char *, size_t, const void *, size_t, int, void **);
    int (*b200_plan_launch_p)(void *, const int64_t *, const b200_field *, int, const b200_scalar *, int, void *, int);
    int (*b200_plan_destroy_p)(void *);
    int (*b200_plan_last_entries_p)(void *, char *, size_t);
    int (*b200_malloc_p)(void **, size_t);
    int (*b200_free_p)(void *);
    int (*b200_memset_p)(void *, int, size_t, void *);
    int (*b200_memcpy_d2h_p)(void *, const void *, size_t, void *);
    int (*b200_fill_synthetic_p)(const b200_field *, const b200_field *, uint64_t, int64_t, int64_t, double, double, void *);
    const char *(*b200_last_error_p)(void);
    LOAD(b200_plan_create) LOAD(b200_plan_launch) LOAD(b200_plan_destroy) LOAD(b200_plan_last_entries) LOAD(b200_malloc)
    LOAD(b200_free) LOAD(b200_memset) LOAD(b200_memcpy_d2h) LOAD(b200_fill_synthetic) LOAD(b200_last_error)

    size_t ndesc = 0, nimage = 0;
    char *desc = slurp(argv[2], &ndesc), *image = slurp(argv[3], &nimage);
    if (!desc || !image) { fprintf(stderr, "cannot read descriptor / image\n"); return 2; }
    const int64_t ni = atoll(argv[4]), nj = atoll(argv[5]), nk = atoll(argv[6]), h = 3;
    void *plan = NULL;
    CHECK(b200_plan_create_p(desc, ndesc, image, nimage, 0, &plan));

    /* fields in the descriptor's order: phi, u, v (halo 3 in I and J), out (no halo); I contiguous, rows padded to 16 */
    const int64_t wi = (ni + 2 * h + 15) / 16 * 16, wj = nj + 2 * h, oi = (ni + 15) / 16 * 16;
    b200_field f[4];
    memset(f, 0, sizeof(f));
    const uint64_t seeds[3] = {91, 92, 93};
    for (int n = 0; n < 4; ++n) {
        const int in = n < 3;
        const int64_t si = in ? wi : oi, sj = in ? wj : nj;
        CHECK(b200_malloc_p(&f[n].ptr, (size_t)(si * sj * nk) * 8));
        f[n].ndim = 3;
        f[n].dtype = B200_FLOAT64;
        f[n].shape[0] = in ? ni + 2 * h : ni; f[n].shape[1] = sj; f[n].shape[2] = nk;
        f[n].stride[0] = 1; f[n].stride[1] = si; f[n].stride[2] = si * sj;
        f[n].origin[0] = in ? (int)h : 0; f[n].origin[1] = in ? (int)h : 0; f[n].origin[2] = 0;
        if (in) { b200_field whole = f[n]; whole.origin[0] = whole.origin[1] = 0;
                  CHECK(b200_fill_synthetic_p(&whole, NULL, seeds[n], -h, -h, -1.0, 1.5, NULL)); }
        else CHECK(b200_memset_p(f[n].ptr, 0, (size_t)(si * sj * nk) * 8, NULL));
    }
    b200_scalar s[2];
    memset(s, 0, sizeof(s));
    s[0].value.f64 = 0.05; s[0].dtype = B200_FLOAT64;   /* dt */
    s[1].value.f64 = 0.03; s[1].dtype = B200_FLOAT64;   /* alpha */
    const int64_t dom[3] = {ni, nj, nk};
    CHECK(b200_plan_launch_p(plan, dom, f, 4, s, 2, NULL, 1));
    CHECK(b200_plan_launch_p(plan, dom, f, 4, s, 2, NULL, 1));  /* a second launch reuses scratch and descriptors */
    char entries[256];
    CHECK(b200_plan_last_entries_p(plan, entries, sizeof(entries)));
    printf("entries %s\n", entries);

    double *host = (double *)malloc((size_t)(oi * nj * nk) * 8), *dense = (double *)malloc((size_t)(ni * nj * nk) * 8);
    CHECK(b200_memcpy_d2h_p(host, f[3].ptr, (size_t)(oi * nj * nk) * 8, NULL));
    for (int64_t k = 0; k < nk; ++k)
        for (int64_t j = 0; j < nj; ++j) memcpy(dense + (k * nj + j) * ni, host + (k * nj + j) * oi, (size_t)ni * 8);
    FILE *out = fopen(argv[7], "wb");
    if (!out || fwrite(dense, 8, (size_t)(ni * nj * nk), out) != (size_t)(ni * nj * nk)) { fprintf(stderr, "cannot write result\n"); return 2; }
    fclose(out);
    /* a wrong call must fail loudly, not crash: one field too few */
    if (b200_plan_launch_p(plan, dom, f, 3, s, 2, NULL, 1) == 0) { fprintf(stderr, "short field list accepted\n"); return 4; }
    CHECK(b200_plan_destroy_p(plan));
    for (int n = 0; n < 4; ++n) b200_free_p(f[n].ptr);
    return 0;
}
