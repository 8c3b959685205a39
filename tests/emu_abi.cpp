// emu_abi.cpp -- TEST INFRASTRUCTURE: a host-memory double of libgt4py_b200.so's C ABI (include/gt4py_b200.h).
//
// Lets the Python host side of the backend -- storage, the StencilObject call path, the gt4py.storage hooks, the
// generated-module seam, plan loading and kernel launching -- run in a container WITHOUT a GPU: "device" memory is
// host memory, streams and events are inert handles, and a "module" is a host shared library that
// tests/host_emulation.py compiled from the very kernel.cu the product would hand to nvcc (CUDA qualifiers defined
// away, one OS thread per CUDA thread for the tile kernels).  The hand-written library kernels (b200_hdiff / vadv /
// tridiag), TMA descriptors, CUDA graphs and the halo exchange have no double: they return B200_EUNSUP, and the
// product's own fallbacks (plain entry points instead of the TMA ones) take over where they exist.
//
// NEVER loaded by the product: gt4py_b200._lib only opens gt4py_b200/lib/libgt4py_b200.so unless a test sets
// GT4PY_B200_LIB explicitly (tests/emu_abi.py).  Results computed through this double prove nothing about the CUDA
// path; the -m gpu tests do that.
#include <dlfcn.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/gt4py_b200.h"
#include "../gt4py_b200/csrc/synthetic.cuh"

namespace {
thread_local char g_err[512] = "";
void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
struct Kernel {
    void *fn;
    bool tile;
    std::string name;
};
struct Module {
    void *dl;
    std::vector<Kernel *> kernels;
};
int g_device = 0;
int g_window_mode = 0;
}  // namespace

#define UNSUP(what)                                                        \
    do {                                                                   \
        set_error("%s: not available in the host double of the C ABI", what); \
        return B200_EUNSUP;                                                \
    } while (0)

extern "C" {
int b200_abi_version(void) { return B200_ABI_VERSION; }
const char *b200_last_error(void) { return g_err; }
const char *b200_last_kernel(void) { return ""; }
int b200_device_count(int *count) { *count = 1; return B200_OK; }
int b200_set_device(int device) { g_device = device; return B200_OK; }
int b200_get_device(int *device) { *device = g_device; return B200_OK; }
int b200_device_sync(void) { return B200_OK; }
int b200_device_props(int64_t props[8])
{
    const int64_t v[8] = {148, 232448, 126 << 20, 10, 0, 180 * 1024, 0, 0};
    memcpy(props, v, sizeof(v));
    return B200_OK;
}
int b200_malloc(void **dptr, size_t bytes)
{
    if (posix_memalign(dptr, 256, bytes ? bytes : 1)) { set_error("out of host memory"); return B200_ENOMEM; }
    memset(*dptr, 0xA5, bytes);  // uninitialised device memory is not zero
    return B200_OK;
}
int b200_free(void *dptr) { free(dptr); return B200_OK; }
int b200_host_alloc(void **hptr, size_t bytes) { return b200_malloc(hptr, bytes); }
int b200_host_free(void *hptr) { free(hptr); return B200_OK; }
static int copy(void *d, const void *s, size_t n) { if (n) memmove(d, s, n); return B200_OK; }
int b200_memcpy_h2d(void *d, const void *s, size_t n, void *) { return copy(d, s, n); }
int b200_memcpy_d2h(void *d, const void *s, size_t n, void *) { return copy(d, s, n); }
int b200_memcpy_d2d(void *d, const void *s, size_t n, void *) { return copy(d, s, n); }
static int copy2d(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h)
{
    for (size_t r = 0; r < h; ++r) memmove((char *)d + r * dp, (const char *)s + r * sp, w);
    return B200_OK;
}
int b200_memcpy2d_h2d(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, void *) { return copy2d(d, dp, s, sp, w, h); }
int b200_memcpy2d_d2h(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, void *) { return copy2d(d, dp, s, sp, w, h); }
int b200_memset(void *dst, int value, size_t bytes, void *) { if (bytes) memset(dst, value, bytes); return B200_OK; }

int b200_fill_synthetic(const b200_field *f, const b200_field *base, uint64_t seed, int64_t gi0, int64_t gj0, double lo,
                        double hi, void *)
{
    double *p = (double *)f->ptr;
    for (int64_t k = 0; k < f->shape[2]; ++k)
        for (int64_t j = 0; j < f->shape[1]; ++j)
            for (int64_t i = 0; i < f->shape[0]; ++i) {
                const double v = b200::synth_value(gi0 + i, gj0 + j, k, seed, lo, hi);
                const double b = base ? ((const double *)base->ptr)[i * base->stride[0] + j * base->stride[1]] : 0.0;
                p[i * f->stride[0] + j * f->stride[1] + k * f->stride[2]] = base ? b + v : v;
            }
    return B200_OK;
}
int b200_stream_create(void **stream) { *stream = malloc(8); return B200_OK; }
int b200_stream_destroy(void *stream) { free(stream); return B200_OK; }
int b200_stream_sync(void *) { return B200_OK; }
int b200_stream_l2_window(void *, const void *, size_t bytes, double, size_t *granted) { if (granted) *granted = bytes; return B200_OK; }
int b200_event_create(void **event) { *event = malloc(8); return B200_OK; }
int b200_event_destroy(void *event) { free(event); return B200_OK; }
int b200_event_record(void *, void *) { return B200_OK; }
int b200_event_sync(void *) { return B200_OK; }
int b200_event_elapsed_ms(void *, void *, float *ms) { *ms = 0.0f; return B200_OK; }
int b200_stream_wait_event(void *, void *) { return B200_OK; }

int b200_graph_begin(void *) { UNSUP("b200_graph_begin"); }
int b200_graph_end(void *, void **) { UNSUP("b200_graph_end"); }
int b200_graph_launch(void *, void *) { UNSUP("b200_graph_launch"); }
int b200_graph_destroy(void *) { return B200_OK; }

int b200_hdiff(const b200_field *, const b200_field *, const b200_field *, const int64_t[3], int, int, void *) { UNSUP("b200_hdiff"); }
int b200_vadv(const b200_field *, const b200_field *, const b200_field *, const b200_field *, const b200_field *, double,
              const int64_t[3], int, void *) { UNSUP("b200_vadv"); }
int b200_vadv_noalloc(const b200_field *, const b200_field *, const b200_field *, const b200_field *, const b200_field *,
                      double, const int64_t[3], void *) { UNSUP("b200_vadv_noalloc"); }
int b200_tridiag(const b200_field *, const b200_field *, const b200_field *, const b200_field *, const b200_field *,
                 const int64_t[3], int, void *) { UNSUP("b200_tridiag"); }

// "image" = the path of the host library tests/host_emulation.py built from kernel.cu
int b200_module_load(const void *image, size_t bytes, void **module)
{
    std::string path((const char *)image, strnlen((const char *)image, bytes));
    void *dl = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!dl) { set_error("host double: cannot dlopen '%s': %s", path.c_str(), dlerror()); return B200_EINVAL; }
    *module = new Module{dl, {}};
    return B200_OK;
}
int b200_module_unload(void *module)
{
    Module *m = (Module *)module;
    for (Kernel *k : m->kernels) delete k;
    dlclose(m->dl);
    delete m;
    return B200_OK;
}
int b200_module_get_kernel(void *module, const char *name, void **kernel)
{
    Module *m = (Module *)module;
    std::string base(name);
    if (base.size() > 2 && (base.compare(base.size() - 2, 2, "_u") == 0)) base.resize(base.size() - 2);
    bool tma = base.size() > 2 && base.compare(base.size() - 2, 2, "_t") == 0;
    if (tma) base.resize(base.size() - 2);
    void *fn = dlsym(m->dl, ("emu_" + base).c_str());
    if (!fn) { set_error("host double: no kernel '%s'", name); return B200_EINVAL; }
    Kernel *k = new Kernel{tma ? nullptr : fn, dlsym(m->dl, ("emu_is_tile_" + base).c_str()) != nullptr, name};
    m->kernels.push_back(k);
    *kernel = k;
    return B200_OK;
}
int b200_kernel_set_smem(void *, int) { return B200_OK; }
int b200_tmap_encode(const b200_field *, const uint32_t[3], void *, int32_t *) { UNSUP("b200_tmap_encode (TMA)"); }
void b200_emu_set_window_mode(int mode) { g_window_mode = mode; }
int b200_launch(void *kernel, const uint32_t grid[3], const uint32_t block[3], uint32_t smem, void *, void **args, int nargs)
{
    Kernel *k = (Kernel *)kernel;
    if (!k || !k->fn || nargs != 1) { set_error("host double: kernel '%s' cannot be launched", k ? k->name.c_str() : "?"); return B200_EUNSUP; }
    if (k->tile) {
        typedef void (*fn_t)(const void *, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, int);
        ((fn_t)k->fn)(args[0], grid[0], grid[1], grid[2], block[0], block[1], block[2], smem, g_window_mode);
    } else {
        typedef void (*fn_t)(const void *, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned);
        ((fn_t)k->fn)(args[0], grid[0], grid[1], grid[2], block[0], block[1], block[2]);
    }
    return B200_OK;
}

int b200_ipc_get_mem_handle(void *, unsigned char[64]) { UNSUP("b200_ipc_get_mem_handle"); }
int b200_ipc_open_mem_handle(const unsigned char[64], void **) { UNSUP("b200_ipc_open_mem_handle"); }
int b200_ipc_close_mem_handle(void *) { UNSUP("b200_ipc_close_mem_handle"); }
int b200_halo_refresh(const b200_halo_box *, int, int, const b200_halo_sync *, uint64_t, unsigned int *, int, void *) { UNSUP("b200_halo_refresh"); }
int b200_halo_wait_done(const b200_halo_sync *, uint64_t, void *) { UNSUP("b200_halo_wait_done"); }
}

// the plan-level entry points: the product's own text (csrc/plan.inl) on top of the doubles above
#define B200_PLAN_SET_ERROR(msg) set_error("%s", msg)
#define B200_PLAN_COLFUSE_SMEM (226 * 1024)
#include "../gt4py_b200/csrc/plan.inl"
