/*
 * gt4py_b200.h -- C ABI of libgt4py_b200.so, the B200 (sm_100a) execution library behind the
 * "b200" backend of gt4py.cartesian.
 *
 * This is the drop-in boundary.  It replaces what the reference reaches through pybind11 in its
 * generated extension modules:
 *   - run_computation(domain, field, origin, ..., exec_info)
 *       /root/reference/src/gt4py/cartesian/backend/gtc_common.py:79-101 (C++ side) and :157-163
 *       (Python call site)  ->  b200_launch() for generated stencils, and the named entry points
 *       b200_hdiff / b200_vadv / b200_tridiag for the library's hand-written hot stencils;
 *   - gt::as_cuda_sid(buffer) + shift_sid_origin(origin)
 *       gtc_common.py:30-62                       ->  b200_field (raw pointer + element strides +
 *                                                     origin, read from __cuda_array_interface__);
 *   - cupy.cuda.Device(0).synchronize()  gtc_common.py:288-296  ->  b200_device_sync / sync flag;
 *   - cupy device allocation used by gt4py.storage
 *       /root/reference/src/gt4py/storage/cartesian/utils.py:258-276 -> b200_malloc / b200_free /
 *                                                     b200_memcpy_*.
 *
 * Conventions: plain C types only; every function returns 0 on success or a negative B200_E* code,
 * with a human-readable message available from b200_last_error() (thread-local).  No function
 * falls back to a CPU implementation: without a CUDA device compute calls fail with
 * B200_ECUDA.
 */
#ifndef GT4PY_B200_H
#define GT4PY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 1

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

/* error codes */
#define B200_OK 0
#define B200_EINVAL -1   /* bad argument (shape, stride, dtype, null pointer)           */
#define B200_ECUDA -2    /* CUDA runtime error (message holds cudaGetErrorString)       */
#define B200_EUNSUP -3   /* valid request the library has no kernel for                 */
#define B200_ENOMEM -4

/* dtype codes (match gt4py.cartesian.gtc.common.DataType ordering, common.py:105-127) */
#define B200_BOOL 1
#define B200_INT8 2
#define B200_INT16 3
#define B200_INT32 4
#define B200_INT64 5
#define B200_FLOAT32 6
#define B200_FLOAT64 7

#define B200_MAX_DIMS 5

/* One field argument: what as_cuda_sid + shift_sid_origin carry in the reference.
 * `ptr` is the address of element (0,0,0,...) of the array (NOT origin-shifted); `stride` is in
 * ELEMENTS per axis, axes ordered I, J, K, data dims; `origin` is the index of the first compute
 * point; `shape` is the array extent (used for bounds validation only). */
typedef struct b200_field {
    void *ptr;
    int64_t stride[B200_MAX_DIMS];
    int64_t shape[B200_MAX_DIMS];
    int32_t origin[B200_MAX_DIMS];
    int32_t ndim;
    int32_t dtype;
} b200_field;

/* ---- library / device ------------------------------------------------------------------- */
B200_API int b200_abi_version(void);
B200_API const char *b200_last_error(void);
/* name of the kernel the last b200_hdiff / b200_vadv / b200_tridiag call of this thread launched (what `variant = 0`
 * resolved to): measurements and tests report what really ran instead of assuming it */
B200_API const char *b200_last_kernel(void);
B200_API int b200_device_count(int *count);
B200_API int b200_set_device(int device);
B200_API int b200_get_device(int *device);
B200_API int b200_device_sync(void);
/* props[0]=SM count, [1]=max smem/block optin, [2]=L2 bytes, [3]=cc major, [4]=cc minor,
 * [5]=total global mem (MiB) */
B200_API int b200_device_props(int64_t props[8]);

/* ---- memory (device allocations are 256 B aligned by cudaMalloc) ---------------------------- */
B200_API int b200_malloc(void **dptr, size_t bytes);
B200_API int b200_free(void *dptr);
B200_API int b200_host_alloc(void **hptr, size_t bytes); /* pinned */
B200_API int b200_host_free(void *hptr);
B200_API int b200_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream);
B200_API int b200_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream);
B200_API int b200_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream);
/* pitched copies (height rows of `width` bytes): a J-slab of an I-contiguous field is one of these */
B200_API int b200_memcpy2d_h2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                      size_t height, void *stream);
B200_API int b200_memcpy2d_d2h(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width,
                      size_t height, void *stream);
/* stream == NULL: legacy default stream and complete on return (like the copies); otherwise asynchronous on `stream` */
B200_API int b200_memset(void *dst, int value, size_t bytes, void *stream);

/* Synthetic data (bench / test plumbing, csrc/synthetic.cu): fill every element of a 3-D float64 field, halo cells
 * included, with  base2d(i, j) + lo + (hi - lo) * u(gi0 + i, gj0 + j, k, seed),  u in [0, 1) a counter-based hash of the
 * GLOBAL coordinates (csrc/synthetic.cuh; numpy restatement oracle/inputs.py:hashed_uniform, bit-identical).
 * base2d may be NULL. */
B200_API int b200_fill_synthetic(const b200_field *field, const b200_field *base2d, uint64_t seed, int64_t gi0, int64_t gj0,
                        double lo, double hi, void *stream);

/* ---- streams / events ------------------------------------------------------------------- */
B200_API int b200_stream_create(void **stream);
B200_API int b200_stream_destroy(void *stream);
B200_API int b200_stream_sync(void *stream);
/* L2 residency of an inter-stencil field (no counterpart in the reference: its GPU backends leave caching to the
 * hardware).  Accesses of work launched on `stream` to [ptr, ptr + bytes) become persisting in
 * the set-aside part of L2 (cudaLimitPersistingL2CacheSize, grown to min(bytes, device maximum); *granted reports it),
 * everything else streams.  ptr == NULL or bytes == 0 removes the window, resets the persisting lines and returns the
 * set-aside (it is taken from the L2 every other kernel uses: do switch it off after use).  `stream`
 * must be a stream of b200_stream_create (the legacy NULL stream has no attributes).  hit_ratio <= 0 means 1. */
B200_API int b200_stream_l2_window(void *stream, const void *ptr, size_t bytes, double hit_ratio, size_t *granted);
B200_API int b200_event_create(void **event);
B200_API int b200_event_destroy(void *event);
B200_API int b200_event_record(void *event, void *stream);
B200_API int b200_event_sync(void *event);
B200_API int b200_event_elapsed_ms(void *start, void *stop, float *ms);
B200_API int b200_stream_wait_event(void *stream, void *event);

/* ---- CUDA graphs: capture a sequence of stencil launches once, replay it with one launch ----------
 * The reference has no counterpart: its callers (model drivers) issue one StencilObject.__call__ per stencil
 * (stencil_object.py:534-605, tens of microseconds of Python each); SURVEY.md 8(f) rank 2.
 * b200_graph_begin puts `stream` into capture mode (relaxed: allocations stay legal); every b200_* launch on that
 * stream is then recorded instead of executed.  b200_graph_end instantiates the executable graph. */
B200_API int b200_graph_begin(void *stream);
B200_API int b200_graph_end(void *stream, void **graph_exec);
B200_API int b200_graph_launch(void *graph_exec, void *stream);
B200_API int b200_graph_destroy(void *graph_exec);

/* ---- hand-written hot stencils ------------------------------------------------------------ */
/* variant: 0 = auto (fastest kernel whose layout preconditions hold), 1 = generic point kernel
 * (any strides), >= 2 = a specific tuned kernel (B200_EUNSUP if its preconditions fail).      */

/* horizontal_diffusion(in_field, out_field, coeff): reference definition
 * tests/cartesian_tests/integration_tests/multi_feature_tests/stencil_definitions.py:316-328.
 * `limiter` = 1 for that stencil, 0 for simple_horizontal_diffusion (:206-216).
 * in_field needs a halo of 2 in I and J around `domain`.  All three fields float64, or all three float32
 * (the definition built with float32 fields and `literal_float_precision=32`: float32 literals and temporaries).
 * variants: 1 point kernel, 2/3 register J-march (prefetch depth 2/4), 4 alternating march directions (float64),
 * 5 TMA-fed shared-memory rings (float64; the `auto` choice when strides/alignment allow and the fields exceed
 * the L2). */
B200_API int b200_hdiff(const b200_field *in_field, const b200_field *out_field, const b200_field *coeff,
               const int64_t domain[3], int limiter, int variant, void *stream);

/* vertical_advection_dycore(utens_stage, u_stage, wcon, u_pos, utens, dtr_stage) with externals
 * BET_M = BET_P = 0.5: stencil_definitions.py:235-313.  wcon is read at i+1.  domain[2] >= 3. */
B200_API int b200_vadv(const b200_field *utens_stage, const b200_field *u_stage, const b200_field *wcon,
              const b200_field *u_pos, const b200_field *utens, double dtr_stage,
              const int64_t domain[3], int variant, void *stream);

/* b200_vadv_noalloc: vertical_advection_dycore for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE:
 * added (code under CUDA-graph capture, tight serving loops and harnesses
 * that forbid cudaMalloc).  Same arguments, checks and results (bit-identical) as b200_vadv with variant 1 -- the
 * thread-per-column kernel whose (ccol, dcol) LIFO lives in global scratch -- but the scratch is a fixed
 * `static __device__` array inside the library instead of a cudaMalloc'd buffer: the call performs no cudaMalloc,
 * cudaFree or any other allocating routine.
 *   limit   domain[0] * domain[1] * domain[2] <= B200_VADV_NOALLOC_MAX_POINTS (2 Mi grid points, e.g. 128 x 128 x 128
 *           or 160 x 160 x 80; 16 bytes of scratch per point = 32 MiB, resident on every GPU that loads the library
 *           until the process exits).  Larger domains are rejected with B200_EINVAL and a message; any strides work.
 *   sharing every call on a device uses the SAME array.  Nothing in it needs clearing (each slot is written by the
 *           forward sweep before the backward sweep reads it), but calls that overlap on the GPU -- different
 *           streams, or graph branches without an edge between them -- race on it: serialise them. */
#define B200_VADV_NOALLOC_MAX_POINTS (2 * 1024 * 1024)
B200_API int b200_vadv_noalloc(const b200_field *utens_stage, const b200_field *u_stage, const b200_field *wcon,
                      const b200_field *u_pos, const b200_field *utens, double dtr_stage,
                      const int64_t domain[3], void *stream);

/* tridiagonal_solver(inf, diag, sup, rhs, out): stencil_definitions.py:219-232 (sup, rhs updated
 * in place). */
B200_API int b200_tridiag(const b200_field *inf, const b200_field *diag, const b200_field *sup,
                 const b200_field *rhs, const b200_field *out, const int64_t domain[3],
                 int variant, void *stream);

/* ---- generated stencils (OIR -> CUDA C -> cubin, built by the Python backend) --------------- */
/* Load a cubin/PTX image, look kernels up by name, launch with a packed argument block.        */
B200_API int b200_module_load(const void *image, size_t bytes, void **module);
B200_API int b200_module_unload(void *module);
B200_API int b200_module_get_kernel(void *module, const char *name, void **kernel);
B200_API int b200_kernel_set_smem(void *kernel, int dynamic_smem_bytes);
/* A 128-byte TMA descriptor (CUtensorMap, cuTensorMapEncodeTiled) over the I, J, K axes of `field` -- the whole
 * allocation, so tile coordinates are array indices (origin + offset) and out-of-range parts of a box are
 * zero-filled by the hardware.  The generated tile kernels take it inside their argument block and stage the
 * halo'd input tiles in shared memory with cp.async.bulk.tensor.  box = tile extent per axis (<= 256 each).
 * Needs unit I stride and J / K strides that are multiples of 16 bytes; otherwise B200_EUNSUP (the caller then
 * launches the non-TMA entry point of the same kernel).  TMA wants a 16-byte aligned base: when the array starts
 * 4 / 8 / 12 bytes off (gt4py.storage aligns the ORIGIN, so an odd halo leaves the base misaligned) the descriptor
 * starts at the aligned address below it and *i_shift receives the number of elements to add to I coordinates. */
B200_API int b200_tmap_encode(const b200_field *field, const uint32_t box[3], void *tmap128, int32_t *i_shift);
/* args: array of nargs pointers to the argument values (cudaLaunchKernel convention). */
B200_API int b200_launch(void *kernel, const uint32_t grid[3], const uint32_t block[3],
                uint32_t dynamic_smem_bytes, void *stream, void **args, int nargs);

/* ---- a generated stencil as one object: the plan-level entry points (csrc/plan.inl) ------------------------------
 * Replace the reference's per-stencil pybind11 entry `run_computation(domain, field, origin, ..., exec_info)`
 * (/root/reference/src/gt4py/cartesian/backend/gtc_common.py:79-101, call site :157-163) for hosts in any language.
 * `desc` is the text the backend writes next to the generated module at build time (`plan.desc`; format: one record
 * per line -- `b200plan 1 <name>`, `field`, `scalar`, `temp`, `kernel`, `member`, optional `library`, `end` -- see
 * gt4py_b200/generated.py:plan_descriptor), `image` the cubin of `kernel.cu` (nvcc -cubin -arch sm_100a -fmad=false).
 * flags bit 0: never dispatch to the hand-written library kernel even if the plan names one; bit 1: never use the
 * TMA-staged entry points (A/B runs).
 * b200_plan_launch: `fields` in the order of the descriptor's `field` records (API order; ptr == NULL for a field the
 * stencil does not reference), `scalars` in the order of its `scalar` records.  The plan packs the argument block,
 * picks the entry point of every kernel (any strides / unit stride + 32-bit indices / TMA-staged), owns the global
 * scratch of the OIR temporaries (per domain and stream), the TMA descriptors and the spill areas of the fused column
 * kernels; all of it is freed by b200_plan_destroy.  Field memory stays the caller's.  sync != 0: return after the
 * stream (or, for a NULL stream, the device) has been synchronised -- the reference's
 * `cupy.cuda.Device(0).synchronize()` (gtc_common.py:288-296).  One launch at a time per plan (calls are serialised
 * by a lock inside the plan). */
typedef struct b200_scalar {
    union {
        double f64;
        float f32;
        int64_t i64;
        int32_t i32;
        int16_t i16;
        int8_t i8;
        uint8_t b;
    } value;
    int32_t dtype; /* B200_* code; must equal the descriptor's */
    int32_t reserved;
} b200_scalar;
B200_API int b200_plan_create(const char *desc, size_t desc_bytes, const void *image, size_t image_bytes, int flags,
                     void **plan);
B200_API int b200_plan_launch(void *plan, const int64_t domain[3], const b200_field *fields, int nfields,
                     const b200_scalar *scalars, int nscalars, void *stream, int sync);
B200_API int b200_plan_destroy(void *plan);
/* comma-separated entry points the last b200_plan_launch of this plan launched (e.g. "k0_0_t"; "library:<kernel>") */
B200_API int b200_plan_last_entries(void *plan, char *buf, size_t n);

/* ---- multi-GPU halo refresh over NVLink peer memory (csrc/halo.cu) ------------------------------
 * New functionality: the reference has no distributed layer.  One process per GPU;
 * a rank maps its neighbours' field allocations with CUDA IPC and pulls the halo boxes with one kernel.
 * Ordering between ranks: step-numbered ready / done flags in peer-visible memory (see halo.cu). */
typedef struct b200_halo_box {
    const double *src;      /* first element of the box in the neighbour's (mapped) memory */
    double *dst;            /* first element of the box in the local halo */
    int64_t src_stride[3];  /* element strides I, J, K */
    int64_t dst_stride[3];
    int32_t extent[3];
    int32_t elem_bytes;     /* element size: 1, 2, 4 or 8 (0 = 8); src / dst are then pointers to that type */
} b200_halo_box;

#define B200_HALO_MAX_PEERS 8
typedef struct b200_halo_sync {
    uint64_t *post_ready[B200_HALO_MAX_PEERS]; /* slot in peer p's flag block that this rank writes */
    uint64_t *post_done[B200_HALO_MAX_PEERS];
    const uint64_t *wait_ready[B200_HALO_MAX_PEERS]; /* local slot that peer p writes */
    const uint64_t *wait_done[B200_HALO_MAX_PEERS];
    int32_t npeers;
} b200_halo_sync;

B200_API int b200_ipc_get_mem_handle(void *dptr, unsigned char handle[64]);
B200_API int b200_ipc_open_mem_handle(const unsigned char handle[64], void **dptr);
B200_API int b200_ipc_close_mem_handle(void *dptr);
/* boxes_dev: device array of nboxes boxes; arrive_dev: zero-initialised device counter.  `use_flags` is the
 * protocol mode: 0 = plain box copy (single-process tests, or the caller orders the ranks itself);
 * 1 = PULL (post ready, wait ready, copy from peer memory, post done); 2 = PUSH (copy into the neighbours'
 * staging buffers, then post ready); 3 = UNPACK (wait ready, copy local staging -> halo); 4 + 8 * nsplit = PUSH of
 * boxes [0, nsplit) and UNPACK of boxes [nsplit, nboxes) in ONE launch -- the fast path: one launch per refresh. */
B200_API int b200_halo_refresh(const b200_halo_box *boxes_dev, int nboxes, int blocks_per_box,
                      const b200_halo_sync *sync, uint64_t step, unsigned int *arrive_dev,
                      int use_flags, void *stream);
/* Block `stream` until every neighbour has finished pulling step `step` out of this rank's fields. */
B200_API int b200_halo_wait_done(const b200_halo_sync *sync, uint64_t step, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GT4PY_B200_H */
