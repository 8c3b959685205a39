// runtime.cu -- device, memory, stream/event and module-launch entry points of the C ABI.
#include <cstring>

#include "common.cuh"

namespace b200 {
static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static thread_local const char *g_kernel = "";
void note_kernel(const char *name) { g_kernel = name; }
}  // namespace b200

extern "C" {

int b200_abi_version(void) { return B200_ABI_VERSION; }
const char *b200_last_error(void) { return b200::g_err; }
const char *b200_last_kernel(void) { return b200::g_kernel; }

int b200_device_count(int *count)
{
    B200_REQUIRE(count, "count is null");
    B200_CUDA_CHECK(cudaGetDeviceCount(count));
    return B200_OK;
}
int b200_set_device(int device)
{
    B200_CUDA_CHECK(cudaSetDevice(device));
    return B200_OK;
}
int b200_get_device(int *device)
{
    B200_CUDA_CHECK(cudaGetDevice(device));
    return B200_OK;
}
int b200_device_sync(void)
{
    B200_CUDA_CHECK(cudaDeviceSynchronize());
    return B200_OK;
}
int b200_device_props(int64_t props[8])
{
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    cudaDeviceProp p;
    B200_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
    props[0] = p.multiProcessorCount;
    props[1] = (int64_t)p.sharedMemPerBlockOptin;
    props[2] = p.l2CacheSize;
    props[3] = p.major;
    props[4] = p.minor;
    props[5] = (int64_t)(p.totalGlobalMem >> 20);
    props[6] = props[7] = 0;
    return B200_OK;
}

int b200_malloc(void **dptr, size_t bytes)
{
    B200_REQUIRE(dptr, "dptr is null");
    cudaError_t e = cudaMalloc(dptr, bytes ? bytes : 1);
    if (e == cudaErrorMemoryAllocation) {
        b200::set_error("cudaMalloc(%zu bytes): out of device memory", bytes);
        return B200_ENOMEM;
    }
    B200_CUDA_CHECK(e);
    return B200_OK;
}
int b200_free(void *dptr)
{
    B200_CUDA_CHECK(cudaFree(dptr));
    return B200_OK;
}
int b200_host_alloc(void **hptr, size_t bytes)
{
    B200_REQUIRE(hptr, "hptr is null");
    B200_CUDA_CHECK(cudaMallocHost(hptr, bytes ? bytes : 1));
    return B200_OK;
}
int b200_host_free(void *hptr)
{
    B200_CUDA_CHECK(cudaFreeHost(hptr));
    return B200_OK;
}
static int copy(void *dst, const void *src, size_t bytes, void *stream, cudaMemcpyKind kind)
{
    if (bytes == 0) return B200_OK;
    B200_REQUIRE(dst && src, "memcpy: null pointer");
    B200_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, kind, (cudaStream_t)stream));
    if (!stream) B200_CUDA_CHECK(cudaStreamSynchronize(0));
    return B200_OK;
}
int b200_memcpy_h2d(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyHostToDevice); }
int b200_memcpy_d2h(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyDeviceToHost); }
int b200_memcpy_d2d(void *d, const void *s, size_t n, void *st) { return copy(d, s, n, st, cudaMemcpyDeviceToDevice); }
static int copy2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                  void *stream, cudaMemcpyKind kind)
{
    if (width == 0 || height == 0) return B200_OK;
    B200_REQUIRE(dst && src, "memcpy2d: null pointer");
    B200_REQUIRE(width <= dpitch && width <= spitch, "memcpy2d: width larger than pitch");
    B200_CUDA_CHECK(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, kind, (cudaStream_t)stream));
    if (!stream) B200_CUDA_CHECK(cudaStreamSynchronize(0));
    return B200_OK;
}
int b200_memcpy2d_h2d(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, void *st)
{
    return copy2d(d, dp, s, sp, w, h, st, cudaMemcpyHostToDevice);
}
int b200_memcpy2d_d2h(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, void *st)
{
    return copy2d(d, dp, s, sp, w, h, st, cudaMemcpyDeviceToHost);
}
int b200_memset(void *dst, int value, size_t bytes, void *stream)
{
    if (bytes == 0) return B200_OK;
    B200_CUDA_CHECK(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
    // null stream: finished on return, like the copies above (the library's own streams are non-blocking and do
    // not order against the legacy default stream)
    if (!stream) B200_CUDA_CHECK(cudaStreamSynchronize(0));
    return B200_OK;
}

int b200_stream_create(void **stream)
{
    cudaStream_t s;
    B200_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = s;
    return B200_OK;
}
int b200_stream_destroy(void *stream)
{
    B200_CUDA_CHECK(cudaStreamDestroy((cudaStream_t)stream));
    return B200_OK;
}
int b200_stream_sync(void *stream)
{
    B200_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return B200_OK;
}
// L2 residency for a field that one stencil writes and the next one reads: the stream's access
// policy window makes accesses to [ptr, ptr + bytes) persisting in the set-aside part of L2, everything else streaming.
int b200_stream_l2_window(void *stream, const void *ptr, size_t bytes, double hit_ratio, size_t *granted)
{
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    int max_persist = 0, max_window = 0;
    B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
    B200_CUDA_CHECK(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev));
    cudaStreamAttrValue attr;
    memset(&attr, 0, sizeof(attr));
    static size_t carved[64] = {0};  // per device: the set-aside only grows while windows are in use
    if (!ptr || !bytes) {
        // switch off: no window, persisting lines back to normal, and the set-aside returned -- it is taken from the L2
        // every other kernel uses (measured: vadv at 2048x2048x80 2.72 -> 4.36 ms with 83 MB set aside)
        attr.accessPolicyWindow.num_bytes = 0;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
        B200_CUDA_CHECK(cudaStreamSetAttribute((cudaStream_t)stream, cudaStreamAttributeAccessPolicyWindow, &attr));
        B200_CUDA_CHECK(cudaCtxResetPersistingL2Cache());
        if (dev < 64 && carved[dev]) {
            B200_CUDA_CHECK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0));
            carved[dev] = 0;
        }
        if (granted) *granted = 0;
        return B200_OK;
    }
    if (max_persist <= 0 || max_window <= 0) {
        b200::set_error("stream_l2_window: this device has no persisting L2");
        return B200_EINVAL;
    }
    const size_t carve = bytes < (size_t)max_persist ? bytes : (size_t)max_persist;
    if (dev < 64 && carve > carved[dev]) {  // (setting the limit is a heavy call: only when it has to grow)
        B200_CUDA_CHECK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
        carved[dev] = carve;
    }
    const size_t window = bytes < (size_t)max_window ? bytes : (size_t)max_window;
    attr.accessPolicyWindow.base_ptr = const_cast<void *>(ptr);
    attr.accessPolicyWindow.num_bytes = window;
    // a window larger than the set-aside: only that fraction of its lines is made persisting, the rest streams
    double ratio = hit_ratio > 0 ? hit_ratio : 1.0;
    if ((double)window * ratio > (double)carve) ratio = (double)carve / (double)window;
    attr.accessPolicyWindow.hitRatio = (float)ratio;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    B200_CUDA_CHECK(cudaStreamSetAttribute((cudaStream_t)stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    if (granted) *granted = carve;
    return B200_OK;
}
int b200_event_create(void **event)
{
    cudaEvent_t e;
    B200_CUDA_CHECK(cudaEventCreate(&e));
    *event = e;
    return B200_OK;
}
int b200_event_destroy(void *event)
{
    B200_CUDA_CHECK(cudaEventDestroy((cudaEvent_t)event));
    return B200_OK;
}
int b200_event_record(void *event, void *stream)
{
    B200_CUDA_CHECK(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
    return B200_OK;
}
int b200_event_sync(void *event)
{
    B200_CUDA_CHECK(cudaEventSynchronize((cudaEvent_t)event));
    return B200_OK;
}
int b200_event_elapsed_ms(void *start, void *stop, float *ms)
{
    B200_CUDA_CHECK(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return B200_OK;
}
int b200_stream_wait_event(void *stream, void *event)
{
    B200_CUDA_CHECK(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0));
    return B200_OK;
}

// ---- CUDA graphs ----------------------------------------------------------------------------
int b200_graph_begin(void *stream)
{
    B200_REQUIRE(stream, "graph_begin: capture needs an explicit (non-default) stream");
    B200_CUDA_CHECK(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeRelaxed));
    return B200_OK;
}
int b200_graph_end(void *stream, void **graph_exec)
{
    B200_REQUIRE(stream && graph_exec, "graph_end: bad arguments");
    cudaGraph_t graph = nullptr;
    B200_CUDA_CHECK(cudaStreamEndCapture((cudaStream_t)stream, &graph));
    cudaGraphExec_t exec = nullptr;
    cudaError_t err = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    B200_CUDA_CHECK(err);
    *graph_exec = exec;
    return B200_OK;
}
int b200_graph_launch(void *graph_exec, void *stream)
{
    B200_REQUIRE(graph_exec, "graph_launch: null graph");
    B200_CUDA_CHECK(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
    return B200_OK;
}
int b200_graph_destroy(void *graph_exec)
{
    if (graph_exec) B200_CUDA_CHECK(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
    return B200_OK;
}

// ---- generated-stencil modules: cubin images built by the Python backend with nvcc ----------
int b200_module_load(const void *image, size_t bytes, void **module)
{
    B200_REQUIRE(image && bytes && module, "module_load: bad arguments");
    cudaLibrary_t lib;
    B200_CUDA_CHECK(cudaLibraryLoadData(&lib, image, nullptr, nullptr, 0, nullptr, nullptr, 0));
    *module = lib;
    return B200_OK;
}
int b200_module_unload(void *module)
{
    B200_CUDA_CHECK(cudaLibraryUnload((cudaLibrary_t)module));
    return B200_OK;
}
int b200_module_get_kernel(void *module, const char *name, void **kernel)
{
    cudaKernel_t k;
    B200_CUDA_CHECK(cudaLibraryGetKernel(&k, (cudaLibrary_t)module, name));
    *kernel = k;
    return B200_OK;
}
int b200_kernel_set_smem(void *kernel, int dynamic_smem_bytes)
{
    B200_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         dynamic_smem_bytes));
    return B200_OK;
}
int b200_launch(void *kernel, const uint32_t grid[3], const uint32_t block[3],
                uint32_t dynamic_smem_bytes, void *stream, void **args, int nargs)
{
    (void)nargs;
    B200_REQUIRE(kernel && grid && block, "launch: bad arguments");
    if (grid[0] == 0 || grid[1] == 0 || grid[2] == 0) return B200_OK;  // empty domain
    B200_CUDA_CHECK(cudaLaunchKernel(kernel, dim3(grid[0], grid[1], grid[2]),
                                     dim3(block[0], block[1], block[2]), args, dynamic_smem_bytes,
                                     (cudaStream_t)stream));
    return B200_OK;
}

}  // extern "C"

// ---- TMA tensor maps (driver entry point resolved through the runtime; libcuda is not linked) ----
#include "tma.cuh"

namespace b200 {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

// The launchers encode 2-7 tensor maps per stencil call; a map depends only on (base, extents, strides, box), which
// repeat from call to call (a model steps through the same few fields).  A small per-thread cache turns the driver
// call into a 64-byte key compare.  Maps hold no resources, so entries never need invalidation: the same key always
// describes the same bytes.
namespace {
struct TmapKey {
    const void *base;
    int64_t ext[3], sj, sk;
    uint32_t box[3], pad;
    bool operator==(const TmapKey &o) const { return memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapEntry {
    TmapKey key;
    CUtensorMap map;
    bool valid;
};
constexpr int TMAP_CACHE = 64;  // direct-mapped
thread_local TmapEntry tmap_cache[TMAP_CACHE];
}  // namespace

int encode_tmap_f64(CUtensorMap *map, const double *base, const int64_t extent[3], int64_t sj, int64_t sk,
                    const uint32_t box[3])
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        B200_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !p) {
            set_error("cuTensorMapEncodeTiled is not available in this driver");
            return B200_ECUDA;
        }
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    TmapKey key;
    memset(&key, 0, sizeof(key));
    key.base = base;
    for (int d = 0; d < 3; ++d) { key.ext[d] = extent[d]; key.box[d] = box[d]; }
    key.sj = sj;
    key.sk = sk;
    uint64_t h = reinterpret_cast<uintptr_t>(base) >> 4;
    h ^= (uint64_t)extent[0] * 0x9E3779B97F4A7C15ull ^ (uint64_t)extent[1] * 0xC2B2AE3D27D4EB4Full ^
         (uint64_t)box[0] * 0x165667B19E3779F9ull ^ (uint64_t)box[1] << 7 ^ (uint64_t)box[2] << 13;
    h ^= h >> 29;
    TmapEntry &e = tmap_cache[h % TMAP_CACHE];
    if (e.valid && e.key == key) {
        *map = e.map;
        return B200_OK;
    }
    cuuint64_t dims[3] = {(cuuint64_t)extent[0], (cuuint64_t)extent[1], (cuuint64_t)extent[2]};
    cuuint64_t strides[2] = {(cuuint64_t)sj * 8, (cuuint64_t)sk * 8};
    cuuint32_t bx[3] = {box[0], box[1], box[2]};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double *>(base), dims, strides, bx, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (extent %lld x %lld x %lld, strides %lld/%lld, "
                  "box %u x %u x %u)", (int)r, (long long)extent[0], (long long)extent[1], (long long)extent[2],
                  (long long)sj, (long long)sk, box[0], box[1], box[2]);
        return B200_ECUDA;
    }
    e.key = key;
    e.map = *map;
    e.valid = true;
    return B200_OK;
}
}  // namespace b200


extern "C" int b200_tmap_encode(const b200_field *f, const uint32_t box[3], void *tmap128, int32_t *i_shift)
{
    using namespace b200;
    B200_REQUIRE(f && f->ptr && box && tmap128 && i_shift, "tmap_encode: bad arguments");
    B200_REQUIRE(f->ndim >= 3, "tmap_encode: field needs the I, J, K axes");
    CUtensorMapDataType dt;
    int item;
    switch (f->dtype) {
    case B200_FLOAT64: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT64; item = 8; break;
    case B200_FLOAT32: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; item = 4; break;
    case B200_INT64: dt = CU_TENSOR_MAP_DATA_TYPE_INT64; item = 8; break;
    case B200_INT32: dt = CU_TENSOR_MAP_DATA_TYPE_INT32; item = 4; break;
    default: set_error("tmap_encode: dtype %d has no TMA element type", f->dtype); return B200_EUNSUP;
    }
    const int64_t sj = f->stride[1] * item, sk = f->stride[2] * item;
    const uintptr_t misalign = (uintptr_t)f->ptr & 15;
    if (f->stride[0] != 1 || misalign % item || sj % 16 || sk % 16 || sj <= 0 || sk <= 0 ||
        (box[0] * item) % 16 || box[0] > 256 || box[1] > 256 || box[2] > 256 || !box[0] || !box[1] || !box[2]) {
        set_error("tmap_encode: layout not TMA-compatible (stride_i %lld, byte strides %lld / %lld, box %u x %u x %u)",
                  (long long)f->stride[0], (long long)sj, (long long)sk, box[0], box[1], box[2]);
        return B200_EUNSUP;
    }
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        B200_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !p) {
            set_error("cuTensorMapEncodeTiled is not available in this driver");
            return B200_ECUDA;
        }
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    const int shift = (int)(misalign / item);  // the bytes below the base belong to the same (256-byte aligned) allocation
    void *base = (char *)f->ptr - misalign;
    *i_shift = shift;
    cuuint64_t dims[3] = {(cuuint64_t)(f->shape[0] + shift), (cuuint64_t)f->shape[1], (cuuint64_t)f->shape[2]};
    cuuint64_t strides[2] = {(cuuint64_t)sj, (cuuint64_t)sk};
    cuuint32_t bx[3] = {box[0], box[1], box[2]};
    cuuint32_t es[3] = {1, 1, 1};
    alignas(64) CUtensorMap map;
    CUresult r = fn(&map, dt, 3, base, dims, strides, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return B200_ECUDA;
    }
    static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap is 128 bytes");
    memcpy(tmap128, &map, 128);
    return B200_OK;
}
