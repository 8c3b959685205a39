// Standalone probe: TMA tile load with the helper PTX of the generated kernels (tools/probe, not part of the product).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)
__device__ __forceinline__ unsigned b200_smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void b200_mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b200_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void b200_mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b200_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void b200_mbar_wait(unsigned long long *bar, unsigned parity)
{
    unsigned done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(b200_smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void b200_tma_load_3d(void *dst, const void *tmap, int c0, int c1, int c2, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(b200_smem_u32(dst)), "l"((unsigned long long)tmap), "r"(c0), "r"(c1), "r"(c2), "r"(b200_smem_u32(bar)) : "memory");
}
struct Args { const void *tm; double *out; int c0, c1, c2, w, h; };
__global__ void __launch_bounds__(288) probe(const Args a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    double *s = (double *)smem;
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0 && threadIdx.y == 0) b200_mbar_init(&bar, 1);
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        b200_mbar_expect_tx(&bar, a.w * a.h * 8);
        b200_tma_load_3d(s, a.tm, a.c0, a.c1, a.c2, &bar);
    }
    b200_mbar_wait(&bar, 0);
    for (int t = threadIdx.y * blockDim.x + threadIdx.x; t < a.w * a.h; t += blockDim.x * blockDim.y) a.out[t] = s[t];
}
struct ArgsG { unsigned char tm[128]; double *out; int c0, c1, c2, w, h; };
__global__ void __launch_bounds__(288) probe_param(const __grid_constant__ ArgsG a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    double *s = (double *)smem;
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0 && threadIdx.y == 0) b200_mbar_init(&bar, 1);
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        b200_mbar_expect_tx(&bar, a.w * a.h * 8);
        b200_tma_load_3d(s, a.tm, a.c0, a.c1, a.c2, &bar);
    }
    b200_mbar_wait(&bar, 0);
    for (int t = threadIdx.y * blockDim.x + threadIdx.x; t < a.w * a.h; t += blockDim.x * blockDim.y) a.out[t] = s[t];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int run(int W, int H, int misalign_elems)
{
    const int NI = 76, NJ = 43, NK = 4, PITCH = 80;
    double *buf; CK(cudaMalloc(&buf, (size_t)PITCH * NJ * NK * 8 + 256));
    std::vector<double> host((size_t)PITCH * NJ * NK + 32);
    for (size_t n = 0; n < host.size(); ++n) host[n] = (double)n;
    CK(cudaMemcpy(buf, host.data(), host.size() * 8, cudaMemcpyHostToDevice));
    double *base = buf + misalign_elems;   // array base (possibly 8 bytes off 16)
    void *p = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    EncodeTiledFn fn = (EncodeTiledFn)p;
    uintptr_t mis = (uintptr_t)base & 15; int shift = (int)(mis / 8);
    cuuint64_t dims[3] = {(cuuint64_t)(NI + shift), NJ, NK};
    cuuint64_t strides[2] = {PITCH * 8, (cuuint64_t)PITCH * NJ * 8};
    cuuint32_t bx[3] = {(cuuint32_t)W, (cuuint32_t)H, 1}, es[3] = {1, 1, 1};
    alignas(64) CUtensorMap map;
    CUresult r = fn(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (char *)base - mis, dims, strides, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("W %d H %d misalign %d: encode %d\n", W, H, misalign_elems, (int)r);
    if (r) return 1;
    void *dmap; CK(cudaMalloc(&dmap, 128)); CK(cudaMemcpy(dmap, &map, 128, cudaMemcpyHostToDevice));
    double *out; CK(cudaMalloc(&out, (size_t)W * H * 8));
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
    CK(cudaFuncSetAttribute(probe_param, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000));
    Args a{dmap, out, 1 + shift, 0, 1, W, H};
    probe<<<1, dim3(72, 4), W * H * 8 + 128>>>(a);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  global-memory descriptor: %s\n", cudaGetErrorString(e));
    if (e == cudaSuccess) {
        std::vector<double> o((size_t)W * H); CK(cudaMemcpy(o.data(), out, o.size() * 8, cudaMemcpyDeviceToHost));
        double want = host[(size_t)misalign_elems + 1 * (size_t)PITCH * NJ + 0 * PITCH + 1];
        printf("  first %.0f want %.0f ; row1 %.0f want %.0f\n", o[0], want, o[W], want + PITCH);
    } else return 2;
    ArgsG g; memcpy(g.tm, &map, 128); g.out = out; g.c0 = 1 + shift; g.c1 = 0; g.c2 = 1; g.w = W; g.h = H;
    probe_param<<<1, dim3(72, 4), W * H * 8 + 128>>>(g);
    e = cudaDeviceSynchronize();
    printf("  param-space descriptor (unsigned char[128] in struct): %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 3;
}
int main(int argc, char **argv)
{
    int W = argc > 1 ? atoi(argv[1]) : 70, H = argc > 2 ? atoi(argv[2]) : 22, m = argc > 3 ? atoi(argv[3]) : 1;
    return run(W, H, m);
}
