// ubench_fp64.cu -- B200 fp64 latency / issue microbenchmark (design input for the vadv kernel).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o ubench_fp64 ubench_fp64.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, int OP>
__global__ void chain(double *out, double seed, int iters, long long *cycles)
{
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = seed + i + threadIdx.x * 1e-3;
    const double a = 1.0000001, b = 1e-9;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == 0) x[i] = fma(x[i], a, b);
            else if (OP == 1) x[i] = x[i] + b;
            else if (OP == 2) x[i] = x[i] * a;
            else if (OP == 3) x[i] = 1.0 / x[i];
            else if (OP == 4) { double y; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x[i])); x[i] = y; }
            else if (OP == 5) { float f = (float)x[i]; f = fmaf(f, 1.0000001f, 1e-9f); x[i] = f; }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int ILP, int OP>
void run(const char *name, int threads)
{
    double *out;
    long long *cyc, h;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&cyc, sizeof(long long));
    const int iters = 2000;
    chain<ILP, OP><<<148, threads>>>(out, 1.5, iters, cyc);
    chain<ILP, OP><<<148, threads>>>(out, 1.5, iters, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-10s ILP=%d threads/SM=%4d : %7.2f cycles per op-instruction per warp-slot (%.2f cycles/iter)\n", name, ILP,
           threads, (double)h / iters / ILP, (double)h / iters);
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    for (int threads : {32, 128, 256, 512, 1024}) {
        run<1, 0>("DFMA", threads);
        run<2, 0>("DFMA", threads);
        run<4, 0>("DFMA", threads);
        run<8, 0>("DFMA", threads);
    }
    run<1, 1>("DADD", 128);
    run<1, 2>("DMUL", 128);
    run<4, 1>("DADD", 128);
    run<1, 3>("1.0/x", 128);
    run<2, 3>("1.0/x", 128);
    run<4, 3>("1.0/x", 128);
    run<4, 3>("1.0/x", 256);
    run<1, 4>("RCP64H", 128);
    run<4, 4>("RCP64H", 128);
    run<1, 5>("FFMA+cvt", 128);
    return 0;
}
