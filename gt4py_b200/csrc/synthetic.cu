// synthetic.cu -- b200_fill_synthetic: fill a float64 field (halo cells included) with counter-based values in
// GLOBAL coordinates, on the device.  Bench / test plumbing ("data: synthetic"): at BASELINE configs[4]'s size a
// rank's eight fields are 21.5 GB -- generated here in milliseconds instead of on the host and over PCIe.
#include "common.cuh"
#include "synthetic.cuh"

namespace b200 {
__global__ void __launch_bounds__(256)
fill_synthetic_kernel(double *__restrict__ p, int64_t si, int64_t sj, int64_t sk, int ni, int nj, int nk,
                      const double *__restrict__ base, int64_t bsi, int64_t bsj, uint64_t seed, int64_t gi0,
                      int64_t gj0, double lo, double hi)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= ni) return;
    const double b = base ? base[i * bsi + j * bsj] : 0.0;
    for (int k = blockIdx.z; k < nk; k += gridDim.z) {
        const double v = synth_value(gi0 + i, gj0 + j, k, seed, lo, hi);
        p[i * si + j * sj + k * sk] = base ? b + v : v;
    }
}
}  // namespace b200

extern "C" int b200_fill_synthetic(const b200_field *f, const b200_field *base2d, uint64_t seed, int64_t gi0, int64_t gj0,
                                   double lo, double hi, void *stream)
{
    using namespace b200;
    B200_REQUIRE(f && f->ptr, "fill_synthetic: field is null");
    B200_REQUIRE(f->ndim == 3 && f->dtype == B200_FLOAT64, "fill_synthetic: a 3-D float64 field is required");
    const double *bp = nullptr;
    int64_t bsi = 0, bsj = 0;
    if (base2d) {
        B200_REQUIRE(base2d->ptr && base2d->ndim == 2 && base2d->dtype == B200_FLOAT64 && base2d->shape[0] >= f->shape[0] &&
                         base2d->shape[1] >= f->shape[1],
                     "fill_synthetic: base2d must be a 2-D float64 field covering the I x J extent");
        bp = static_cast<const double *>(base2d->ptr);
        bsi = base2d->stride[0];
        bsj = base2d->stride[1];
    }
    const int ni = (int)f->shape[0], nj = (int)f->shape[1], nk = (int)f->shape[2];
    if (ni <= 0 || nj <= 0 || nk <= 0) return B200_OK;
    dim3 grid((ni + 255) / 256, nj, nk < 8 ? nk : 8);
    fill_synthetic_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(static_cast<double *>(f->ptr), f->stride[0], f->stride[1],
                                                                  f->stride[2], ni, nj, nk, bp, bsi, bsj, seed, gi0, gj0, lo, hi);
    B200_CUDA_CHECK(cudaGetLastError());
    return B200_OK;
}
