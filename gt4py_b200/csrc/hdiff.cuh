// hdiff.cuh -- launch entry points of the horizontal-diffusion kernels living outside hdiff.cu.
#pragma once
#include "common.cuh"

namespace b200 {

// hdiff_tma.cu: TMA-ring kernel (variant 5).  *handled = false with B200_OK: preconditions not met.
int launch_hdiff_tma(bool limiter, const View3<const double> &in, const View3<double> &out,
                     const View3<const double> &co, int ni, int nj, int nk, cudaStream_t stream, bool *handled);

}  // namespace b200
