// synthetic.cuh -- counter-based synthetic field values (bench / test inputs), usable from host and device code.
//
// value(gi, gj, k) = base2d(i, j) + lo + (hi - lo) * u,  u = (mix(gi, gj, k, seed) >> 11) * 2^-53  in [0, 1)
// with GLOBAL coordinates (gi, gj): every rank of an IJ decomposition -- and the CPU oracle, through the numpy
// restatement oracle/inputs.py:hashed_uniform -- computes the same value for the same global point, so a
// decomposed run can be compared with a single-domain one without ever materialising the global field.
// Integer arithmetic is modulo 2^64 (splitmix64 finaliser); the floating-point part is one multiply and one add,
// each correctly rounded (no FMA: the library is built with -fmad=false), hence bit-identical on both sides.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif

namespace b200 {
B200_HD uint64_t synth_mix(int64_t gi, int64_t gj, int64_t k, uint64_t seed)
{
    uint64_t x = seed + (uint64_t)(gi + (1 << 20)) * 0x9E3779B97F4A7C15ull + (uint64_t)(gj + (1 << 20)) * 0xC2B2AE3D27D4EB4Full +
                 (uint64_t)(k + (1 << 20)) * 0x165667B19E3779F9ull;
    x ^= x >> 30;
    x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}
B200_HD double synth_value(int64_t gi, int64_t gj, int64_t k, uint64_t seed, double lo, double hi)
{
    const double u = (double)(synth_mix(gi, gj, k, seed) >> 11) * (1.0 / 9007199254740992.0);
    const double span = hi - lo;
    const double t = span * u;
    return lo + t;
}
}  // namespace b200
