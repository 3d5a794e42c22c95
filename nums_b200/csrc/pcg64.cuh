// PCG64 (128-bit setseq LCG + XSL-RR output) pieces shared by the block-RNG kernels.  NumPy's PCG64 bit generator --
// the stream the reference draws every block from (nums/core/kernel/numpy_kernel.py:31-32) -- is this generator;
// word k (1-based) of a stream is xsl_rr(state after k steps).
#pragma once
#include "common.cuh"

namespace nb {

typedef unsigned __int128 u128;

__host__ __device__ __forceinline__ u128 make128(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | (u128)lo; }

#define NB_PCG_MULT make128(0x2360ED051FC65DA4ull, 0x4385DF649FCCF645ull)   // PCG_DEFAULT_MULTIPLIER_128

// (mult, plus) such that state_{k} = mult * state_0 + plus  (pcg_advance_lcg_128)
__device__ __forceinline__ void lcg_advance(u128 a, u128 c, uint64_t delta, u128& mult, u128& plus) {
  u128 acc_m = 1, acc_p = 0, cur_m = a, cur_p = c;
  while (delta > 0) {
    if (delta & 1) { acc_m *= cur_m; acc_p = acc_p * cur_m + cur_p; }
    cur_p = (cur_m + 1) * cur_p;
    cur_m *= cur_m;
    delta >>= 1;
  }
  mult = acc_m; plus = acc_p;
}

__device__ __forceinline__ uint64_t xsl_rr(u128 s) {
  const uint64_t hi = (uint64_t)(s >> 64), lo = (uint64_t)s;
  const uint64_t x = hi ^ lo;
  const unsigned rot = (unsigned)(hi >> 58);
  return (x >> rot) | (x << ((64u - rot) & 63u));
}

}  // namespace nb
