// Block RNG on the device, bit-compatible with the reference's host stream for uniform doubles:
// nums/core/kernel/numpy_kernel.py:31-32,103-110 draws block j from Generator(PCG64(seed).jumped(j)), and
// Generator.random() is (next_uint64 >> 11) * 2^-53 with PCG64 = 128-bit LCG (setseq) + XSL-RR output.
// Seeding and the per-block jump stay on the host (NumPy's SeedSequence); the host hands over the block's
// 128-bit (state, inc) and every thread jumps to its own position with the O(log k) LCG advance, then
// strides through the sequence with the grid-wide jump (a^T, c_T): element i is produced by exactly the
// i-th step of the stream, whatever the launch geometry.
#include "pcg64.cuh"

namespace nb {

__global__ void __launch_bounds__(256)
k_pcg64_random(int64_t n, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo, double* __restrict__ out) {
  const u128 a = NB_PCG_MULT;
  const u128 inc = make128(i_hi, i_lo);
  const uint64_t T = (uint64_t)gridDim.x * 256;
  const uint64_t t = (uint64_t)blockIdx.x * 256 + threadIdx.x;
  u128 m, p;
  lcg_advance(a, inc, t + 1, m, p);                 // element t is the output of step t + 1
  u128 s = make128(s_hi, s_lo) * m + p;
  u128 mt, pt;
  lcg_advance(a, inc, T, mt, pt);
  for (uint64_t i = t; i < (uint64_t)n; i += T) {
    out[i] = (double)(xsl_rr(s) >> 11) * (1.0 / 9007199254740992.0);
    s = s * mt + pt;
  }
}

}  // namespace nb

using namespace nb;

extern "C" int nb_pcg64_random(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                               double* out_d, void* stream) {
  NB_ENTER(device);
  if (n <= 0) return 0;
  int grid = grid_for(device, n, 256, 4);
  k_pcg64_random<<<grid, 256, 0, st>>>(n, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1], out_d);
  NB_LAUNCH_CHECK();
  return 0;
}
