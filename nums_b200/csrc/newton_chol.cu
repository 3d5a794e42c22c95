// The Newton step of glms.py:660-676, beta - H^-1 g, by a blocked Cholesky solve instead of the explicit inverse.
//
// The reference forms inv(H) @ g (numpy_kernel.py:471, LAPACK getrf/getri); for a GLM Hessian with an L2 term H is
// symmetric positive definite, so the same step is H = L L^T, L z = g, L^T delta = z -- a third of the flops and,
// what matters at p = 257, 9 panel steps of 3 grid barriers instead of 257 pivoted columns of one barrier each
// (k_inv: 1.47 ms of a 34 ms step).  One cooperative launch:
//   per 32-column panel: CTA 0 factorises the diagonal block in shared memory (one warp) | every thread solves one row
//   of the panel against it | the trailing lower triangle takes its rank-32 update;
//   then CTA 0 runs the two blocked substitutions and writes beta_out.
// If a pivot is not safely positive (H not positive definite: unpenalised separable data, a singular design) the kernel sets
// *flag and leaves everything else untouched; the caller's pivoted inverse path is gated on that flag, so the result
// is then exactly the old one.  Sums run in a fixed order: deterministic.
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace nb {

struct CholArgs {
  int p;
  const double* h;       // p x p, symmetric (lower triangle read)
  double* l;             // p x p scratch: the factor
  const double* g;
  const double* beta;
  double* beta_out;
  double* z;             // p doubles of scratch
  int* flag;             // 0: solved; k + 1: pivot k not positive
  int32_t* info;         // set to 0 on success (the inverse path sets it otherwise)
};

constexpr int CH_T = 256;
constexpr int CH_B = 32;

__global__ void __launch_bounds__(CH_T)
k_chol_solve(const CholArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double sb[CH_B][CH_B + 1];
  const int p = a.p;
  const int tid = threadIdx.x, lane = tid & 31;
  const int gthreads = gridDim.x * CH_T, gtid = blockIdx.x * CH_T + tid;
  for (int e = gtid; e < p * p; e += gthreads) {
    const int i = e / p, j = e - i * p;
    a.l[e] = (j <= i) ? a.h[e] : 0.0;
  }
  if (gtid == 0) *a.flag = 0;
  grid.sync();
  for (int k0 = 0; k0 < p; k0 += CH_B) {
    const int nb = (p - k0 < CH_B) ? (p - k0) : CH_B;
    // ---- diagonal block: CTA 0, one warp, lane = row
    if (blockIdx.x == 0) {
      for (int e = tid; e < nb * nb; e += CH_T) {
        const int r = e / nb, c = e - r * nb;
        sb[r][c] = a.l[(size_t)(k0 + r) * p + k0 + c];
      }
      __syncthreads();
      if (tid < 32) {
        int bad = 0;
        for (int j = 0; j < nb; ++j) {
          const double djj = sb[j][j];
          // a pivot that cancelled down to rounding noise (exactly singular H: LU would meet a zero pivot) is not
          // trusted either: the pivoted inverse decides such cases, as it did before.  Uniform over the warp.
          const double orig = a.h[(size_t)(k0 + j) * p + k0 + j];
          if (!(djj > 1e-13 * fabs(orig))) { bad = j + 1; break; }
          const double r = sqrt(djj);
          __syncwarp();
          if (lane == j) sb[j][j] = r;
          if (lane > j && lane < nb) sb[lane][j] /= r;
          __syncwarp();
          if (lane > j && lane < nb) {
            const double lij = sb[lane][j];
            for (int c = j + 1; c <= lane; ++c) sb[lane][c] -= lij * sb[c][j];
          }
          __syncwarp();
        }
        if (bad && lane == 0) *a.flag = k0 + bad;
      }
      __syncthreads();
      for (int e = tid; e < nb * nb; e += CH_T) {
        const int r = e / nb, c = e - r * nb;
        if (c <= r) a.l[(size_t)(k0 + r) * p + k0 + c] = sb[r][c];
      }
    }
    __threadfence();
    grid.sync();
    if (*(volatile int*)a.flag != 0) return;                   // not positive definite: the caller's inverse path runs
    if (k0 + nb >= p) break;
    // ---- panel below the block: one row per thread, x L_kk^T = a_row  (nb == 32 here)
    for (int e = tid; e < CH_B * CH_B; e += CH_T) {
      const int r = e / CH_B, c = e - r * CH_B;
      sb[r][c] = a.l[(size_t)(k0 + r) * p + k0 + c];
    }
    __syncthreads();
    for (int i = k0 + CH_B + gtid; i < p; i += gthreads) {
      double* row = a.l + (size_t)i * p + k0;
      double x[CH_B];
#pragma unroll
      for (int c = 0; c < CH_B; ++c) x[c] = row[c];
#pragma unroll
      for (int c = 0; c < CH_B; ++c) {
        double s = x[c];
#pragma unroll
        for (int m = 0; m < c; ++m) s -= x[m] * sb[c][m];
        x[c] = s / sb[c][c];
      }
#pragma unroll
      for (int c = 0; c < CH_B; ++c) row[c] = x[c];
    }
    __threadfence();
    grid.sync();
    // ---- trailing lower triangle: a_ij -= sum_m l_im l_jm over the panel
    const int t = p - k0 - CH_B;
    for (int e = gtid; e < t * t; e += gthreads) {
      const int r = e / t, c = e - r * t;
      if (c > r) continue;
      const double* li = a.l + (size_t)(k0 + CH_B + r) * p + k0;
      const double* lj = a.l + (size_t)(k0 + CH_B + c) * p + k0;
      double s = 0.0;
#pragma unroll
      for (int m = 0; m < CH_B; ++m) s = fma(li[m], lj[m], s);
      a.l[(size_t)(k0 + CH_B + r) * p + k0 + CH_B + c] -= s;
    }
    __threadfence();
    grid.sync();
  }
  if (blockIdx.x != 0) return;
  // ---- L z = g, L^T delta = z: blocked substitutions by CTA 0
  for (int i = tid; i < p; i += CH_T) a.z[i] = a.g[i];
  __syncthreads();
  for (int k0 = 0; k0 < p; k0 += CH_B) {
    const int nb = (p - k0 < CH_B) ? (p - k0) : CH_B;
    if (tid < 32) {
      double v = lane < nb ? a.z[k0 + lane] : 0.0;
      for (int j = 0; j < nb; ++j) {
        if (lane == j) v /= a.l[(size_t)(k0 + j) * p + k0 + j];
        const double zj = __shfl_sync(0xffffffffu, v, j);
        if (lane > j && lane < nb) v -= a.l[(size_t)(k0 + lane) * p + k0 + j] * zj;
      }
      if (lane < nb) a.z[k0 + lane] = v;
    }
    __syncthreads();
    for (int i = k0 + nb + tid; i < p; i += CH_T) {
      const double* li = a.l + (size_t)i * p + k0;
      double s = 0.0;
      for (int m = 0; m < nb; ++m) s = fma(li[m], a.z[k0 + m], s);
      a.z[i] -= s;
    }
    __syncthreads();
  }
  for (int k0 = ((p - 1) / CH_B) * CH_B; k0 >= 0; k0 -= CH_B) {
    const int nb = (p - k0 < CH_B) ? (p - k0) : CH_B;
    if (tid < 32) {
      double v = lane < nb ? a.z[k0 + lane] : 0.0;
      for (int j = nb - 1; j >= 0; --j) {
        if (lane == j) v /= a.l[(size_t)(k0 + j) * p + k0 + j];
        const double dj = __shfl_sync(0xffffffffu, v, j);
        if (lane < j) v -= a.l[(size_t)(k0 + j) * p + k0 + lane] * dj;
      }
      if (lane < nb) a.z[k0 + lane] = v;
    }
    __syncthreads();
    for (int i = tid; i < k0; i += CH_T) {
      double s = 0.0;
      for (int m = 0; m < nb; ++m) s = fma(a.l[(size_t)(k0 + m) * p + i], a.z[k0 + m], s);
      a.z[i] -= s;
    }
    __syncthreads();
  }
  for (int i = tid; i < p; i += CH_T) a.beta_out[i] = a.beta[i] - a.z[i];
  if (tid == 0) *a.info = 0;
}

// scratch: l (p*p doubles, may alias the inverse path's output buffer), z (p doubles), flag (one int)
int newton_chol_solve(int device, int64_t p, const double* h, double* l, const double* g, const double* beta,
                      double* beta_out, double* z, int* flag, int32_t* info, cudaStream_t st) {
  CholArgs a;
  a.p = (int)p; a.h = h; a.l = l; a.g = g; a.beta = beta; a.beta_out = beta_out; a.z = z; a.flag = flag; a.info = info;
  int G = dev_info(device).sms < 32 ? dev_info(device).sms : 32;       // few CTAs: the barriers are what costs
  if ((int64_t)G * CH_T > p * p) G = (int)((p * p + CH_T - 1) / CH_T);
  if (G < 1) G = 1;
  void* args[] = {(void*)&a};
  NB_CHECK(cudaLaunchCooperativeKernel((void*)k_chol_solve, dim3(G), dim3(CH_T), args, 0, st));
  return 0;
}

}  // namespace nb
