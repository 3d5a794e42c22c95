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

// Everything that is serial here is kept short: the diagonal block is factorised by the whole CTA (one element per thread
// and step, three block barriers per column), its inverse W = L_kk^-1 is formed once (one column per lane, in registers)
// and stored in the unused strict upper triangle of the block, and from then on the panel below the block (X = A W^T), and
// both substitutions at the end (z_k = W_k t_k), are products with no dependent chain of loads or divisions.  The first
// version solved every row of a panel and every step of the substitutions by forward substitution against L_kk (32
// dependent steps each, the substitutions with a trip to L2 per step): 0.79 ms at p = 257 against 28 grid barriers' worth
// (~0.1 ms) of unavoidable latency.
__global__ void __launch_bounds__(CH_T)
k_chol_solve(const CholArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double sb[CH_B][CH_B + 1];      // lower triangle: L_kk; strict upper triangle: W^T (sb[m][c] = W[c][m], m < c)
  __shared__ double sorig[CH_B];
  __shared__ int sbad;
  const int p = a.p;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gthreads = gridDim.x * CH_T, gtid = blockIdx.x * CH_T + tid;
  for (int e = gtid; e < p * p; e += gthreads) {
    const int i = e / p, j = e - i * p;
    a.l[e] = (j <= i) ? a.h[e] : 0.0;
  }
  if (gtid == 0) *a.flag = 0;
  grid.sync();
  for (int k0 = 0; k0 < p; k0 += CH_B) {
    const int nb = (p - k0 < CH_B) ? (p - k0) : CH_B;
    if (blockIdx.x == 0) {
      // ---- diagonal block: factorise, invert, store both
      for (int e = tid; e < CH_B * CH_B; e += CH_T) {
        const int r = e >> 5, c = e & 31;
        sb[r][c] = (r < nb && c <= r) ? a.l[(size_t)(k0 + r) * p + k0 + c] : 0.0;
      }
      if (tid < nb) sorig[tid] = a.h[(size_t)(k0 + tid) * p + k0 + tid];
      if (tid == 0) sbad = 0;
      __syncthreads();
      for (int j = 0; j < nb; ++j) {
        const double djj = sb[j][j];
        // a pivot that cancelled down to rounding noise (exactly singular H: LU would meet a zero pivot) is not
        // trusted either: the pivoted inverse decides such cases.  Uniform over the CTA.
        if (!(djj > 1e-13 * fabs(sorig[j]))) { if (tid == 0) sbad = j + 1; break; }
        const double r = sqrt(djj);
        __syncthreads();
        if (tid == j) sb[j][j] = r;
        if (tid > j && tid < nb) sb[tid][j] /= r;
        __syncthreads();
        for (int e = tid; e < CH_B * CH_B; e += CH_T) {
          const int i = e >> 5, c = e & 31;
          if (c > j && c <= i && i < nb) sb[i][c] -= sb[i][j] * sb[c][j];
        }
        __syncthreads();
      }
      __syncthreads();
      if (sbad == 0) {
        if (warp == 0 && lane < nb) {
          // column `lane` of W = L_kk^-1 by forward substitution, kept in registers
          double w[CH_B];
#pragma unroll
          for (int i = 0; i < CH_B; ++i) {
            double v = 0.0;
            if (i < nb && i >= lane) {
              if (i == lane) v = 1.0 / sb[i][i];
              else {
                double acc = 0.0;
#pragma unroll
                for (int m = 0; m < i; ++m) if (m >= lane) acc = fma(sb[i][m], w[m], acc);
                v = -acc / sb[i][i];
              }
            }
            w[i] = v;
          }
#pragma unroll
          for (int i = 0; i < CH_B; ++i) if (i > lane && i < nb) sb[lane][i] = w[i];
        }
        __syncthreads();
        for (int e = tid; e < nb * nb; e += CH_T) {
          const int r = e / nb, c = e - r * nb;
          a.l[(size_t)(k0 + r) * p + k0 + c] = sb[r][c];
        }
      } else if (tid == 0) {
        *a.flag = k0 + sbad;
      }
    }
    __threadfence();
    grid.sync();
    if (*(volatile int*)a.flag != 0) return;                   // not positive definite: the caller's inverse path runs
    if (k0 + nb >= p) break;
    // ---- panel below the block (nb == 32 here): X = A W^T, one warp per row, lane = column
    for (int e = tid; e < CH_B * CH_B; e += CH_T) {
      const int r = e >> 5, c = e & 31;
      sb[r][c] = a.l[(size_t)(k0 + r) * p + k0 + c];
    }
    __syncthreads();
    {
      const double wdiag = 1.0 / sb[lane][lane];
      const int gwarps = gridDim.x * (CH_T / 32);
      for (int i = k0 + CH_B + blockIdx.x * (CH_T / 32) + warp; i < p; i += gwarps) {
        double* row = a.l + (size_t)i * p + k0;
        const double av = row[lane];
        double acc = 0.0;
#pragma unroll
        for (int m = 0; m < CH_B; ++m) {
          const double am = __shfl_sync(0xffffffffu, av, m);
          if (m < lane) acc = fma(am, sb[m][lane], acc);
        }
        row[lane] = fma(av, wdiag, acc);
      }
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
  // ---- L z = g, L^T delta = z by CTA 0: per block a product with W_k (warp 0), then the update of the rows it feeds
  for (int i = tid; i < p; i += CH_T) a.z[i] = a.g[i];
  __syncthreads();
  for (int k0 = 0; k0 < p; k0 += CH_B) {
    const int nb = (p - k0 < CH_B) ? (p - k0) : CH_B;
    if (tid < 32) {
      // z_k = W_k t: lane r needs W[r][m], m <= r: column r of the block's upper storage, the diagonal inverted
      double wr[CH_B];
#pragma unroll
      for (int m = 0; m < CH_B; ++m)
        wr[m] = (lane < nb && m <= lane) ? a.l[(size_t)(k0 + m) * p + k0 + lane] : 0.0;
      const double tv = lane < nb ? a.z[k0 + lane] : 0.0;
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < CH_B; ++m) {
        const double tm = __shfl_sync(0xffffffffu, tv, m);
        if (m < lane) acc = fma(wr[m], tm, acc);
      }
      if (lane < nb) a.z[k0 + lane] = acc + tv / wr[lane];
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
      // delta_k = W_k^T z_k: lane c needs W[m][c], m >= c: row c of the block's upper storage (contiguous), diagonal inverted
      double wc[CH_B];
#pragma unroll
      for (int m = 0; m < CH_B; ++m)
        wc[m] = (lane < nb && m < nb && m >= lane) ? a.l[(size_t)(k0 + lane) * p + k0 + m] : 0.0;
      const double zv = lane < nb ? a.z[k0 + lane] : 0.0;
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < CH_B; ++m) {
        const double zm = __shfl_sync(0xffffffffu, zv, m);
        if (m > lane) acc = fma(wc[m], zm, acc);
      }
      if (lane < nb) a.z[k0 + lane] = acc + zv / wc[lane];
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
