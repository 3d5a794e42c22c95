// Small dense factorisations behind KernelCls.cholesky / qr / svd (nums/core/kernel/numpy_kernel.py:455-474), the
// d x d ends of the tall-skinny linear algebra of nums/core/linalg.py.  The n x d work of
// those routines (Gram matrices, X @ R^-1) runs on the GEMM kernels of matmul.cu / glm_fused.cu; what is here is the
// latency-bound p x p part, p <= 1024, kept on the device so a QR or SVD never leaves HBM:
//   k_cholesky   lower Cholesky factor (left-looking, one warp per row dot, fixed summation order)
//   k_qr_signs   the signs LAPACK's Householder QR would give the diagonal of R, recovered from the top p x p block of
//                an orthonormal factor by the sign-choosing LU of Householder reconstruction (Ballard et al. 2014)
//   k_jacobi_svd one-sided Jacobi (Hestenes) with round-robin pair ordering, one warp per row pair, cooperative grid
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace nb {

constexpr int FACT_MAX_P = 1024;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- Cholesky: a = l l^T, l lower triangular (np.linalg.cholesky reads the lower triangle) -------------
__global__ void __launch_bounds__(1024)
k_cholesky(int p, const double* __restrict__ a, double* __restrict__ l, int32_t* __restrict__ info) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int e = tid; e < p * p; e += 1024) {
    const int i = e / p, j = e - i * p;
    l[e] = (j <= i) ? a[e] : 0.0;
  }
  __syncthreads();
  for (int j = 0; j < p; ++j) {
    const double* lj = l + (size_t)j * p;
    for (int i = j + warp; i < p; i += 32) {
      const double* li = l + (size_t)i * p;
      double s = 0.0;
      for (int k = lane; k < j; k += 32) s = fma(li[k], lj[k], s);
      s = warp_sum(s);
      if (lane == 0) l[(size_t)i * p + j] = li[j] - s;
    }
    __syncthreads();
    const double d = l[(size_t)j * p + j];
    if (!(d > 0.0)) {                    // also catches NaN
      if (tid == 0) *info = j + 1;
      return;
    }
    const double r = sqrt(d);
    __syncthreads();
    for (int i = j + 1 + tid; i < p; i += 1024) l[(size_t)i * p + j] /= r;
    if (tid == 0) l[(size_t)j * p + j] = r;
    __syncthreads();
  }
  if (tid == 0) *info = 0;
}

// ---- Householder signs -----------------------------------------------------------------------------------
// LAPACK's dgeqrf gives R(j,j) = -sign(alpha_j) * norm, alpha_j the diagonal entry the j-th reflector sees.  For an
// orthonormal factor Q+ with R+ > 0 on the diagonal (what a Cholesky QR produces), the Householder factor is Q+ S and
// S is fixed by an LU of Q+(0:p, 0:p) that picks S(i,i) = -sgn(current pivot) and eliminates with pivot - S(i,i)
// (|pivot - S| >= 1, so no row exchanges are needed).  A square input ends with the identity reflector: its last
// sign keeps R(p,p) = alpha_p.
__global__ void __launch_bounds__(1024)
k_qr_signs(int p, const double* __restrict__ q, int64_t ldq, int square, double* __restrict__ w,
           double* __restrict__ signs) {
  const int tid = threadIdx.x;
  for (int e = tid; e < p * p; e += 1024) {
    const int i = e / p, j = e - i * p;
    w[e] = q[(size_t)i * ldq + j];
  }
  for (int i = 0; i < p; ++i) {
    __syncthreads();
    const double qi = w[(size_t)i * p + i];
    double s = (qi >= 0.0) ? -1.0 : 1.0;
    if (square && i == p - 1) s = -s;
    if (tid == 0) signs[i] = s;
    const double piv = qi - s;
    const int t = p - 1 - i;                       // trailing block is t x t
    const double* wi = w + (size_t)i * p;
    for (int e = tid; e < t * t; e += 1024) {
      const int r = i + 1 + e / t, c = i + 1 + e % t;
      const double m = w[(size_t)r * p + i] / piv;
      w[(size_t)r * p + c] = fma(-m, wi[c], w[(size_t)r * p + c]);
    }
  }
}

// ---- one-sided Jacobi SVD of a square matrix ---------------------------------------------------------------
// wt holds the columns of A as rows (wt = (A V)^T), vt the columns of V as rows (= V^T, the third output).  A
// round of the round-robin tournament pairs every row with one other; each warp orthogonalises one pair.
struct SvdArgs {
  int p;
  const double* a;     // p x p row-major input
  double* wt;          // scratch p x p
  double* vt_work;     // scratch p x p
  double* norms;       // scratch p
  int* rotated;        // scratch: [0] rotations in the current sweep (device-wide counter)
  double* u;
  double* s;
  double* vt;
  int32_t* info;       // sweeps used, or -1 when max sweeps hit
  int max_sweeps;
  double tol;
};

__global__ void __launch_bounds__(256)
k_jacobi_svd(SvdArgs a) {
  cg::grid_group grid = cg::this_grid();
  const int p = a.p;
  const int tid = threadIdx.x, lane = tid & 31;
  const int gwarp = blockIdx.x * 8 + (tid >> 5), nwarps = gridDim.x * 8;
  const int64_t gthreads = (int64_t)gridDim.x * 256, gtid = (int64_t)blockIdx.x * 256 + tid;
  for (int64_t e = gtid; e < (int64_t)p * p; e += gthreads) {
    const int i = (int)(e / p), j = (int)(e - (int64_t)i * p);
    a.wt[e] = a.a[(size_t)j * p + i];
    a.vt_work[e] = (i == j) ? 1.0 : 0.0;
  }
  if (gtid == 0) a.rotated[0] = 0;
  grid.sync();
  const int P = (p + 1) & ~1;
  const int pairs = P / 2;
  int sweeps = 0;
  bool converged = (p <= 1);
  while (!converged && sweeps < a.max_sweeps) {
    for (int round = 0; round < P - 1; ++round) {
      for (int k = gwarp; k < pairs; k += nwarps) {
        int i, j;
        if (k == 0) { i = P - 1; j = round; }
        else { i = (round + k) % (P - 1); j = (round - k + (P - 1)) % (P - 1); }
        if (i >= p || j >= p) continue;            // dummy player of an odd tournament
        if (i > j) { const int t = i; i = j; j = t; }
        double* wi = a.wt + (size_t)i * p;
        double* wj = a.wt + (size_t)j * p;
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int c = lane; c < p; c += 32) {
          const double x = wi[c], y = wj[c];
          al = fma(x, x, al); be = fma(y, y, be); ga = fma(x, y, ga);
        }
        al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
        if (fabs(ga) > a.tol * sqrt(al * be) && ga != 0.0) {
          const double zeta = (be - al) / (2.0 * ga);
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
          for (int c = lane; c < p; c += 32) {
            const double x = wi[c], y = wj[c];
            wi[c] = cs * x - sn * y;
            wj[c] = sn * x + cs * y;
          }
          double* vi = a.vt_work + (size_t)i * p;
          double* vj = a.vt_work + (size_t)j * p;
          for (int c = lane; c < p; c += 32) {
            const double x = vi[c], y = vj[c];
            vi[c] = cs * x - sn * y;
            vj[c] = sn * x + cs * y;
          }
          if (lane == 0) atomicAdd(a.rotated, 1);
        }
      }
      grid.sync();
    }
    ++sweeps;
    const int rot = *(volatile int*)a.rotated;
    grid.sync();
    if (gtid == 0) a.rotated[0] = 0;
    grid.sync();
    converged = (rot == 0);
  }
  // singular values = row norms of wt
  for (int r = gwarp; r < p; r += nwarps) {
    const double* wr = a.wt + (size_t)r * p;
    double n2 = 0.0;
    for (int c = lane; c < p; c += 32) n2 = fma(wr[c], wr[c], n2);
    n2 = warp_sum(n2);
    if (lane == 0) a.norms[r] = sqrt(n2);
  }
  grid.sync();
  // descending order (ties by original index), then scatter U columns / VT rows
  for (int r = gwarp; r < p; r += nwarps) {
    const double nr = a.norms[r];
    int rank = 0;
    for (int c = lane; c < p; c += 32) {
      const double nc = a.norms[c];
      rank += (nc > nr || (nc == nr && c < r)) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rank += __shfl_xor_sync(0xffffffffu, rank, o);
    const double* wr = a.wt + (size_t)r * p;
    const double* vr = a.vt_work + (size_t)r * p;
    const double inv = nr > 0.0 ? 1.0 / nr : 0.0;
    for (int c = lane; c < p; c += 32) {
      a.u[(size_t)c * p + rank] = wr[c] * inv;
      a.vt[(size_t)rank * p + c] = vr[c];
    }
    if (lane == 0) a.s[rank] = nr;
  }
  if (gtid == 0) *a.info = converged ? sweeps : -1;
}

static inline size_t fact_align(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace nb

using namespace nb;

extern "C" int nb_cholesky_f64(int device, int64_t p, const double* a_d, double* l_d, int32_t* info_d, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(p >= 1 && p <= FACT_MAX_P, "nb_cholesky_f64: p=%lld outside 1..%d", (long long)p, FACT_MAX_P);
  NB_REQUIRE(a_d != l_d, "nb_cholesky_f64: the factor cannot overwrite its input");
  k_cholesky<<<1, 1024, 0, st>>>((int)p, a_d, l_d, info_d);
  NB_LAUNCH_CHECK();
  return 0;
}

extern "C" int64_t nb_qr_signs_scratch_bytes(int64_t p) {
  return p > 0 ? (int64_t)fact_align((size_t)p * p * sizeof(double)) : 0;
}

extern "C" int nb_qr_signs_f64(int device, int64_t p, const double* qtop_d, int64_t ldq, int square, double* signs_d,
                               void* scratch_d, int64_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(p >= 1 && p <= FACT_MAX_P, "nb_qr_signs_f64: p=%lld outside 1..%d", (long long)p, FACT_MAX_P);
  NB_REQUIRE(ldq >= p, "nb_qr_signs_f64: ldq < p");
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_qr_signs_scratch_bytes(p), "nb_qr_signs_f64: scratch too small");
  k_qr_signs<<<1, 1024, 0, st>>>((int)p, qtop_d, ldq, square, (double*)scratch_d, signs_d);
  NB_LAUNCH_CHECK();
  return 0;
}

extern "C" int64_t nb_svd_scratch_bytes(int64_t p) {
  if (p <= 0) return 0;
  return (int64_t)(2 * fact_align((size_t)p * p * sizeof(double)) + fact_align((size_t)p * sizeof(double)) + 256);
}

extern "C" int nb_svd_f64(int device, int64_t p, const double* a_d, double* u_d, double* s_d, double* vt_d,
                          int32_t* info_d, void* scratch_d, int64_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(p >= 1 && p <= FACT_MAX_P, "nb_svd_f64: p=%lld outside 1..%d", (long long)p, FACT_MAX_P);
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_svd_scratch_bytes(p), "nb_svd_f64: scratch too small");
  SvdArgs a;
  a.p = (int)p;
  a.a = a_d;
  char* sp = (char*)scratch_d;
  a.wt = (double*)sp; sp += fact_align((size_t)p * p * sizeof(double));
  a.vt_work = (double*)sp; sp += fact_align((size_t)p * p * sizeof(double));
  a.norms = (double*)sp; sp += fact_align((size_t)p * sizeof(double));
  a.rotated = (int*)sp;
  a.u = u_d; a.s = s_d; a.vt = vt_d; a.info = info_d;
  a.max_sweeps = 60;
  a.tol = 2.220446049250313e-16 * sqrt((double)p);
  const int pairs = (int)((p + 1) / 2);
  int G = (pairs + 7) / 8;
  const int sms = dev_info(device).sms;
  if (G > sms) G = sms;
  if (G < 1) G = 1;
  void* args[] = {(void*)&a};
  NB_CHECK(cudaLaunchCooperativeKernel((void*)k_jacobi_svd, dim3(G), dim3(256), args, 0, st));
  return 0;
}
