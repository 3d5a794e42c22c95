// Dense inverse (KernelCls.inv, nums/core/kernel/numpy_kernel.py:471-472 -> np.linalg.inv) and the
// Newton update of nums/models/glms.py:660-676, which calls it through linalg.inv (linalg.py:246-261).
//
// In-place Gauss-Jordan with partial pivoting.  Row swaps are virtual (a logical->physical row map
// kept by every CTA), the scaled pivot row is written back one step late by its owner, and the pivot
// search for column j+1 is fused into the elimination of column j, so one grid-wide barrier per
// column suffices.  p <= 32 runs in one CTA; larger systems use a cooperative launch with rows dealt
// round-robin to one CTA per SM.
#include "common.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>
namespace cg = cooperative_groups;

namespace nb {

constexpr int INV_THREADS = 256;
constexpr int64_t INV_MAX_P = 4096;

struct InvArgs {
  int64_t p;
  double* a;        // p x p working copy (scratch), destroyed
  double* out;      // p x p result
  double* slot_v;   // [2][G] pivot candidates: |value|
  int* slot_r;      // [2][G] pivot candidates: logical row
  int32_t* info;
  const int* gate;  // null: always run; else run only when *gate != 0 (fallback of the Cholesky Newton step)
};

template <bool COOP>
__device__ __forceinline__ void inv_barrier() {
  if constexpr (COOP) cg::this_grid().sync();
  else __syncthreads();
}

// candidate ordering: larger |v| wins, NaN loses, ties -> lower logical row
__device__ __forceinline__ bool piv_better(double va, int ra, double vb, int rb) {
  if (rb < 0) return ra >= 0;
  if (ra < 0) return false;
  if (va > vb) return true;
  if (va == vb) return ra < rb;
  return false;
}

__device__ __forceinline__ void piv_block_reduce(double& v, int& r, double* smv, int* smr) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double ov = __shfl_down_sync(0xffffffffu, v, o);
    int orr = __shfl_down_sync(0xffffffffu, r, o);
    if (piv_better(ov, orr, v, r)) { v = ov; r = orr; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) { smv[warp] = v; smr[warp] = r; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < INV_THREADS / 32; w++)
      if (piv_better(smv[w], smr[w], v, r)) { v = smv[w]; r = smr[w]; }
    smv[0] = v; smr[0] = r;
  }
  __syncthreads();
  v = smv[0]; r = smr[0];
  __syncthreads();
}

template <bool COOP>
__global__ void __launch_bounds__(INV_THREADS)
k_inv(const InvArgs a) {
  extern __shared__ __align__(16) unsigned char inv_smem[];
  if (a.gate != nullptr && *a.gate == 0) return;        // uniform over the grid, before any barrier
  const int p = (int)a.p;
  const int G = COOP ? (int)gridDim.x : 1;
  const int b = COOP ? (int)blockIdx.x : 0;
  const int tid = threadIdx.x;
  const int own_max = (p + G - 1) / G;
  double* prow = (double*)inv_smem;          // p   current pivot row (unscaled)
  double* keep = prow + p;                   // p   scaled pivot row awaiting write-back
  double* fcol = keep + p;                   // own_max  multipliers of the owned rows
  int* rowmap = (int*)(fcol + own_max + (own_max & 1));   // logical -> physical
  int* logof = rowmap + p;                   // physical -> logical
  int* perm = logof + p;                     // recorded swaps
  __shared__ double smv[INV_THREADS / 32];
  __shared__ int smr[INV_THREADS / 32];
  __shared__ int s_keep_row;

  for (int i = tid; i < p; i += INV_THREADS) { rowmap[i] = i; logof[i] = i; }
  if (tid == 0) s_keep_row = -1;
  __syncthreads();

  // initial pivot search in column 0
  {
    double bv = 0.0; int br = -1;
    for (int r = b + tid * G; r < p; r += G * INV_THREADS) {
      double v = fabs(a.a[(size_t)r * p]);
      if (piv_better(v, r, bv, br)) { bv = v; br = r; }
    }
    piv_block_reduce(bv, br, smv, smr);
    if (tid == 0) { a.slot_v[b] = bv; a.slot_r[b] = br; }
  }
  inv_barrier<COOP>();

  bool singular = false;
  for (int j = 0; j < p; j++) {
    const int par = j & 1;
    // deferred write-back of the previous pivot row (only its owner has one pending)
    if (s_keep_row >= 0) {
      double* dst = a.a + (size_t)s_keep_row * p;
      for (int c = tid; c < p; c += INV_THREADS) dst[c] = keep[c];
    }
    // every CTA picks the same pivot from the published candidates
    double bv = 0.0; int br = -1;
    for (int q = tid; q < G; q += INV_THREADS) {
      double v = a.slot_v[par * G + q]; int r = a.slot_r[par * G + q];
      if (piv_better(v, r, bv, br)) { bv = v; br = r; }
    }
    piv_block_reduce(bv, br, smv, smr);
    if (br < 0 || !(bv > 0.0)) {   // exactly singular (or NaN column)
      if (b == 0 && tid == 0) *a.info = j + 1;
      singular = true;
      break;
    }
    if (tid == 0) {
      perm[j] = br;
      int pj = rowmap[j], pb = rowmap[br];
      rowmap[j] = pb; rowmap[br] = pj;
      logof[pb] = j; logof[pj] = br;
      s_keep_row = -1;
    }
    __syncthreads();
    const int pr = rowmap[j];
    const double* src = a.a + (size_t)pr * p;
    for (int c = tid; c < p; c += INV_THREADS) prow[c] = src[c];
    // multipliers of the owned rows (read before column j is overwritten)
    for (int s = tid; s < own_max; s += INV_THREADS) {
      int r = b + s * G;
      fcol[s] = r < p ? a.a[(size_t)r * p + j] : 0.0;
    }
    __syncthreads();
    const double dinv = 1.0 / prow[j];
    if ((pr % G) == b) {
      for (int c = tid; c < p; c += INV_THREADS) keep[c] = (c == j) ? dinv : prow[c] * dinv;
      if (tid == 0) s_keep_row = pr;
    }
    // eliminate column j from the owned rows; track the best pivot candidate for column j+1
    double nv = 0.0; int nr = -1;
    const int64_t work = (int64_t)own_max * p;
    for (int64_t e = tid; e < work; e += INV_THREADS) {
      int s = (int)(e / p), c = (int)(e - (int64_t)s * p);
      int r = b + s * G;
      if (r >= p || r == pr) continue;
      double f = fcol[s];
      double* cell = a.a + (size_t)r * p + c;
      double v = (c == j) ? -f * dinv : *cell - f * (prow[c] * dinv);
      *cell = v;
      if (c == j + 1) {
        int lg = logof[r];
        if (lg > j) { double av = fabs(v); if (piv_better(av, lg, nv, nr)) { nv = av; nr = lg; } }
      }
    }
    piv_block_reduce(nv, nr, smv, smr);
    if (tid == 0) { a.slot_v[(par ^ 1) * G + b] = nv; a.slot_r[(par ^ 1) * G + b] = nr; }
    inv_barrier<COOP>();
  }
  if (singular) return;   // uniform across the grid: every CTA saw the same candidates
  if (s_keep_row >= 0) {
    double* dst = a.a + (size_t)s_keep_row * p;
    for (int c = tid; c < p; c += INV_THREADS) dst[c] = keep[c];
  }
  inv_barrier<COOP>();
  // undo the row swaps as column swaps, in reverse order; reuse logof as the source-column table
  if (tid == 0) {
    for (int c = 0; c < p; c++) logof[c] = c;
    for (int j = p - 1; j >= 0; j--) { int x = logof[j]; logof[j] = logof[perm[j]]; logof[perm[j]] = x; }
  }
  __syncthreads();
  const int64_t total = (int64_t)p * p;
  for (int64_t e = (int64_t)b * INV_THREADS + tid; e < total; e += (int64_t)G * INV_THREADS) {
    int i = (int)(e / p), c = (int)(e - (int64_t)i * p);
    a.out[e] = a.a[(size_t)rowmap[i] * p + logof[c]];
  }
  if (b == 0 && tid == 0) *a.info = 0;
}

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static size_t inv_scratch(int64_t p) {
  // working copy + pivot slots (2 x up to 148... use 2 x 1024 entries of double + int)
  return align256((size_t)p * p * sizeof(double)) + align256(2 * 1024 * sizeof(double)) + align256(2 * 1024 * sizeof(int));
}

static int inv_f64(int device, int64_t p, const double* a_in, double* out, int32_t* info, void* scratch,
                   size_t scratch_bytes, cudaStream_t st, const int* gate = nullptr) {
  NB_REQUIRE(p >= 1 && p <= INV_MAX_P, "inv: p=%lld outside 1..%lld", (long long)p, (long long)INV_MAX_P);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= inv_scratch(p), "inv: scratch too small (%zu < %zu)",
             scratch_bytes, inv_scratch(p));
  InvArgs a;
  a.p = p;
  a.a = (double*)scratch;
  a.slot_v = (double*)((char*)scratch + align256((size_t)p * p * sizeof(double)));
  a.slot_r = (int*)((char*)a.slot_v + align256(2 * 1024 * sizeof(double)));
  a.out = out;
  a.info = info;
  a.gate = gate;
  NB_CHECK(cudaMemcpyAsync(a.a, a_in, (size_t)p * p * sizeof(double), cudaMemcpyDeviceToDevice, st));
  const bool coop = p > 32;
  int G = 1;
  if (coop) {
    // measured at p = 257: 8 CTAs 4.9 ms, 33 CTAs 2.2 ms, 148 CTAs 1.5 ms — the per-column cost is the chain of
    // global round trips inside a CTA, not the grid barrier, so use every SM
    G = dev_info(device).sms;
    if (G > p) G = (int)p;
    if (G < 1) G = 1;
  }
  int own_max = (int)((p + G - 1) / G);
  size_t smem = (size_t)(2 * p + own_max + (own_max & 1)) * sizeof(double) + (size_t)3 * p * sizeof(int);
  if (coop) {
    NB_CHECK(cudaFuncSetAttribute(k_inv<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    void* args[] = {(void*)&a};
    NB_CHECK(cudaLaunchCooperativeKernel((void*)k_inv<true>, dim3(G), dim3(INV_THREADS), args, smem, st));
  } else {
    NB_CHECK(cudaFuncSetAttribute(k_inv<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_inv<false><<<1, INV_THREADS, smem, st>>>(a);
    NB_LAUNCH_CHECK();
  }
  return 0;
}

// g += l2vec*beta ; H += diag(l2vec) ; gmax = max|g| (NaN propagates like np.max)
__global__ void __launch_bounds__(256)
k_newton_prep(int64_t p, double* __restrict__ stats, const double* __restrict__ l2vec,
              const double* __restrict__ beta, double* __restrict__ gmax) {
  __shared__ double sm[8];
  double* H = stats;
  double* g = stats + p * p;
  double m = 0.0;
  bool has_nan = false;
  for (int64_t i = threadIdx.x; i < p; i += 256) {
    double l = l2vec ? l2vec[i] : 0.0;
    double gv = g[i];
    if (l2vec) { gv += l * beta[i]; g[i] = gv; H[i * p + i] += l; }
    double av = fabs(gv);
    if (isnan(av)) has_nan = true;
    else if (av > m) m = av;
  }
  if (has_nan) m = nan("");
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double ov = __shfl_down_sync(0xffffffffu, m, o);
    m = (isnan(m) || isnan(ov)) ? nan("") : (ov > m ? ov : m);
  }
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    double r = sm[0];
    for (int w = 1; w < 8; w++) r = (isnan(r) || isnan(sm[w])) ? nan("") : (sm[w] > r ? sm[w] : r);
    *gmax = r;
  }
}

// beta_out = beta - Hinv @ g, one warp per row
__global__ void __launch_bounds__(256)
k_newton_apply(int64_t p, const double* __restrict__ hinv, const double* __restrict__ g,
               const double* __restrict__ beta, double* __restrict__ beta_out, const int* __restrict__ gate) {
  if (gate != nullptr && *gate == 0) return;
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * 8;
  for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < p; i += warps) {
    double s = 0.0;
    for (int64_t c = lane; c < p; c += 32) s = fma(hinv[i * p + c], g[c], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) beta_out[i] = beta[i] - s;
  }
}

// The whole Newton update for p <= 32 in one warp (the README fit has p = 3): penalty, max|g|, the solve of
// H delta = g by Gaussian elimination with partial pivoting in shared memory, beta_out = beta - delta.  At this size the
// three-kernel path (prep, cooperative Cholesky with grid barriers, gated inverse + apply) is four launches and ~60 us
// of latency for a few hundred flops.  info = k + 1 when column k has no non-zero pivot (np.linalg.inv raises then).
__global__ void __launch_bounds__(32)
k_newton_small(int p, double* __restrict__ stats, const double* __restrict__ l2vec, const double* __restrict__ beta,
               double* __restrict__ beta_out, double* __restrict__ gmax, int32_t* __restrict__ info) {
  __shared__ double A[32][33];
  __shared__ double rhs[32];
  const int lane = threadIdx.x;
  const double* H = stats;
  const double* g = stats + (size_t)p * p;
  double gv = 0.0;
  if (lane < p) {
    const double l = l2vec ? l2vec[lane] : 0.0;
    gv = g[lane] + (l2vec ? l * beta[lane] : 0.0);
    for (int c = 0; c < p; c++) A[lane][c] = H[(size_t)lane * p + c] + (c == lane ? l : 0.0);
    rhs[lane] = gv;
  }
  // max |g| with NumPy's rule: a NaN propagates
  double m = lane < p ? fabs(gv) : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, m, o);
    m = (isnan(m) || isnan(ov)) ? nan("") : (ov > m ? ov : m);
  }
  __syncwarp();
  int bad = 0;
  for (int k = 0; k < p; k++) {
    // pivot: largest |A[r][k]|, r >= k (first such row on ties, like getrf)
    double best = (lane >= k && lane < p) ? fabs(A[lane][k]) : -1.0;
    int who = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int ow = __shfl_xor_sync(0xffffffffu, who, o);
      if (ob > best || (ob == best && ow < who)) { best = ob; who = ow; }
    }
    if (!(best > 0.0)) { bad = k + 1; break; }
    if (who != k) {               // swap rows k and who (lane = column)
      if (lane < p) { const double t = A[k][lane]; A[k][lane] = A[who][lane]; A[who][lane] = t; }
      if (lane == 0) { const double t = rhs[k]; rhs[k] = rhs[who]; rhs[who] = t; }
    }
    __syncwarp();
    const double piv = A[k][k];
    if (lane > k && lane < p) {
      const double f = A[lane][k] / piv;
      for (int c = k + 1; c < p; c++) A[lane][c] = fma(-f, A[k][c], A[lane][c]);
      rhs[lane] = fma(-f, rhs[k], rhs[lane]);
    }
    __syncwarp();
  }
  if (!bad) {
    // back substitution: lane 0 walks the rows (p <= 32)
    if (lane == 0) {
      for (int r = p - 1; r >= 0; r--) {
        double s = rhs[r];
        for (int c = r + 1; c < p; c++) s = fma(-A[r][c], rhs[c], s);
        rhs[r] = s / A[r][r];
      }
    }
    __syncwarp();
    if (lane < p) beta_out[lane] = beta[lane] - rhs[lane];
  }
  if (lane == 0) { *gmax = m; *info = bad; }
}

// newton_chol.cu
int newton_chol_solve(int device, int64_t p, const double* h, double* l, const double* g, const double* beta,
                      double* beta_out, double* z, int* flag, int32_t* info, cudaStream_t st);

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_inv_scratch_bytes(int64_t p) { return inv_scratch(p); }

int nb_inv_f64(int device, int64_t p, const double* a_d, double* out_d, int32_t* info_d, void* scratch_d,
               size_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  return inv_f64(device, p, a_d, out_d, info_d, scratch_d, scratch_bytes, st);
}

size_t nb_glm_newton_scratch_bytes(int64_t p) {
  return inv_scratch(p) + align256((size_t)p * p * sizeof(double)) + align256((size_t)p * sizeof(double)) + 256;
}

int nb_glm_newton_update(int device, int64_t p, double* stats_d, const double* l2vec_d, const double* beta_d,
                         double* beta_out_d, double* gmax_d, int32_t* info_d, void* scratch_d, size_t scratch_bytes,
                         void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_glm_newton_scratch_bytes(p),
             "nb_glm_newton_update: scratch too small (%zu < %zu)", scratch_bytes, nb_glm_newton_scratch_bytes(p));
  // scratch: [hinv / Cholesky factor p*p | z p | flag | inverse scratch]
  double* hinv = (double*)scratch_d;
  char* sp = (char*)scratch_d + align256((size_t)p * p * sizeof(double));
  double* z = (double*)sp; sp += align256((size_t)p * sizeof(double));
  int* flag = (int*)sp; sp += 256;
  void* inv_sc = sp;
  const size_t inv_bytes = scratch_bytes - (size_t)(sp - (char*)scratch_d);
  if (p <= 32 && !NB_ENV_FLAG("NUMS_B200_NEWTON_INVERSE")) {
    k_newton_small<<<1, 32, 0, st>>>((int)p, stats_d, l2vec_d, beta_d, beta_out_d, gmax_d, info_d);
    NB_LAUNCH_CHECK();
    return 0;
  }
  k_newton_prep<<<1, 256, 0, st>>>(p, stats_d, l2vec_d, beta_d, gmax_d);
  NB_LAUNCH_CHECK();
  const int* gate = nullptr;
  if (p >= 2 && !NB_ENV_FLAG("NUMS_B200_NEWTON_INVERSE")) {
    // H = L L^T, two substitutions (newton_chol.cu); the pivoted inverse below runs only if H is not positive definite
    int rc = newton_chol_solve(device, p, stats_d, hinv, stats_d + p * p, beta_d, beta_out_d, z, flag, info_d, st);
    if (rc) return rc;
    gate = flag;
  }
  int rc = inv_f64(device, p, stats_d, hinv, info_d, inv_sc, inv_bytes, st, gate);
  if (rc) return rc;
  int64_t blocks = (p + 7) / 8;
  k_newton_apply<<<(unsigned)(blocks < 1184 ? blocks : 1184), 256, 0, st>>>(p, hinv, stats_d + p * p, beta_d, beta_out_d,
                                                                           gate);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
