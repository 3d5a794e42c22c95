// The GLM inner loop of nums/models/glms.py (newton :660-676, LBFGS lbfgs.py:102-155) as device
// passes over one row block of X.
//
// Reference per row block and Newton iteration: ~19 NumPy calls, 5 reads + 1 write
// of X-sized data (X@beta, X^T r, s*X materialised, X^T(sX)).  Here:
//   d <= 7   one fused pass (k_glm_small): eta, link, objective, gradient and the full Hessian are
//            accumulated in registers per thread; X and y are read exactly once.
//   d >= 8   pass 1 (k_glm_rows, HBM-bound): eta, link, objective, g = [X,1]^T r, u = [X,1]^T w and the
//            weights w[n] (scratch);  pass 2 (gemm_f64, FP64 DMMA, compute-bound for d >~ 40):
//            G = X^T diag(w) X without materialising w*X.  H = [[G, u], [u^T, sum w]].
// The intercept column of GLM.fit (glms.py:124-135) is implicit: it is never stored.
// All cross-thread sums are fixed-order (per-CTA partials in scratch, then an ordered second stage).
#include "gemm.cuh"
#include "glm_common.cuh"
#include <stdlib.h>
#include <string.h>

namespace nb {

constexpr int GLM_THREADS = 256;

// ------------------------------------------------------------------------------------------------
// d <= 7: fully fused, one thread per row, everything in registers.
// Per-CTA partial layout: [H upper triangle row-major: P*(P+1)/2][g: P][obj: 1]
// ------------------------------------------------------------------------------------------------
template <int FAMILY, int D>
__global__ void __launch_bounds__(GLM_THREADS)
k_glm_small(const double* __restrict__ x, int64_t n, int64_t ldx, const void* __restrict__ y, int ydt,
            const double* __restrict__ beta, int want_hess, double* __restrict__ partials) {
  constexpr int P = D + 1;
  constexpr int NH = P * (P + 1) / 2;
  constexpr int NV = NH + P + 1;
  double b[P];
#pragma unroll
  for (int j = 0; j < P; j++) b[j] = beta[j];
  double acc[NV];
#pragma unroll
  for (int j = 0; j < NV; j++) acc[j] = 0.0;

  const int64_t stride = (int64_t)gridDim.x * GLM_THREADS;
  for (int64_t r = (int64_t)blockIdx.x * GLM_THREADS + threadIdx.x; r < n; r += stride) {
    double xv[P];
    if constexpr (D == 2) {
      if ((ldx & 1) == 0 && (((uintptr_t)x) & 15) == 0) {
        double2 v = *reinterpret_cast<const double2*>(x + r * ldx);
        xv[0] = v.x; xv[1] = v.y;
      } else {
        xv[0] = x[r * ldx]; xv[1] = x[r * ldx + 1];
      }
    } else {
#pragma unroll
      for (int j = 0; j < D; j++) xv[j] = x[r * ldx + j];
    }
    xv[D] = 1.0;
    double yv = load_y(y, ydt, r);
    double eta = b[D];
#pragma unroll
    for (int j = 0; j < D; j++) eta = fma(xv[j], b[j], eta);
    LinkOut lo = glm_link<FAMILY>(eta, yv);
    acc[NH + P] += lo.obj;
#pragma unroll
    for (int j = 0; j < P; j++) acc[NH + j] = fma(lo.resid, xv[j], acc[NH + j]);
    if (want_hess) {
      int q = 0;
#pragma unroll
      for (int i = 0; i < P; i++) {
        double wi = lo.wgt * xv[i];
#pragma unroll
        for (int j = i; j < P; j++) { acc[q] = fma(wi, xv[j], acc[q]); q++; }
      }
    }
  }
  // block reduction, fixed order: lanes (tree), then warps
  __shared__ double sm[GLM_THREADS / 32][NV];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < NV; j++) {
    double v = acc[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) sm[warp][j] = v;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < NV; j += GLM_THREADS) {
    double s = sm[0][j];
    for (int w = 1; w < GLM_THREADS / 32; w++) s += sm[w][j];
    partials[(size_t)blockIdx.x * NV + j] = s;
  }
}

// Second stage for k_glm_small: sums the CTA partials in a fixed order (one warp per statistic: a single thread walking
// ~600 partials was 50 us of a 10 us pass) and unpacks into the stats buffer.
__global__ void __launch_bounds__(256)
k_glm_small_finish(int P, int nparts, const double* __restrict__ partials, int want_hess, double* __restrict__ stats) {
  const int NH = P * (P + 1) / 2;
  const int NV = NH + P + 1;
  const int lane = threadIdx.x & 31;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < NV; j += gridDim.x * 8) {
    const double s = sum_partials_warp(partials + j, (size_t)NV, nparts);
    if (lane != 0) continue;
    if (j < NH) {
      if (want_hess) {
        int i = 0, rem = j;                       // j -> (i, c) of the row-major upper triangle
        while (rem >= P - i) { rem -= P - i; i++; }
        const int c = i + rem;
        stats[i * P + c] = s; stats[c * P + i] = s;
      }
    } else if (j < NH + P) {
      stats[P * P + (j - NH)] = s;
    } else {
      stats[P * P + P] = s;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// d >= 8, pass 1: eta / link / objective / g / u / weights.  HBM-bound: X is read once, straight
// into registers with 16-byte loads; each warp owns R rows x (64*NJ) columns per step.
// WPR warps share one row when d > 1024 (partial dot products meet in shared memory).
// ------------------------------------------------------------------------------------------------
struct GlmRowsArgs {
  const double* x; int64_t n, d, ldx;
  const void* y; int ydt;
  const double* beta;
  int family, want_hess;
  int skip_obj;        // Newton does not read the objective: leave out its two logarithms per row
  double* wbuf;        // n weights for pass 2 (null: not needed)
  double* partials;    // [gridDim.x][2*d + 3]
};

// R rows per warp step.  NJ = 4 (d <= 256) runs 2 rows x 2 CTAs per SM rather than 4 x 1: the same bytes
// in flight, but twice the warps to overlap the link's exp/log chain with the next loads.
// Without the Hessian (L-BFGS / gradient passes) the u accumulators are not needed, which buys a second row
// per step for the wide configurations: the link (~220 FP64 instructions, SIMT) is then paid once per 2-4
// rows instead of once per row.
template <int NJ, bool HESS>
struct RowsCfg {
  static constexpr int R = NJ <= 4 ? 4 : (NJ <= 8 ? (HESS ? 2 : 4) : (HESS ? 1 : 2));
  static constexpr int MIN_CTAS = NJ <= 4 ? 2 : 1;
};

template <int NJ, int WPR, bool VEC, bool HESS>
__global__ void __launch_bounds__(GLM_THREADS, RowsCfg<NJ, HESS>::MIN_CTAS)
k_glm_rows(const GlmRowsArgs a) {
  constexpr int R = RowsCfg<NJ, HESS>::R;
  constexpr int NE = 2 * NJ;
  constexpr int NW = GLM_THREADS / 32;
  constexpr int RG = NW / WPR;
  constexpr int CW = 64 * NJ;
  extern __shared__ __align__(16) double dyn[];
  double* s_beta = dyn;                         // WPR * CW: beta[0..d) then zeros (so no column guard is needed)
  double* s_acc = dyn + WPR * CW;               // 2 * WPR * CW  (g then u), then 4 scalars
  __shared__ double s_eta[2][RG][WPR][R];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rg = warp / WPR, wc = warp % WPR;
  const int64_t d = a.d, n = a.n;
  for (int j = tid; j < WPR * CW; j += GLM_THREADS) s_beta[j] = j < d ? a.beta[j] : 0.0;
  for (int j = tid; j < 2 * WPR * CW + 4; j += GLM_THREADS) s_acc[j] = 0.0;
  __syncthreads();

  // column owned by this lane for element slot e (compile-time e): cheap to recompute, so not kept in registers
  const int c0 = wc * CW + (VEC ? 2 * lane : lane);
#define COL(e) (c0 + (VEC ? (64 * ((e) >> 1) + ((e) & 1)) : 32 * (e)))
#define CV(e) ((int64_t)COL(e) < d)
  double bet[NJ <= 2 ? NE : 1];
  if constexpr (NJ <= 2) {
#pragma unroll
    for (int e = 0; e < NE; e++) bet[e] = CV(e) ? s_beta[COL(e)] : 0.0;
  }
  const double beta0 = a.beta[d];

  double g[NE], u[HESS ? NE : 1];
#pragma unroll
  for (int e = 0; e < NE; e++) g[e] = 0.0;
#pragma unroll
  for (int e = 0; e < (HESS ? NE : 1); e++) u[e] = 0.0;
  double gi = 0.0, ui = 0.0, ob = 0.0;

  const int64_t nbatches = (n + R - 1) / R;
  const int64_t per_it = (int64_t)gridDim.x * RG;
  const int64_t iters = (nbatches + per_it - 1) / per_it;
  int buf = 0;
  for (int64_t it = 0; it < iters; it++) {
    const int64_t batch = it * per_it + (int64_t)blockIdx.x * RG + rg;
    const int64_t row0 = batch * R;
    double xv[R][NE];
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int64_t row = row0 + r;
      const double* xr = a.x + row * a.ldx;
      if constexpr (VEC) {
#pragma unroll
        for (int j = 0; j < NJ; j++) {
          double2 v = make_double2(0.0, 0.0);
          if (row < n && CV(2 * j)) {
            if (CV(2 * j + 1)) v = *reinterpret_cast<const double2*>(xr + COL(2 * j));
            else v.x = xr[COL(2 * j)];
          }
          xv[r][2 * j] = v.x; xv[r][2 * j + 1] = v.y;
        }
      } else {
#pragma unroll
        for (int e = 0; e < NE; e++) xv[r][e] = (row < n && CV(e)) ? xr[COL(e)] : 0.0;
      }
    }
    // eta partials: beta comes from shared memory once per column slot and serves all R rows; four independent
    // sums per row keep the FMA chain short (the kernel runs at one or two warps per scheduler)
    double part[R];
    {
      constexpr int NP = R >= 4 ? (HESS ? 1 : 2) : 4;   // fewer partial sums where four rows already give the ILP
      double ps[R][NP];
#pragma unroll
      for (int r = 0; r < R; r++)
#pragma unroll
        for (int k = 0; k < NP; k++) ps[r][k] = 0.0;
      if constexpr (NJ <= 2) {
#pragma unroll
        for (int e = 0; e < NE; e++)
#pragma unroll
          for (int r = 0; r < R; r++) ps[r][e & (NP - 1)] = fma(xv[r][e], bet[e], ps[r][e & (NP - 1)]);
      } else if constexpr (VEC) {
#pragma unroll
        for (int j = 0; j < NJ; j++) {
          const double2 b = *reinterpret_cast<const double2*>(s_beta + COL(2 * j));
#pragma unroll
          for (int r = 0; r < R; r++) {
            ps[r][(2 * j) & (NP - 1)] = fma(xv[r][2 * j], b.x, ps[r][(2 * j) & (NP - 1)]);
            ps[r][(2 * j + 1) & (NP - 1)] = fma(xv[r][2 * j + 1], b.y, ps[r][(2 * j + 1) & (NP - 1)]);
          }
        }
      } else {
#pragma unroll
        for (int e = 0; e < NE; e++) {
          const double bv = s_beta[COL(e)];
#pragma unroll
          for (int r = 0; r < R; r++) ps[r][e & (NP - 1)] = fma(xv[r][e], bv, ps[r][e & (NP - 1)]);
        }
      }
#pragma unroll
      for (int r = 0; r < R; r++) {
        double s = ps[r][0];
        if constexpr (NP >= 2) s += ps[r][1];
        if constexpr (NP == 4) s += ps[r][2] + ps[r][3];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        part[r] = s;
      }
    }
    if constexpr (WPR > 1) {
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; r++) s_eta[buf][rg][wc][r] = part[r];
      }
      __syncthreads();
#pragma unroll
      for (int r = 0; r < R; r++) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < WPR; w++) s += s_eta[buf][rg][w][r];
        part[r] = s;
      }
      buf ^= 1;
    }
    // lane l < R evaluates the link for row l of the batch (other lanes mirror lane l % R)
    double eta = part[0];
#pragma unroll
    for (int r = 1; r < R; r++) if ((lane % R) == r) eta = part[r];
    eta += beta0;
    const int64_t myrow = row0 + (lane % R);
    const bool inb = myrow < n;
    const double yv = inb ? load_y(a.y, a.ydt, myrow) : 0.0;
    LinkOut lo = a.skip_obj ? glm_link_noobj_rt(a.family, eta, yv) : glm_link_rt(a.family, eta, yv);
    if (!inb) { lo.resid = 0.0; lo.wgt = 0.0; lo.obj = 0.0; }
    if (wc == 0 && lane < R && inb) {
      ob += lo.obj; gi += lo.resid; ui += lo.wgt;
      if (a.wbuf) a.wbuf[myrow] = lo.wgt;
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
      const double rr = __shfl_sync(0xffffffffu, lo.resid, r);
#pragma unroll
      for (int e = 0; e < NE; e++) g[e] = fma(rr, xv[r][e], g[e]);
      if constexpr (HESS) {
        const double ww = __shfl_sync(0xffffffffu, lo.wgt, r);
#pragma unroll
        for (int e = 0; e < NE; e++) u[e] = fma(ww, xv[r][e], u[e]);
      }
    }
  }

  // CTA reduction over the row groups, in row-group order (deterministic)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gi += __shfl_down_sync(0xffffffffu, gi, o);
    ui += __shfl_down_sync(0xffffffffu, ui, o);
    ob += __shfl_down_sync(0xffffffffu, ob, o);
  }
  double* s_g = s_acc;
  double* s_u = s_acc + WPR * CW;
  double* s_sc = s_acc + 2 * WPR * CW;
  for (int q = 0; q < RG; q++) {
    if (rg == q) {
#pragma unroll
      for (int e = 0; e < NE; e++) {
        if (CV(e)) { s_g[COL(e)] += g[e]; if constexpr (HESS) s_u[COL(e)] += u[e]; }
      }
      if (wc == 0 && lane == 0) { s_sc[0] += gi; s_sc[1] += ui; s_sc[2] += ob; }
    }
    __syncthreads();
  }
  double* out = a.partials + (size_t)blockIdx.x * (2 * d + 3);
  for (int64_t j = tid; j < d; j += GLM_THREADS) { out[j] = s_g[j]; out[d + j] = s_u[j]; }
  if (tid == 0) { out[2 * d] = s_sc[0]; out[2 * d + 1] = s_sc[1]; out[2 * d + 2] = s_sc[2]; }
}

#undef COL
#undef CV

// Sums the CTA partials (in CTA order) and writes g, obj and the intercept row/column of H.
__global__ void __launch_bounds__(256)
k_glm_rows_finish(int64_t d, int nparts, const double* __restrict__ partials, int want_hess,
                  double* __restrict__ stats) {
  const int64_t p = d + 1, ps = 2 * d + 3;
  for (int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); j < ps; j += (int64_t)gridDim.x * 8) {   // one warp per statistic
    const double s = sum_partials_warp(partials + j, (size_t)ps, nparts);
    if ((threadIdx.x & 31) != 0) continue;
    if (j < d) {
      stats[p * p + j] = s;                               // g[j]
    } else if (j < 2 * d) {
      if (want_hess) { int64_t i = j - d; stats[i * p + d] = s; stats[d * p + i] = s; }   // u
    } else if (j == 2 * d) {
      stats[p * p + d] = s;                               // g[intercept] = sum r
    } else if (j == 2 * d + 1) {
      if (want_hess) stats[d * p + d] = s;                // sum w
    } else {
      stats[p * p + p] = s;                               // objective
    }
  }
}

// mu[i] = link_inv(mu[i] + beta0)
__global__ void __launch_bounds__(256)
k_glm_link_inplace(int family, int64_t n, const double* __restrict__ beta0, double* __restrict__ mu) {
  const double b0 = *beta0;
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const double eta = mu[i] + b0;
    mu[i] = (family == NB_GLM_LOGISTIC || family == NB_GLM_LOGISTIC_IRLS) ? glm_mu<NB_GLM_LOGISTIC>(eta)
          : (family == NB_GLM_POISSON ? glm_mu<NB_GLM_POISSON>(eta) : eta);
  }
}

// mu = link_inv(X beta + beta0), one warp per row
__global__ void __launch_bounds__(GLM_THREADS)
k_glm_predict(int family, const double* __restrict__ x, int64_t n, int64_t d, int64_t ldx,
              const double* __restrict__ beta, double* __restrict__ mu) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (GLM_THREADS / 32);
  const double b0 = beta[d];
  for (int64_t r = (int64_t)blockIdx.x * (GLM_THREADS / 32) + (threadIdx.x >> 5); r < n; r += warps) {
    const double* row = x + r * ldx;
    double s = 0.0;
    for (int64_t j = lane; j < d; j += 32) s = fma(row[j], beta[j], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      double eta = s + b0;
      mu[r] = (family == NB_GLM_LOGISTIC || family == NB_GLM_LOGISTIC_IRLS) ? glm_mu<NB_GLM_LOGISTIC>(eta)
            : (family == NB_GLM_POISSON ? glm_mu<NB_GLM_POISSON>(eta) : eta);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// glm_fused.cu
bool glm_regs_applicable(int64_t d, int want_hess);
size_t glm_regs_scratch_bytes(int device, int64_t d, int want_hess);
int glm_regs_stats(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                   const double* beta, int want_hess, int skip_obj, double* stats, void* scratch, size_t scratch_bytes,
                   cudaStream_t st);
bool glm_regs_pass1_applicable(int64_t d);
int glm_regs_pass1(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                   const double* beta, int skip_obj, double* wbuf, double* stats, void* scratch, size_t scratch_bytes,
                   cudaStream_t st);
bool fused_applicable(int64_t n, int64_t d, int64_t ldx, const double* x);
size_t fused_scratch_bytes(int device, int64_t n, int64_t d);
int tri_gram(int device, const double* x, int64_t n, int64_t d, int64_t ldx, const double* wext, double* out,
             int64_t ldc, void* scratch, size_t scratch_bytes, cudaStream_t st);

// glm_ozaki.cu: X^T diag(w) X on the int8 tensor path (tcgen05), 128 < d <= 256
bool oz_applicable(int family, int64_t n, int64_t d, int64_t ldx, const double* x);
size_t oz_scratch_bytes(int device, int64_t n);
bool oz_rows_applicable(int64_t n, int64_t d, int64_t ldx, const double* x);
int oz_gradient_rows(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                     const double* beta, int skip_obj, double* parts, int max_parts, int* nparts, cudaStream_t st);
int oz_glm_stats(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                 const double* beta, int skip_obj, const void* prep, double* stats, double* parts, int max_parts,
                 int* nparts, void* scratch, size_t scratch_bytes, int phases, int scheme, cudaStream_t st);

struct GlmLayout {
  size_t off_w, off_parts, off_gemm, off_oz, total;
  int max_parts;
};

static GlmLayout glm_layout(int device, int64_t n, int64_t d, int want_hess, bool int8_hess = false) {
  GlmLayout L;
  const int sms = dev_info(device).sms;
  L.max_parts = sms * 4;
  size_t cur = 0;
  L.off_w = cur;
  if (want_hess && d >= 8) cur += align256((size_t)n * sizeof(double));
  L.off_parts = cur;
  size_t per = d >= 8 ? (size_t)(2 * d + 3) : 64;
  size_t parts_bytes = (size_t)L.max_parts * per * sizeof(double);
  if (want_hess && glm_regs_pass1_applicable(d)) {
    const size_t r = glm_regs_scratch_bytes(device, d, 0);
    if (r > parts_bytes) parts_bytes = r;
  }
  cur += align256(parts_bytes);
  L.off_gemm = cur;
  if (want_hess && d >= 8) cur += align256(gemm_scratch_bytes(device, d, d, n));
  if (want_hess && d >= 8 && d <= 256) {   // the triangular kernel takes their partials from the GEMM region
    size_t f = L.off_gemm + align256(fused_scratch_bytes(device, n, d));
    if (f > cur) cur = f;
  }
  if (glm_regs_applicable(d, want_hess)) {  // register-resident single pass (glm_regs.cu) takes its partials from offset 0
    size_t r = align256(glm_regs_scratch_bytes(device, d, want_hess));
    if (r > cur) cur = r;
  }
  L.off_oz = cur;
  if (int8_hess && want_hess && d > 128 && d <= 256) cur += oz_scratch_bytes(device, n);
  L.total = cur;
  return L;
}

template <int FAMILY, int D>
static int launch_small(int device, const double* x, int64_t n, int64_t ldx, const void* y, int ydt,
                        const double* beta, int want_hess, double* stats, double* parts, int max_parts,
                        cudaStream_t st) {
  int64_t blocks = (n + GLM_THREADS * 2 - 1) / (GLM_THREADS * 2);
  if (blocks > max_parts) blocks = max_parts;
  if (blocks < 1) blocks = 1;
  k_glm_small<FAMILY, D><<<(unsigned)blocks, GLM_THREADS, 0, st>>>(x, n, ldx, y, ydt, beta, want_hess, parts);
  NB_LAUNCH_CHECK();
  constexpr int NVS = (D + 1) * (D + 2) / 2 + D + 2;
  k_glm_small_finish<<<(NVS + 7) / 8, 256, 0, st>>>(D + 1, (int)blocks, parts, want_hess, stats);
  NB_LAUNCH_CHECK();
  return 0;
}

template <int FAMILY>
static int launch_small_d(int device, int64_t d, const double* x, int64_t n, int64_t ldx, const void* y, int ydt,
                          const double* beta, int want_hess, double* stats, double* parts, int max_parts,
                          cudaStream_t st) {
  switch (d) {
#define NB_SMALL(DD) case DD: return launch_small<FAMILY, DD>(device, x, n, ldx, y, ydt, beta, want_hess, stats, parts, max_parts, st);
    NB_SMALL(1) NB_SMALL(2) NB_SMALL(3) NB_SMALL(4) NB_SMALL(5) NB_SMALL(6) NB_SMALL(7)
#undef NB_SMALL
  }
  set_error("glm small path: d=%lld out of range", (long long)d);
  return 2;
}

template <int NJ, int WPR, bool VEC, bool HESS>
static int launch_rows_h(int device, const GlmRowsArgs& a, int max_parts, int* nparts, cudaStream_t st) {
  auto kern = k_glm_rows<NJ, WPR, VEC, HESS>;
  size_t smem = (3 * (size_t)WPR * 64 * NJ + 4) * sizeof(double);
  static KernelSetup setup;          // the shared-memory size depends on the template arguments only
  int occ = 1;
  if (int rc = setup.get(device, kern, GLM_THREADS, smem, &occ)) return rc;
  constexpr int R = RowsCfg<NJ, HESS>::R;
  constexpr int RG = (GLM_THREADS / 32) / WPR;
  int64_t batches = (a.n + R - 1) / R;
  int64_t want = (batches + RG - 1) / RG;
  int64_t grid = (int64_t)dev_info(device).sms * occ;
  if (grid > max_parts) grid = max_parts;
  if (grid > want) grid = want;
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, GLM_THREADS, smem, st>>>(a);
  NB_LAUNCH_CHECK();
  *nparts = (int)grid;
  return 0;
}

template <int NJ, int WPR, bool VEC>
static int launch_rows(int device, const GlmRowsArgs& a, int max_parts, int* nparts, cudaStream_t st) {
  return a.want_hess ? launch_rows_h<NJ, WPR, VEC, true>(device, a, max_parts, nparts, st)
                     : launch_rows_h<NJ, WPR, VEC, false>(device, a, max_parts, nparts, st);
}

template <bool VEC>
static int launch_rows_cfg(int device, const GlmRowsArgs& a, int max_parts, int* nparts, cudaStream_t st) {
  const int64_t d = a.d;
  if (d <= 64) return launch_rows<1, 1, VEC>(device, a, max_parts, nparts, st);
  if (d <= 128) return launch_rows<2, 1, VEC>(device, a, max_parts, nparts, st);
  if (d <= 256) return launch_rows<4, 1, VEC>(device, a, max_parts, nparts, st);
  if (d <= 512) return launch_rows<8, 1, VEC>(device, a, max_parts, nparts, st);
  if (d <= 1024) return launch_rows<16, 1, VEC>(device, a, max_parts, nparts, st);
  if (d <= 2048) return launch_rows<16, 2, VEC>(device, a, max_parts, nparts, st);
  if (d <= 4096) return launch_rows<16, 4, VEC>(device, a, max_parts, nparts, st);
  if (d <= 8192) return launch_rows<16, 8, VEC>(device, a, max_parts, nparts, st);
  set_error("nb_glm_stats: d=%lld > 8192 is not supported", (long long)d);
  return 2;
}

}  // namespace nb

using namespace nb;

extern "C" {

int64_t nb_glm_stats_len(int64_t d) { int64_t p = d + 1; return p * p + p + 1; }

size_t nb_glm_scratch_bytes(int device, int64_t n, int64_t d, int want_hess) {
  return glm_layout(device, n, d, want_hess & 1, (want_hess & NB_GLM_INT8_HESSIAN) != 0).total;
}

int nb_glm_stats(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx, const void* y_d,
                 int ydt, const double* beta_d, int want_hess, double* stats_d, void* scratch_d,
                 size_t scratch_bytes, void* stream) {
  return nb_glm_stats_prepared(device, family, x_d, n, d, ldx, y_d, ydt, beta_d, want_hess, nullptr, stats_d, scratch_d,
                               scratch_bytes, stream);
}

int nb_glm_stats_prepared(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx, const void* y_d,
                          int ydt, const double* beta_d, int want_hess, const void* prep_d, double* stats_d,
                          void* scratch_d, size_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(family >= NB_GLM_LINEAR && family <= NB_GLM_LOGISTIC_IRLS, "nb_glm_stats: unknown family %d", family);
  NB_REQUIRE(d >= 1 && n >= 0 && ldx >= d, "nb_glm_stats: bad shape n=%lld d=%lld ldx=%lld", (long long)n,
             (long long)d, (long long)ldx);
  NB_REQUIRE(ydt == NB_F64 || ydt == NB_I64, "nb_glm_stats: y dtype %d not supported", ydt);
  const int skip_obj = (want_hess & 2) ? 1 : 0;      // NB_GLM_NO_OBJECTIVE
  const bool int8_hess = (want_hess & NB_GLM_INT8_HESSIAN) != 0;
  const int oz_phases = (want_hess & NB_GLM_INT8_ROWS_ONLY) ? 1 : ((want_hess & NB_GLM_INT8_PRODUCTS_ONLY) ? 2 : 3);
  const int oz_scheme = (want_hess & NB_GLM_INT8_FOUR_PLANES) ? 2 : ((want_hess & NB_GLM_INT8_NO_LEVEL5) ? 1 : 0);
  want_hess &= 1;
  NB_REQUIRE(!int8_hess || prep_d != nullptr, "nb_glm_stats: NB_GLM_INT8_HESSIAN needs the block's nb_glm_prepare output");
  GlmLayout L = glm_layout(device, n, d, want_hess, int8_hess);
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= L.total, "nb_glm_stats: scratch too small (%zu < %zu)",
             scratch_bytes, L.total);
  char* sc = (char*)scratch_d;
  double* parts = (double*)(sc + L.off_parts);
  if (d <= 7) {
    switch (family) {
      case NB_GLM_LOGISTIC_IRLS: return launch_small_d<NB_GLM_LOGISTIC_IRLS>(device, d, x_d, n, ldx, y_d, ydt, beta_d, want_hess, stats_d, parts, L.max_parts, st);
      case NB_GLM_LOGISTIC: return launch_small_d<NB_GLM_LOGISTIC>(device, d, x_d, n, ldx, y_d, ydt, beta_d, want_hess, stats_d, parts, L.max_parts, st);
      case NB_GLM_POISSON: return launch_small_d<NB_GLM_POISSON>(device, d, x_d, n, ldx, y_d, ydt, beta_d, want_hess, stats_d, parts, L.max_parts, st);
      default: return launch_small_d<NB_GLM_LINEAR>(device, d, x_d, n, ldx, y_d, ydt, beta_d, want_hess, stats_d, parts, L.max_parts, st);
    }
  }
  if (glm_regs_applicable(d, want_hess))
    return glm_regs_stats(device, family, x_d, n, d, ldx, y_d, ydt, beta_d, want_hess, skip_obj, stats_d, sc, L.total, st);
  if (int8_hess && want_hess && oz_applicable(family, n, d, ldx, x_d)) {
    // pass A (link, g, u, digit planes) + pass B (int8 tensor products) + finish; the vector statistics go through
    // the same second stage as k_glm_rows
    int np = 0;
    int rc = oz_glm_stats(device, family, x_d, n, d, ldx, y_d, ydt, beta_d, skip_obj, prep_d, stats_d, parts, L.max_parts,
                          &np, sc + L.off_oz, L.total - L.off_oz, oz_phases, oz_scheme, st);
    if (rc) return rc;
    if (np > 0) {
      k_glm_rows_finish<<<(unsigned)((2 * d + 3 + 7) / 8), 256, 0, st>>>(d, np, parts, want_hess, stats_d);
      NB_LAUNCH_CHECK();
    }
    return 0;
  }
  if (!want_hess && oz_rows_applicable(n, d, ldx, x_d)) {
    // gradient / objective pass for 128 < d <= 256: the bulk-copy ring of the int8 path's row kernel (k_oz_rows<0>)
    int np = 0;
    int rc = oz_gradient_rows(device, family, x_d, n, d, ldx, y_d, ydt, beta_d, skip_obj, parts, L.max_parts, &np, st);
    if (rc) return rc;
    k_glm_rows_finish<<<(unsigned)((2 * d + 3 + 7) / 8), 256, 0, st>>>(d, np, parts, want_hess, stats_d);
    NB_LAUNCH_CHECK();
    return 0;
  }
  const bool tri_ok = want_hess && fused_applicable(n, d, ldx, x_d);
  GlmRowsArgs a;
  a.x = x_d; a.n = n; a.d = d; a.ldx = ldx; a.y = y_d; a.ydt = ydt; a.beta = beta_d;
  a.family = family; a.want_hess = want_hess; a.skip_obj = skip_obj;
  const bool need_w = want_hess && family != NB_GLM_LINEAR;
  a.wbuf = need_w ? (double*)(sc + L.off_w) : nullptr;
  a.partials = parts;
  int nparts = 0;
  const bool vec = ((ldx & 1) == 0) && ((((uintptr_t)x_d) & 15) == 0);
  int rc;
  if (tri_ok && glm_regs_pass1_applicable(d)) {
    // 64 < d <= 192: the register-resident kernel is the faster first pass (its partials live in the parts region)
    rc = glm_regs_pass1(device, family, x_d, n, d, ldx, y_d, ydt, beta_d, skip_obj, a.wbuf, stats_d, parts,
                        L.off_gemm - L.off_parts, st);
    if (rc) return rc;
    return tri_gram(device, x_d, n, d, ldx, a.wbuf, stats_d, d + 1, sc + L.off_gemm, L.total - L.off_gemm, st);
  }
  rc = vec ? launch_rows_cfg<true>(device, a, L.max_parts, &nparts, st)
           : launch_rows_cfg<false>(device, a, L.max_parts, &nparts, st);
  if (rc) return rc;
  k_glm_rows_finish<<<(unsigned)((2 * d + 3 + 7) / 8), 256, 0, st>>>(d, nparts, parts, want_hess, stats_d);
  NB_LAUNCH_CHECK();
  if (tri_ok) {
    // pass 2: upper-triangular blocks of X^T diag(w) X by FP64 DMMA, mirrored -> exactly symmetric
    return tri_gram(device, x_d, n, d, ldx, a.wbuf, stats_d, d + 1, sc + L.off_gemm, L.total - L.off_gemm, st);
  }
  if (want_hess) {
    // G = X^T diag(w) X into the top-left d x d of the p x p Hessian (row stride p)
    rc = gemm_f64(device, d, d, n, x_d, 1, ldx, x_d, ldx, 1, a.wbuf, stats_d, d + 1, sc + L.off_gemm,
                  L.total - L.off_gemm, st);
    if (rc) return rc;
  }
  return 0;
}

int nb_glm_predict(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx,
                   const double* beta_d, double* mu_d, void* stream) {
  NB_ENTER(device);
  if (n == 0) return 0;
  if (d <= 512 && (d & 1) == 0 && (ldx & 1) == 0 && n >= 256 && ((((uintptr_t)x_d) | ((uintptr_t)beta_d)) & 15) == 0) {
    // narrow rows: eta by the sub-warp mat-vec (16-byte loads, several rows per warp), then the link in place
    int rc = gemm_f64(device, n, 1, d, x_d, ldx, 1, beta_d, 1, 1, nullptr, mu_d, 1, nullptr, 0, st);
    if (rc) return rc;
    const int grid = grid_for(device, n, 256, 4);
    k_glm_link_inplace<<<grid, 256, 0, st>>>(family, n, beta_d + d, mu_d);
    NB_LAUNCH_CHECK();
    return 0;
  }
  int64_t blocks = (n + 7) / 8;
  int64_t cap = (int64_t)dev_info(device).sms * 8;
  k_glm_predict<<<(unsigned)(blocks < cap ? blocks : cap), GLM_THREADS, 0, st>>>(family, x_d, n, d, ldx, beta_d, mu_d);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// The whole Newton loop (glms.py:660-676) driven from native code
// ------------------------------------------------------------------------------------------------
// newton() is a host loop: statistics per row block, the FIFO sum of the partials (blockarray.py:643-689), the update,
// and one scalar comparison that needs the device's answer.  Driven from Python every one of those steps costs tens of
// microseconds of interpreter and allocator time; on the README problem (10^6 x 2: 24 MB, ~5 us per pass) that is all
// of the run time.  Here the same launches are issued back to back from C++, and the 16-byte verdict of an iteration is the
// only thing read back (one cudaMemcpyAsync + stream synchronise per iteration, as the reference's `<= tol` test implies).
namespace nb {
static inline size_t fit_align(size_t v) { return (v + 255) & ~(size_t)255; }
}

extern "C" {

size_t nb_glm_newton_fit_scratch_bytes(int device, int nblocks, int64_t max_rows, int64_t d, int want_hess_flags) {
  const int64_t p = d + 1;
  const size_t len = (size_t)nb_glm_stats_len(d) * sizeof(double);
  return nb::fit_align(len) * (size_t)(2 * nblocks) + nb::fit_align((size_t)p * sizeof(double)) + 256 +
         nb::fit_align(nb_glm_newton_scratch_bytes(p)) + nb::fit_align(nb_glm_scratch_bytes(device, max_rows, d, want_hess_flags));
}

int nb_glm_newton_fit(int device, int family, int nblocks, const double* const* x_blocks, const int64_t* rows, int64_t d,
                      const int64_t* ldx, const void* const* y_blocks, int ydt, const void* const* preps,
                      int want_hess_flags, const double* l2vec_d, double* beta_d, double tol, int max_iter,
                      int32_t* iters_out, double* gmax_out, int32_t* info_out, void* scratch_d, size_t scratch_bytes,
                      void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(nblocks >= 1 && max_iter >= 0, "nb_glm_newton_fit: bad nblocks=%d max_iter=%d", nblocks, max_iter);
  int64_t max_rows = 0;
  for (int b = 0; b < nblocks; b++) if (rows[b] > max_rows) max_rows = rows[b];
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_glm_newton_fit_scratch_bytes(device, nblocks, max_rows, d, want_hess_flags),
             "nb_glm_newton_fit: scratch too small");
  const int64_t p = d + 1;
  const int64_t len = nb_glm_stats_len(d);
  const size_t slot = nb::fit_align((size_t)len * sizeof(double));
  char* sp = (char*)scratch_d;
  char* stats0 = sp; sp += slot * (size_t)(2 * nblocks);
  double* beta_new = (double*)sp; sp += nb::fit_align((size_t)p * sizeof(double));
  double* verdict = (double*)sp; sp += 256;                  // [max|g| (f64) | info (i32)]
  void* nsc = sp; const size_t nsc_bytes = nb::fit_align(nb_glm_newton_scratch_bytes(p)); sp += nsc_bytes;
  void* gsc = sp; const size_t gsc_bytes = scratch_bytes - (size_t)(sp - (char*)scratch_d);
  auto buf = [&](int i) { return (double*)(stats0 + slot * (size_t)i); };
  *iters_out = 0; *gmax_out = 0.0; *info_out = 0;
  // the verdict lands in page-locked host memory (one small buffer per calling thread, allocated on first use): a copy
  // into pageable memory goes through the driver's staging buffer and costs ~10 us more per iteration
  static thread_local double* host_verdict = nullptr;
  if (host_verdict == nullptr) NB_CHECK(cudaHostAlloc((void**)&host_verdict, 64, cudaHostAllocDefault));
  double* cur = beta_d;                          // iterates ping-pong between beta_d and beta_new
  double* nxt = beta_new;
  for (int it = 0; it < max_iter; it++) {
    for (int b = 0; b < nblocks; b++) {
      const void* prep = preps ? preps[b] : nullptr;
      const int mode = prep ? want_hess_flags : (want_hess_flags & ~(NB_GLM_INT8_HESSIAN | NB_GLM_INT8_NO_LEVEL5 | NB_GLM_INT8_FOUR_PLANES));
      int rc = nb_glm_stats_prepared(device, family, x_blocks[b], rows[b], d, ldx[b], y_blocks[b], ydt, cur, mode, prep,
                                     buf(b), gsc, gsc_bytes, stream);
      if (rc) return rc;
    }
    // the reference's FIFO tree: pop two, push their sum at the back
    int head = 0, tail = nblocks;                // live partials are buffers head .. tail-1 of a queue that only grows
    while (tail - head > 1) {
      int rc = nb_bop_reduce(device, NB_RSUM, len, buf(head), buf(head + 1), buf(tail), NB_F64, stream);
      if (rc) return rc;
      head += 2; tail += 1;
    }
    int rc = nb_glm_newton_update(device, p, buf(head), l2vec_d, cur, nxt, verdict, (int32_t*)(verdict + 1), nsc,
                                  nsc_bytes, stream);
    if (rc) return rc;
    NB_CHECK(cudaMemcpyAsync(host_verdict, verdict, 16, cudaMemcpyDeviceToHost, st));
    NB_CHECK(cudaStreamSynchronize(st));
    int32_t info = 0;
    memcpy(&info, &host_verdict[1], sizeof(int32_t));
    *iters_out = it + 1; *gmax_out = host_verdict[0]; *info_out = info;
    if (info != 0) break;                        // singular / non-finite: `cur` is the last good iterate, the caller decides
    double* t = cur; cur = nxt; nxt = t;
    if (host_verdict[0] <= tol) break;           // glms.py:674 (the gradient from before the update)
  }
  if (cur != beta_d) NB_CHECK(cudaMemcpyAsync(beta_d, cur, (size_t)p * sizeof(double), cudaMemcpyDeviceToDevice, st));
  return 0;
}

}  // extern "C"
