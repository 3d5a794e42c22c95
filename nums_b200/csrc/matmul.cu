// FP64 contractions: KernelCls.bop "tensordot" branch (nums/core/kernel/numpy_kernel.py:419-425)
// and the weighted Gram X^T.diag(w).X of the GLM Hessian (nums/models/glms.py:491-499, 548).
//
// Matrix.matrix: FP64 tensor-core tiles.  sm_100a's FP64 tensor instruction is DMMA.8x8x4
// (mma.sync.aligned.m8n8k4.f64; tcgen05 has no f64 kind), fed from shared memory through a
// multi-stage cp.async pipeline with padded (stride = 4 mod 16 doubles) tiles so every fragment
// load is bank-conflict free.  Tall-skinny contractions (huge k, small m,n) are split along k over
// the whole grid; partial tiles go to caller-provided scratch and are summed in a fixed order
// (deterministic, no atomics).  Matrix.vector / dot are HBM-bound streaming kernels.
#include "gemm.cuh"
#include <stdlib.h>

namespace nb {

// ------------------------------------------------------------------------------------------------
// PTX helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N)); }

// ------------------------------------------------------------------------------------------------
// GEMM
// ------------------------------------------------------------------------------------------------
constexpr int GEMM_BK = 16;
constexpr int GEMM_PAD = 4;

struct GemmArgs {
  int64_t m, n, k;
  const double* a; int64_t lda;   // A_MC: element (k,m) at a[k*lda+m];  else (m,k) at a[m*lda+k]
  const double* b; int64_t ldb;   // B_NC: element (k,n) at b[k*ldb+n];  else (n,k) at b[n*ldb+k]
  const double* w;                // optional weights along k
  double* c; int64_t ldc;
  double* part;                   // [splits][ntiles][BM*BN] when splits > 1
  int tiles_m, tiles_n, ntiles, splits, sym;
  int64_t kchunk;                 // k range per split (multiple of GEMM_BK)
};

template <int BM, int BN, bool A_MC, bool B_NC>
struct GemmSmem {
  static constexpr int A_ROWS = A_MC ? GEMM_BK : BM;
  static constexpr int A_COLS = A_MC ? BM : GEMM_BK;
  static constexpr int B_ROWS = B_NC ? GEMM_BK : BN;
  static constexpr int B_COLS = B_NC ? BN : GEMM_BK;
  static constexpr int A_LD = A_COLS + GEMM_PAD;
  static constexpr int B_LD = B_COLS + GEMM_PAD;
  static constexpr int A_ELEMS = A_ROWS * A_LD;
  static constexpr int B_ELEMS = B_ROWS * B_LD;
  static constexpr int STAGE_ELEMS = A_ELEMS + B_ELEMS + GEMM_BK;   // + weights
};

// Copies a ROWS x COLS tile (COLS contiguous in global memory) into padded shared memory,
// zero-filling everything outside [row_lim, col_lim).
template <int ROWS, int COLS, int LD, int NT, bool AL>
__device__ __forceinline__ void load_tile(double* sm, const double* __restrict__ g, int64_t ld, int64_t row0,
                                          int64_t col0, int64_t row_lim, int64_t col_lim, int tid) {
  if constexpr (AL) {
    constexpr int CPR = COLS / 2;  // 16-byte chunks per row
    constexpr int TOTAL = ROWS * CPR;
#pragma unroll
    for (int c = tid; c < TOTAL; c += NT) {
      int r = c / CPR, cc = (c % CPR) * 2;
      int64_t gr = row0 + r, gc = col0 + cc;
      int bytes = 0;
      if (gr < row_lim) {
        int64_t rem = col_lim - gc;
        bytes = rem >= 2 ? 16 : (rem == 1 ? 8 : 0);
      }
      const double* src = bytes ? g + gr * ld + gc : g;
      cp_async16(sm + r * LD + cc, src, bytes);
    }
  } else {
    constexpr int TOTAL = ROWS * COLS;
#pragma unroll 4
    for (int c = tid; c < TOTAL; c += NT) {
      int r = c / COLS, cc = c % COLS;
      int64_t gr = row0 + r, gc = col0 + cc;
      int bytes = (gr < row_lim && gc < col_lim) ? 8 : 0;
      const double* src = bytes ? g + gr * ld + gc : g;
      cp_async8(sm + r * LD + cc, src, bytes);
    }
  }
}

template <int BM, int BN, int WM, int WN, int WK, int STAGES, bool A_MC, bool B_NC, bool AL>
__global__ void __launch_bounds__((BM / WM) * (BN / WN) * WK * 32)
k_gemm(const GemmArgs g) {
  using SM = GemmSmem<BM, BN, A_MC, B_NC>;
  constexpr int WARPS_M = BM / WM, WARPS_N = BN / WN;
  constexpr int NT = WARPS_M * WARPS_N * WK * 32;
  constexpr int FM = WM / 8, FN = WN / 8;
  constexpr int KSTEPS = GEMM_BK / 4;
  static_assert(KSTEPS % WK == 0, "k-split must divide the k steps of a stage");
  extern __shared__ __align__(16) double smem[];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wk = warp / (WARPS_M * WARPS_N);
  const int wmn = warp % (WARPS_M * WARPS_N);
  const int wm = wmn / WARPS_N, wn = wmn % WARPS_N;
  const int lq = lane >> 2, lr = lane & 3;   // groupID, threadID_in_group

  // which tile / split
  const int t = blockIdx.x % g.ntiles;
  const int z = blockIdx.x / g.ntiles;
  int tm, tn;
  if (g.sym) {          // upper-triangular tile list, row by row
    int rem = t; tm = 0;
    while (rem >= g.tiles_n - tm) { rem -= g.tiles_n - tm; tm++; }
    tn = tm + rem;
  } else {
    tm = t / g.tiles_n; tn = t % g.tiles_n;
  }
  const int64_t m0 = (int64_t)tm * BM, n0 = (int64_t)tn * BN;
  const int64_t kbeg = (int64_t)z * g.kchunk;
  const int64_t kend = (kbeg + g.kchunk < g.k) ? kbeg + g.kchunk : g.k;
  const int ktiles = kend > kbeg ? (int)((kend - kbeg + GEMM_BK - 1) / GEMM_BK) : 0;

  auto load_stage = [&](int slot, int kt) {
    double* sA = smem + (size_t)slot * SM::STAGE_ELEMS;
    double* sB = sA + SM::A_ELEMS;
    double* sW = sB + SM::B_ELEMS;
    const int64_t k0 = kbeg + (int64_t)kt * GEMM_BK;
    if constexpr (A_MC) load_tile<SM::A_ROWS, SM::A_COLS, SM::A_LD, NT, AL>(sA, g.a, g.lda, k0, m0, kend, g.m, tid);
    else                load_tile<SM::A_ROWS, SM::A_COLS, SM::A_LD, NT, AL>(sA, g.a, g.lda, m0, k0, g.m, kend, tid);
    if constexpr (B_NC) load_tile<SM::B_ROWS, SM::B_COLS, SM::B_LD, NT, AL>(sB, g.b, g.ldb, k0, n0, kend, g.n, tid);
    else                load_tile<SM::B_ROWS, SM::B_COLS, SM::B_LD, NT, AL>(sB, g.b, g.ldb, n0, k0, g.n, kend, tid);
    if (g.w != nullptr && tid < GEMM_BK) {
      int64_t kk = k0 + tid;
      cp_async8(sW + tid, kk < kend ? g.w + kk : g.w, kk < kend ? 8 : 0);
    }
  };

  double acc[FM][FN][2];
#pragma unroll
  for (int i = 0; i < FM; i++)
#pragma unroll
    for (int j = 0; j < FN; j++) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < ktiles) load_stage(s, s);
    cp_async_commit();
  }
  const bool has_w = g.w != nullptr;
  for (int kt = 0; kt < ktiles; kt++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    {
      int nk = kt + STAGES - 1;
      if (nk < ktiles) load_stage(nk % STAGES, nk);
      cp_async_commit();
    }
    const double* sA = smem + (size_t)(kt % STAGES) * SM::STAGE_ELEMS;
    const double* sB = sA + SM::A_ELEMS;
    const double* sW = sB + SM::B_ELEMS;
#pragma unroll
    for (int ks = 0; ks < KSTEPS / WK; ks++) {
      const int kk = (ks * WK + wk) * 4 + lr;
      double af[FM], bf[FN];
#pragma unroll
      for (int i = 0; i < FM; i++) {
        int mm = wm * WM + i * 8 + lq;
        af[i] = A_MC ? sA[kk * SM::A_LD + mm] : sA[mm * SM::A_LD + kk];
      }
#pragma unroll
      for (int j = 0; j < FN; j++) {
        int nn = wn * WN + j * 8 + lq;
        bf[j] = B_NC ? sB[kk * SM::B_LD + nn] : sB[nn * SM::B_LD + kk];
      }
      if (has_w) {
        double wv = sW[kk];
#pragma unroll
        for (int j = 0; j < FN; j++) bf[j] *= wv;
      }
#pragma unroll
      for (int i = 0; i < FM; i++)
#pragma unroll
        for (int j = 0; j < FN; j++) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
  }
  cp_async_wait<0>();
  __syncthreads();

  // intra-CTA k-split: groups wk > 0 hand their accumulators to group 0 through shared memory
  if constexpr (WK > 1) {
    constexpr int PER_WARP = FM * FN * 2 * 32;
    for (int src = 1; src < WK; src++) {
      if (wk == src) {
        double* dst = smem + (size_t)wmn * PER_WARP;
#pragma unroll
        for (int i = 0; i < FM; i++)
#pragma unroll
          for (int j = 0; j < FN; j++) {
            dst[((i * FN + j) * 2 + 0) * 32 + lane] = acc[i][j][0];
            dst[((i * FN + j) * 2 + 1) * 32 + lane] = acc[i][j][1];
          }
      }
      __syncthreads();
      if (wk == 0) {
        const double* s2 = smem + (size_t)wmn * PER_WARP;
#pragma unroll
        for (int i = 0; i < FM; i++)
#pragma unroll
          for (int j = 0; j < FN; j++) {
            acc[i][j][0] += s2[((i * FN + j) * 2 + 0) * 32 + lane];
            acc[i][j][1] += s2[((i * FN + j) * 2 + 1) * 32 + lane];
          }
      }
      __syncthreads();
    }
    if (wk != 0) return;
  }

  if (g.splits > 1) {
    double* p = g.part + ((size_t)z * g.ntiles + t) * (size_t)(BM * BN);
#pragma unroll
    for (int i = 0; i < FM; i++)
#pragma unroll
      for (int j = 0; j < FN; j++) {
        int r = wm * WM + i * 8 + lq, c = wn * WN + j * 8 + 2 * lr;
        *reinterpret_cast<double2*>(p + r * BN + c) = make_double2(acc[i][j][0], acc[i][j][1]);
      }
  } else {
#pragma unroll
    for (int i = 0; i < FM; i++)
#pragma unroll
      for (int j = 0; j < FN; j++) {
        int64_t r = m0 + wm * WM + i * 8 + lq, c = n0 + wn * WN + j * 8 + 2 * lr;
        if (r < g.m) {
#pragma unroll
          for (int e = 0; e < 2; e++) {
            int64_t cc = c + e;
            if (cc >= g.n) continue;
            if (g.sym) {
              // only the upper triangle is authoritative; the lower one is its mirror image, so the
              // result is exactly symmetric
              if (r <= cc) { g.c[r * g.ldc + cc] = acc[i][j][e]; if (r != cc) g.c[cc * g.ldc + r] = acc[i][j][e]; }
            } else {
              g.c[r * g.ldc + cc] = acc[i][j][e];
            }
          }
        }
      }
  }
}

// Sums split-k partial tiles in split order and scatters them (with mirroring in sym mode) into C.
template <int BM, int BN>
__global__ void __launch_bounds__(256)
k_gemm_reduce(const GemmArgs g) {
  const int64_t total = g.m * g.n;
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += stride) {
    int64_t r = e / g.n, c = e - r * g.n;
    int tm = (int)(r / BM), tn = (int)(c / BN);
    int lr_ = (int)(r - (int64_t)tm * BM), lc = (int)(c - (int64_t)tn * BN);
    int t;
    if (g.sym) {
      // read every element from the upper triangle (r <= c): exact symmetry
      if (c < r) { int x = tm; tm = tn; tn = x; x = lr_; lr_ = lc; lc = x; }
      // index of (tm, tn) in the row-by-row upper-triangular list
      t = tm * g.tiles_n - (tm * (tm - 1)) / 2 + (tn - tm);
    } else {
      t = tm * g.tiles_n + tn;
    }
    const double* p = g.part + (size_t)t * (BM * BN) + lr_ * BN + lc;
    double s = 0.0;
    for (int z = 0; z < g.splits; z++) s += p[(size_t)z * g.ntiles * (BM * BN)];
    g.c[r * g.ldc + c] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// launch plan
// ------------------------------------------------------------------------------------------------
struct GemmPlan {
  int cfg;            // 0: 128x128 tiles, 1: 64x64, 2: 32x32 (4-way in-CTA k split)
  int bm, bn;
  int tiles_m, tiles_n, ntiles, splits, sym;
  int64_t kchunk;
  size_t scratch;
};

static GemmPlan plan_gemm(int device, int64_t m, int64_t n, int64_t k, bool sym) {
  GemmPlan p;
  int64_t big = m > n ? m : n;
  p.cfg = big <= 32 ? 2 : (big <= 64 ? 1 : 0);
  p.bm = p.bn = p.cfg == 0 ? 128 : (p.cfg == 1 ? 64 : 32);
  p.tiles_m = (int)((m + p.bm - 1) / p.bm);
  p.tiles_n = (int)((n + p.bn - 1) / p.bn);
  p.sym = sym ? 1 : 0;
  p.ntiles = sym ? p.tiles_m * (p.tiles_m + 1) / 2 : p.tiles_m * p.tiles_n;
  const int sms = dev_info(device).sms;
  const int64_t target = (int64_t)sms * (p.cfg == 0 ? 1 : (p.cfg == 1 ? 3 : 4));
  int64_t splits = target / p.ntiles;
  int64_t max_splits = (k + GEMM_BK * 8 - 1) / (GEMM_BK * 8);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int64_t kchunk = (k + splits - 1) / splits;
  kchunk = (kchunk + GEMM_BK - 1) / GEMM_BK * GEMM_BK;
  if (kchunk < GEMM_BK) kchunk = GEMM_BK;
  splits = (k + kchunk - 1) / kchunk;
  if (splits < 1) splits = 1;
  p.splits = (int)splits;
  p.kchunk = kchunk;
  p.scratch = splits > 1 ? (size_t)splits * p.ntiles * p.bm * p.bn * sizeof(double) : 0;
  return p;
}

template <int BM, int BN, int WM, int WN, int WK, int STAGES, bool A_MC, bool B_NC, bool AL>
static int launch_gemm(const GemmArgs& g, cudaStream_t st) {
  using SM = GemmSmem<BM, BN, A_MC, B_NC>;
  constexpr int NT = (BM / WM) * (BN / WN) * WK * 32;
  size_t smem = (size_t)STAGES * SM::STAGE_ELEMS * sizeof(double);
  auto kern = k_gemm<BM, BN, WM, WN, WK, STAGES, A_MC, B_NC, AL>;
  NB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)(g.ntiles * g.splits), NT, smem, st>>>(g);
  NB_LAUNCH_CHECK();
  if (g.splits > 1) {
    int64_t total = g.m * g.n;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_gemm_reduce<BM, BN><<<(unsigned)blocks, 256, 0, st>>>(g);
    NB_LAUNCH_CHECK();
  }
  return 0;
}

template <bool A_MC, bool B_NC, bool AL>
static int launch_gemm_cfg(int cfg, const GemmArgs& g, cudaStream_t st) {
  switch (cfg) {
    case 0: return launch_gemm<128, 128, 64, 32, 1, 4, A_MC, B_NC, AL>(g, st);
    case 1: return launch_gemm<64, 64, 32, 32, 1, 4, A_MC, B_NC, AL>(g, st);
    default: return launch_gemm<32, 32, 32, 32, 4, 4, A_MC, B_NC, AL>(g, st);
  }
}

template <bool A_MC, bool B_NC>
static int launch_gemm_al(bool al, int cfg, const GemmArgs& g, cudaStream_t st) {
  return al ? launch_gemm_cfg<A_MC, B_NC, true>(cfg, g, st) : launch_gemm_cfg<A_MC, B_NC, false>(cfg, g, st);
}

// ------------------------------------------------------------------------------------------------
// matrix.vector / dot (HBM-bound)
// ------------------------------------------------------------------------------------------------
constexpr int GV_THREADS = 256;
constexpr int64_t GV_TARGET_BLOCKS = 592;

// y[r*ys] = sum_k A[r*lda + k] * w[k] * x[k*xs], one warp per row (short rows, many of them)
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_rows_warp(int64_t m, int64_t k, const double* __restrict__ a, int64_t lda, const double* __restrict__ x,
                 int64_t xs, const double* __restrict__ w, double* __restrict__ y, int64_t ys) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (GV_THREADS / 32);
  for (int64_t r = (int64_t)blockIdx.x * (GV_THREADS / 32) + (threadIdx.x >> 5); r < m; r += warps) {
    const double* row = a + r * lda;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    int64_t i = lane;
    for (; i + 96 < k; i += 128) {
      double v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) v[u] = row[i + 32 * u];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        double xv = x[(i + 32 * u) * xs];
        if (w) xv *= w[i + 32 * u];
        acc[u] = fma(v[u], xv, acc[u]);
      }
    }
    for (; i < k; i += 32) {
      double xv = x[i * xs];
      if (w) xv *= w[i];
      acc[0] = fma(row[i], xv, acc[0]);
    }
    double s = (acc[0] + acc[1]) + (acc[2] + acc[3]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) y[r * ys] = s;
  }
}

// Short aligned rows (k <= 512, even, 16-byte aligned): LPR lanes per row, 32 / LPR rows per warp and two such row
// groups per step, every lane loading double2 -- a warp per 512-byte row left most of the load slots idle
// (X @ b at d = 64 ran at 39 % of the HBM rate).
template <int LPR>
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_rows_sub(int64_t m, int64_t k, const double* __restrict__ a, int64_t lda, const double* __restrict__ x,
                const double* __restrict__ w, double* __restrict__ y, int64_t ys) {
  constexpr int RPW = 32 / LPR;                      // rows per warp per group
  const int lane = threadIdx.x & 31;
  const int sub = lane / LPR, pos = lane % LPR;
  const int64_t warps = (int64_t)gridDim.x * (GV_THREADS / 32);
  const int64_t pairs = k >> 1;
  const int64_t groups = (m + RPW - 1) / RPW;
  const double2* x2 = reinterpret_cast<const double2*>(x);
  const double2* w2 = reinterpret_cast<const double2*>(w);
  for (int64_t g0 = ((int64_t)blockIdx.x * (GV_THREADS / 32) + (threadIdx.x >> 5)) * 2; g0 < groups; g0 += 2 * warps) {
    double acc[2] = {0.0, 0.0};
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int64_t r = (g0 + h) * RPW + sub;
      if (g0 + h < groups && r < m) {
        const double2* row = reinterpret_cast<const double2*>(a + r * lda);
        double s0 = 0.0, s1 = 0.0;
        for (int64_t c = pos; c < pairs; c += LPR) {
          const double2 v = row[c];
          double2 xv = x2[c];
          if (w) { const double2 wv = w2[c]; xv.x *= wv.x; xv.y *= wv.y; }
          s0 = fma(v.x, xv.x, s0);
          s1 = fma(v.y, xv.y, s1);
        }
        acc[h] = s0 + s1;
      }
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
      double s = acc[h];
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const int64_t r = (g0 + h) * RPW + sub;
      if (pos == 0 && g0 + h < groups && r < m) y[r * ys] = s;
    }
  }
}

// part[r*S + s] = sum over the s-th k-chunk of A[r*lda + k] * w[k] * x[k*xs]  (few, long rows; dot)
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_rows_split(int64_t m, int64_t k, int64_t S, int64_t chunk, const double* __restrict__ a, int64_t lda,
                  const double* __restrict__ x, int64_t xs, const double* __restrict__ w, double* __restrict__ part) {
  __shared__ double sm[8];
  const int64_t s = blockIdx.x;
  for (int64_t r = blockIdx.y; r < m; r += gridDim.y) {
    const int64_t lo = s * chunk, hi = (lo + chunk < k) ? lo + chunk : k;
    const double* row = a + r * lda;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    int64_t i = lo + threadIdx.x;
    for (; i + 3 * GV_THREADS < hi; i += 4 * GV_THREADS) {
      double v[4], xv[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { v[u] = row[i + u * GV_THREADS]; xv[u] = x[(i + u * GV_THREADS) * xs]; }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (w) xv[u] *= w[i + u * GV_THREADS];
        acc[u] = fma(v[u], xv[u], acc[u]);
      }
    }
    for (; i < hi; i += GV_THREADS) {
      double xv = x[i * xs];
      if (w) xv *= w[i];
      acc[0] = fma(row[i], xv, acc[0]);
    }
    double v = (acc[0] + acc[1]) + (acc[2] + acc[3]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = sm[0];
      for (int q = 1; q < GV_THREADS / 32; q++) t += sm[q];
      part[r * S + s] = t;
    }
  }
}

// The same with 16-byte loads: unit-stride x, everything 16-byte aligned, even chunk (four double2 of each operand in
// flight per thread: 128 B, against 64 B of the scalar variant -- the dot of two 10^8 vectors is a pure stream)
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_rows_split_v2(int64_t m, int64_t k, int64_t S, int64_t chunk, const double* __restrict__ a, int64_t lda,
                     const double* __restrict__ x, const double* __restrict__ w, double* __restrict__ part) {
  __shared__ double sm[8];
  const int64_t s = blockIdx.x;
  for (int64_t r = blockIdx.y; r < m; r += gridDim.y) {
    const int64_t lo = s * chunk, hi = (lo + chunk < k) ? lo + chunk : k;
    const double* row = a + r * lda;
    const int64_t pairs = (hi - lo) / 2;
    const double2* r2 = reinterpret_cast<const double2*>(row + lo);
    const double2* x2 = reinterpret_cast<const double2*>(x + lo);
    const double2* w2 = w ? reinterpret_cast<const double2*>(w + lo) : nullptr;
    double acc[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    int64_t i = threadIdx.x;
    for (; i + 3 * GV_THREADS < pairs; i += 4 * GV_THREADS) {
      double2 v[4], xv[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { v[u] = r2[i + u * GV_THREADS]; xv[u] = x2[i + u * GV_THREADS]; }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (w2) { const double2 wv = w2[i + u * GV_THREADS]; xv[u].x *= wv.x; xv[u].y *= wv.y; }
        acc[2 * u] = fma(v[u].x, xv[u].x, acc[2 * u]);
        acc[2 * u + 1] = fma(v[u].y, xv[u].y, acc[2 * u + 1]);
      }
    }
    for (; i < pairs; i += GV_THREADS) {
      double2 xv = x2[i];
      if (w2) { const double2 wv = w2[i]; xv.x *= wv.x; xv.y *= wv.y; }
      const double2 v = r2[i];
      acc[0] = fma(v.x, xv.x, acc[0]);
      acc[1] = fma(v.y, xv.y, acc[1]);
    }
    if (threadIdx.x == 0 && ((hi - lo) & 1)) {          // odd tail element of the last chunk
      double xv = x[hi - 1];
      if (w) xv *= w[hi - 1];
      acc[2] = fma(row[hi - 1], xv, acc[2]);
    }
    double v = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = sm[0];
      for (int q = 1; q < GV_THREADS / 32; q++) t += sm[q];
      part[r * S + s] = t;
    }
  }
}

// part[s*m + c] = sum over the s-th k-chunk of A[k*lda + c] * w[k] * x[k*xs]   (X^T v, tall-skinny)
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_cols(int64_t m, int64_t k, int64_t chunk, int tile, const double* __restrict__ a, int64_t lda,
            const double* __restrict__ x, int64_t xs, const double* __restrict__ w, double* __restrict__ part) {
  __shared__ double sm[GV_THREADS];
  const int kpar = GV_THREADS / tile;
  const int t = threadIdx.x;
  const bool active = t < kpar * tile;
  const int ti = t % tile, tk = t / tile;
  const int64_t col = (int64_t)blockIdx.x * tile + ti;
  const int64_t s = blockIdx.y;
  const int64_t lo = s * chunk, hi = (lo + chunk < k) ? lo + chunk : k;
  double acc0 = 0.0, acc1 = 0.0;
  if (active && col < m) {
    int64_t kk = lo + tk;
    double acc2 = 0.0, acc3 = 0.0;
    for (; kk + 3 * (int64_t)kpar < hi; kk += 4 * kpar) {
      double v0 = a[kk * lda + col], v1 = a[(kk + kpar) * lda + col];
      double v2 = a[(kk + 2 * kpar) * lda + col], v3 = a[(kk + 3 * kpar) * lda + col];
      double x0 = x[kk * xs], x1 = x[(kk + kpar) * xs], x2 = x[(kk + 2 * kpar) * xs], x3 = x[(kk + 3 * kpar) * xs];
      if (w) { x0 *= w[kk]; x1 *= w[kk + kpar]; x2 *= w[kk + 2 * kpar]; x3 *= w[kk + 3 * kpar]; }
      acc0 = fma(v0, x0, acc0);
      acc1 = fma(v1, x1, acc1);
      acc2 = fma(v2, x2, acc2);
      acc3 = fma(v3, x3, acc3);
    }
    for (; kk < hi; kk += kpar) {
      double x0 = x[kk * xs];
      if (w) x0 *= w[kk];
      acc0 = fma(a[kk * lda + col], x0, acc0);
    }
    acc0 += acc2;
    acc1 += acc3;
  }
  sm[t] = acc0 + acc1;
  __syncthreads();
  if (t < tile && col < m) {
    double r = sm[t];
    for (int j = 1; j < kpar; j++) r += sm[t + j * tile];
    part[s * m + col] = r;
  }
}

// X^T v with 16-byte loads: a thread owns two adjacent columns and walks the rows of its chunk eight at a time
// (128 B in flight per thread); needs an even, 16-byte aligned leading dimension and an even column count.
__global__ void __launch_bounds__(GV_THREADS)
k_gemv_cols_v2(int64_t m, int64_t k, int64_t chunk, int tile2, const double* __restrict__ a, int64_t lda,
               const double* __restrict__ x, int64_t xs, const double* __restrict__ w, double* __restrict__ part) {
  __shared__ double2 sm[GV_THREADS];
  const int kpar = GV_THREADS / tile2;
  const int t = threadIdx.x;
  const bool active = t < kpar * tile2;
  const int ti = t % tile2, tk = t / tile2;
  const int64_t col = ((int64_t)blockIdx.x * tile2 + ti) * 2;
  const int64_t s = blockIdx.y;
  const int64_t lo = s * chunk, hi = (lo + chunk < k) ? lo + chunk : k;
  double2 acc[4];
#pragma unroll
  for (int u = 0; u < 4; u++) acc[u] = make_double2(0.0, 0.0);
  if (active && col < m) {
    int64_t kk = lo + tk;
    for (; kk + 7 * (int64_t)kpar < hi; kk += 8 * kpar) {
      double2 v[8];
      double xv[8];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        v[u] = *reinterpret_cast<const double2*>(a + (kk + (int64_t)u * kpar) * lda + col);
        xv[u] = x[(kk + (int64_t)u * kpar) * xs];
      }
      if (w) {
#pragma unroll
        for (int u = 0; u < 8; u++) xv[u] *= w[kk + (int64_t)u * kpar];
      }
#pragma unroll
      for (int u = 0; u < 8; u++) {
        acc[u & 3].x = fma(v[u].x, xv[u], acc[u & 3].x);
        acc[u & 3].y = fma(v[u].y, xv[u], acc[u & 3].y);
      }
    }
    for (; kk < hi; kk += kpar) {
      double x0 = x[kk * xs];
      if (w) x0 *= w[kk];
      const double2 v = *reinterpret_cast<const double2*>(a + kk * lda + col);
      acc[0].x = fma(v.x, x0, acc[0].x);
      acc[0].y = fma(v.y, x0, acc[0].y);
    }
  }
  sm[t] = make_double2((acc[0].x + acc[1].x) + (acc[2].x + acc[3].x), (acc[0].y + acc[1].y) + (acc[2].y + acc[3].y));
  __syncthreads();
  if (t < tile2 && col < m) {
    double2 r = sm[t];
    for (int j = 1; j < kpar; j++) { r.x += sm[t + j * tile2].x; r.y += sm[t + j * tile2].y; }
    part[s * m + col] = r.x;
    part[s * m + col + 1] = r.y;
  }
}

// out[j*os] = sum_s part[s*ss + j*js], in split order
__global__ void __launch_bounds__(256)
k_sum_parts(int64_t count, int64_t S, const double* __restrict__ part, int64_t ss, int64_t js,
            double* __restrict__ out, int64_t os) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < count; j += stride) {
    double s = 0.0;
    for (int64_t q = 0; q < S; q++) s += part[q * ss + j * js];
    out[j * os] = s;
  }
}

struct GemvPlan { int kind; int64_t S, chunk; int tile; size_t scratch; };

static GemvPlan plan_gemv(int64_t m, int64_t k, bool k_contig) {
  GemvPlan p;
  p.tile = 1; p.S = 1; p.chunk = k; p.scratch = 0;
  if (k_contig) {
    if (m >= 256 && k <= 16384) { p.kind = 0; return p; }
    p.kind = 1;
    int64_t want = GV_TARGET_BLOCKS / (m < GV_TARGET_BLOCKS ? (m > 0 ? m : 1) : GV_TARGET_BLOCKS);
    if (want < 1) want = 1;
    int64_t max_s = (k + 4095) / 4096;
    p.S = want < max_s ? want : max_s;
  } else {
    p.kind = 2;
    p.tile = m >= GV_THREADS ? GV_THREADS : (int)m;
    int64_t col_blocks = (m + p.tile - 1) / p.tile;
    int64_t want = col_blocks >= GV_TARGET_BLOCKS ? 1 : GV_TARGET_BLOCKS / col_blocks;
    int64_t kpar = GV_THREADS / p.tile;
    int64_t max_s = (k + 16 * kpar - 1) / (16 * kpar);
    p.S = want < max_s ? want : max_s;
  }
  if (p.S < 1) p.S = 1;
  p.chunk = (k + p.S - 1) / p.S;
  if (k_contig) p.chunk = (p.chunk + 1) & ~(int64_t)1;      // even chunks keep every split 16-byte aligned
  if (p.chunk < 1) p.chunk = 1;
  p.S = (k + p.chunk - 1) / p.chunk;
  if (p.S < 1) p.S = 1;
  p.scratch = (size_t)p.S * m * sizeof(double);
  return p;
}

// y[r*ys] = sum_k A(r,k) * w[k] * x[k*xs];  A(r,k) at a[r*a_rs + k*a_cs], one of a_rs/a_cs == 1.
static int gemv_f64(int device, int64_t m, int64_t k, const double* a, int64_t a_rs, int64_t a_cs,
                    const double* x, int64_t xs, const double* w, double* y, int64_t ys, void* scratch,
                    size_t scratch_bytes, cudaStream_t st) {
  const bool k_contig = (a_cs == 1) || (m == 1);
  if (m == 1 && a_cs != 1) { /* single row with stride a_cs: handled as a column walk below */ }
  if (k_contig && a_cs == 1) {
    GemvPlan p = plan_gemv(m, k, true);
    if (p.kind == 0 && k <= 512 && (k & 1) == 0 && (a_rs & 1) == 0 && xs == 1 &&
        ((((uintptr_t)a) | ((uintptr_t)x) | ((uintptr_t)w)) & 15) == 0) {
      const int lpr = k <= 32 ? 4 : (k <= 64 ? 8 : (k <= 128 ? 16 : 32));
      const int64_t rows_per_block = (GV_THREADS / 32) * (32 / lpr) * 2;
      int64_t blocks = (m + rows_per_block - 1) / rows_per_block;
      const int64_t cap = (int64_t)dev_info(device).sms * 8;
      const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
      switch (lpr) {
        case 4: k_gemv_rows_sub<4><<<grid, GV_THREADS, 0, st>>>(m, k, a, a_rs, x, w, y, ys); break;
        case 8: k_gemv_rows_sub<8><<<grid, GV_THREADS, 0, st>>>(m, k, a, a_rs, x, w, y, ys); break;
        case 16: k_gemv_rows_sub<16><<<grid, GV_THREADS, 0, st>>>(m, k, a, a_rs, x, w, y, ys); break;
        default: k_gemv_rows_sub<32><<<grid, GV_THREADS, 0, st>>>(m, k, a, a_rs, x, w, y, ys); break;
      }
      NB_LAUNCH_CHECK();
      return 0;
    }
    if (p.kind == 0) {
      int64_t blocks = (m + 7) / 8;
      int64_t cap = (int64_t)dev_info(device).sms * 8;
      k_gemv_rows_warp<<<(unsigned)(blocks < cap ? blocks : cap), GV_THREADS, 0, st>>>(m, k, a, a_rs, x, xs, w, y, ys);
      NB_LAUNCH_CHECK();
      return 0;
    }
    NB_REQUIRE(scratch != nullptr && scratch_bytes >= p.scratch, "matmul: scratch too small (%zu < %zu)",
               scratch_bytes, p.scratch);
    dim3 grid((unsigned)p.S, (unsigned)(m < 65535 ? m : 65535));
    const bool al16 = ((((uintptr_t)a) | ((uintptr_t)x) | ((uintptr_t)w)) & 15) == 0;
    if (xs == 1 && al16 && (p.chunk & 1) == 0 && (m == 1 || (a_rs & 1) == 0))
      k_gemv_rows_split_v2<<<grid, GV_THREADS, 0, st>>>(m, k, p.S, p.chunk, a, a_rs, x, w, (double*)scratch);
    else
      k_gemv_rows_split<<<grid, GV_THREADS, 0, st>>>(m, k, p.S, p.chunk, a, a_rs, x, xs, w, (double*)scratch);
    NB_LAUNCH_CHECK();
    int64_t blocks = (m + 255) / 256;
    k_sum_parts<<<(unsigned)(blocks < 1184 ? blocks : 1184), 256, 0, st>>>(m, p.S, (const double*)scratch, 1, p.S, y, ys);
    NB_LAUNCH_CHECK();
    return 0;
  }
  // m contiguous (a_rs == 1), or a single strided row (m == 1): walk k with stride a_cs
  int64_t lda = a_cs;
  GemvPlan p = plan_gemv(m, k, false);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= p.scratch, "matmul: scratch too small (%zu < %zu)",
             scratch_bytes, p.scratch);
  if ((m & 1) == 0 && (lda & 1) == 0 && (((uintptr_t)a) & 15) == 0 && m >= 2 && a_rs == 1) {
    // two columns per thread; the split count of the plan (made for one column per thread) is kept
    const int tile2 = (int)(m / 2 >= GV_THREADS / 2 ? GV_THREADS / 2 : m / 2);
    dim3 grid2((unsigned)((m / 2 + tile2 - 1) / tile2), (unsigned)p.S);
    k_gemv_cols_v2<<<grid2, GV_THREADS, 0, st>>>(m, k, p.chunk, tile2, a, lda, x, xs, w, (double*)scratch);
    NB_LAUNCH_CHECK();
  } else {
    dim3 grid((unsigned)((m + p.tile - 1) / p.tile), (unsigned)p.S);
    k_gemv_cols<<<grid, GV_THREADS, 0, st>>>(m, k, p.chunk, p.tile, a, lda, x, xs, w, (double*)scratch);
    NB_LAUNCH_CHECK();
  }
  int64_t blocks = (m + 255) / 256;
  k_sum_parts<<<(unsigned)(blocks < 1184 ? blocks : 1184), 256, 0, st>>>(m, p.S, (const double*)scratch, m, 1, y, ys);
  NB_LAUNCH_CHECK();
  return 0;
}

static size_t gemv_scratch(int64_t m, int64_t k) {
  size_t a = plan_gemv(m, k, true).scratch, b = plan_gemv(m, k, false).scratch;
  return a > b ? a : b;
}

// ------------------------------------------------------------------------------------------------
// int64 contraction (NumPy has no BLAS path for integers; wrapping arithmetic)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_matmul_i64(int64_t m, int64_t n, int64_t k, const int64_t* __restrict__ a, int64_t a_rs, int64_t a_cs,
             const int64_t* __restrict__ b, int64_t b_rs, int64_t b_cs, int64_t* __restrict__ c, int64_t ldc) {
  // one block per output element, strided over k, fixed-order block reduction
  __shared__ uint64_t sm[8];
  const int64_t total = m * n;
  for (int64_t e = blockIdx.x; e < total; e += gridDim.x) {
    int64_t r = e / n, cc = e - r * n;
    uint64_t acc = 0;
    for (int64_t i = threadIdx.x; i < k; i += 256)
      acc += (uint64_t)a[r * a_rs + i * a_cs] * (uint64_t)b[i * b_rs + cc * b_cs];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint64_t t = 0;
      for (int q = 0; q < 8; q++) t += sm[q];
      c[r * ldc + cc] = (int64_t)t;
    }
  }
}

__global__ void __launch_bounds__(256)
k_matmul_i64_thread(int64_t m, int64_t n, int64_t k, const int64_t* __restrict__ a, int64_t a_rs, int64_t a_cs,
                    const int64_t* __restrict__ b, int64_t b_rs, int64_t b_cs, int64_t* __restrict__ c, int64_t ldc) {
  const int64_t total = m * n;
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += stride) {
    int64_t r = e / n, cc = e - r * n;
    uint64_t acc = 0;
    for (int64_t i = 0; i < k; i++) acc += (uint64_t)a[r * a_rs + i * a_cs] * (uint64_t)b[i * b_rs + cc * b_cs];
    c[r * ldc + cc] = (int64_t)acc;
  }
}

// ------------------------------------------------------------------------------------------------
// dispatcher
// ------------------------------------------------------------------------------------------------
// src[r*s_rs + c] (c contiguous) -> dst[c*d_cs + r] (r contiguous), through a padded shared-memory tile.  Used to put a
// k-contiguous operand of a large square product into the [k][m] layout the TMA kernel streams in 1 KB rows: one bulk
// copy per 256-byte tile row sustains only ~1 copy per 57 cycles per SM (measured: A @ B^T ran at 19.8 TFLOP/s against
// 34.8 for A^T @ B), while the transpose costs 0.3 ms of a 57 ms product.
__global__ void __launch_bounds__(256)
k_mm_transpose(int64_t rows, int64_t cols, const double* __restrict__ src, int64_t s_rs, double* __restrict__ dst,
               int64_t d_cs) {
  __shared__ double tile[32][33];
  const int64_t tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int64_t t = blockIdx.x; t < tiles_c * tiles_r; t += gridDim.x) {
    const int64_t tr = t / tiles_c, tc = t - tr * tiles_c;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const int64_t r = tr * 32 + ty + j, c = tc * 32 + tx;
      if (r < rows && c < cols) tile[ty + j][tx] = src[r * s_rs + c];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const int64_t c = tc * 32 + ty + j, r = tr * 32 + tx;
      if (r < rows && c < cols) dst[c * d_cs + r] = tile[tx][ty + j];
    }
    __syncthreads();
  }
}

constexpr size_t WS_TRANSPOSE_CAP = (size_t)2 << 30;      // per operand
static inline size_t mm_up(size_t v) { return (v + 255) & ~(size_t)255; }

// a k-contiguous operand is worth transposing when the product runs in direct mode (enough output tiles) and the copy fits
static bool ws_transpose_ok(const WsPlan& wp, int64_t m, int64_t n, int64_t k) {
  return wp.ok && wp.splits == 1 && wp.sym == 0 && (size_t)m * k * 8 <= WS_TRANSPOSE_CAP &&
         (size_t)k * n * 8 <= WS_TRANSPOSE_CAP;
}

// matmul_ws.cu (declared in gemm.cuh): warp-specialised, TMA-fed kernel for large m, n -- direct for square
// products, split along k for tall-skinny ones
// glm_fused.cu: warp-specialised upper-triangular X^T diag(w) X for 8 <= d <= 256
bool fused_applicable(int64_t n, int64_t d, int64_t ldx, const double* x);
size_t fused_scratch_bytes(int device, int64_t n, int64_t d);
int tri_gram(int device, const double* x, int64_t n, int64_t d, int64_t ldx, const double* wext, double* out,
             int64_t ldc, void* scratch, size_t scratch_bytes, cudaStream_t st);

size_t gemm_scratch_bytes(int device, int64_t m, int64_t n, int64_t k) {
  if (m <= 0 || n <= 0 || k <= 0) return 0;
  size_t s = 0;
  if (n == 1) s = gemv_scratch(m, k);
  else if (m == 1) s = gemv_scratch(n, k);
  else {
    size_t a = plan_gemm(device, m, n, k, false).scratch;
    size_t b = (m == n) ? plan_gemm(device, m, n, k, true).scratch : 0;
    s = a > b ? a : b;
    if (m == n && n >= 8 && n <= 256) { size_t f = fused_scratch_bytes(device, k, n); if (f > s) s = f; }
    for (int sy = 0; sy < (m == n ? 2 : 1); sy++) {
      const WsPlan wp = ws_plan(device, m, n, k, sy == 1, true);
      if (wp.ok && wp.scratch > s) s = wp.scratch;
      if (sy == 0 && ws_transpose_ok(wp, m, n, k)) {
        const size_t t = mm_up((size_t)m * k * 8) + mm_up((size_t)k * n * 8);
        if (t > s) s = t;
      }
    }
  }
  return s;
}

int gemm_f64(int device, int64_t m, int64_t n, int64_t k, const double* a, int64_t a_rs, int64_t a_cs,
             const double* b, int64_t b_rs, int64_t b_cs, const double* w, double* c, int64_t ldc,
             void* scratch, size_t scratch_bytes, cudaStream_t st) {
  if (m <= 0 || n <= 0) return 0;
  if (k <= 0) {  // empty contraction: zeros
    for (int64_t r = 0; r < m; r++) NB_CHECK(cudaMemsetAsync(c + r * ldc, 0, (size_t)n * sizeof(double), st));
    return 0;
  }
  if (n == 1) return gemv_f64(device, m, k, a, a_rs, a_cs, b, b_rs, w, c, ldc, scratch, scratch_bytes, st);
  if (m == 1) return gemv_f64(device, n, k, b, b_cs, b_rs, a, a_cs, w, c, 1, scratch, scratch_bytes, st);
  NB_REQUIRE(a_rs == 1 || a_cs == 1, "matmul: operand A must be contiguous along one axis (strides %lld,%lld)",
             (long long)a_rs, (long long)a_cs);
  NB_REQUIRE(b_rs == 1 || b_cs == 1, "matmul: operand B must be contiguous along one axis (strides %lld,%lld)",
             (long long)b_rs, (long long)b_cs);
  // a_rs == 1 -> m contiguous (A stored [k][m]); prefer the k-contiguous reading when both are 1
  const bool a_mc = (a_rs == 1) && (a_cs != 1);
  const bool b_nc = (b_cs == 1) && (b_rs != 1 || k == 1);
  GemmArgs g;
  g.m = m; g.n = n; g.k = k;
  g.a = a; g.lda = a_mc ? a_cs : a_rs;
  g.b = b; g.ldb = b_nc ? b_rs : b_cs;
  g.w = w; g.c = c; g.ldc = ldc;
  const bool sym = (a == b) && (m == n) && a_mc && b_nc && (a_cs == b_rs);
  if (sym && fused_applicable(k, n, a_cs, a) && scratch != nullptr && scratch_bytes >= fused_scratch_bytes(device, k, n))
    return tri_gram(device, a, k, n, a_cs, w, c, ldc, scratch, scratch_bytes, st);
  if (!NB_ENV_FLAG("NUMS_B200_NO_WS") && !(a_mc && b_nc) && !sym && !NB_ENV_FLAG("NUMS_B200_NO_WS_TRANSPOSE")) {
    // large product with a k-contiguous operand: copy it into the [k][.] layout, then take the fast variant
    const WsPlan wp = ws_plan(device, m, n, k, false, true);
    const size_t need = mm_up((size_t)m * k * 8) + mm_up((size_t)k * n * 8);
    if (ws_transpose_ok(wp, m, n, k) && scratch != nullptr && scratch_bytes >= need) {
      const double* a2 = a;
      const double* b2 = b;
      int64_t lda2 = g.lda, ldb2 = g.ldb;
      char* sp = (char*)scratch;
      const int tgrid = dev_info(device).sms * 8;
      if (!a_mc) {
        double* at = (double*)sp;
        k_mm_transpose<<<tgrid, 256, 0, st>>>(m, k, a, g.lda, at, m);      // A(m,k) at a[m*lda+k] -> at[k*m + m]
        NB_LAUNCH_CHECK();
        a2 = at; lda2 = m;
      }
      if (!b_nc) {
        double* bt = (double*)(sp + mm_up((size_t)m * k * 8));
        k_mm_transpose<<<tgrid, 256, 0, st>>>(n, k, b, g.ldb, bt, n);      // B(k,n) at b[n*ldb+k] -> bt[k*n + n]
        NB_LAUNCH_CHECK();
        b2 = bt; ldb2 = n;
      }
      if (gemm_ws_aligned(a2, lda2, b2, ldb2, w, c, ldc))
        return gemm_ws_f64(wp, m, n, k, a2, lda2, true, b2, ldb2, true, w, c, ldc, nullptr, st);
    }
  }
  if (!NB_ENV_FLAG("NUMS_B200_NO_WS") && gemm_ws_aligned(a, g.lda, b, g.ldb, w, c, ldc)) {
    const WsPlan wp = ws_plan(device, m, n, k, sym, a_mc && b_nc);
    if (wp.ok && (wp.scratch == 0 || (scratch != nullptr && scratch_bytes >= wp.scratch))) {
      int rc = gemm_ws_f64(wp, m, n, k, a, g.lda, a_mc, b, g.ldb, b_nc, w, c, ldc, (double*)scratch, st);
      if (rc || wp.scratch == 0) return rc;
      g.tiles_m = wp.tiles_m; g.tiles_n = wp.tiles_n; g.ntiles = wp.ntiles; g.splits = wp.splits; g.sym = wp.sym;
      g.kchunk = wp.kchunk;
      g.part = (double*)scratch;
      int64_t blocks = (m * n + 255) / 256;
      if (blocks > 148 * 8) blocks = 148 * 8;
      k_gemm_reduce<128, 128><<<(unsigned)blocks, 256, 0, st>>>(g);
      NB_LAUNCH_CHECK();
      return 0;
    }
  }
  GemmPlan p = plan_gemm(device, m, n, k, sym);
  g.tiles_m = p.tiles_m; g.tiles_n = p.tiles_n; g.ntiles = p.ntiles; g.splits = p.splits; g.sym = p.sym;
  g.kchunk = p.kchunk;
  g.part = (double*)scratch;
  if (p.splits > 1)
    NB_REQUIRE(scratch != nullptr && scratch_bytes >= p.scratch, "matmul: scratch too small (%zu < %zu)",
               scratch_bytes, p.scratch);
  const bool al = (((uintptr_t)a | (uintptr_t)b) & 15) == 0 && (g.lda % 2 == 0) && (g.ldb % 2 == 0);
  if (a_mc && b_nc) return launch_gemm_al<true, true>(al, p.cfg, g, st);
  if (a_mc && !b_nc) return launch_gemm_al<true, false>(al, p.cfg, g, st);
  if (!a_mc && b_nc) return launch_gemm_al<false, true>(al, p.cfg, g, st);
  return launch_gemm_al<false, false>(al, p.cfg, g, st);
}

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_matmul_scratch_bytes(int device, int64_t m, int64_t n, int64_t k) {
  return gemm_scratch_bytes(device, m, n, k);
}

int nb_matmul_f64(int device, int64_t m, int64_t n, int64_t k, const double* a_d, int64_t a_rs, int64_t a_cs,
                  const double* b_d, int64_t b_rs, int64_t b_cs, const double* w_d, double* c_d, int64_t ldc,
                  void* scratch_d, size_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  return gemm_f64(device, m, n, k, a_d, a_rs, a_cs, b_d, b_rs, b_cs, w_d, c_d, ldc, scratch_d, scratch_bytes, st);
}

int nb_matmul_i64(int device, int64_t m, int64_t n, int64_t k, const int64_t* a_d, int64_t a_rs, int64_t a_cs,
                  const int64_t* b_d, int64_t b_rs, int64_t b_cs, int64_t* c_d, int64_t ldc, void* stream) {
  NB_ENTER(device);
  if (m <= 0 || n <= 0) return 0;
  int64_t total = m * n;
  if (total <= 4096 && k >= 1024) {
    k_matmul_i64<<<(unsigned)total, 256, 0, st>>>(m, n, k, a_d, a_rs, a_cs, b_d, b_rs, b_cs, c_d, ldc);
  } else {
    int grid = grid_for(device, total, 256, 1);
    k_matmul_i64_thread<<<grid, 256, 0, st>>>(m, n, k, a_d, a_rs, a_cs, b_d, b_rs, b_cs, c_d, ldc);
  }
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
