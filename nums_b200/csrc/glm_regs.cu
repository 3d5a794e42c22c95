// GLM statistics for narrow designs: one pass over X with everything in registers (Newton statistics for
// 8 <= d <= 64, the gradient/objective evaluation of L-BFGS for 8 <= d <= 192).
//
// For d <= 64 the whole upper triangle of X^T diag(w) X fits in a warp's registers as 8 x 8 DMMA tiles (36 tiles x 2
// doubles at d = 64), so no shared-memory staging, no barriers and no second pass are needed: a warp streams four
// rows per step straight from HBM, lane (q, g) = (lane & 3, lane >> 2) loading X[row q][8 I + g] for every column
// block I -- that one value is at once the A fragment (after scaling by w) and the B fragment of the tile products,
// its contribution to eta = x.beta (reduced over g by three shuffles), to the gradient and to the intercept column.
// The link functions (glms.py:457-548) are evaluated once per row: a batch of up to 32 rows is loaded, every lane
// takes one row's eta, and the residual / weight come back to the lanes that hold that row's x by shuffles (FP64
// exp/log issued per four rows would cost more than the tile products).  Reference path:
// glms.py:660-676 (newton: gradient + hessian) and lbfgs.py:32-35 (objective + gradient); the two-pass kernels of
// glm.cu / glm_fused.cu serve d > 64, where the triangle no longer fits.
//
// Roofline: X is read exactly once (8 d bytes / row); at d = 64 the 36 tile products per four rows make the kernel
// FP64-tensor-bound at ~60 % of the HBM rate, below that it is HBM-bound.  Reductions are deterministic: warps add
// into shared memory in warp order, CTAs write partials, k_regs_finish sums them in CTA order and mirrors H.
#include "glm_common.cuh"

namespace nb {

constexpr int RG_WARPS = 4;
constexpr int RG_THREADS = RG_WARPS * 32;

__device__ __forceinline__ void rg_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

struct RegsArgs {
  const double* x; int64_t n; int64_t d; int64_t ldx;
  const void* y; int ydt;
  const double* beta;
  int family;
  int skip_obj;       // the objective is not needed (Newton): its two logarithms per row are left out
  double* wbuf;       // mode 2: the Hessian weights w (n doubles) for the second pass, or null
  double* parts;      // per CTA: [dd*dd tiles | g dd | u dd | sum r | sum w | obj]
};

// per-CTA partial (and shared-memory reduction buffer): [DD*DD tiles (Hessian variant only) | g DD | u DD | 3 sums]
// MODE 0: gradient + objective; 1: Newton statistics with the tile products; 2: first pass of the two-pass Newton
// (g, the intercept column u, sum w, objective, and the weights written out for glm_fused.cu)
template <int NB, int MODE> struct RegsLayout {
  static constexpr int DD = 8 * NB;
  static constexpr int G_OFF = MODE == 1 ? DD * DD : 0, U_OFF = G_OFF + DD, S_OFF = U_OFF + DD, LEN = S_OFF + 3;
};

// rows per batch = 4 * BATCH: the link function of a batch is evaluated once, lane (q, g) taking row 4 (g % BATCH) + q,
// and handed to the lanes that need it by shuffles; BATCH is bounded by the registers the x values of a batch take
template <int NB, int MODE> struct RegsBatch {
  static constexpr int value = MODE == 0 ? (NB <= 8 ? 8 : (NB <= 16 ? 4 : 2))
                             : MODE == 1 ? (NB <= 4 ? 8 : (NB <= 6 ? 4 : 2))
                                         : (NB <= 8 ? 8 : (NB <= 12 ? 4 : 2));
};

template <int NB, int MODE>
__global__ void __launch_bounds__(RG_THREADS)
k_glm_regs(RegsArgs a) {
  using L = RegsLayout<NB, MODE>;
  constexpr int DD = L::DD;
  constexpr int BATCH = RegsBatch<NB, MODE>::value;
  constexpr bool HESS = MODE == 1;                // tile products
  constexpr bool WANT_U = MODE >= 1;              // intercept column and sum of weights
  constexpr bool BET_SMEM = NB > 16;              // wide gradient-only variants keep beta in shared memory
  extern __shared__ double red[];                 // L::LEN doubles (+ DD of beta)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q = lane & 3, g = lane >> 2;
  const unsigned full = 0xffffffffu;
  double bet[BET_SMEM ? 1 : NB];
  double* s_bet = red + L::LEN;
  if constexpr (BET_SMEM) {
    for (int j = threadIdx.x; j < DD; j += RG_THREADS) s_bet[j] = j < a.d ? a.beta[j] : 0.0;
    __syncthreads();
  } else {
#pragma unroll
    for (int I = 0; I < NB; ++I) bet[I] = (8 * I + g < a.d) ? a.beta[8 * I + g] : 0.0;
  }
  const double b0 = a.beta[a.d];
  double acc[HESS ? NB * (NB + 1) / 2 : 1][2];
#pragma unroll
  for (int t = 0; t < (HESS ? NB * (NB + 1) / 2 : 1); ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
  double gacc[NB], uacc[WANT_U ? NB : 1];
#pragma unroll
  for (int I = 0; I < NB; ++I) gacc[I] = 0.0;
#pragma unroll
  for (int I = 0; I < (WANT_U ? NB : 1); ++I) uacc[I] = 0.0;
  double sr = 0.0, sw = 0.0, sobj = 0.0;

  constexpr int ROWS = 4 * BATCH;
  const int64_t batches = (a.n + ROWS - 1) / ROWS;
  const int64_t stride = (int64_t)gridDim.x * RG_WARPS;
  const int gl = g % BATCH;                        // the group whose row (gl, q) this lane evaluates the link for
  for (int64_t bt = (int64_t)blockIdx.x * RG_WARPS + warp; bt < batches; bt += stride) {
    const int64_t row0 = bt * ROWS;
    double xv[BATCH][NB];
#pragma unroll
    for (int u = 0; u < BATCH; ++u) {
      const int64_t r = row0 + 4 * u + q;
      const double* row = a.x + r * a.ldx;
#pragma unroll
      for (int I = 0; I < NB; ++I) xv[u][I] = (r < a.n && 8 * I + g < a.d) ? __ldg(row + 8 * I + g) : 0.0;
    }
    const int64_t rl = row0 + 4 * gl + q;
    const bool live = rl < a.n && g < BATCH;       // lanes g >= BATCH would repeat a row: they contribute nothing
    const double yv = rl < a.n ? load_y(a.y, a.ydt, rl) : 0.0;
    double my_eta = 0.0;
#pragma unroll
    for (int u = 0; u < BATCH; ++u) {
      double eta = 0.0;
#pragma unroll
      for (int I = 0; I < NB; ++I) eta = fma(xv[u][I], BET_SMEM ? s_bet[8 * I + g] : bet[BET_SMEM ? 0 : I], eta);
      eta += __shfl_xor_sync(full, eta, 4);
      eta += __shfl_xor_sync(full, eta, 8);
      eta += __shfl_xor_sync(full, eta, 16);
      if (u == gl) my_eta = eta;
    }
    LinkOut lo = a.skip_obj ? glm_link_noobj_rt(a.family, my_eta + b0, yv) : glm_link_rt(a.family, my_eta + b0, yv);
    if (!live) { lo.resid = 0.0; lo.wgt = 0.0; lo.obj = 0.0; }
    sr += lo.resid; sobj += lo.obj;
    if constexpr (WANT_U) sw += lo.wgt;
    if constexpr (MODE == 2) { if (live && a.wbuf) a.wbuf[rl] = lo.wgt; }
#pragma unroll
    for (int u = 0; u < BATCH; ++u) {
      const double ru = __shfl_sync(full, lo.resid, (u << 2) | q);
#pragma unroll
      for (int I = 0; I < NB; ++I) gacc[I] = fma(ru, xv[u][I], gacc[I]);
      if constexpr (WANT_U) {
        const double wu = __shfl_sync(full, lo.wgt, (u << 2) | q);
        double af[NB];
#pragma unroll
        for (int I = 0; I < NB; ++I) { af[I] = wu * xv[u][I]; uacc[I] += af[I]; }
        if constexpr (HESS) {
          int t = 0;
#pragma unroll
          for (int I = 0; I < NB; ++I)
#pragma unroll
            for (int J = I; J < NB; ++J) { rg_dmma(acc[t][0], acc[t][1], af[I], xv[u][J]); ++t; }
        }
      }
    }
  }
  // per-feature sums over the four rows a lane group covers (q); row sums over all lanes (each live lane = one row)
#pragma unroll
  for (int I = 0; I < NB; ++I) {
    gacc[I] += __shfl_xor_sync(full, gacc[I], 1);
    gacc[I] += __shfl_xor_sync(full, gacc[I], 2);
    if constexpr (WANT_U) {
      uacc[I] += __shfl_xor_sync(full, uacc[I], 1);
      uacc[I] += __shfl_xor_sync(full, uacc[I], 2);
    }
  }
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    sr += __shfl_xor_sync(full, sr, o);
    sw += __shfl_xor_sync(full, sw, o);
    sobj += __shfl_xor_sync(full, sobj, o);
  }
  // warps add into shared memory in warp order
  for (int w = 0; w < RG_WARPS; ++w) {
    if (warp == w) {
      const bool first = (w == 0);
      if constexpr (HESS) {
        int t = 0;
#pragma unroll
        for (int I = 0; I < NB; ++I)
#pragma unroll
          for (int J = I; J < NB; ++J) {
            double* p = red + (8 * I + g) * DD + 8 * J + 2 * q;
            if (first) { p[0] = acc[t][0]; p[1] = acc[t][1]; } else { p[0] += acc[t][0]; p[1] += acc[t][1]; }
            ++t;
          }
      }
      if (q == 0) {
#pragma unroll
        for (int I = 0; I < NB; ++I) {
          double* pg = red + L::G_OFF + 8 * I + g;
          if (first) *pg = gacc[I]; else *pg += gacc[I];
          if constexpr (WANT_U) {
            double* pu = red + L::U_OFF + 8 * I + g;
            if (first) *pu = uacc[I]; else *pu += uacc[I];
          }
        }
      }
      if (lane == 0) {
        double* ps = red + L::S_OFF;
        if (first) { ps[0] = sr; ps[1] = sw; ps[2] = sobj; } else { ps[0] += sr; ps[1] += sw; ps[2] += sobj; }
      }
    }
    __syncthreads();
  }
  double* out = a.parts + (size_t)blockIdx.x * L::LEN;
  for (int e = threadIdx.x; e < L::LEN; e += RG_THREADS) out[e] = red[e];
}

// stats = [H p x p | g p | obj] from the CTA partials, summed in CTA order; H mirrored from its upper triangle
template <int NB, int MODE>
__global__ void __launch_bounds__(256)
k_regs_finish(int64_t d, int nparts, const double* __restrict__ parts, double* __restrict__ stats) {
  using L = RegsLayout<NB, MODE>;
  constexpr bool HESS = MODE >= 1;                // the intercept column and sum w are part of the result
  constexpr int DD = L::DD;
  const int64_t p = d + 1;
  // short vectors: one warp per statistic; with the tile area: one thread per statistic
  constexpr bool WARP_EACH = L::LEN <= 2048;
  const int first = WARP_EACH ? blockIdx.x * 8 + (threadIdx.x >> 5) : blockIdx.x * 256 + threadIdx.x;
  const int step = WARP_EACH ? gridDim.x * 8 : gridDim.x * 256;
  for (int e = first; e < L::LEN; e += step) {
    int i = 0, j = 0;
    if (e < L::G_OFF) {
      // tiles on the diagonal hold both halves (computed with different roundings): only the upper one is used
      i = e / DD; j = e % DD;
      if (i > j || j >= d) continue;
    } else if (e < L::S_OFF) {
      if ((e - L::G_OFF) % DD >= d) continue;
      if (e >= L::U_OFF && !HESS) continue;
    } else if (e == L::S_OFF + 1 && !HESS) {
      continue;
    }
    const double s = WARP_EACH ? sum_partials_warp(parts + e, (size_t)L::LEN, nparts)
                               : sum_partials(parts + e, (size_t)L::LEN, nparts);
    if (WARP_EACH && (threadIdx.x & 31) != 0) continue;
    if (e < L::G_OFF) {
      stats[i * p + j] = s;
      stats[j * p + i] = s;
    } else if (e < L::U_OFF) {
      stats[p * p + (e - L::G_OFF)] = s;
    } else if (e < L::S_OFF) {
      const int64_t k = e - L::U_OFF;
      stats[k * p + d] = s;
      stats[d * p + k] = s;
    } else if (e == L::S_OFF) {
      stats[p * p + d] = s;
    } else if (e == L::S_OFF + 1) {
      stats[d * p + d] = s;
    } else {
      stats[p * p + p] = s;
    }
  }
}

// Newton statistics fit a warp's registers up to d = 64; the gradient-only pass (no tiles) up to d = 192
bool glm_regs_applicable(int64_t d, int want_hess) {
  return d >= 8 && d <= (want_hess ? 64 : 192) && !NB_ENV_FLAG("NUMS_B200_NO_REGS");
}

constexpr int RG_MAX_CTAS_PER_SM = 8;
static int regs_ctas(int device) { return dev_info(device).sms * RG_MAX_CTAS_PER_SM; }   // scratch is sized for this

static int regs_nb(int64_t d) {        // column blocks of the instantiated variant serving d
  const int nb = (int)((d + 7) / 8);
  if (nb <= 8) return nb;
  if (nb <= 16) return (nb + 1) & ~1;
  return (nb + 3) & ~3;
}

size_t glm_regs_scratch_bytes(int device, int64_t d, int want_hess) {
  const size_t dd = (size_t)regs_nb(d) * 8;
  return (size_t)regs_ctas(device) * ((want_hess ? dd * dd : 0) + 2 * dd + 3) * sizeof(double);
}

template <int NB, int MODE>
static int launch_regs(int device, const RegsArgs& a, double* stats, cudaStream_t st) {
  using L = RegsLayout<NB, MODE>;
  const int64_t rows_per_batch = 4 * RegsBatch<NB, MODE>::value;
  int64_t batches = (a.n + rows_per_batch - 1) / rows_per_batch;
  int64_t want = (batches + RG_WARPS - 1) / RG_WARPS;
  const size_t smem = (size_t)(L::LEN + (NB > 16 ? L::DD : 0)) * sizeof(double);
  // one resident wave: as many CTAs per SM as the kernel's registers allow (the narrow variants need the extra
  // warps to keep enough loads in flight)
  int per_sm = 1;
  static KernelSetup setup;
  if (int rc = setup.get(device, k_glm_regs<NB, MODE>, RG_THREADS, smem, &per_sm)) return rc;
  if (per_sm > RG_MAX_CTAS_PER_SM) per_sm = RG_MAX_CTAS_PER_SM;
  int ctas = dev_info(device).sms * per_sm;
  if (want < ctas) ctas = (int)(want < 1 ? 1 : want);
  k_glm_regs<NB, MODE><<<ctas, RG_THREADS, smem, st>>>(a);
  NB_LAUNCH_CHECK();
  k_regs_finish<NB, MODE><<<L::LEN <= 2048 ? (L::LEN + 7) / 8 : (L::LEN + 255) / 256, 256, 0, st>>>(a.d, ctas, a.parts, stats);
  NB_LAUNCH_CHECK();
  return 0;
}

int glm_regs_stats(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                   const double* beta, int want_hess, int skip_obj, double* stats, void* scratch, size_t scratch_bytes,
                   cudaStream_t st) {
  NB_REQUIRE(glm_regs_applicable(d, want_hess), "glm_regs_stats: d=%lld outside the register-resident range", (long long)d);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= glm_regs_scratch_bytes(device, d, want_hess),
             "glm_regs_stats: scratch too small");
  RegsArgs a;
  a.x = x; a.n = n; a.d = d; a.ldx = ldx; a.y = y; a.ydt = ydt; a.beta = beta; a.family = family;
  a.wbuf = nullptr;
  a.skip_obj = skip_obj;
  a.parts = (double*)scratch;
  const int nb = regs_nb(d);
#define RG_CASE(N) case N: return want_hess ? launch_regs<(N <= 8 ? N : 8), 1>(device, a, stats, st) \
                                            : launch_regs<N, 0>(device, a, stats, st)
  switch (nb) {
    RG_CASE(1); RG_CASE(2); RG_CASE(3); RG_CASE(4); RG_CASE(5); RG_CASE(6); RG_CASE(7); RG_CASE(8);
    RG_CASE(10); RG_CASE(12); RG_CASE(14); RG_CASE(16); RG_CASE(20); RG_CASE(24);
  }
#undef RG_CASE
  set_error("glm_regs_stats: no variant for d=%lld", (long long)d);
  return 2;
}

// First pass of the two-pass Newton statistics for 64 < d <= 160 (beyond that the accumulators spill): g, u, sum r, sum w, objective into `stats`, the
// weights into wbuf (null for the linear family, whose weights are 1); glm_fused.cu adds X^T diag(w) X.
bool glm_regs_pass1_applicable(int64_t d) { return d > 64 && d <= 160 && !NB_ENV_FLAG("NUMS_B200_NO_REGS"); }

int glm_regs_pass1(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                   const double* beta, int skip_obj, double* wbuf, double* stats, void* scratch, size_t scratch_bytes,
                   cudaStream_t st) {
  NB_REQUIRE(glm_regs_pass1_applicable(d), "glm_regs_pass1: d=%lld outside 65..160", (long long)d);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= glm_regs_scratch_bytes(device, d, 0), "glm_regs_pass1: scratch too small");
  RegsArgs a;
  a.x = x; a.n = n; a.d = d; a.ldx = ldx; a.y = y; a.ydt = ydt; a.beta = beta; a.family = family;
  a.wbuf = wbuf;
  a.skip_obj = skip_obj;
  a.parts = (double*)scratch;
  switch (regs_nb(d)) {
    case 10: return launch_regs<10, 2>(device, a, stats, st);
    case 12: return launch_regs<12, 2>(device, a, stats, st);
    case 14: return launch_regs<14, 2>(device, a, stats, st);
    case 16: return launch_regs<16, 2>(device, a, stats, st);
    case 20: return launch_regs<20, 2>(device, a, stats, st);
  }
  set_error("glm_regs_pass1: no variant for d=%lld", (long long)d);
  return 2;
}

}  // namespace nb
