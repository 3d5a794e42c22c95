// Pass 1 of the GLM statistics (and the whole L-BFGS / gradient evaluation): eta = X.beta + beta0, link,
// objective, g = [X,1]^T r, u = [X,1]^T w, weights w -> scratch.   HBM-bound: X is read exactly once.
// Reference: nums/models/glms.py:457-489 (logistic), :283-316 (linear), :516-544 (poisson).
//
// Why shared-memory staging: the link (exp, divide, two logs ~ 220 FP64 instructions) is SIMT work, so it
// should be evaluated for as many rows per instruction as possible.  Rows are staged whole by TMA bulk
// copies (one per row, issued by the lanes of a producer warp, completing on the slot's mbarrier), and a
// stage of FK rows gets ONE link evaluation with a row per lane, on a warp that rotates from stage to stage so
// that its ~2000-cycle dependent chain overlaps three later stages.  Compared with the register-resident
// k_glm_rows (link once per 1-2 rows) this cuts the FP64 instruction count per byte ~4x, which is what kept
// that kernel at ~60 % of the HBM peak.
//
// Roles per CTA: warp 8 = producer; warps 0-7 = workers.  Per stage kt every worker
//   (1) computes its share of eta for stage kt+3 from shared memory          -> etaDone[slot] (8 arrivals)
//   (2) if it is the stage's link warp ((kt+3) % 8): waits etaDone, evaluates the link for the FK rows,
//       stores r, w (shared) and w (global scratch)                           -> linkDone[slot] (FK arrivals)
//   (3) waits linkDone[kt] and accumulates g, u for its columns (thread = column) over the FK rows
//   (4) releases the slot                                                     -> empty[slot] (8 arrivals)
// No block-wide barrier in steady state.  Per-CTA partials go to scratch and are summed in CTA order by
// k_glm_rows_finish (deterministic).
#include "glm_common.cuh"

namespace nb {

__device__ __forceinline__ void r_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void r_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void r_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void r_mbar_wait(uint64_t* bar, unsigned parity) {
  unsigned addr = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(addr), "r"(parity) : "memory");
}
__device__ __forceinline__ void r_bulk_g2s(void* smem, const void* gmem, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem), "r"(bytes),
                  "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

constexpr int RW = 8;            // worker warps
constexpr int RS = 6;            // stage slots
constexpr int RAHEAD = 3;        // eta / link run this many stages ahead of the g,u accumulation
constexpr int RMAXC = 8;         // columns per thread (d <= 2048)

struct RowsTmaArgs {
  const double* x; int64_t n, d, ldx;
  const void* y; int ydt;
  const double* beta;
  int family, want_hess;
  int fk;                  // rows per stage (1, 2, 4, 8 or 16)
  int64_t rows_per_cta;    // multiple of fk
  double* wbuf;            // n weights for pass 2 (null: not needed)
  double* partials;        // [gridDim.x][2*d + 3]
};

__global__ void __launch_bounds__((RW + 1) * 32)
k_glm_rows_tma(const RowsTmaArgs a) {
  extern __shared__ __align__(16) double rsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t d = a.d;
  const int fk = a.fk;
  const int ld = (int)d + 2;                        // row stride in shared memory (16-byte aligned rows)
  const int64_t row_beg = (int64_t)blockIdx.x * a.rows_per_cta;
  const int64_t row_end = (row_beg + a.rows_per_cta < a.n) ? row_beg + a.rows_per_cta : a.n;
  const int ktiles = row_end > row_beg ? (int)((row_end - row_beg + fk - 1) / fk) : 0;
  // eta work split: `nch` column chunks per row so that all 8 workers are busy when fk < 8
  const int nch = fk >= RW ? 1 : RW / fk;

  double* sX = rsm;                                 // RS * fk * ld
  double* sBeta = sX + (size_t)RS * fk * ld;        // d + 2
  double* sEta = sBeta + ((d + 2 + 1) & ~(int64_t)1);   // RS * 16 * 8   (row, chunk) partial dot products
  double* sRes = sEta + RS * 16 * 8;                // RS * 16
  double* sWgt = sRes + RS * 16;                    // RS * 16
  double* sFin = sWgt + RS * 16;                    // RW * 4 final scalars
  uint64_t* fullx = reinterpret_cast<uint64_t*>(sFin + RW * 4);
  uint64_t* etaDone = fullx + RS;
  uint64_t* linkDone = etaDone + RS;
  uint64_t* empty = linkDone + RS;

  for (int64_t j = tid; j <= d; j += (RW + 1) * 32) sBeta[j] = a.beta[j];
  if (tid == 0) {
    for (int s = 0; s < RS; s++) {
      r_mbar_init(&fullx[s], fk);
      r_mbar_init(&etaDone[s], RW);
      r_mbar_init(&linkDone[s], fk);
      r_mbar_init(&empty[s], RW);
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  if (warp == RW) {
    // ---------------------------------------------------------------- producer
    const unsigned row_bytes = (unsigned)(d * sizeof(double));
    for (int kt = 0; kt < ktiles; kt++) {
      const int slot = kt % RS;
      if (kt >= RS) r_mbar_wait(&empty[slot], (unsigned)(((kt / RS) - 1) & 1));
      if (lane < fk) {
        const int64_t gr = row_beg + (int64_t)kt * fk + lane;
        if (gr < row_end) {
          r_mbar_expect_tx(&fullx[slot], row_bytes);
          r_bulk_g2s(sX + ((size_t)slot * fk + lane) * ld, a.x + gr * a.ldx, row_bytes, &fullx[slot]);
        } else {
          r_mbar_arrive(&fullx[slot]);
        }
      }
    }
    return;
  }

  // ---------------------------------------------------------------- workers
  const double beta0 = sBeta[d];
  const int wtid = tid;                              // 0 .. 255: column owner
  double g[RMAXC], u[RMAXC];
#pragma unroll
  for (int j = 0; j < RMAXC; j++) { g[j] = 0.0; u[j] = 0.0; }
  double acc_obj = 0.0, acc_r = 0.0, acc_w = 0.0;    // lanes < fk of the link warps

  // eta share of this warp for one stage: units (row, chunk), unit index = warp + RW * q
  auto eta_stage = [&](int kt) {
    const int slot = kt % RS;
    r_mbar_wait(&fullx[slot], (unsigned)((kt / RS) & 1));
    const double* xs = sX + (size_t)slot * fk * ld;
    const int units = fk * nch;
    const int64_t cw = (d + nch - 1) / nch;          // columns per chunk
    for (int un = warp; un < units; un += RW) {
      const int r = un % fk, ch = un / fk;
      const int64_t c0 = ch * cw, c1 = (c0 + cw < d) ? c0 + cw : d;
      double s0 = 0.0, s1 = 0.0;
      int64_t cidx = c0 + lane;
      for (; cidx + 32 < c1; cidx += 64) {
        s0 = fma(xs[r * ld + cidx], sBeta[cidx], s0);
        s1 = fma(xs[r * ld + cidx + 32], sBeta[cidx + 32], s1);
      }
      if (cidx < c1) s0 = fma(xs[r * ld + cidx], sBeta[cidx], s0);
      double s = s0 + s1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) sEta[(slot * 16 + r) * 8 + ch] = s;
    }
    __syncwarp();
    if (lane == 0) r_mbar_arrive(&etaDone[slot]);
  };
  auto link_stage = [&](int kt) {
    const int slot = kt % RS;
    r_mbar_wait(&etaDone[slot], (unsigned)((kt / RS) & 1));
    if (lane < fk) {
      const int64_t gr = row_beg + (int64_t)kt * fk + lane;
      LinkOut lo;
      lo.resid = 0.0; lo.wgt = 0.0; lo.obj = 0.0;
      if (gr < row_end) {
        double eta = beta0;
        for (int ch = 0; ch < nch; ch++) eta += sEta[(slot * 16 + lane) * 8 + ch];
        lo = glm_link_rt(a.family, eta, load_y(a.y, a.ydt, gr));
        if (a.wbuf) a.wbuf[gr] = lo.wgt;
      }
      sRes[slot * 16 + lane] = lo.resid;
      sWgt[slot * 16 + lane] = lo.wgt;
      acc_obj += lo.obj; acc_r += lo.resid; acc_w += lo.wgt;
      r_mbar_arrive(&linkDone[slot]);
    }
  };

  // prologue: eta for stages 0 .. RAHEAD-1, link for those stages
  for (int s = 0; s < RAHEAD && s < ktiles; s++) eta_stage(s);
  for (int s = 0; s < RAHEAD && s < ktiles; s++)
    if (warp == (s % RW)) link_stage(s);

  for (int kt = 0; kt < ktiles; kt++) {
    const int slot = kt % RS;
    if (kt + RAHEAD < ktiles) {
      eta_stage(kt + RAHEAD);
      if (warp == ((kt + RAHEAD) % RW)) link_stage(kt + RAHEAD);
    }
    r_mbar_wait(&linkDone[slot], (unsigned)((kt / RS) & 1));
    const double* xs = sX + (size_t)slot * fk * ld;
    const double* rs = sRes + slot * 16;
    const double* ws = sWgt + slot * 16;
    const int64_t valid = (row_end - (row_beg + (int64_t)kt * fk) < fk) ? row_end - (row_beg + (int64_t)kt * fk) : fk;
    for (int r = 0; r < (int)valid; r++) {
      const double rr = rs[r], ww = ws[r];
#pragma unroll
      for (int j = 0; j < RMAXC; j++) {
        const int64_t col = wtid + 256 * j;
        if (col < d) {
          const double xv = xs[r * ld + col];
          g[j] = fma(rr, xv, g[j]);
          if (a.want_hess) u[j] = fma(ww, xv, u[j]);
        }
      }
    }
    __syncwarp();
    if (lane == 0) r_mbar_arrive(&empty[slot]);
  }

  // ---- CTA partials (worker threads own disjoint columns)
  double* out = a.partials + (size_t)blockIdx.x * (2 * d + 3);
#pragma unroll
  for (int j = 0; j < RMAXC; j++) {
    const int64_t col = wtid + 256 * j;
    if (col < d) { out[col] = g[j]; out[d + col] = u[j]; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc_obj += __shfl_down_sync(0xffffffffu, acc_obj, o);
    acc_r += __shfl_down_sync(0xffffffffu, acc_r, o);
    acc_w += __shfl_down_sync(0xffffffffu, acc_w, o);
  }
  if (lane == 0) { sFin[warp * 4] = acc_r; sFin[warp * 4 + 1] = acc_w; sFin[warp * 4 + 2] = acc_obj; }
  asm volatile("bar.sync 1, %0;" :: "n"(RW * 32));
  if (tid == 0) {
    double sr = 0.0, sw = 0.0, so = 0.0;
    for (int w = 0; w < RW; w++) { sr += sFin[w * 4]; sw += sFin[w * 4 + 1]; so += sFin[w * 4 + 2]; }
    out[2 * d] = sr; out[2 * d + 1] = sw; out[2 * d + 2] = so;
  }
}

bool rows_tma_applicable(int64_t n, int64_t d, int64_t ldx, const double* x) {
  return d >= 8 && d <= 2048 && n > 0 && (d % 2 == 0) && (ldx % 2 == 0) && ((((uintptr_t)x) & 15) == 0);
}

// Launches pass 1; *nparts = number of per-CTA partial vectors written (layout of k_glm_rows_finish).
int glm_rows_tma(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                 const double* beta, int want_hess, double* wbuf, double* partials, int max_parts, int* nparts,
                 cudaStream_t st) {
  RowsTmaArgs a;
  a.x = x; a.n = n; a.d = d; a.ldx = ldx; a.y = y; a.ydt = ydt; a.beta = beta; a.family = family;
  a.want_hess = want_hess; a.wbuf = wbuf; a.partials = partials;
  const int ld = (int)d + 2;
  int fk = 16;
  while (fk > 1 && (size_t)fk * ld * sizeof(double) > (size_t)34 * 1024) fk >>= 1;
  a.fk = fk;
  int grid = dev_info(device).sms;
  if (grid > max_parts) grid = max_parts;
  int64_t tiles = (n + fk - 1) / fk;
  if (grid > tiles) grid = (int)tiles;
  int64_t tpc = (tiles + grid - 1) / grid;
  a.rows_per_cta = tpc * fk;
  grid = (int)((n + a.rows_per_cta - 1) / a.rows_per_cta);
  size_t smem = ((size_t)RS * fk * ld + ((d + 2 + 1) & ~(int64_t)1) + RS * 16 * 8 + 2 * RS * 16 + RW * 4) * sizeof(double) +
                4 * RS * sizeof(uint64_t);
  NB_CHECK(cudaFuncSetAttribute(k_glm_rows_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_glm_rows_tma<<<grid, (RW + 1) * 32, smem, st>>>(a);
  NB_LAUNCH_CHECK();
  *nparts = grid;
  return 0;
}

}  // namespace nb
