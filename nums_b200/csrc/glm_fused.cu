// Single-pass fused Newton statistics for 8 <= d <= 256 (nums/models/glms.py:457-499 as called from
// newton() :660-676): one read of X and y yields the objective, g = [X,1]^T r and
// H = [X,1]^T diag(w) [X,1].
//
// A group of C CTAs walks one contiguous range of rows, 16 rows per pipeline stage (cp.async, S stages,
// full-width padded rows in shared memory, stride = 4 mod 16 doubles -> conflict-free DMMA fragments).
// Per stage: (1) every warp computes eta = x.beta for its rows from shared memory; (2) one warp
// (rotating) evaluates the link for the 16 rows -> residuals r and weights w; (3) CTA 0 of the group
// accumulates g and u = X^T w with plain FMAs, one column per thread; (4) every warp owns ONE 32x32
// block of the UPPER TRIANGLE of X^T W X and accumulates it with FP64 DMMA.8x8x4 (A fragments scaled by
// w).  Only d/32*(d/32+1)/2 blocks exist (36 at d = 256, spread over 3 CTAs x 12 warps), so the tensor
// pipe does 56 % of the work of a full d x d product.  The steps are software-pipelined over three
// consecutive stages so that one __syncthreads per stage suffices.
// Partial blocks / vectors go to scratch per group and are summed in group order by k_fused_finish
// (deterministic; lower triangle mirrored from the upper one -> exactly symmetric H).
#include "glm_common.cuh"

namespace nb {

constexpr int FK = 16;          // rows per stage
constexpr int FB = 32;          // block edge
constexpr int FPAD = 4;

__device__ __forceinline__ void f_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void f_cp16(void* smem, const void* gmem, int src_bytes) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(s), "l"(gmem), "r"(src_bytes));
}
// ---- TMA bulk copies (cp.async.bulk -> SASS UBLKCP) completing on an mbarrier ---------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
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
__device__ __forceinline__ void bulk_g2s(void* smem, const void* gmem, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem), "r"(bytes),
                  "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct FusedArgs {
  const double* x; int64_t n, d, ldx;
  const void* y; int ydt;
  const double* beta;
  const double* wext;      // precomputed weights (null = all ones) when the link is not fused
  int family;
  int nblk;                // ceil(d / 32)
  int T;                   // nblk (nblk + 1) / 2 upper-triangular blocks
  int C;                   // CTAs per group
  int groups;
  int64_t rows_per_group;  // multiple of FK
  int64_t ldc;             // row stride of the output matrix (d + 1 inside the packed stats buffer)
  double* part_h;          // [groups][T][32*32]
  double* part_v;          // [groups][2*dp + 3]   g[dp] | u[dp] | sum r | sum w | objective
};

template <int W, int S, bool LINK>
__global__ void __launch_bounds__(W * 32)
k_glm_fused(const FusedArgs a) {
  constexpr int NT = W * 32;
  extern __shared__ __align__(16) double fsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int lq = lane >> 2, lr = lane & 3;
  const int dp = a.nblk * FB;
  const int ld = dp + FPAD;
  const int64_t d = a.d;
  const int group = blockIdx.x / a.C, c = blockIdx.x % a.C;
  const int64_t row_beg = (int64_t)group * a.rows_per_group;
  const int64_t row_end = (row_beg + a.rows_per_group < a.n) ? row_beg + a.rows_per_group : a.n;
  const int ktiles = row_end > row_beg ? (int)((row_end - row_beg + FK - 1) / FK) : 0;

  // shared memory carve-up
  double* sX = fsm;                               // S * FK * ld
  double* sEta = sX + (size_t)S * FK * ld;        // S * FK
  double* sWgt = sEta + S * FK;                   // S * FK
  double* sRes = sWgt + S * FK;                   // S * FK
  double* sRed = sRes + S * FK;                   // W * 4 (final scalar reduction)
  uint64_t* full = reinterpret_cast<uint64_t*>(sRed + 12 * 4);   // S mbarriers: stage landed

  // this warp's block of the upper triangle
  const int t = c * W + warp;
  const bool has_block = t < a.T;
  int bi = 0, bj = 0;
  if (has_block) {
    int rem = t;
    while (rem >= a.nblk - bi) { rem -= a.nblk - bi; bi++; }
    bj = bi + rem;
  }
  // beta for the eta phase: lane owns columns lane + 32*j
  double bet[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    int col = lane + 32 * j;
    bet[j] = (j < a.nblk && col < d) ? a.beta[col] : 0.0;
  }
  const double beta0 = a.beta[d];

  // Stage loads: ONE thread issues 16 row-sized bulk copies (2 KB each at d = 256) that complete on the
  // slot's mbarrier; nobody else spends an instruction on data movement.  Rows past the end of the range
  // are simply not copied: their weights are zero and shared memory never holds non-finite stale data
  // (it is zero-filled once below), so they contribute exactly zero.
  const unsigned row_bytes = (unsigned)(d * sizeof(double));
  auto load_stage = [&](int kt) {
    const int slot = kt % S;
    const int64_t r0 = row_beg + (int64_t)kt * FK;
    if (tid == 0) {
      int64_t valid = row_end - r0;
      if (valid > FK) valid = FK;
      mbar_expect_tx(&full[slot], (unsigned)valid * row_bytes);
      double* dst = sX + (size_t)slot * FK * ld;
      for (int r = 0; r < (int)valid; r++) bulk_g2s(dst + r * ld, a.x + (r0 + r) * a.ldx, row_bytes, &full[slot]);
    }
    if constexpr (!LINK) {
      if (tid < FK) {
        int64_t gr = r0 + tid;
        double wv = 0.0;
        if (gr < row_end) wv = a.wext ? a.wext[gr] : 1.0;
        sWgt[slot * FK + tid] = wv;          // plain store: made visible by the stage's barrier
      }
    }
  };
  auto eta_phase = [&](int kt) {                 // rows warp, warp + W, ... of stage kt
    const double* src = sX + (size_t)(kt % S) * FK * ld;
    for (int r = warp; r < FK; r += W) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < 8; j++)
        if (j < a.nblk) s = fma(src[r * ld + lane + 32 * j], bet[j], s);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) sEta[(kt % S) * FK + r] = s + beta0;
    }
  };
  double acc_obj = 0.0, acc_r = 0.0, acc_w = 0.0;   // lanes < FK of whichever warp ran the link
  auto link_phase = [&](int kt) {                // one warp: lane r handles row r of stage kt
    if (lane < FK) {
      const int64_t gr = row_beg + (int64_t)kt * FK + lane;
      LinkOut lo;
      lo.resid = 0.0; lo.wgt = 0.0; lo.obj = 0.0;
      if (gr < row_end) lo = glm_link_rt(a.family, sEta[(kt % S) * FK + lane], load_y(a.y, a.ydt, gr));
      sWgt[(kt % S) * FK + lane] = lo.wgt;
      sRes[(kt % S) * FK + lane] = lo.resid;
      acc_obj += lo.obj; acc_r += lo.resid; acc_w += lo.wgt;
    }
  };

  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  double gacc = 0.0, uacc = 0.0;                  // column `tid` (CTA 0 only)

  // ---- prologue: zero the staging buffers, arm the barriers, put stages 0 .. S-2 in flight
  for (int i = tid; i < S * FK * ld; i += NT) sX[i] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < S; s++) mbar_init(&full[s], 1);
  }
  fence_async_proxy();                            // generic-proxy zero fill before async-proxy writes
  __syncthreads();
  for (int s = 0; s < S - 1; s++)
    if (s < ktiles) load_stage(s);
  if constexpr (LINK) {
    if (ktiles > 0) mbar_wait(&full[0], 0);
    if (ktiles > 1) mbar_wait(&full[1 % S], 0);
    if (ktiles > 0) eta_phase(0);
    if (ktiles > 1) eta_phase(1);
    __syncthreads();
    if (warp == 0 && ktiles > 0) link_phase(0);
  }

  for (int kt = 0; kt < ktiles; kt++) {
    mbar_wait(&full[kt % S], (unsigned)((kt / S) & 1));                 // stage kt has landed
    if constexpr (LINK) {
      if (kt + 2 < ktiles) mbar_wait(&full[(kt + 2) % S], (unsigned)(((kt + 2) / S) & 1));
    }
    __syncthreads();                              // everyone is done with stage kt - 1: its slot is free
    {
      int nk = kt + S - 1;
      if (nk < ktiles) load_stage(nk);
    }
    if constexpr (LINK) {
      if (kt + 2 < ktiles) eta_phase(kt + 2);
      if (kt + 1 < ktiles && warp == ((kt + 1) % W)) link_phase(kt + 1);
    }

    const double* xs = sX + (size_t)(kt % S) * FK * ld;
    const double* ws = sWgt + (kt % S) * FK;
    if (LINK && c == 0 && tid < dp) {             // g and u, one column per thread
      const double* rs = sRes + (kt % S) * FK;
#pragma unroll
      for (int r = 0; r < FK; r++) {
        double xv = xs[r * ld + tid];
        gacc = fma(rs[r], xv, gacc);
        uacc = fma(ws[r], xv, uacc);
      }
    }
    if (has_block) {
      const double* pa = xs + bi * FB + lq;
      const double* pb = xs + bj * FB + lq;
      double af[2][4], bf[2][4];
#pragma unroll
      for (int i = 0; i < 4; i++) { af[0][i] = pa[lr * ld + i * 8] * ws[lr]; bf[0][i] = pb[lr * ld + i * 8]; }
#pragma unroll
      for (int ks = 0; ks < FK / 4; ks++) {
        const int cur = ks & 1, nxt = cur ^ 1;
        if (ks + 1 < FK / 4) {
          const int rr = (ks + 1) * 4 + lr;
          const double wv = ws[rr];
#pragma unroll
          for (int i = 0; i < 4; i++) { af[nxt][i] = pa[rr * ld + i * 8] * wv; bf[nxt][i] = pb[rr * ld + i * 8]; }
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) f_dmma(acc[i][j][0], acc[i][j][1], af[cur][i], bf[cur][j]);
      }
    }
  }
  __syncthreads();

  // ---- epilogue: block partial, column partials, scalars
  if (has_block) {
    double* p = a.part_h + ((size_t)group * a.T + t) * (FB * FB);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        *reinterpret_cast<double2*>(p + (i * 8 + lq) * FB + j * 8 + 2 * lr) = make_double2(acc[i][j][0], acc[i][j][1]);
  }
  if (LINK && c == 0) {
    double* pv = a.part_v + (size_t)group * (2 * dp + 3);
    if (tid < dp) { pv[tid] = gacc; pv[dp + tid] = uacc; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      acc_obj += __shfl_down_sync(0xffffffffu, acc_obj, o);
      acc_r += __shfl_down_sync(0xffffffffu, acc_r, o);
      acc_w += __shfl_down_sync(0xffffffffu, acc_w, o);
    }
    if (lane == 0) { sRed[warp * 4 + 0] = acc_r; sRed[warp * 4 + 1] = acc_w; sRed[warp * 4 + 2] = acc_obj; }
    __syncthreads();
    if (tid == 0) {
      double sr = 0.0, sw = 0.0, so = 0.0;
      for (int w = 0; w < W; w++) { sr += sRed[w * 4]; sw += sRed[w * 4 + 1]; so += sRed[w * 4 + 2]; }
      pv[2 * dp] = sr; pv[2 * dp + 1] = sw; pv[2 * dp + 2] = so;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Pass 2 of the two-pass path: upper-triangular X^T diag(w) X only, warp-specialised.
// Warps 0..W-1 are consumers (one 32x32 block each, FP64 DMMA); warp W is the producer: for every
// stage lane r stores the weight of row r and issues the TMA bulk copy of row r, each lane arriving on
// the stage's "full" mbarrier with its own expected byte count.  Consumers release a stage by arriving
// on its "empty" mbarrier, which the producer waits on before refilling the slot.  There is no
// block-wide barrier in the steady state, so the tensor pipe never drains between stages.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

template <int W, int S>
__global__ void __launch_bounds__((W + 1) * 32)
k_tri_gram(const FusedArgs a) {
  constexpr int NT = (W + 1) * 32;
  extern __shared__ __align__(16) double fsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int lq = lane >> 2, lr = lane & 3;
  const int dp = a.nblk * FB;
  const int ld = dp + FPAD;
  const int64_t d = a.d;
  const int group = blockIdx.x / a.C, c = blockIdx.x % a.C;
  const int64_t row_beg = (int64_t)group * a.rows_per_group;
  const int64_t row_end = (row_beg + a.rows_per_group < a.n) ? row_beg + a.rows_per_group : a.n;
  const int ktiles = row_end > row_beg ? (int)((row_end - row_beg + FK - 1) / FK) : 0;

  double* sX = fsm;                               // S * FK * ld
  double* sWgt = sX + (size_t)S * FK * ld;        // S * FK
  uint64_t* full = reinterpret_cast<uint64_t*>(sWgt + S * FK);
  uint64_t* empty = full + S;

  for (int i = tid; i < S * FK * ld; i += NT) sX[i] = 0.0;     // stale rows must be finite (weight 0)
  if (tid == 0) {
    for (int s = 0; s < S; s++) { mbar_init(&full[s], FK); mbar_init(&empty[s], W); }
  }
  fence_async_proxy();
  __syncthreads();

  if (warp == W) {
    // ---------------- producer
    const unsigned row_bytes = (unsigned)(d * sizeof(double));
    for (int kt = 0; kt < ktiles; kt++) {
      const int slot = kt % S;
      if (kt >= S) mbar_wait(&empty[slot], (unsigned)(((kt / S) - 1) & 1));
      if (lane < FK) {
        const int64_t gr = row_beg + (int64_t)kt * FK + lane;
        if (gr < row_end) {
          sWgt[slot * FK + lane] = a.wext ? a.wext[gr] : 1.0;
          mbar_expect_tx(&full[slot], row_bytes);              // this lane's arrival + its bytes
          bulk_g2s(sX + (size_t)slot * FK * ld + lane * ld, a.x + gr * a.ldx, row_bytes, &full[slot]);
        } else {
          sWgt[slot * FK + lane] = 0.0;
          mbar_arrive(&full[slot]);
        }
      }
    }
    return;
  }

  // ---------------- consumers
  const int t = c * W + warp;
  const bool has_block = t < a.T;
  int bi = 0, bj = 0;
  if (has_block) {
    int rem = t;
    while (rem >= a.nblk - bi) { rem -= a.nblk - bi; bi++; }
    bj = bi + rem;
  }
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

  for (int kt = 0; kt < ktiles; kt++) {
    const int slot = kt % S;
    // every consumer waits (also a warp without a block): it keeps a fast warp from arriving twice on
    // the same "empty" phase
    mbar_wait(&full[slot], (unsigned)((kt / S) & 1));
    if (has_block) {
      const double* xs = sX + (size_t)slot * FK * ld;
      const double* ws = sWgt + slot * FK;
      const double* pa = xs + bi * FB + lq;
      const double* pb = xs + bj * FB + lq;
      double af[2][4], bf[2][4];
#pragma unroll
      for (int i = 0; i < 4; i++) { af[0][i] = pa[lr * ld + i * 8] * ws[lr]; bf[0][i] = pb[lr * ld + i * 8]; }
#pragma unroll
      for (int ks = 0; ks < FK / 4; ks++) {
        const int cur = ks & 1, nxt = cur ^ 1;
        if (ks + 1 < FK / 4) {
          const int rr = (ks + 1) * 4 + lr;
          const double wv = ws[rr];
#pragma unroll
          for (int i = 0; i < 4; i++) { af[nxt][i] = pa[rr * ld + i * 8] * wv; bf[nxt][i] = pb[rr * ld + i * 8]; }
        }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) f_dmma(acc[i][j][0], acc[i][j][1], af[cur][i], bf[cur][j]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[slot]);     // this warp no longer reads the slot
  }
  if (has_block) {
    double* p = a.part_h + ((size_t)group * a.T + t) * (FB * FB);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        *reinterpret_cast<double2*>(p + (i * 8 + lq) * FB + j * 8 + 2 * lr) = make_double2(acc[i][j][0], acc[i][j][1]);
  }
}

// Sums the group partials in group order and assembles the packed [H | g | obj] buffer.
__global__ void __launch_bounds__(256)
k_fused_finish(const FusedArgs a, double* __restrict__ stats, int h_only) {
  const int64_t d = a.d, p = a.ldc;
  const int dp = a.nblk * FB;
  const int64_t nh = d * (d + 1) / 2;                // upper triangle of the d x d part
  const int64_t total = h_only ? nh : nh + 2 * d + 3;
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (int64_t)gridDim.x * 256) {
    if (e < nh) {
      // e -> (i, j), i <= j, row by row
      int64_t i = 0, rem = e;
      // closed form would need a sqrt; d <= 256 so a short search is fine
      while (rem >= d - i) { rem -= d - i; i++; }
      int64_t j = i + rem;
      int bi = (int)(i / FB), bj = (int)(j / FB);
      int t = bi * a.nblk - (bi * (bi - 1)) / 2 + (bj - bi);
      const double* src = a.part_h + (size_t)t * (FB * FB) + (i - bi * FB) * FB + (j - bj * FB);
      double s = 0.0;
      for (int g = 0; g < a.groups; g++) s += src[(size_t)g * a.T * (FB * FB)];
      stats[i * p + j] = s;
      stats[j * p + i] = s;
    } else {
      int64_t q = e - nh;
      int64_t idx;                                   // index into the per-group vector
      if (q < d) idx = q;                            // g[q]
      else if (q < 2 * d) idx = dp + (q - d);        // u[q - d]
      else idx = 2 * dp + (q - 2 * d);               // sum r, sum w, objective
      double s = 0.0;
      for (int g = 0; g < a.groups; g++) s += a.part_v[(size_t)g * (2 * dp + 3) + idx];
      if (q < d) stats[p * p + q] = s;
      else if (q < 2 * d) { int64_t i = q - d; stats[i * p + d] = s; stats[d * p + i] = s; }
      else if (q == 2 * d) stats[p * p + d] = s;
      else if (q == 2 * d + 1) stats[d * p + d] = s;
      else stats[p * p + p] = s;
    }
  }
}

struct FusedPlan { int W, S, nblk, T, C, groups; int64_t rows_per_group; size_t smem, scratch, off_v; };

static FusedPlan plan_fused(int device, int64_t n, int64_t d) {
  FusedPlan f;
  f.nblk = (int)((d + FB - 1) / FB);
  f.T = f.nblk * (f.nblk + 1) / 2;
  f.W = f.T <= 4 ? 4 : (f.T <= 8 ? 8 : 12);
  f.C = (f.T + f.W - 1) / f.W;
  const int sms = dev_info(device).sms;
  f.groups = sms / f.C;
  if (f.groups < 1) f.groups = 1;
  int64_t tiles = (n + FK - 1) / FK;
  if (f.groups > tiles) f.groups = (int)(tiles > 0 ? tiles : 1);
  int64_t tpg = (tiles + f.groups - 1) / f.groups;
  f.rows_per_group = tpg * FK;
  f.groups = (int)((n + f.rows_per_group - 1) / f.rows_per_group);
  if (f.groups < 1) f.groups = 1;
  const int dp = f.nblk * FB, ld = dp + FPAD;
  const size_t stage = (size_t)FK * ld * sizeof(double);
  f.S = 5;
  while (f.S > 4 && (size_t)f.S * stage + 3 * f.S * FK * 8 + 12 * 4 * 8 + 64 > (size_t)200 * 1024) f.S--;
  f.smem = (size_t)f.S * stage + (size_t)3 * f.S * FK * 8 + (size_t)12 * 4 * 8 + 8 * 8;
  size_t ph = (size_t)f.groups * f.T * FB * FB * sizeof(double);
  f.off_v = (ph + 255) & ~(size_t)255;
  f.scratch = f.off_v + (((size_t)f.groups * (2 * dp + 3) * sizeof(double) + 255) & ~(size_t)255);
  return f;
}

bool fused_applicable(int64_t n, int64_t d, int64_t ldx, const double* x) {
  // bulk copies need 16-byte aligned rows whose length is a multiple of 16 bytes
  return d >= 8 && d <= 256 && n > 0 && (d % 2 == 0) && (ldx % 2 == 0) && ((((uintptr_t)x) & 15) == 0);
}

size_t fused_scratch_bytes(int device, int64_t n, int64_t d) {
  // upper bound over n: the group count only shrinks for tiny n
  FusedPlan f = plan_fused(device, n > 0 ? n : 1, d);
  FusedPlan g = plan_fused(device, (int64_t)1 << 40, d);
  return f.scratch > g.scratch ? f.scratch : g.scratch;
}

template <int W, int S, bool LINK>
static int launch_fused(const FusedArgs& a, const FusedPlan& f, cudaStream_t st) {
  auto kern = k_glm_fused<W, S, LINK>;
  NB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f.smem));
  kern<<<(unsigned)(f.groups * f.C), W * 32, f.smem, st>>>(a);
  NB_LAUNCH_CHECK();
  return 0;
}

template <int W, int S>
static int launch_tri(const FusedArgs& a, const FusedPlan& f, cudaStream_t st) {
  auto kern = k_tri_gram<W, S>;
  size_t smem = (size_t)S * FK * (f.nblk * FB + FPAD) * sizeof(double) + (size_t)S * FK * 8 + 2 * S * 8;
  NB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)(f.groups * f.C), (W + 1) * 32, smem, st>>>(a);
  NB_LAUNCH_CHECK();
  return 0;
}

template <bool LINK>
static int run_fused(int device, const FusedArgs& a0, const FusedPlan& f, double* stats, cudaStream_t st) {
  int rc;
  if constexpr (!LINK) {
    if (f.S == 5) rc = f.W == 4 ? launch_tri<4, 5>(a0, f, st) : (f.W == 8 ? launch_tri<8, 5>(a0, f, st) : launch_tri<12, 5>(a0, f, st));
    else          rc = f.W == 4 ? launch_tri<4, 4>(a0, f, st) : (f.W == 8 ? launch_tri<8, 4>(a0, f, st) : launch_tri<12, 4>(a0, f, st));
  } else if (f.S == 5) {
    rc = f.W == 4 ? launch_fused<4, 5, LINK>(a0, f, st)
                  : (f.W == 8 ? launch_fused<8, 5, LINK>(a0, f, st) : launch_fused<12, 5, LINK>(a0, f, st));
  } else {
    rc = f.W == 4 ? launch_fused<4, 4, LINK>(a0, f, st)
                  : (f.W == 8 ? launch_fused<8, 4, LINK>(a0, f, st) : launch_fused<12, 4, LINK>(a0, f, st));
  }
  if (rc) return rc;
  const int64_t d = a0.d;
  int64_t total = d * (d + 1) / 2 + (LINK ? 2 * d + 3 : 0);
  int64_t blocks = (total + 255) / 256;
  k_fused_finish<<<(unsigned)(blocks < 592 ? blocks : 592), 256, 0, st>>>(a0, stats, LINK ? 0 : 1);
  NB_LAUNCH_CHECK();
  return 0;
}

// link == true : everything in one pass.
// link == false: only the upper-triangular X^T diag(w) X (w = wext, null = ones) into the d x d corner of stats.
int glm_fused_stats(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y,
                    int ydt, const double* beta, const double* wext, bool link, double* stats, int64_t ldc,
                    void* scratch, size_t scratch_bytes, cudaStream_t st) {
  FusedPlan f = plan_fused(device, n, d);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= f.scratch, "glm fused: scratch too small (%zu < %zu)",
             scratch_bytes, f.scratch);
  FusedArgs a;
  a.x = x; a.n = n; a.d = d; a.ldx = ldx; a.y = y; a.ydt = ydt; a.beta = beta; a.wext = wext; a.family = family;
  a.nblk = f.nblk; a.T = f.T; a.C = f.C; a.groups = f.groups; a.rows_per_group = f.rows_per_group;
  a.ldc = ldc;
  a.part_h = (double*)scratch;
  a.part_v = (double*)((char*)scratch + f.off_v);
  return link ? run_fused<true>(device, a, f, stats, st) : run_fused<false>(device, a, f, stats, st);
}

}  // namespace nb
