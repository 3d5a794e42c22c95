// Square / large block matmul (BASELINE config 3: 10^4 x 10^4 @ 10^4 x 10^4 float64), warp-specialised:
// FP64 DMMA.8x8x4 tiles fed by TMA bulk copies.  Reference: KernelCls.bop "tensordot" -> np.matmul
// (nums/core/kernel/numpy_kernel.py:419-425), OpenBLAS dgemm on the CPU.
//
// CTA tile 128 x 128, k step 32.  A producer warp streams the two operand tiles of a stage into padded shared
// memory (row stride = 4 mod 16 doubles: every DMMA fragment load is bank-conflict free) with one
// cp.async.bulk per contiguous tile row (SASS UBLKCP), completing on the slot's "full" mbarrier; eight
// consumer warps (2 x 4, 64 x 32 each, 64 accumulator registers per thread) wait on it, run 256 DMMAs per
// stage and release the slot through its "empty" mbarrier.  No block-wide barrier in the main loop.
// All four operand layouts are supported (a tile row is whichever index is contiguous in memory).
// Requirements (else the cp.async kernel of matmul.cu is used): 16-byte aligned operands, even leading
// dimensions, even m / n / k so that every partial row copy is a multiple of 16 bytes.
#include "gemm.cuh"

namespace nb {

constexpr int WS_BM = 128, WS_BN = 128, WS_BK = 32, WS_STAGES = 3, WS_PAD = 4;
constexpr int WS_CONSUMERS = 8;

__device__ __forceinline__ void ws_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void ws_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void ws_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ws_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void ws_mbar_wait(uint64_t* bar, unsigned parity) {
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
__device__ __forceinline__ void ws_bulk_g2s(void* smem, const void* gmem, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem), "r"(bytes),
                  "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

struct WsArgs {
  int64_t m, n, k;
  const double* a; int64_t lda;    // A_MC: (k,m) at a[k*lda+m]; else (m,k) at a[m*lda+k]
  const double* b; int64_t ldb;    // B_NC: (k,n) at b[k*ldb+n]; else (n,k) at b[n*ldb+k]
  const double* w;                 // optional weights along k: C = op(A) diag(w) op(B)
  double* c; int64_t ldc;
  int tiles_m, tiles_n;
  // tall-skinny products (few output tiles, long k): the k range is cut into `splits` chunks, one CTA per
  // (tile, chunk), partial tiles go to `part` and k_gemm_reduce sums them in chunk order; `sym` computes the
  // upper-triangular tiles only (the reduce mirrors them)
  int ntiles, splits, sym;
  int64_t kchunk;
  double* part;
};

template <bool A_MC, bool B_NC>
struct WsSmem {
  static constexpr int A_ROWS = A_MC ? WS_BK : WS_BM;
  static constexpr int A_COLS = A_MC ? WS_BM : WS_BK;
  static constexpr int B_ROWS = B_NC ? WS_BK : WS_BN;
  static constexpr int B_COLS = B_NC ? WS_BN : WS_BK;
  static constexpr int A_LD = A_COLS + WS_PAD;
  static constexpr int B_LD = B_COLS + WS_PAD;
  static constexpr int A_ELEMS = A_ROWS * A_LD;
  static constexpr int B_ELEMS = B_ROWS * B_LD;
  static constexpr int STAGE = A_ELEMS + B_ELEMS + WS_BK;      // + the weights of the stage's k slice
};

// Producer helpers for a ROWS x COLS tile (COLS contiguous in memory).  The part of the k range that lies
// beyond the matrix is zero-filled so that stale shared memory never enters a sum (m / n tails only feed
// outputs that are discarded).  Zero fill (plain stores) comes BEFORE the lane's mbarrier arrive, which
// releases it to the consumers; the bulk copies come after it.
template <int ROWS, int COLS, int LD, bool K_IS_ROW>
__device__ __forceinline__ void ws_zero_tail(double* sm, int64_t vr, int64_t vc, int lane) {
  if (K_IS_ROW) {
    for (int64_t e = vr * LD + lane; e < (int64_t)ROWS * LD; e += 32) sm[e] = 0.0;
  } else if (vc < COLS) {
    const int w = (int)(COLS - vc);
    for (int64_t e = lane; e < vr * w; e += 32) sm[(e / w) * LD + vc + (e % w)] = 0.0;
  }
}
template <int LD>
__device__ __forceinline__ void ws_copy_rows(double* sm, const double* g, int64_t ld, int64_t row0, int64_t col0,
                                             int64_t vr, int64_t vc, int lane, uint64_t* bar) {
  const unsigned bytes = (unsigned)(vc * sizeof(double));
  for (int64_t r = lane; r < vr; r += 32) ws_bulk_g2s(sm + r * LD, g + (row0 + r) * ld + col0, bytes, bar);
}
__device__ __forceinline__ int64_t ws_clip(int64_t lim, int64_t origin, int64_t extent) {
  int64_t v = lim - origin;
  return v < 0 ? 0 : (v < extent ? v : extent);
}

template <bool A_MC, bool B_NC, bool HAS_W, bool PART>
__global__ void __launch_bounds__((WS_CONSUMERS + 1) * 32)
k_gemm_ws(const WsArgs g) {
  using SM = WsSmem<A_MC, B_NC>;
  extern __shared__ __align__(16) double wsm[];
  uint64_t* full = reinterpret_cast<uint64_t*>(wsm + (size_t)WS_STAGES * SM::STAGE);
  uint64_t* empty = full + WS_STAGES;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < WS_STAGES; s++) { ws_mbar_init(&full[s], 32); ws_mbar_init(&empty[s], WS_CONSUMERS); }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  int tm, tn;
  const int z = blockIdx.x / g.ntiles;            // k chunk
  const int tile = blockIdx.x - z * g.ntiles;
  if (g.sym) {
    // upper-triangular tiles, row by row: tile = tm * tiles_n - tm (tm - 1) / 2 + (tn - tm)
    int t = tile;
    tm = 0;
    while (t >= g.tiles_n - tm) { t -= g.tiles_n - tm; tm++; }
    tn = tm + t;
  } else {
    // rasterisation in bands of 8 tile columns: the ~148 co-resident CTAs cover ~18 x 8 tiles, so every A panel
    // is shared by 8 CTAs and every B panel by ~18 through L2
    const int band_w = 8;
    const int per_band = g.tiles_m * band_w;
    const int band = tile / per_band;
    const int in_band = tile - band * per_band;
    const int w = (g.tiles_n - band * band_w) < band_w ? (g.tiles_n - band * band_w) : band_w;   // last band may be narrower
    tm = in_band / w;
    tn = band * band_w + in_band % w;
  }
  const int64_t m0 = (int64_t)tm * WS_BM, n0 = (int64_t)tn * WS_BN;
  const int64_t k_lo = (int64_t)z * g.kchunk;
  const int64_t k_hi = (k_lo + g.kchunk < g.k) ? k_lo + g.kchunk : g.k;
  const int ktiles = (int)((k_hi - k_lo + WS_BK - 1) / WS_BK);

  if (warp == WS_CONSUMERS) {
    // ---------------------------------------------------------------- producer
    for (int kt = 0; kt < ktiles; kt++) {
      const int slot = kt % WS_STAGES;
      if (kt >= WS_STAGES) ws_mbar_wait(&empty[slot], (unsigned)(((kt / WS_STAGES) - 1) & 1));
      double* sA = wsm + (size_t)slot * SM::STAGE;
      double* sB = sA + SM::A_ELEMS;
      double* sW = sB + SM::B_ELEMS;
      const int64_t k0 = k_lo + (int64_t)kt * WS_BK;
      const int64_t a_vr = A_MC ? ws_clip(k_hi, k0, SM::A_ROWS) : ws_clip(g.m, m0, SM::A_ROWS);
      const int64_t a_vc = A_MC ? ws_clip(g.m, m0, SM::A_COLS) : ws_clip(k_hi, k0, SM::A_COLS);
      const int64_t b_vr = B_NC ? ws_clip(k_hi, k0, SM::B_ROWS) : ws_clip(g.n, n0, SM::B_ROWS);
      const int64_t b_vc = B_NC ? ws_clip(g.n, n0, SM::B_COLS) : ws_clip(k_hi, k0, SM::B_COLS);
      ws_zero_tail<SM::A_ROWS, SM::A_COLS, SM::A_LD, A_MC>(sA, a_vr, a_vc, lane);
      ws_zero_tail<SM::B_ROWS, SM::B_COLS, SM::B_LD, B_NC>(sB, b_vr, b_vc, lane);
      // weights of the k slice: one bulk copy for a full slice, plain stores (before the arrive) for the ragged last one
      const int64_t w_v = ws_clip(k_hi, k0, WS_BK);
      const bool w_bulk = HAS_W && w_v == WS_BK;
      if (HAS_W && !w_bulk) sW[lane] = lane < w_v ? g.w[k0 + lane] : 0.0;
      unsigned mine = 0;
      for (int64_t r = lane; r < a_vr; r += 32) mine += (unsigned)(a_vc * sizeof(double));
      for (int64_t r = lane; r < b_vr; r += 32) mine += (unsigned)(b_vc * sizeof(double));
      if (w_bulk && lane == 0) mine += (unsigned)(WS_BK * sizeof(double));
      if (mine) ws_mbar_expect_tx(&full[slot], mine); else ws_mbar_arrive(&full[slot]);
      if (w_bulk && lane == 0) ws_bulk_g2s(sW, g.w + k0, (unsigned)(WS_BK * sizeof(double)), &full[slot]);
      if (A_MC) ws_copy_rows<SM::A_LD>(sA, g.a, g.lda, k0, m0, a_vr, a_vc, lane, &full[slot]);
      else      ws_copy_rows<SM::A_LD>(sA, g.a, g.lda, m0, k0, a_vr, a_vc, lane, &full[slot]);
      if (B_NC) ws_copy_rows<SM::B_LD>(sB, g.b, g.ldb, k0, n0, b_vr, b_vc, lane, &full[slot]);
      else      ws_copy_rows<SM::B_LD>(sB, g.b, g.ldb, n0, k0, b_vr, b_vc, lane, &full[slot]);
    }
    return;
  }

  // ---------------------------------------------------------------- consumers: 2 x 4 warps of 64 x 32
  const int wm = warp / 4, wn = warp % 4;
  const int lq = lane >> 2, lr = lane & 3;
  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

  for (int kt = 0; kt < ktiles; kt++) {
    const int slot = kt % WS_STAGES;
    ws_mbar_wait(&full[slot], (unsigned)((kt / WS_STAGES) & 1));
    const double* sA = wsm + (size_t)slot * SM::STAGE;
    const double* sB = sA + SM::A_ELEMS;
    const double* sW = sB + SM::B_ELEMS;
#pragma unroll
    for (int ks = 0; ks < WS_BK / 4; ks++) {
      const int kk = ks * 4 + lr;
      double af[8], bf[4];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int mm = wm * 64 + i * 8 + lq;
        af[i] = A_MC ? sA[kk * SM::A_LD + mm] : sA[mm * SM::A_LD + kk];
      }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int nn = wn * 32 + j * 8 + lq;
        bf[j] = B_NC ? sB[kk * SM::B_LD + nn] : sB[nn * SM::B_LD + kk];
      }
      if (HAS_W) {                 // the narrower fragment carries the weights
        const double wk = sW[kk];
#pragma unroll
        for (int j = 0; j < 4; j++) bf[j] *= wk;
      }
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) ws_dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
    __syncwarp();
    if (lane == 0) ws_mbar_arrive(&empty[slot]);
  }

  if constexpr (PART) {
    // partial tile of this k chunk (whole tile: rows / columns beyond the matrix are never read back)
    double* pt = g.part + ((size_t)z * g.ntiles + tile) * (WS_BM * WS_BN);
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int r = wm * 64 + i * 8 + lq, c = wn * 32 + j * 8 + 2 * lr;
        *reinterpret_cast<double2*>(pt + r * WS_BN + c) = make_double2(acc[i][j][0], acc[i][j][1]);
      }
  } else {
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int64_t r = m0 + wm * 64 + i * 8 + lq, c = n0 + wn * 32 + j * 8 + 2 * lr;
        if (r < g.m && c < g.n) {
          if (c + 1 < g.n && ((g.ldc & 1) == 0)) {
            *reinterpret_cast<double2*>(g.c + r * g.ldc + c) = make_double2(acc[i][j][0], acc[i][j][1]);
          } else {
            g.c[r * g.ldc + c] = acc[i][j][0];
            if (c + 1 < g.n) g.c[r * g.ldc + c + 1] = acc[i][j][1];
          }
        }
      }
  }
}

// Launch plan.  Enough output tiles for the machine -> one CTA per tile, direct stores.  Few tiles and a long k
// (tall-skinny X^T diag(w) X for d > 256, or any A^T B of two tall blocks) -> split k so that tiles x splits fills
// one to four waves of SMs as evenly as possible.
WsPlan ws_plan(int device, int64_t m, int64_t n, int64_t k, bool sym, bool k_rows) {
  WsPlan p{};
  p.ok = false;
  if (m < 256 || n < 256 || k < 64) return p;
  // every bulk copy must be a multiple of 16 bytes: the contiguous extent of a tile row is m / n when both operands
  // are stored [k][.] (k_rows), else it is a k slice
  if ((m | n) & 1) return p;
  if (!k_rows && (k & 1)) return p;
  p.tiles_m = (int)((m + WS_BM - 1) / WS_BM);
  p.tiles_n = (int)((n + WS_BN - 1) / WS_BN);
  p.sym = sym ? 1 : 0;
  p.ntiles = sym ? p.tiles_m * (p.tiles_m + 1) / 2 : p.tiles_m * p.tiles_n;
  const int sms = dev_info(device).sms;
  p.splits = 1;
  p.kchunk = k;
  if (p.ntiles < sms || sym) {
    const int64_t max_splits = k / (WS_BK * 16) > 0 ? k / (WS_BK * 16) : 1;
    double best = -1.0;
    int64_t best_s = 1;
    for (int wv = 1; wv <= 4; wv++) {
      int64_t sp = (int64_t)sms * wv / p.ntiles;
      if (sp < 1) sp = 1;
      if (sp > max_splits) sp = max_splits;
      const int64_t ctas = sp * p.ntiles;
      const double eff = (double)ctas / (double)(((ctas + sms - 1) / sms) * sms);
      if (eff > best + 1e-9) { best = eff; best_s = sp; }
    }
    if (best_s * p.ntiles * 10 < (int64_t)sms * 6) return p;       // k too short to fill the machine: cp.async split-k kernel
    int64_t kchunk = (k + best_s - 1) / best_s;
    kchunk = (kchunk + WS_BK - 1) / WS_BK * WS_BK;
    p.kchunk = kchunk;
    p.splits = (int)((k + kchunk - 1) / kchunk);
  }
  p.scratch = (p.splits > 1 || sym) ? (size_t)p.splits * p.ntiles * WS_BM * WS_BN * sizeof(double) : 0;
  p.ok = true;
  return p;
}

bool gemm_ws_aligned(const double* a, int64_t lda, const double* b, int64_t ldb, const double* w, const double* c,
                     int64_t ldc) {
  if ((lda | ldb) & 1) return false;
  if ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)w) & 15) != 0) return false;
  if ((ldc & 1) == 0 && (((uintptr_t)c) & 15) != 0) return false;
  return true;
}

template <bool A_MC, bool B_NC, bool HAS_W, bool PART>
static int launch_ws_k(const WsArgs& g, cudaStream_t st) {
  using SM = WsSmem<A_MC, B_NC>;
  size_t smem = (size_t)WS_STAGES * SM::STAGE * sizeof(double) + 2 * WS_STAGES * sizeof(uint64_t);
  auto kern = k_gemm_ws<A_MC, B_NC, HAS_W, PART>;
  NB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)(g.ntiles * g.splits), (WS_CONSUMERS + 1) * 32, smem, st>>>(g);
  NB_LAUNCH_CHECK();
  return 0;
}

template <bool A_MC, bool B_NC>
static int launch_ws(const WsArgs& g, cudaStream_t st) {
  if (g.part) return g.w ? launch_ws_k<A_MC, B_NC, true, true>(g, st) : launch_ws_k<A_MC, B_NC, false, true>(g, st);
  return g.w ? launch_ws_k<A_MC, B_NC, true, false>(g, st) : launch_ws_k<A_MC, B_NC, false, false>(g, st);
}

// Runs the plan.  With plan.scratch != 0 the result is left as partial tiles in `part` (layout of k_gemm_reduce<128,128>,
// which the caller launches).
int gemm_ws_f64(const WsPlan& plan, int64_t m, int64_t n, int64_t k, const double* a, int64_t lda, bool a_mc,
                const double* b, int64_t ldb, bool b_nc, const double* w, double* c, int64_t ldc, double* part,
                cudaStream_t st) {
  WsArgs g;
  g.m = m; g.n = n; g.k = k; g.a = a; g.lda = lda; g.b = b; g.ldb = ldb; g.w = w; g.c = c; g.ldc = ldc;
  g.tiles_m = plan.tiles_m; g.tiles_n = plan.tiles_n;
  g.ntiles = plan.ntiles; g.splits = plan.splits; g.sym = plan.sym; g.kchunk = plan.kchunk;
  g.part = plan.scratch ? part : nullptr;
  if (a_mc && b_nc) return launch_ws<true, true>(g, st);
  if (a_mc && !b_nc) return launch_ws<true, false>(g, st);
  if (!a_mc && b_nc) return launch_ws<false, true>(g, st);
  return launch_ws<false, false>(g, st);
}

}  // namespace nb
