// The Hessian X^T diag(w) X of the Newton step (nums/models/glms.py:491-499) for 128 < d <= 256 on the INTEGER tensor
// path of sm_100a (tcgen05.mma kind::i8, int32 accumulators in tensor memory), exact to ~2^-40 of the largest
// entry of each column: an Ozaki-style error-free split.
//
// Why: FP64 has no tcgen05 kind; the DMMA pipe peaks at 37 TFLOP/s, which holds X^T W X at d = 256 to 28 ms per
// 1.25e7-row shard (k_tri_gram, 88 % of that peak) while the bytes take 4 ms.  The int8 path is ~100x faster per
// multiply, so a handful of exact integer products replace one FP64 product.
//
// Scheme.  With per-column powers of two 2^q_i (from the column maxima of |X|, computed once per block:
// nb_glm_prepare) every a_ki = x_ki * sqrt(w_k) is rounded to a 39-bit integer v_ki = rint(a_ki 2^q_i) and written as
// five balanced base-256 digits v = sum_p d_p 256^(4-p), d_p in [-128, 127] ("planes" p = 0..4 of int8, plane 0 most
// significant).  Then sum_k v_ki v_kj = sum_L 256^(8-L) S_L with S_L = sum over ordered digit pairs p + p' = L of the
// plane products P_pp' = plane_p^T plane_p'.  Every P is an exact int8 x int8 -> int32 tensor product (at most 32768
// rows per accumulation: 2 * 2^14 * 2^15 < 2^31).  Levels L <= 5 are kept (11 products, using P_p'p = P_pp'^T); the
// dropped levels are below 2^-48 of the column scale squared.  Rounding to v is the approximation that matters:
// half a unit in 2^-38 of the column maximum per element.
//
// Blocks of millions of rows (NB_GLM_INT8_FOUR_PLANES, scheme 2 below): the rounding errors of the rows cancel as 1 / sqrt(n), so
// 31 bits per element are enough for the 1e-10 bar: FOUR planes, eight products (levels p + p' <= 4), one per accumulator.
// The rounding to 31 bits is dithered (see k_oz_rows) so that the cancellation holds for any data, repeated rows included.
//
//   pass A  k_oz_rows    reads X once (HBM-bound): eta, link, g, u, sum w as k_glm_rows does, and writes the digit
//                        planes [plane][row][256] (4 or 5 bytes per element instead of the 8 of w or of X*sqrt(w)).
//   pass B  k_oz_gemm    CTA pairs (cta_group::2, M = 256, N = 256): TMA tensor copies (128-byte swizzle) bring
//                        [64 or 128 rows][128 bytes] tiles of the planes a pair needs, one elected thread issues
//                        tcgen05.mma kind::i8 on MN-major operands, the accumulators of two "slots" fill the 512
//                        tensor-memory columns.  Eight slots are needed (a slot holds same-level products of one
//                        kind: P_pp' with p < p' enter H as P + P^T, P_pp once), so four groups of pairs work on the
//                        same rows at about the same time and share most plane reads in L2 (measured DRAM traffic:
//                        1.4x the planes; holding the pairs in lock step through global progress counters brought it to
//                        1.0x and cost 2.5x in time, and was removed).  Epilogue warps read the
//                        accumulators (tcgen05.ld) after each range of rows and store them as int32 partials.
//   finish  k_oz_reduce  sums the partials exactly (int64);  k_oz_finish applies 256^(8-L) and 2^(-q_i-q_j), mirrors.
//
// A value that does not fit the 39 bits (the prepared column maxima do not belong to this X, or X holds Inf/NaN)
// raises a flag that turns H into NaN: the caller's solve reports it; nothing is silently wrong.
#include "glm_common.cuh"
#include <cuda.h>
#include <mutex>

namespace nb {

constexpr int OZ_S = 5;                 // digit planes
constexpr int OZ_COLS = 256;            // plane width (columns of X, zero padded)
constexpr int OZ_HALF = 128;
constexpr int OZ_KC = 64;               // rows per pipeline stage (schemes 0, 1; scheme 2 runs 128-row stages)
constexpr int OZ_KC_MAX = 128;
constexpr int OZ_SMEM_TILES = 26 * 64 * OZ_HALF;   // 208 KB of plane tiles per CTA: (tiles - 1) / np stages, the last tile stays zero
constexpr int OZ_MAX_STAGES = 12;
constexpr int OZ_MAX_RANGE = 32768;     // rows per accumulation: two products of digits in [-128, 127] share an int32 accumulator
constexpr int OZ_MAX_RANGE_SINGLE = 65536;      // scheme 2: one product per accumulator, |sum| <= 2^30
constexpr int OZ_GROUPS = 4;
constexpr int OZ_SLOTS = 2;
constexpr int OZ_ROWS_THREADS = 256;
// the gradient-only instantiation needs 168 registers: nine warps fit a scheduler's 16 K registers (three warps on one of the four),
// and their 216 KB of ring the shared memory; the plane kernels (215 registers) spill at nine
constexpr int OZ_GRAD_THREADS = 288;
__host__ __device__ constexpr int oz_rows_threads(int nplanes) { return nplanes == 0 ? OZ_GRAD_THREADS : OZ_ROWS_THREADS; }
constexpr int OZ_R = 4;                 // rows per warp step in pass A
// |v| < 2^vbits: five planes carry 39-bit values, four planes 31-bit values
__host__ __device__ constexpr int oz_vbits(int nplanes) { return nplanes == 5 ? 38 : 30; }

// products of each group: {slot, plane of A, plane of B}; planes the group stages.  Levels 0..5 (11 products, 8 slots:
// the first dropped terms, level 6, are 2^-48 of the column scale squared).
struct OzProduct { int slot, pa, pb; };
// Three schemes.  2: four planes, see the table.  0: levels 0..5, 11 products (any n).  1: levels 0..4, 9 products: the level-5 terms (1,4), (2,3) are sums of
// n products of independent, zero-mean low digits, i.e. a relative contribution ~ 2^(2r - 38.6) / sqrt(n) to H (r = range_log2
// of nb_glm_prepare); the caller drops them (NB_GLM_INT8_NO_LEVEL5) when that is below 1e-11 — at 1.25e7 Gaussian rows it is
// 3e-13 — which removes four of the sixteen plane streams (ncu on scheme 2: the tensor pipe is busy 92 % of the elapsed cycles, the
// kernel follows the clock the board grants it).
// Grouping rule of scheme 0: no group stages more than four planes; every group stages four planes and issues three products per
// stage, so the four pairs of a lane advance at the same pace and stay close: the pair that asks for a tile first pays the
// trip to HBM, the others find it in L2.  Group 3 has only two useful products on three planes; its third is plane 3 against a
// tile of zeros (pb = -1, adds nothing) -- the idle tensor time buys the lock step (DRAM re-reads 2.1x -> 1.66x).
constexpr int OZ_SCHEMES = 3;
__constant__ OzProduct c_oz_products[OZ_SCHEMES][OZ_GROUPS][3] = {
    {{{0, 0, 3}, {0, 1, 2}, {1, 0, 0}},        // level 3 | level 0
     {{0, 0, 4}, {0, 1, 3}, {1, 0, 1}},        // level 4, pairs p < p' | level 1
     {{0, 1, 4}, {0, 2, 3}, {1, 2, 2}},        // level 5 | level 4, (2,2)
     {{0, 0, 2}, {1, 1, 1}, {0, 3, -1}}},      // level 2: (0,2) | (1,1); plane 3 x zeros
    {{{0, 0, 3}, {0, 1, 2}, {1, 0, 0}},        // level 3 | level 0
     {{0, 0, 4}, {0, 1, 3}, {1, 0, 1}},        // level 4, pairs p < p' | level 1
     {{0, 2, 2}, {1, 1, 1}, {-1, 0, 0}},       // level 4, (2,2) | level 2, (1,1)
     {{0, 0, 2}, {-1, 0, 0}, {-1, 0, 0}}},     // level 2, (0,2) | (slot 1 unused)
    // scheme 2: FOUR planes (31-bit values), levels 0..4: eight products, two per group, every accumulator slot in use.
    // Three groups pair a p = p' product with a p < p' one on two planes; the fourth needs three planes: nine plane reads
    // per stage for eight products
    {{{0, 0, 3}, {1, 1, 3}, {-1, 0, 0}},       // level 3, (0,3) | level 4, (1,3)
     {{0, 0, 1}, {1, 0, 0}, {-1, 0, 0}},       // level 1, (0,1) | level 0
     {{0, 0, 2}, {1, 2, 2}, {-1, 0, 0}},       // level 2, (0,2) | level 4, (2,2)
     {{0, 1, 2}, {1, 1, 1}, {-1, 0, 0}}},      // level 3, (1,2) | level 2, (1,1)
};
__constant__ int c_oz_nplanes[OZ_SCHEMES][OZ_GROUPS] = {{4, 4, 4, 4}, {4, 4, 2, 2}, {3, 2, 2, 2}};
__constant__ int c_oz_planes[OZ_SCHEMES][OZ_GROUPS][4] = {{{0, 1, 2, 3}, {0, 1, 3, 4}, {1, 2, 3, 4}, {0, 1, 2, 3}},
                                                          {{0, 1, 2, 3}, {0, 1, 3, 4}, {1, 2, 0, 0}, {0, 2, 0, 0}},
                                                          {{0, 1, 3, 0}, {0, 1, 0, 0}, {0, 2, 0, 0}, {1, 2, 0, 0}}};
// slot (group, s) -> level (-1: unused) and kind (1 = pairs p < p': enters as P + P^T; 0 = p == p')
constexpr int OZ_LEVELS = 6;
__constant__ int c_oz_slot_level[OZ_SCHEMES][OZ_GROUPS][OZ_SLOTS] = {{{3, 0}, {4, 1}, {5, 4}, {2, 2}}, {{3, 0}, {4, 1}, {4, 2}, {2, -1}},
                                                                     {{3, 4}, {1, 0}, {2, 4}, {3, 2}}};
__constant__ int c_oz_slot_off[OZ_SCHEMES][OZ_GROUPS][OZ_SLOTS] = {{{1, 0}, {1, 1}, {1, 0}, {1, 0}}, {{1, 0}, {1, 1}, {0, 0}, {1, 0}},
                                                                   {{1, 1}, {1, 0}, {1, 0}, {1, 0}}};
static const int h_oz_scheme_planes[OZ_SCHEMES] = {5, 5, 4};      // digit planes of each scheme
static const int h_oz_scheme_kc[OZ_SCHEMES] = {64, 64, 128};        // rows per pipeline stage

struct OzPrep {               // nb_glm_prepare output (device memory)
  int32_t colexp[OZ_COLS];    // max|x[:, i]| < 2^colexp[i];  INT32_MIN for an all-zero column
  int32_t bad;                // non-finite entry seen
  int32_t range_log2;         // ceil(log2(max_i max|x[:, i]| / rms(x[:, i]))): how many leading bits the largest entries cost
  int32_t pad[62];
};
constexpr int OZ_PREP_CTAS = 592;

// ------------------------------------------------------------------------------------------------ prepare
// Column maxima of |X| (-> the fixed-point scale of each column) and column sums of squares (-> how far the typical
// entry sits below that scale; fixed summation order, so the verdict does not change from run to run).
__device__ __forceinline__ void oz_ld256(const double* p, double& a, double& b, double& c, double& d) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

// 64 threads cover a row with one 32-byte load each (columns 4t .. 4t+3); the four row groups of a CTA take rows
// 4 apart, four loads in flight per thread.  VEC = false: 8-byte loads for blocks that are not 32-byte aligned.
template <bool VEC>
__global__ void __launch_bounds__(256)
k_oz_colmax(const double* __restrict__ x, int64_t n, int64_t d, int64_t ldx, unsigned long long* __restrict__ maxbits,
            double* __restrict__ sumsq) {
  __shared__ unsigned long long s_m[4][OZ_COLS];
  __shared__ double s_q[4][OZ_COLS];
  const int t = threadIdx.x & 63, rg = threadIdx.x >> 6;
  const int c0 = 4 * t;
  unsigned long long m[4] = {0, 0, 0, 0};
  double q[4] = {0.0, 0.0, 0.0, 0.0};
  const int64_t stride = (int64_t)gridDim.x * 4;
  auto take = [&](double v, int e) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(fabs(v));
    m[e] = bits > m[e] ? bits : m[e];
    q[e] = fma(v, v, q[e]);
  };
  if (c0 < d) {
    int64_t r = (int64_t)blockIdx.x * 4 + rg;
    if constexpr (VEC) {
      for (; r + 3 * stride < n; r += 4 * stride) {
        double v[4][4];
#pragma unroll
        for (int k = 0; k < 4; k++) oz_ld256(x + (r + k * stride) * ldx + c0, v[k][0], v[k][1], v[k][2], v[k][3]);
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
          for (int e = 0; e < 4; e++) take(v[k][e], e);
      }
      for (; r < n; r += stride) {
        double v0, v1, v2, v3;
        oz_ld256(x + r * ldx + c0, v0, v1, v2, v3);
        take(v0, 0); take(v1, 1); take(v2, 2); take(v3, 3);
      }
    } else {
      for (; r < n; r += stride)
#pragma unroll
        for (int e = 0; e < 4; e++) if (c0 + e < d) take(x[r * ldx + c0 + e], e);
    }
  }
#pragma unroll
  for (int e = 0; e < 4; e++) { s_m[rg][c0 + e] = m[e]; s_q[rg][c0 + e] = q[e]; }
  __syncthreads();
  const int c = threadIdx.x;
  unsigned long long mm = max(max(s_m[0][c], s_m[1][c]), max(s_m[2][c], s_m[3][c]));
  if (c < d && mm) atomicMax(&maxbits[c], mm);      // non-negative doubles order like their bit patterns (NaN above Inf)
  sumsq[(size_t)blockIdx.x * OZ_COLS + c] = (s_q[0][c] + s_q[1][c]) + (s_q[2][c] + s_q[3][c]);
}

__global__ void k_oz_colexp(const unsigned long long* __restrict__ maxbits, const double* __restrict__ sumsq, int nparts,
                            int64_t n, int64_t d, OzPrep* __restrict__ prep) {
  const int c = threadIdx.x;
  int bad = 0, range = 0;
  if (c < OZ_COLS) {
    int e = INT32_MIN;
    if (c < d) {
      const unsigned long long b = maxbits[c];
      if (b >= 0x7ff0000000000000ull) bad = 1;
      else if (b != 0) {
        const double m = __longlong_as_double((long long)b);
        int ex;
        frexp(m, &ex);            // m = f 2^ex, f in [0.5, 1)  =>  m < 2^ex
        e = ex;
        double ss = 0.0;
        for (int k = 0; k < nparts; k++) ss += sumsq[(size_t)k * OZ_COLS + c];
        const double rms = sqrt(ss / (double)(n > 0 ? n : 1));
        if (rms > 0.0 && isfinite(rms)) range = (int)ceil(log2(m / rms));
      }
    }
    prep->colexp[c] = e;
  }
  bad = __syncthreads_or(bad);
  __shared__ int s_range;
  if (c == 0) s_range = 0;
  __syncthreads();
  atomicMax(&s_range, range);
  __syncthreads();
  if (c == 0) { prep->bad = bad; prep->range_log2 = s_range; }
}

// ------------------------------------------------------------------------------------------------ pass A
struct OzRowsArgs {
  const double* x; int64_t n, d, ldx;
  const void* y; int ydt;
  const double* beta;
  int family, skip_obj;
  int ew;                       // sqrt(w) <= 2^ew for this family
  const OzPrep* prep;
  int8_t* planes; int64_t plane_stride;     // plane p at planes + p * plane_stride, row k at + k * 256
  double* partials;             // [gridDim.x][2*d + 3]  (layout of k_glm_rows, summed by k_glm_rows_finish)
  int* flag;                    // set when a value leaves the 39-bit range
  uint32_t seed;                // four planes: seed of the rounding dither (the chunk's first row)
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

constexpr int OZ_ROWS_STAGES = 3;        // batches of OZ_R rows in flight per warp (bulk copies into its own ring)

__device__ __forceinline__ void oz_rows_fetch(const OzRowsArgs& a, int64_t batch, double* stage, uint64_t* bar) {
  // one lane: rows [batch*R, +R) of X -> stage[R][d] (compact rows), completion on `bar`
  const int64_t row0 = batch * OZ_R;
  const int64_t left = a.n - row0;
  const int rows = left < OZ_R ? (int)left : OZ_R;
  const unsigned row_bytes = (unsigned)(a.d * sizeof(double));
  const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_s), "r"(row_bytes * (unsigned)rows) : "memory");
  if (a.ldx == a.d) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"((unsigned)__cvta_generic_to_shared(stage)), "l"(a.x + row0 * a.ldx), "r"(row_bytes * (unsigned)rows), "r"(bar_s)
                 : "memory");
  } else {
    for (int r = 0; r < rows; r++)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"((unsigned)__cvta_generic_to_shared(stage + (size_t)r * a.d)), "l"(a.x + (row0 + r) * a.ldx), "r"(row_bytes),
                      "r"(bar_s) : "memory");
  }
}

// Latency, not bandwidth, held the first version (direct 32-byte loads, 8 warps of 228 registers per SM, each warp
// alternating between waiting for its rows and a ~600-instruction batch) to 4.6 TB/s of combined traffic.  Each warp now
// keeps OZ_ROWS_STAGES batches in flight as bulk copies into its own shared-memory ring (lane 0 issues the copy for batch
// i + STAGES - 1 before the warp starts on batch i); no producer warp and no "empty" barriers are needed because the
// warp that refills a stage is the one that has just finished reading it.
template <int NP>
__global__ void __launch_bounds__(oz_rows_threads(NP), 1)
k_oz_rows(const OzRowsArgs a) {
  constexpr int THREADS = oz_rows_threads(NP);
  constexpr int NW = THREADS / 32;
  constexpr int R = OZ_R;
  constexpr int NS = OZ_ROWS_STAGES;
  extern __shared__ __align__(128) double oz_ring[];          // [NW][NS][R][d]
  __shared__ __align__(8) uint64_t ring_full[NW][NS];
  __shared__ double s_g[OZ_COLS], s_u[OZ_COLS], s_sc[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t d = a.d, n = a.n;
  for (int j = tid; j < OZ_COLS; j += THREADS) { s_g[j] = 0.0; s_u[j] = 0.0; }
  if (tid < 4) s_sc[tid] = 0.0;
  // The inner loop reads whole 256-wide rows of its ring without guards: columns beyond d and rows beyond n take part in
  // the arithmetic with zero factors (beta, the column scale, the row's residual and weight are 0 there), so what they
  // read only has to be finite: the ring starts out as zeros and afterwards holds rows of X.
  for (int j = tid; j < NW * NS * R * OZ_COLS; j += THREADS) oz_ring[j] = 0.0;
  if (tid < NW * NS) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((unsigned)__cvta_generic_to_shared(&ring_full[tid / NS][tid % NS])));
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  // this lane owns columns cA..cA+3 and cB..cB+3 (element slots 0..3 and 4..7)
  const int cA = 4 * lane, cB = OZ_HALF + 4 * lane;
  double bet[8], cs[8];
#pragma unroll
  for (int e = 0; e < 8; e++) {
    const int col = (e < 4 ? cA : cB) + (e & 3);
    bet[e] = col < d ? a.beta[col] : 0.0;
    if constexpr (NP == 0) { cs[e] = 0.0; continue; }
    const int ce = a.prep->colexp[col];
    // v = x sqrt(w) 2^q with q = 38 - colexp - ew, so that |v| < 2^38 (four planes drop the lowest byte after a dither)
    int q = oz_vbits(5) - ce - a.ew;
    cs[e] = (col < d && ce != INT32_MIN && q < 1000 && q > -1000) ? scalbn(1.0, q) : 0.0;
  }
  const double beta0 = a.beta[d];
  // 1.5 * 2^52 + sum_s 128 * 256^s: after fma(x, scale, MAGIC) the low 40 mantissa bits are v + bias, i.e. the
  // bytes d_s + 128 of the balanced base-256 digits of v (round to nearest even comes with the fma)
  // Four planes: the same 39-bit value, but the lowest byte carries no bias and receives a uniform 8-bit dither; the four
  // upper bytes are then floor((v + u) / 256) + bias, a rounding of v / 256 whose error has zero mean whatever v is.
  const double MAGIC = 6755399441055744.0 + (NP == 5 ? 551911719040.0 : 551911718912.0);
  constexpr uint32_t HI_OK = 0x43380000u, HI_MASK = 0xffffff00u;
  constexpr uint32_t DITHER_LO = 0x80808000u;        // low word of MAGIC (four planes) without the dither byte

  double g[8], u[8];
#pragma unroll
  for (int e = 0; e < 8; e++) { g[e] = 0.0; u[e] = 0.0; }
  double gi = 0.0, ui = 0.0, ob = 0.0;
  uint32_t bad = 0;
  // Four planes keep 31 bits of x sqrt(w), and the 1e-10 bar on H then rests on the rounding errors of the n rows
  // cancelling (~ 1 / sqrt(n)).  Round-to-nearest errors do not cancel over REPEATED rows (one-hot data, tiled test
  // matrices): the same row rounds the same way every time.  The dither byte is hash(row) ^ (a constant per lane): a
  // given column sees independent uniform dithers down the rows, so its rounding errors are zero-mean and independent
  // of the data.  It depends on the row's position only: the same block gives the same bits.
  const uint32_t lane_key = ((uint32_t)lane * 0x9E3779B1u) >> 24;

  const int64_t nbatches = (n + R - 1) / R;
  const int64_t per_it = (int64_t)gridDim.x * NW;
  const int64_t first = (int64_t)blockIdx.x * NW + warp;
  double* my_ring = oz_ring + (size_t)warp * NS * R * d;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NS - 1; s++)
      if (first + s * per_it < nbatches) oz_rows_fetch(a, first + s * per_it, my_ring + (size_t)s * R * d, &ring_full[warp][s]);
  }
  int stage = 0;
  unsigned phase = 0;
  long long ynext = 0;                      // y (float64 or int64: 8 bytes either way) of this lane's row in the next batch
  // after the transposed reduction below, the eight lanes with (lane >> 3) == r hold eta of row r of the batch
  const int myr = lane >> 3;
  if (first < nbatches && first * R + myr < n) ynext = reinterpret_cast<const long long*>(a.y)[first * R + myr];
  // this lane's byte position in every plane for the first row of the batch: advanced once per batch, so that the forty
  // plane stores of a batch need no address arithmetic beyond constant offsets
  int8_t* pl[NP > 0 ? NP : 1];
#pragma unroll
  for (int p = 0; p < NP; p++) pl[p] = a.planes + p * a.plane_stride + first * R * OZ_COLS + cA;
  const int64_t pl_step = per_it * R * OZ_COLS;
  for (int64_t batch = first; batch < nbatches; batch += per_it) {
    const int64_t row0 = batch * R;
    // y of the row this lane evaluates was asked for one batch ago (as raw bits: converting an int64 at once would make
    // the warp wait for the load right there -- a fifth of all stall samples in the first version); ask for the next one
    const int64_t myrow = row0 + myr;
    const bool inb = myrow < n;
    const long long ybits = ynext;
    {
      const int64_t nrow = myrow + per_it * R;
      ynext = nrow < n ? reinterpret_cast<const long long*>(a.y)[nrow] : 0;
    }
    // refill the stage this warp finished reading in the previous iteration
    if (lane == 0) {
      const int64_t ahead = batch + (NS - 1) * per_it;
      const int sa = stage == 0 ? NS - 1 : stage - 1;
      if (ahead < nbatches) oz_rows_fetch(a, ahead, my_ring + (size_t)sa * R * d, &ring_full[warp][sa]);
    }
    {
      const unsigned bar_s = (unsigned)__cvta_generic_to_shared(&ring_full[warp][stage]);
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "WAIT_%=:\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
          "@p bra DONE_%=;\n\t"
          "bra WAIT_%=;\n\t"
          "DONE_%=:\n\t}"
          :: "r"(bar_s), "r"(phase) : "memory");
    }
    const double* st_rows = my_ring + (size_t)stage * R * d;
    double xv[R][8];
#pragma unroll
    for (int r = 0; r < R; r++) {
      const double* xr = st_rows + (size_t)r * d;
      const double2 p0 = *reinterpret_cast<const double2*>(xr + cA), p1 = *reinterpret_cast<const double2*>(xr + cA + 2);
      const double2 p2 = *reinterpret_cast<const double2*>(xr + cB), p3 = *reinterpret_cast<const double2*>(xr + cB + 2);
      xv[r][0] = p0.x; xv[r][1] = p0.y; xv[r][2] = p1.x; xv[r][3] = p1.y;
      xv[r][4] = p2.x; xv[r][5] = p2.y; xv[r][6] = p3.x; xv[r][7] = p3.y;
    }
    __syncwarp();                                  // every lane holds its part of the stage in registers: it may be refilled
    if (++stage == NS) { stage = 0; phase ^= 1u; }
    double part[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int e = 0; e < 4; e++) { s0 = fma(xv[r][e], bet[e], s0); s1 = fma(xv[r][4 + e], bet[4 + e], s1); }
      part[r] = s0 + s1;
    }
    // four row sums over 32 lanes with 6 exchanges instead of 20: halves of the warp trade the rows they will not keep
    double eta;
    {
      const bool up16 = (lane & 16) != 0, up8 = (lane & 8) != 0;
      // lanes 0-15 keep rows 0,1 and lanes 16-31 rows 2,3
      const double k0 = up16 ? part[2] : part[0], k1 = up16 ? part[3] : part[1];
      const double g0 = up16 ? part[0] : part[2], g1 = up16 ? part[1] : part[3];
      const double a0 = k0 + __shfl_xor_sync(0xffffffffu, g0, 16);
      const double a1 = k1 + __shfl_xor_sync(0xffffffffu, g1, 16);
      // bit 3 of the lane picks one of the two
      const double keep = up8 ? a1 : a0, give = up8 ? a0 : a1;
      double v = keep + __shfl_xor_sync(0xffffffffu, give, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      eta = v + beta0;
    }
    const double yv = !inb ? 0.0 : (a.ydt == NB_F64 ? __longlong_as_double(ybits) : (double)ybits);
    LinkOut lo = a.skip_obj ? glm_link_noobj_rt(a.family, eta, yv) : glm_link_rt(a.family, eta, yv);
    if (!inb) { lo.resid = 0.0; lo.wgt = 0.0; lo.obj = 0.0; }
    const double sw = sqrt(lo.wgt);
    uint32_t rowkey = 0;
    if constexpr (NP == 4) {
      uint32_t h = ((uint32_t)myrow + a.seed) * 0x9E3779B1u;
      h ^= h >> 15; h *= 0x85EBCA77u;
      rowkey = h >> 24;
    }
    if ((lane & 7) == 0 && inb) { ob += lo.obj; gi += lo.resid; ui += lo.wgt; }
#pragma unroll
    for (int r = 0; r < R; r++) {
      const double rr = __shfl_sync(0xffffffffu, lo.resid, 8 * r);
      const double ww = __shfl_sync(0xffffffffu, lo.wgt, 8 * r);
      const double sq = __shfl_sync(0xffffffffu, sw, 8 * r);
      uint32_t lo32[8], hi32[8];
      uint32_t dlo = 0;
      if constexpr (NP == 4) dlo = (__shfl_sync(0xffffffffu, rowkey, 8 * r) ^ lane_key) | DITHER_LO;
      // four planes: one dither byte per (row, lane) serves the lane's eight columns -- what matters is that a COLUMN sees
      // independent uniform bytes down the rows; columns sharing a byte only correlate their errors with each other, which
      // enters H at second order (2^-57)
      const double add = NP == 5 ? MAGIC : __hiloint2double((int)0x43380080u, (int)dlo);
#pragma unroll
      for (int e = 0; e < 8; e++) {
        g[e] = fma(rr, xv[r][e], g[e]);
        if constexpr (NP == 0) continue;          // gradient / objective pass: no u, no digits
        u[e] = fma(ww, xv[r][e], u[e]);
        const double t = fma(xv[r][e], __dmul_rn(sq, cs[e]), add);
        lo32[e] = (uint32_t)__double2loint(t);
        hi32[e] = (uint32_t)__double2hiint(t);
        bad |= hi32[e] ^ HI_OK;                     // (the digit byte of the high word is masked out at the end)
      }
      // (rows beyond n, at most three in the last batch, store zero digits: the plane buffer is a whole number of 128-row
      // stages long, and pass B's tensor map ends at row n anyway)
      if constexpr (NP > 0) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const uint32_t* L = lo32 + 4 * h;
          const uint32_t* H = hi32 + 4 * h;
          // 4x4 byte transpose of the low words: w_s = {byte s of element 0, 1, 2, 3}
          const uint32_t t01a = prmt(L[0], L[1], 0x5140), t01b = prmt(L[0], L[1], 0x7362);
          const uint32_t t23a = prmt(L[2], L[3], 0x5140), t23b = prmt(L[2], L[3], 0x7362);
          const uint32_t w1 = prmt(t01a, t23a, 0x7632);
          const uint32_t w2 = prmt(t01b, t23b, 0x5410), w3 = prmt(t01b, t23b, 0x7632);
          const uint32_t w4 = prmt(prmt(H[0], H[1], 0x0040), prmt(H[2], H[3], 0x0040), 0x5410);
          const int off = r * OZ_COLS + h * OZ_HALF;
          // plane p holds digit s = 4 - p (byte 4 - p of the 40-bit value); signed digit = byte ^ 0x80; columns beyond d have
          // scale 0, i.e. digits 0 (the planes are read 256 wide)
          *reinterpret_cast<uint32_t*>(pl[0] + off) = w4 ^ 0x80808080u;
          *reinterpret_cast<uint32_t*>(pl[1] + off) = w3 ^ 0x80808080u;
          *reinterpret_cast<uint32_t*>(pl[2] + off) = w2 ^ 0x80808080u;
          *reinterpret_cast<uint32_t*>(pl[3] + off) = w1 ^ 0x80808080u;
          if constexpr (NP == 5) {
            const uint32_t w0 = prmt(t01a, t23a, 0x5410);
            *reinterpret_cast<uint32_t*>(pl[4] + off) = w0 ^ 0x80808080u;
          }
        }
      }
    }
#pragma unroll
    for (int p = 0; p < NP; p++) pl[p] += pl_step;
  }
  if (bad & HI_MASK) atomicExch(a.flag, 1);

  // CTA reduction in warp order (deterministic)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gi += __shfl_down_sync(0xffffffffu, gi, o);
    ui += __shfl_down_sync(0xffffffffu, ui, o);
    ob += __shfl_down_sync(0xffffffffu, ob, o);
  }
  for (int q = 0; q < NW; q++) {
    if (warp == q) {
#pragma unroll
      for (int e = 0; e < 8; e++) {
        const int col = (e < 4 ? cA : cB) + (e & 3);
        s_g[col] += g[e]; s_u[col] += u[e];
      }
      if (lane == 0) { s_sc[0] += gi; s_sc[1] += ui; s_sc[2] += ob; }
    }
    __syncthreads();
  }
  double* out = a.partials + (size_t)blockIdx.x * (2 * d + 3);
  for (int64_t j = tid; j < d; j += THREADS) { out[j] = s_g[j]; out[d + j] = s_u[j]; }
  if (tid == 0) { out[2 * d] = s_sc[0]; out[2 * d + 1] = s_sc[1]; out[2 * d + 2] = s_sc[2]; }
}

// ------------------------------------------------------------------------------------------------ pass B
__device__ __forceinline__ uint32_t oz_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void oz_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(oz_smem(bar)), "r"(count));
}
__device__ __forceinline__ void oz_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(oz_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void oz_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(oz_smem(bar)) : "memory");
}
// arrive on the same barrier of CTA `rank` of the cluster
__device__ __forceinline__ void oz_mbar_arrive_remote(uint64_t* bar, uint32_t rank) {
  asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\tmbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
               :: "r"(oz_smem(bar)), "r"(rank) : "memory");
}
__device__ __forceinline__ void oz_mbar_wait(uint64_t* bar, unsigned parity) {
  const uint32_t addr = oz_smem(bar);
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(addr), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t oz_cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void oz_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared-memory matrix descriptor, MN-major operand with the 128-byte swizzle: atoms of 8 rows x 128 bytes, groups of
// 8 rows (K) are 1024 bytes apart; one atom along MN (the CTA's 128 columns)
__device__ __forceinline__ uint64_t oz_desc(uint32_t saddr) {
  uint64_t dsc = 0;
  dsc |= (uint64_t)((saddr >> 4) & 0x3fff);
  dsc |= (uint64_t)1 << 16;                              // leading byte offset (unused: one atom along MN)
  dsc |= (uint64_t)((8 * OZ_HALF) >> 4) << 32;           // stride byte offset: next group of 8 rows
  dsc |= (uint64_t)1 << 46;                              // descriptor version
  dsc |= (uint64_t)2 << 61;                              // SWIZZLE_128B
  return dsc;
}
// instruction descriptor, kind::i8: D = s32, A = B = signed 8 bit, both MN-major, M = 256 (pair), N = 256
constexpr uint32_t OZ_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((256u >> 3) << 17) | ((256u >> 4) << 24);

// one accumulator row (256 int32: this thread's lane of tensor memory, columns taddr .. taddr + 255) -> dst
__device__ __forceinline__ void oz_drain_row(uint32_t taddr0, int32_t* dst) {
  for (int c0 = 0; c0 < 256; c0 += 32) {
    uint32_t v[32];
    const uint32_t taddr = taddr0 + (uint32_t)c0;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      *reinterpret_cast<uint4*>(dst + c0 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
  }
}

struct OzGemmArgs {
  int64_t n;
  int range_rows;           // rows per accumulation item (multiple of OZ_KC_MAX)
  int nranges;
  int lanes;                // pairs per group
  int32_t* part;            // [nranges][OZ_GROUPS][OZ_SLOTS][256][256]
  int scheme;               // which product table
  int debug;                // NUMS_B200_OZ_DEBUG: pair 0 prints the time from a stage's copies to its arrival
};

// Pair p works for group p % 4 on the row ranges p / 4, p / 4 + lanes, ...: the four pairs of a lane walk the same rows;
// whichever asks for a tile first brings it into L2 and the others mostly find it there.
template <int KC>
__global__ void __launch_bounds__(192, 1)
k_oz_gemm(const __grid_constant__ CUtensorMap tmap, const OzGemmArgs a) {
  constexpr uint32_t OZ_TILE = KC * OZ_HALF;                   // one plane, one column half, one stage: 8 or 16 KB
  constexpr int OZ_TILES = OZ_SMEM_TILES / (int)OZ_TILE;
  extern __shared__ __align__(1024) uint8_t oz_raw[];
  uint8_t* tiles = (uint8_t*)(((uintptr_t)oz_raw + 1023) & ~(uintptr_t)1023);       // [stage][plane slot][64][128]
  uint8_t* zero_tile = tiles + (size_t)(OZ_TILES - 1) * OZ_TILE;                      // the last tile stays zero
  __shared__ __align__(8) uint64_t full[OZ_MAX_STAGES], empty[OZ_MAX_STAGES], acc_full, acc_empty;
  __shared__ uint32_t tmem_base_s;
  __shared__ long long dbg_issue[OZ_MAX_STAGES];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = oz_cluster_rank();
  const int pair = blockIdx.x >> 1;
  const int group = pair % OZ_GROUPS, lane_id = pair / OZ_GROUPS;
  const int np = c_oz_nplanes[a.scheme][group];
  const int nst = (OZ_TILES - 1) / np < OZ_MAX_STAGES ? (OZ_TILES - 1) / np : OZ_MAX_STAGES;     // pipeline stages of this group

  for (int i = threadIdx.x; i < (int)(OZ_TILE / 16); i += blockDim.x) reinterpret_cast<uint4*>(zero_tile)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // the tensor core reads it through the async proxy
  if (threadIdx.x == 0) {
    for (int s = 0; s < OZ_MAX_STAGES; s++) { oz_mbar_init(&full[s], 2); oz_mbar_init(&empty[s], 1); }

    oz_mbar_init(&acc_full, 1);
    oz_mbar_init(&acc_empty, 8);          // four epilogue warps (2..5) of each CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(oz_smem(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  oz_cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  const int chunks_per_range = a.range_rows / KC;

  if (warp == 0) {
    // ------------------------------------------------------------ producer (both CTAs): TMA tensor copies
    if (lane == 0) {
      const uint32_t lead_full0 = oz_smem(&full[0]) & 0xFEFFFFFFu;      // the leader's copy of the barrier
      int it = 0, st = 0;
      unsigned ph = 1;                        // parity that lets the first pass over the ring through
      for (int range = lane_id; range < a.nranges; range += a.lanes) {
        const int64_t r0 = (int64_t)range * a.range_rows;
        int64_t rem = a.n - r0;
        const int nch = (int)((rem < a.range_rows ? (rem + KC - 1) / KC : chunks_per_range));
        for (int ch = 0; ch < nch; ch++, it++) {
          oz_mbar_wait(&empty[st], ph);
          if (rank == 0) oz_mbar_expect_tx(&full[st], 2u * (uint32_t)np * OZ_TILE);
          else oz_mbar_arrive_remote(&full[st], 0);
          if (a.debug) dbg_issue[st] = clock64();
          const uint32_t bar = lead_full0 + (uint32_t)st * 8u;
          const int c0 = (int)rank * OZ_HALF;
          const int c1 = (int)(r0 + (int64_t)ch * KC);
          for (int q = 0; q < np; q++) {
            const uint32_t dst = oz_smem(tiles + ((size_t)st * np + q) * OZ_TILE);
            const int c2 = c_oz_planes[a.scheme][group][q];
            asm volatile("cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
                         " [%0], [%1, {%3, %4, %5}], [%2];"
                         :: "r"(dst), "l"(&tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
          }
          if (++st == nst) { st = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA, one thread)
    if (rank == 0 && lane == 0) {
      // Everything the issue loop needs sits in registers before it starts: per product the operands' offsets inside a stage
      // (16-byte units), the accumulator's tensor-memory address and whether it is the first product into its slot.  One thread
      // has 128 cycles per MMA to build two descriptors and issue; constant-memory and local-memory reads in that path (the
      // first version indexed the tables per MMA) held the tensor pipe at ~55 %.
      uint32_t a_off[3], b_off[3], dcol[3], first[3];
      bool b_fixed[3], valid[3];
      {
        uint32_t seen = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const OzProduct pr = c_oz_products[a.scheme][group][k];
          valid[k] = pr.slot >= 0;
          int pa = 0, pb = 0;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (q < np && c_oz_planes[a.scheme][group][q] == pr.pa) pa = q;
            if (q < np && c_oz_planes[a.scheme][group][q] == pr.pb) pb = q;
          }
          a_off[k] = (uint32_t)pa * (OZ_TILE >> 4);
          b_fixed[k] = pr.pb < 0;
          b_off[k] = pr.pb < 0 ? (oz_smem(zero_tile) >> 4) : (uint32_t)pb * (OZ_TILE >> 4);
          const int slot = valid[k] ? pr.slot : 0;
          dcol[k] = tmem_base + (uint32_t)slot * 256u;
          first[k] = ((seen >> slot) & 1u) ? 0u : 1u;
          if (valid[k]) seen |= 1u << slot;
        }
      }
      const uint64_t desc_hi = oz_desc(0);                       // everything but the start address
      const uint32_t tiles_u = oz_smem(tiles) >> 4;
      const uint32_t stage_u = (uint32_t)np * (OZ_TILE >> 4);
      constexpr uint32_t KSTEP = (32 * OZ_HALF) >> 4;            // 32 rows of a tile, in 16-byte units
      int item = 0, st = 0;
      unsigned ph = 0;
      long long dbg_lat = 0, dbg_wait = 0, dbg_n = 0, dbg_t0 = clock64();
      for (int range = lane_id; range < a.nranges; range += a.lanes, item++) {
        const int64_t r0 = (int64_t)range * a.range_rows;
        int64_t rem = a.n - r0;
        const int nch = (int)((rem < a.range_rows ? (rem + KC - 1) / KC : chunks_per_range));
        if (item > 0) oz_mbar_wait(&acc_empty, (unsigned)((item - 1) & 1));      // the epilogue has drained the accumulators
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int ch = 0; ch < nch; ch++) {
          const long long w0 = a.debug ? clock64() : 0;
          oz_mbar_wait(&full[st], ph);
          if (a.debug) { const long long w1 = clock64(); dbg_wait += w1 - w0; dbg_lat += w1 - dbg_issue[st]; dbg_n++; }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t base_u = tiles_u + (uint32_t)st * stage_u;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            if (!valid[k]) continue;
            const uint32_t au = base_u + a_off[k];
            const uint32_t bu = b_fixed[k] ? b_off[k] : base_u + b_off[k];
#pragma unroll
            for (int ks = 0; ks < KC / 32; ks++) {
              const uint64_t ad = desc_hi | (uint64_t)((au + ks * KSTEP) & 0x3fffu);
              const uint64_t bd = desc_hi | (uint64_t)((bu + ks * KSTEP) & 0x3fffu);
              const uint32_t acc = (ch > 0 || ks > 0 || !first[k]) ? 1u : 0u;
              asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                           "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
                           :: "r"(dcol[k]), "l"(ad), "l"(bd), "r"(OZ_IDESC), "r"(acc), "r"(0u) : "memory");
            }
          }
          // the stage may be refilled (both CTAs) once these MMAs have read it
          asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                       :: "r"(oz_smem(&empty[st])), "h"((uint16_t)3) : "memory");
          if (++st == nst) { st = 0; ph ^= 1u; }
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     :: "r"(oz_smem(&acc_full)), "h"((uint16_t)3) : "memory");
      }
      if (a.debug && pair < 4 && dbg_n > 0)
        printf("k_oz_gemm pair %d (group %d, %d planes, %d stages): %lld stages, %.0f clk per stage, copy->arrival %.0f clk, "
               "issuer waited %.0f clk per stage\n", pair, group, np, nst, dbg_n, (double)(clock64() - dbg_t0) / dbg_n,
               (double)dbg_lat / dbg_n, (double)dbg_wait / dbg_n);
    }
  } else {
    // ------------------------------------------------------------ epilogue warps 2..5 (both CTAs)
    const int q = warp & 3;                          // lane quarter of tensor memory this warp may read
    int item = 0;
    for (int range = lane_id; range < a.nranges; range += a.lanes, item++) {
      oz_mbar_wait(&acc_full, (unsigned)(item & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row = (int)rank * 128 + q * 32 + lane;
      for (int s = 0; s < OZ_SLOTS; s++) {
        int32_t* dst = a.part + ((((size_t)range * OZ_GROUPS + group) * OZ_SLOTS + s) * 256 + row) * 256;
        oz_drain_row(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * 256), dst);
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) oz_mbar_arrive(&acc_empty);
        else oz_mbar_arrive_remote(&acc_empty, 0);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  oz_cluster_sync();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

// ------------------------------------------------------------------------------------------------ finish
// step 1: exact sums over the ranges, one thread per accumulator entry (coalesced)
__global__ void __launch_bounds__(256)
k_oz_reduce(const int32_t* __restrict__ part, int nranges, long long* __restrict__ sums) {
  const size_t total = (size_t)OZ_GROUPS * OZ_SLOTS * 65536;
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (size_t)gridDim.x * 256) {
    long long a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    int r = 0;
    for (; r + 4 <= nranges; r += 4) {
      a0 += part[(size_t)r * total + e];
      a1 += part[(size_t)(r + 1) * total + e];
      a2 += part[(size_t)(r + 2) * total + e];
      a3 += part[(size_t)(r + 3) * total + e];
    }
    for (; r < nranges; r++) a0 += part[(size_t)r * total + e];
    sums[e] = (a0 + a1) + (a2 + a3);
  }
}

struct OzFinishArgs {
  int64_t d, ldc;
  const long long* sums;    // [OZ_GROUPS][OZ_SLOTS][256][256]
  const OzPrep* prep;
  int ew;
  const int* flag;
  int scheme, nplanes;
  double* out;              // d x d corner, row stride ldc
};

// step 2: levels, scales, mirror
__global__ void __launch_bounds__(256)
k_oz_finish(const OzFinishArgs a) {
  const int64_t d = a.d;
  const int64_t nh = d * (d + 1) / 2;
  const bool poisoned = (*a.flag != 0) || (a.prep->bad != 0);
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < nh; e += (int64_t)gridDim.x * 256) {
    int64_t i = 0, rem = e;
    while (rem >= d - i) { rem -= d - i; i++; }
    const int64_t j = i + rem;
    long long lev[OZ_LEVELS] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int g = 0; g < OZ_GROUPS; g++)
#pragma unroll
      for (int s = 0; s < OZ_SLOTS; s++) {
        const long long* t = a.sums + ((size_t)g * OZ_SLOTS + s) * 65536;
        const int level = c_oz_slot_level[a.scheme][g][s];
        if (level < 0) continue;                 // a slot the scheme does not use holds whatever tensor memory held
        long long v = t[i * 256 + j];
        if (c_oz_slot_off[a.scheme][g][s]) v += t[j * 256 + i];
        lev[level] += v;
      }
    // sum_L lev[L] 256^(2 (nplanes - 1) - L), smallest terms first; every term is exact in double
    double acc = 0.0;
    const int top = 2 * (a.nplanes - 1);
#pragma unroll
    for (int L = OZ_LEVELS - 1; L >= 0; L--) acc += scalbn((double)lev[L], 8 * (top - L));
    const int vb = oz_vbits(a.nplanes);
    const int qi = vb - a.prep->colexp[i] - a.ew, qj = vb - a.prep->colexp[j] - a.ew;
    double h = (a.prep->colexp[i] == INT32_MIN || a.prep->colexp[j] == INT32_MIN) ? 0.0 : scalbn(scalbn(acc, -qi), -qj);
    if (poisoned) h = __longlong_as_double(0x7ff8000000000000ll);
    a.out[i * a.ldc + j] = h;
    a.out[j * a.ldc + i] = h;
  }
}

// ------------------------------------------------------------------------------------------------ host side
static inline size_t oz_align(size_t v) { return (v + 1023) & ~(size_t)1023; }

// The digit planes are 5/8 of the size of X: rows are therefore processed in up to OZ_MAX_CHUNKS chunks through one
// plane buffer of at most ~16 GB (pass A of a chunk, then pass B of the same chunk), so the scratch of the call does
// not grow with the block beyond that.
constexpr int OZ_MAX_CHUNKS = 4;                       // the row pass writes per-CTA partials per chunk (max_parts = 4 per SM)
constexpr int64_t OZ_CHUNK_ROWS = 12517376;            // 382 ranges of 32768 rows (>= the 1.25e7-row benchmark shard)

struct OzPlan {
  int range_rows, lanes, pairs;
  int nchunks; int64_t chunk_rows;      // rows per chunk (the last one may be shorter); a multiple of range_rows
  int nranges;                          // over all chunks
  size_t off_planes, off_part, off_sums, off_flag, total;
};

static OzPlan oz_plan(int device, int64_t n, int scheme) {
  OzPlan p;
  const int sms = dev_info(device).sms;
  p.pairs = ((sms / 2) / OZ_GROUPS) * OZ_GROUPS;
  if (p.pairs < OZ_GROUPS) p.pairs = OZ_GROUPS;
  p.lanes = p.pairs / OZ_GROUPS;
  if (n < 1) n = 1;
  // as few rounds of ranges per lane as the int32 accumulators allow, and ranges of equal length over those rounds
  const int64_t max_range = scheme == 2 ? OZ_MAX_RANGE_SINGLE : OZ_MAX_RANGE;
  const int64_t rounds = (n + (int64_t)p.lanes * max_range - 1) / ((int64_t)p.lanes * max_range);
  int64_t per_lane = (n + p.lanes * rounds - 1) / (p.lanes * rounds);
  int64_t rr = ((per_lane + OZ_KC_MAX - 1) / OZ_KC_MAX) * OZ_KC_MAX;
  if (rr < OZ_KC_MAX) rr = OZ_KC_MAX;
  if (rr > max_range) rr = max_range;
  p.range_rows = (int)rr;
  int64_t chunk = OZ_CHUNK_ROWS;
  if (const char* e = getenv("NUMS_B200_OZ_CHUNK_ROWS")) { const long long v = atoll(e); if (v >= OZ_KC_MAX) chunk = v; }   // tests
  if ((n + chunk - 1) / chunk > OZ_MAX_CHUNKS) chunk = (n + OZ_MAX_CHUNKS - 1) / OZ_MAX_CHUNKS;
  chunk = ((chunk + rr - 1) / rr) * rr;
  if (chunk > n) chunk = ((n + rr - 1) / rr) * rr;
  p.chunk_rows = chunk;
  p.nchunks = (int)((n + chunk - 1) / chunk);
  p.nranges = (int)((n + rr - 1) / rr);
  size_t cur = 0;
  p.off_planes = cur; cur += oz_align((size_t)OZ_S * (size_t)p.chunk_rows * OZ_COLS);
  p.off_part = cur; cur += oz_align((size_t)p.nranges * OZ_GROUPS * OZ_SLOTS * 65536 * sizeof(int32_t));
  p.off_sums = cur; cur += oz_align((size_t)OZ_GROUPS * OZ_SLOTS * 65536 * sizeof(long long));
  p.off_flag = cur; cur += 1024;
  p.total = cur;
  return p;
}

bool oz_applicable(int family, int64_t n, int64_t d, int64_t ldx, const double* x) {
  if (!(family == NB_GLM_LOGISTIC || family == NB_GLM_LOGISTIC_IRLS || family == NB_GLM_LINEAR)) return false;
  return n > 0 && d > 128 && d <= 256 && (d % 4 == 0) && (ldx % 4 == 0) && ((((uintptr_t)x) & 31) == 0) &&
         n < ((int64_t)1 << 31);
}

size_t oz_scratch_bytes(int device, int64_t n) {
  const size_t a = oz_plan(device, n, 0).total, b = oz_plan(device, n, 2).total;
  return a > b ? a : b;
}

typedef CUresult (*OzEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static OzEncodeFn oz_encode_fn() {
  static OzEncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = (OzEncodeFn)p;
  });
  return fn;
}

// pass A + pass B + finish.  `stats` receives g, objective, the intercept row/column of H (through
// k_glm_rows_finish, launched by the caller's usual path) and the d x d corner of H.
// phases: bit 0 = pass A, bit 1 = pass B + finish (both for a normal call; one at a time lets a benchmark time them apart,
// pass B then works on the planes the previous call left in the scratch).
int oz_glm_stats(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                 const double* beta, int skip_obj, const void* prep, double* stats, double* parts, int max_parts,
                 int* nparts, void* scratch, size_t scratch_bytes, int phases, int scheme, cudaStream_t st) {
  const OzPlan p = oz_plan(device, n, scheme);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= p.total, "int8 Hessian: scratch too small (%zu < %zu)",
             scratch_bytes, p.total);
  OzEncodeFn enc = oz_encode_fn();
  NB_REQUIRE(enc != nullptr, "int8 Hessian: cuTensorMapEncodeTiled is not available from this driver");
  char* sc = (char*)scratch;
  int8_t* planes = (int8_t*)(sc + p.off_planes);
  int32_t* part = (int32_t*)(sc + p.off_part);
  int* flag = (int*)(sc + p.off_flag);
  const int ew = family == NB_GLM_LINEAR ? 0 : -1;       // sqrt(w) <= 1 (linear), <= 1/2 (logistic: w = mu (1 - mu))
  *nparts = 0;
  if (phases & 1) NB_CHECK(cudaMemsetAsync(flag, 0, sizeof(int), st));
  const int nplanes = h_oz_scheme_planes[scheme], kc = h_oz_scheme_kc[scheme];
  static KernelSetup rows_setup5, rows_setup4;
  int occ = 1;
  const size_t rows_smem = (size_t)(OZ_ROWS_THREADS / 32) * OZ_ROWS_STAGES * OZ_R * OZ_COLS * sizeof(double);
  if (int rc = nplanes == 5 ? rows_setup5.get(device, k_oz_rows<5>, OZ_ROWS_THREADS, rows_smem, &occ)
                            : rows_setup4.get(device, k_oz_rows<4>, OZ_ROWS_THREADS, rows_smem, &occ)) return rc;
  const size_t smem = (size_t)OZ_SMEM_TILES + 1024;
  static bool gemm_ready[NB_MAX_DEVICES] = {};
  if (device < 0 || device >= NB_MAX_DEVICES || !gemm_ready[device]) {
    NB_CHECK(cudaFuncSetAttribute(k_oz_gemm<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NB_CHECK(cudaFuncSetAttribute(k_oz_gemm<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (device >= 0 && device < NB_MAX_DEVICES) gemm_ready[device] = true;
  }
  const int item = 8;                    // y is float64 or int64
  int ranges_done = 0;
  for (int c = 0; c < p.nchunks; c++) {
    const int64_t c0 = (int64_t)c * p.chunk_rows;
    const int64_t cn = (n - c0 < p.chunk_rows) ? n - c0 : p.chunk_rows;
    const int cranges = (int)((cn + p.range_rows - 1) / p.range_rows);
    if (phases & 1) {
      OzRowsArgs ra;
      ra.x = x + c0 * ldx; ra.n = cn; ra.d = d; ra.ldx = ldx; ra.y = (const char*)y + c0 * item; ra.ydt = ydt; ra.beta = beta;
      ra.family = family; ra.skip_obj = skip_obj; ra.ew = ew; ra.prep = (const OzPrep*)prep; ra.planes = planes;
      ra.plane_stride = p.chunk_rows * OZ_COLS; ra.flag = flag;
      int64_t want = ((cn + OZ_R - 1) / OZ_R + (OZ_ROWS_THREADS / 32) - 1) / (OZ_ROWS_THREADS / 32);
      int64_t grid = (int64_t)dev_info(device).sms * occ;
      if (grid > max_parts / p.nchunks) grid = max_parts / p.nchunks;
      if (grid > want) grid = want;
      if (grid < 1) grid = 1;
      ra.partials = parts + (size_t)(*nparts) * (2 * d + 3);
      ra.seed = (uint32_t)(c0 / 64);
      if (nplanes == 5) k_oz_rows<5><<<(unsigned)grid, OZ_ROWS_THREADS, rows_smem, st>>>(ra);
      else k_oz_rows<4><<<(unsigned)grid, OZ_ROWS_THREADS, rows_smem, st>>>(ra);
      NB_LAUNCH_CHECK();
      *nparts += (int)grid;
    }
    if (phases & 2) {
      CUtensorMap tm;
      cuuint64_t dims[3] = {(cuuint64_t)OZ_COLS, (cuuint64_t)cn, (cuuint64_t)OZ_S};
      cuuint64_t strides[2] = {(cuuint64_t)OZ_COLS, (cuuint64_t)p.chunk_rows * OZ_COLS};
      cuuint32_t box[3] = {OZ_HALF, (cuuint32_t)kc, 1};
      cuuint32_t estr[3] = {1, 1, 1};
      CUresult cr = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, planes, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      NB_REQUIRE(cr == CUDA_SUCCESS, "int8 Hessian: cuTensorMapEncodeTiled failed (%d)", (int)cr);
      OzGemmArgs ga;
      ga.n = cn; ga.range_rows = p.range_rows; ga.nranges = cranges; ga.lanes = p.lanes; ga.scheme = scheme;
      ga.debug = NB_ENV_FLAG("NUMS_B200_OZ_DEBUG") ? 1 : 0;
      if (ga.debug) fprintf(stderr, "int8 Hessian: scheme %d, %d lanes, ranges of %d rows, %d ranges\n", scheme, p.lanes, p.range_rows, cranges);
      ga.part = part + (size_t)ranges_done * OZ_GROUPS * OZ_SLOTS * 65536;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)(2 * p.pairs));
      cfg.blockDim = dim3(192);
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      if (kc == 64) NB_CHECK(cudaLaunchKernelEx(&cfg, k_oz_gemm<64>, tm, ga));
      else NB_CHECK(cudaLaunchKernelEx(&cfg, k_oz_gemm<128>, tm, ga));
    }
    ranges_done += cranges;
  }
  if (!(phases & 2)) return 0;

  long long* sums = (long long*)(sc + p.off_sums);
  k_oz_reduce<<<dev_info(device).sms * 4, 256, 0, st>>>(part, ranges_done, sums);
  NB_LAUNCH_CHECK();
  OzFinishArgs fa;
  fa.d = d; fa.ldc = d + 1; fa.sums = sums; fa.prep = (const OzPrep*)prep; fa.ew = ew;
  fa.flag = flag; fa.out = stats; fa.scheme = scheme; fa.nplanes = nplanes;
  const int64_t blocks = (d * (d + 1) / 2 + 255) / 256;
  k_oz_finish<<<(unsigned)blocks, 256, 0, st>>>(fa);
  NB_LAUNCH_CHECK();
  return 0;
}

// The gradient / objective pass (no Hessian: an L-BFGS evaluation, a line-search trial) through the row kernel's bulk-copy
// ring, without digit planes: k_oz_rows<0>.  Same partials as k_glm_rows (the caller runs k_glm_rows_finish).
bool oz_rows_applicable(int64_t n, int64_t d, int64_t ldx, const double* x) {
  return n > 0 && d > 128 && d <= 256 && (d % 4 == 0) && (ldx % 4 == 0) && ((((uintptr_t)x) & 31) == 0);
}

int oz_gradient_rows(int device, int family, const double* x, int64_t n, int64_t d, int64_t ldx, const void* y, int ydt,
                     const double* beta, int skip_obj, double* parts, int max_parts, int* nparts, cudaStream_t st) {
  static KernelSetup setup;
  int occ = 1;
  const size_t rows_smem = (size_t)(OZ_GRAD_THREADS / 32) * OZ_ROWS_STAGES * OZ_R * OZ_COLS * sizeof(double);
  if (int rc = setup.get(device, k_oz_rows<0>, OZ_GRAD_THREADS, rows_smem, &occ)) return rc;
  OzRowsArgs ra;
  ra.x = x; ra.n = n; ra.d = d; ra.ldx = ldx; ra.y = y; ra.ydt = ydt; ra.beta = beta;
  ra.family = family; ra.skip_obj = skip_obj; ra.ew = 0; ra.prep = nullptr; ra.planes = nullptr; ra.plane_stride = 0;
  ra.flag = nullptr; ra.seed = 0; ra.partials = parts;
  int64_t want = ((n + OZ_R - 1) / OZ_R + (OZ_GRAD_THREADS / 32) - 1) / (OZ_GRAD_THREADS / 32);
  int64_t grid = (int64_t)dev_info(device).sms * occ;
  if (grid > max_parts) grid = max_parts;
  if (grid > want) grid = want;
  if (grid < 1) grid = 1;
  k_oz_rows<0><<<(unsigned)grid, OZ_GRAD_THREADS, rows_smem, st>>>(ra);
  NB_LAUNCH_CHECK();
  *nparts = (int)grid;
  return 0;
}

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_glm_prep_bytes(int64_t d) {
  (void)d;
  return sizeof(OzPrep) + OZ_COLS * sizeof(unsigned long long) + (size_t)OZ_PREP_CTAS * OZ_COLS * sizeof(double);
}

int nb_glm_prepare(int device, const double* x_d, int64_t n, int64_t d, int64_t ldx, void* prep_d, size_t prep_bytes,
                   void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(d >= 1 && d <= OZ_COLS && n >= 0 && ldx >= d, "nb_glm_prepare: bad shape n=%lld d=%lld ldx=%lld",
             (long long)n, (long long)d, (long long)ldx);
  NB_REQUIRE(prep_d != nullptr && prep_bytes >= nb_glm_prep_bytes(d), "nb_glm_prepare: buffer too small");
  unsigned long long* maxbits = (unsigned long long*)((char*)prep_d + sizeof(OzPrep));
  double* sumsq = (double*)(maxbits + OZ_COLS);
  NB_CHECK(cudaMemsetAsync(maxbits, 0, OZ_COLS * sizeof(unsigned long long), st));
  int64_t grid = OZ_PREP_CTAS;
  if (grid > (n + 3) / 4) grid = (n + 3) / 4;
  if (n > 0) {
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((((uintptr_t)x_d) & 31) == 0);
    if (vec) k_oz_colmax<true><<<(unsigned)grid, 256, 0, st>>>(x_d, n, d, ldx, maxbits, sumsq);
    else k_oz_colmax<false><<<(unsigned)grid, 256, 0, st>>>(x_d, n, d, ldx, maxbits, sumsq);
    NB_LAUNCH_CHECK();
  }
  k_oz_colexp<<<1, 256, 0, st>>>(maxbits, sumsq, n > 0 ? (int)grid : 0, n, d, (OzPrep*)prep_d);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
