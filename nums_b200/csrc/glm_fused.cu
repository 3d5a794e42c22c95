// The Hessian X^T diag(w) X for 8 <= d <= 256 (nums/models/glms.py:491-499 as called from newton() :660-676; also
// the plain symmetric Gram X^T X) on the FP64 tensor pipe, warp-specialised.
//
// A group of C CTAs walks one contiguous range of rows, 16 rows per pipeline stage.  Rows are staged whole
// (all d columns, padded stride = 4 mod 16 doubles -> conflict-free DMMA fragment loads) so that every
// 32x32 block of the UPPER TRIANGLE of X^T diag(w) X can be formed from one copy of the data: only
// (d/32)(d/32+1)/2 blocks exist (36 at d = 256, one per consumer warp, 12 warps x 3 CTAs), 56 % of the tensor
// work of a full d x d product.
//
// Roles inside a CTA (no block-wide barrier in steady state, only mbarriers per stage slot):
//   producer warp   lane r issues the TMA bulk copy (cp.async.bulk, SASS UBLKCP) of row r of the stage and
//                   arrives on the slot's "fullx" barrier with its own byte count;
//   consumer warps  wait fullx, accumulate their block with DMMA.8x8x4 (A fragments scaled by w), arrive on "empty".
// The weights come from the first pass (k_glm_rows / k_glm_regs); the producer lanes store them next to the rows.
// (A single-pass variant that evaluated the link in extra warps of this kernel was measured at 48 ms against 33 ms
// for two passes -- scalar FP64 and DMMA share one pipe -- and was removed; see DESIGN.md.)
// Partials go to scratch per group and are summed in a fixed order by k_fused_finish; the lower triangle is
// the mirror image of the upper one, so H is exactly symmetric.
#include "glm_common.cuh"
#include <stdlib.h>

namespace nb {

constexpr int FK = 16;          // rows per stage
constexpr int FB = 32;          // block edge
constexpr int FPAD = 4;

__device__ __forceinline__ void f_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
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
  const double* wext;      // precomputed weights (null = all ones)
  int nblk;                // ceil(d / 32)
  int T;                   // nblk (nblk + 1) / 2 upper-triangular blocks
  int C;                   // CTAs per group
  int KG;                  // warps sharing one block (k-split inside the CTA) when there are few blocks
  int groups;
  int64_t rows_per_group;  // multiple of FK
  int64_t ldc;             // row stride of the output matrix (d + 1 inside the packed stats buffer)
  double* part_h;          // [groups * KG][T][32*32]
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

template <int W, int S>
__global__ void __launch_bounds__((W + 1) * 32)
k_tri_gram(const FusedArgs a) {
  constexpr int NT = (W + 1) * 32;
  constexpr int PRODUCER = W;
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
  uint64_t* fullx = reinterpret_cast<uint64_t*>(sWgt + S * FK);
  uint64_t* empty = fullx + S;

  for (int i = tid; i < S * FK * ld; i += NT) sX[i] = 0.0;     // stale rows must be finite (their weight is 0)
  if (tid == 0) {
    for (int s = 0; s < S; s++) {
      mbar_init(&fullx[s], FK);
      mbar_init(&empty[s], W);
    }
  }
  fence_async_proxy();
  __syncthreads();

  if (warp == PRODUCER) {
    // ---------------------------------------------------------------- producer
    const unsigned row_bytes = (unsigned)(d * sizeof(double));
    for (int kt = 0; kt < ktiles; kt++) {
      const int slot = kt % S;
      if (kt >= S) mbar_wait(&empty[slot], (unsigned)(((kt / S) - 1) & 1));
      if (lane < FK) {
        const int64_t gr = row_beg + (int64_t)kt * FK + lane;
        if (gr < row_end) {
          sWgt[slot * FK + lane] = a.wext ? a.wext[gr] : 1.0;
          mbar_expect_tx(&fullx[slot], row_bytes);             // this lane's arrival + its bytes
          bulk_g2s(sX + (size_t)slot * FK * ld + lane * ld, a.x + gr * a.ldx, row_bytes, &fullx[slot]);
        } else {
          sWgt[slot * FK + lane] = 0.0;
          mbar_arrive(&fullx[slot]);
        }
      }
    }
    return;
  }

  // ---------------------------------------------------------------- consumers
  const int bpc = W / a.KG;                        // blocks per CTA
  const int kg = warp / bpc;                       // which k4-steps of a stage this warp takes
  // Blocks are dealt to the C CTAs of a group from the list [diagonal blocks, then the others row by row], round robin,
  // so that the diagonal blocks -- which skip their lower 8 x 8 tiles, 10 of 16 tile products -- are spread evenly.
  const int u = (warp % bpc) * a.C + c;
  const bool has_block = u < a.T;
  int bi = 0, bj = 0;
  if (has_block) {
    if (u < a.nblk) {
      bi = bj = u;
    } else {
      int rem = u - a.nblk;
      while (rem >= a.nblk - 1 - bi) { rem -= a.nblk - 1 - bi; bi++; }
      bj = bi + 1 + rem;
    }
  }
  const int t = bi * a.nblk - (bi * (bi - 1)) / 2 + (bj - bi);      // index in the row-by-row upper-triangular list
  const bool diag = bi == bj;
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

  for (int kt = 0; kt < ktiles; kt++) {
    const int slot = kt % S;
    // every consumer waits (also a warp without a block): it keeps a fast warp from arriving twice on
    // the same "empty" phase
    mbar_wait(&fullx[slot], (unsigned)((kt / S) & 1));
    if (has_block && a.KG == 1) {
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
          for (int j = 0; j < 4; j++)
            if (j >= i || !diag) f_dmma(acc[i][j][0], acc[i][j][1], af[cur][i], bf[cur][j]);
      }
    } else if (has_block) {
      // few blocks (small d): KG warps share a block, each taking every KG-th group of four rows
      const double* xs = sX + (size_t)slot * FK * ld;
      const double* ws = sWgt + slot * FK;
      const double* pa = xs + bi * FB + lq;
      const double* pb = xs + bj * FB + lq;
      for (int ks = kg; ks < FK / 4; ks += a.KG) {
        const int rr = ks * 4 + lr;
        const double wv = ws[rr];
        double af[4], bf[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { af[i] = pa[rr * ld + i * 8] * wv; bf[i] = pb[rr * ld + i * 8]; }
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (j >= i || !diag) f_dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[slot]);     // this warp no longer reads the slot
  }
  if (has_block) {
    double* p = a.part_h + (((size_t)group * a.KG + kg) * a.T + t) * (FB * FB);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++)
        *reinterpret_cast<double2*>(p + (i * 8 + lq) * FB + j * 8 + 2 * lr) = make_double2(acc[i][j][0], acc[i][j][1]);
  }
}

// Sums the group partials in group order into the upper triangle of the d x d corner of `out` and mirrors it.
__global__ void __launch_bounds__(256)
k_fused_finish(const FusedArgs a, double* __restrict__ out) {
  const int64_t d = a.d, p = a.ldc;
  const int64_t nh = d * (d + 1) / 2;                // upper triangle of the d x d part
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < nh; e += (int64_t)gridDim.x * 256) {
    // e -> (i, j), i <= j, row by row (a closed form would need a sqrt; d <= 256 so a short search is fine)
    int64_t i = 0, rem = e;
    while (rem >= d - i) { rem -= d - i; i++; }
    int64_t j = i + rem;
    int bi = (int)(i / FB), bj = (int)(j / FB);
    int t = bi * a.nblk - (bi * (bi - 1)) / 2 + (bj - bi);
    const double* src = a.part_h + (size_t)t * (FB * FB) + (i - bi * FB) * FB + (j - bj * FB);
    const double s = sum_partials(src, (size_t)a.T * (FB * FB), a.groups * a.KG);
    out[i * p + j] = s;
    out[j * p + i] = s;
  }
}

struct FusedPlan { int W, S, nblk, T, C, KG, groups; int64_t rows_per_group; size_t smem, scratch; };

static FusedPlan plan_fused(int device, int64_t n, int64_t d) {
  FusedPlan f;
  f.nblk = (int)((d + FB - 1) / FB);
  f.T = f.nblk * (f.nblk + 1) / 2;
  // consumer warps W, of which KG share one block: T = 1 -> 4 warps x 1 block; 3 -> 12 warps, 4 per block;
  // 6 -> 12 warps, 2 per block; >= 10 -> 12 warps, one block each over C CTAs
  f.W = f.T == 1 ? 4 : 12;
  f.KG = f.T == 1 ? 4 : (f.T == 3 ? 4 : (f.T == 6 ? 2 : 1));
  f.C = (f.T + (f.W / f.KG) - 1) / (f.W / f.KG);
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
  const size_t fixed = 2 * 8 * 8 + 64;
  f.S = 5;
  while (f.S > 4 && (size_t)f.S * (stage + FK * 8) + fixed > (size_t)220 * 1024) f.S--;
  f.smem = (size_t)f.S * (stage + FK * 8) + fixed;
  f.scratch = ((size_t)f.groups * f.KG * f.T * FB * FB * sizeof(double) + 255) & ~(size_t)255;
  return f;
}

bool fused_applicable(int64_t n, int64_t d, int64_t ldx, const double* x) {
  // bulk copies need 16-byte aligned rows whose length is a multiple of 16 bytes
  return d >= 8 && d <= 256 && n > 0 && (d % 2 == 0) && (ldx % 2 == 0) && ((((uintptr_t)x) & 15) == 0);
}

size_t fused_scratch_bytes(int device, int64_t n, int64_t d) {
  // upper bound over n (the group count only shrinks for tiny n)
  const FusedPlan f = plan_fused(device, n > 0 ? n : 1, d);
  const FusedPlan g = plan_fused(device, (int64_t)1 << 40, d);
  return f.scratch > g.scratch ? f.scratch : g.scratch;
}

template <int W, int S>
static int launch_tri(int device, const FusedArgs& a, const FusedPlan& f, cudaStream_t st) {
  auto kern = k_tri_gram<W, S>;
  static KernelSetup setup;
  int occ = 1;
  if (int rc = setup.get(device, kern, (W + 1) * 32, (size_t)220 * 1024, &occ)) return rc;
  kern<<<(unsigned)(f.groups * f.C), (W + 1) * 32, f.smem, st>>>(a);
  NB_LAUNCH_CHECK();
  return 0;
}

template <int S>
static int launch_tri_w(int device, const FusedArgs& a, const FusedPlan& f, cudaStream_t st) {
  return f.W == 4 ? launch_tri<4, S>(device, a, f, st) : launch_tri<12, S>(device, a, f, st);
}

// Upper-triangular X^T diag(w) X (w = wext, null = ones), mirrored, into the d x d corner of `out` (row stride ldc).
int tri_gram(int device, const double* x, int64_t n, int64_t d, int64_t ldx, const double* wext, double* out,
             int64_t ldc, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  FusedPlan f = plan_fused(device, n, d);
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= f.scratch, "tri_gram: scratch too small (%zu < %zu)",
             scratch_bytes, f.scratch);
  FusedArgs a;
  a.x = x; a.n = n; a.d = d; a.ldx = ldx; a.wext = wext;
  a.nblk = f.nblk; a.T = f.T; a.C = f.C; a.KG = f.KG; a.groups = f.groups; a.rows_per_group = f.rows_per_group;
  a.ldc = ldc;
  a.part_h = (double*)scratch;
  int rc = f.S == 5 ? launch_tri_w<5>(device, a, f, st) : launch_tri_w<4>(device, a, f, st);
  if (rc) return rc;
  const int64_t blocks = (d * (d + 1) / 2 + 255) / 256;
  k_fused_finish<<<(unsigned)(blocks < 592 ? blocks : 592), 256, 0, st>>>(a, out);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // namespace nb
