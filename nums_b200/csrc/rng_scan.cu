// Block RNG draws whose word consumption varies per output -- Generator.normal (ziggurat) and Generator.integers
// (Lemire rejection) -- produced on the device in the reference's host order (SURVEY.md §8(f) rank 1;
// nums/core/kernel/numpy_kernel.py:31-32,103-110 calls Generator(PCG64(seed).jumped(j)).<rfunc>(*args)).
//
// A sequential sampler reads the 64-bit word stream left to right and emits an output every so often; which word an
// output comes from depends on every rejection before it.  Both samplers are finite-state transducers over the word
// stream, so the dependence is a prefix scan of state maps:
//   pass 1 (k_*_map)   each thread owns a chunk of L = 64 consecutive words (reached by the O(log k) LCG jump) and
//                      computes, for every state the chunk could be entered in, the exit state and the number of
//                      outputs emitted -- one simulation from the start state plus short ones that merge into it;
//   pass 2 (k_scan_maps) one CTA composes the chunk maps in stream order -> true entry state and output offset per chunk;
//   pass 3 (k_*_emit)  each thread re-runs its chunk from the true entry state and writes its outputs at their offset.
// Element i of the result is therefore exactly the i-th value the host generator would return, independent of launch
// geometry.  Ziggurat states: S (fresh word), W (wedge test pending on the previous word), T1/T2 (tail loop, first /
// second uniform of a pair) with the sign bit of the word that entered the tail.
#include "pcg64.cuh"
#define NB_ZIG_QUAL static __device__
#include "zig_tables.h"

namespace nb {

constexpr int ZL = 64;             // words per chunk
constexpr int ZS = 0, ZW = 1, ZT1 = 2, ZT2 = 3;
constexpr int ZSIGN = 4;           // tail sign bit (set = negative)
constexpr int ZINHERIT = 8;        // tail sign comes from the entry state (tail entered before this chunk)

struct ZigTables {
  const uint64_t* ki;
  const double* wi;
  const double* fi;
};

__device__ __forceinline__ double u53(uint64_t w) { return (double)(w >> 11) * (1.0 / 9007199254740992.0); }

// One step of NumPy's random_standard_normal as a transducer.  `st` is updated; returns true when a value is emitted.
__device__ __forceinline__ bool zig_step(int& st, uint64_t w, uint64_t prev, double& val, const ZigTables& T) {
  const int base = st & 3;
  if (base == ZS) {
    const int idx = (int)(w & 0xff);
    const uint64_t r = w >> 8;
    const uint64_t rabs = (r >> 1) & 0x000fffffffffffffull;
    double x = (double)rabs * T.wi[idx];
    if (r & 1) x = -x;
    if (rabs < T.ki[idx]) { val = x; return true; }
    if (idx == 0) st = ZT1 | (((rabs >> 8) & 1) ? ZSIGN : 0);
    else st = ZW;
    return false;
  }
  if (base == ZW) {
    const int idx = (int)(prev & 0xff);
    const uint64_t r = prev >> 8;
    const uint64_t rabs = (r >> 1) & 0x000fffffffffffffull;
    double x = (double)rabs * T.wi[idx];
    if (r & 1) x = -x;
    st = ZS;
    const double lhs = __dadd_rn(__dmul_rn(T.fi[idx - 1] - T.fi[idx], u53(w)), T.fi[idx]);
    const double rhs = exp(__dmul_rn(__dmul_rn(-0.5, x), x));
    if (lhs < rhs) { val = x; return true; }
    return false;
  }
  if (base == ZT1) { st = (st & ~3) | ZT2; return false; }
  // ZT2: the pair (prev, w)
  const double xx = __dmul_rn(-NB_ZIG_INV_R, log1p(-u53(prev)));
  const double yy = -log1p(-u53(w));
  if (__dadd_rn(yy, yy) > __dmul_rn(xx, xx)) {
    const double t = __dadd_rn(NB_ZIG_R, xx);
    val = (st & ZSIGN) ? -t : t;
    st = ZS;
    return true;
  }
  st = (st & ~3) | ZT1;
  return false;
}

// chunk map: byte b (b = entry base state 0..3) of the low word = exit code (base | sign | inherit), byte b of the
// high word = outputs emitted when entered in b
__device__ __forceinline__ uint64_t pack_map(const int exits[4], const int cnt[4]) {
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) { lo |= (uint32_t)(exits[b] & 0xff) << (8 * b); hi |= (uint32_t)(cnt[b] & 0xff) << (8 * b); }
  return ((uint64_t)hi << 32) | lo;
}

__device__ __forceinline__ void load_tables(uint64_t* ski, double* swi, double* sfi) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    ski[i] = zig_ki[i];
    swi[i] = __longlong_as_double((long long)zig_wi_bits[i]);
    sfi[i] = __longlong_as_double((long long)zig_fi_bits[i]);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(128)
k_zig_map(int64_t nchunks, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo,
          uint64_t* __restrict__ maps, ulonglong2* __restrict__ lcg) {
  __shared__ uint64_t ski[256];
  __shared__ double swi[256], sfi[256];
  load_tables(ski, swi, sfi);
  const ZigTables T{ski, swi, sfi};
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  u128 m, p;
  lcg_advance(a, inc, (uint64_t)c * ZL, m, p);
  const u128 s_start = make128(s_hi, s_lo) * m + p;      // state whose output is word c*L - 1
  lcg[c] = make_ulonglong2((uint64_t)s_start, (uint64_t)(s_start >> 64));
  const uint64_t prev0 = xsl_rr(s_start);
  // the path entered in S
  uint64_t emit_mask = 0, s_mask = 0;
  int st = ZS;
  {
    u128 s = s_start;
    uint64_t prev = prev0;
#pragma unroll 4
    for (int j = 0; j < ZL; ++j) {
      if (st == ZS) s_mask |= 1ull << j;
      s = s * a + inc;
      const uint64_t w = xsl_rr(s);
      double v;
      if (zig_step(st, w, prev, v, T)) emit_mask |= 1ull << j;
      prev = w;
    }
  }
  int exits[4], cnt[4];
  exits[ZS] = st;
  cnt[ZS] = __popcll(emit_mask);
  // the paths entered mid-draw: simulate until they meet the S path in state S
  for (int b = ZW; b <= ZT2; ++b) {
    int sb = (b == ZW) ? ZW : (b | ZINHERIT);
    int n = 0;
    bool merged = false;
    u128 s = s_start;
    uint64_t prev = prev0;
    for (int j = 0; j < ZL; ++j) {
      if (sb == ZS && ((s_mask >> j) & 1)) {
        n += __popcll(emit_mask >> j);
        merged = true;
        break;
      }
      s = s * a + inc;
      const uint64_t w = xsl_rr(s);
      double v;
      if (zig_step(sb, w, prev, v, T)) ++n;
      prev = w;
    }
    exits[b] = merged ? st : sb;
    cnt[b] = n;
  }
  maps[c] = pack_map(exits, cnt);
}

// Composes the chunk maps in stream order (one CTA).  entry[c] = state code chunk c is entered in (base | sign),
// offs[c] = outputs emitted before chunk c, header[0] = outputs of the whole scan.
struct RangeMap {
  int ex[4];
  long long cnt[4];
};

__device__ __forceinline__ int apply_exit(int cur, int exit_code) {
  // exit_code may inherit the tail sign of the state the chunk was entered in
  if (exit_code & ZINHERIT) return (exit_code & 3) | (cur & (ZSIGN | ZINHERIT));
  return exit_code & 7;
}

constexpr int SCAN_T = 512;

__global__ void __launch_bounds__(SCAN_T)
k_scan_maps(int64_t nchunks, const uint64_t* __restrict__ maps, unsigned char* __restrict__ entry,
            long long* __restrict__ offs, long long* __restrict__ header) {
  __shared__ RangeMap rm[SCAN_T];
  __shared__ int ent_state[SCAN_T];
  __shared__ long long ent_off[SCAN_T];
  const int t = threadIdx.x;
  const int64_t q = (nchunks + SCAN_T - 1) / SCAN_T;
  const int64_t lo = min(nchunks, (int64_t)t * q), hi = min(nchunks, lo + q);
  RangeMap r;
#pragma unroll
  for (int b = 0; b < 4; ++b) { r.ex[b] = (b >= ZT1) ? (b | ZINHERIT) : b; r.cnt[b] = 0; }
  for (int64_t c = lo; c < hi; ++c) {
    const uint64_t mp = maps[c];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int cur = r.ex[b];
      const int base = cur & 3;
      r.cnt[b] += (long long)((mp >> (32 + 8 * base)) & 0xff);
      r.ex[b] = apply_exit(cur, (int)((mp >> (8 * base)) & 0xff));
    }
  }
  rm[t] = r;
  __syncthreads();
  if (t == 0) {
    int cur = ZS;
    long long off = 0;
    for (int i = 0; i < SCAN_T; ++i) {
      ent_state[i] = cur;
      ent_off[i] = off;
      const int base = cur & 3;
      off += rm[i].cnt[base];
      cur = apply_exit(cur, rm[i].ex[base]) & 7;
    }
    header[0] = off;
    header[1] = cur;
  }
  __syncthreads();
  int cur = ent_state[t];
  long long off = ent_off[t];
  for (int64_t c = lo; c < hi; ++c) {
    const uint64_t mp = maps[c];
    entry[c] = (unsigned char)cur;
    offs[c] = off;
    const int base = cur & 3;
    off += (long long)((mp >> (32 + 8 * base)) & 0xff);
    cur = apply_exit(cur, (int)((mp >> (8 * base)) & 0xff)) & 7;
  }
}

__global__ void __launch_bounds__(128)
k_zig_emit(int64_t nchunks, int64_t n, uint64_t i_hi, uint64_t i_lo, const ulonglong2* __restrict__ lcg,
           const unsigned char* __restrict__ entry, const long long* __restrict__ offs,
           double loc, double scale, int affine, double* __restrict__ out) {
  __shared__ uint64_t ski[256];
  __shared__ double swi[256], sfi[256];
  load_tables(ski, swi, sfi);
  const ZigTables T{ski, swi, sfi};
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  long long off = offs[c];
  if (off >= n) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  const ulonglong2 ls = lcg[c];
  u128 s = make128(ls.y, ls.x);
  uint64_t prev = xsl_rr(s);
  int st = entry[c];
  for (int j = 0; j < ZL; ++j) {
    s = s * a + inc;
    const uint64_t w = xsl_rr(s);
    double v;
    if (zig_step(st, w, prev, v, T)) {
      if (off < n) out[off] = affine ? __dadd_rn(loc, __dmul_rn(scale, v)) : v;
      ++off;
    }
    prev = w;
  }
}

// ---- Generator.integers (int64): Lemire's multiply-shift with rejection, 32-bit samples when the range fits ----
// mode 0: 32-bit samples (low half of a word first, then the high half -- PCG64's buffered next_uint32)
// mode 1: 64-bit samples
__device__ __forceinline__ bool lemire32(uint32_t x, uint32_t rng_excl, uint32_t threshold, uint64_t& v) {
  const uint64_t m = (uint64_t)x * rng_excl;
  v = m >> 32;
  return (uint32_t)m >= threshold;
}
__device__ __forceinline__ bool lemire64(uint64_t x, uint64_t rng_excl, uint64_t threshold, uint64_t& v) {
  v = __umul64hi(x, rng_excl);
  return x * rng_excl >= threshold;
}

template <int MODE>
__global__ void __launch_bounds__(128)
k_int_map(int64_t nchunks, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo, uint64_t rng_excl,
          uint64_t threshold, uint64_t* __restrict__ maps, ulonglong2* __restrict__ lcg) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  u128 m, p;
  lcg_advance(a, inc, (uint64_t)c * ZL, m, p);
  u128 s = make128(s_hi, s_lo) * m + p;
  lcg[c] = make_ulonglong2((uint64_t)s, (uint64_t)(s >> 64));
  int n = 0;
#pragma unroll 4
  for (int j = 0; j < ZL; ++j) {
    s = s * a + inc;
    const uint64_t w = xsl_rr(s);
    uint64_t v;
    if (MODE == 0) {
      n += lemire32((uint32_t)w, (uint32_t)rng_excl, (uint32_t)threshold, v) ? 1 : 0;
      n += lemire32((uint32_t)(w >> 32), (uint32_t)rng_excl, (uint32_t)threshold, v) ? 1 : 0;
    } else {
      n += lemire64(w, rng_excl, threshold, v) ? 1 : 0;
    }
  }
  int exits[4] = {ZS, ZS, ZS, ZS}, cnt[4] = {n, n, n, n};
  maps[c] = pack_map(exits, cnt);
}

template <int MODE>
__global__ void __launch_bounds__(128)
k_int_emit(int64_t nchunks, int64_t n, uint64_t i_hi, uint64_t i_lo, const ulonglong2* __restrict__ lcg,
           const long long* __restrict__ offs, uint64_t rng_excl, uint64_t threshold, uint64_t low,
           long long* __restrict__ out) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  long long off = offs[c];
  if (off >= n) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  const ulonglong2 ls = lcg[c];
  u128 s = make128(ls.y, ls.x);
  for (int j = 0; j < ZL; ++j) {
    s = s * a + inc;
    const uint64_t w = xsl_rr(s);
    uint64_t v;
    if (MODE == 0) {
      if (lemire32((uint32_t)w, (uint32_t)rng_excl, (uint32_t)threshold, v)) {
        if (off < n) out[off] = (long long)(low + v);
        ++off;
      }
      if (lemire32((uint32_t)(w >> 32), (uint32_t)rng_excl, (uint32_t)threshold, v)) {
        if (off < n) out[off] = (long long)(low + v);
        ++off;
      }
    } else {
      if (lemire64(w, rng_excl, threshold, v)) {
        if (off < n) out[off] = (long long)(low + v);
        ++off;
      }
    }
  }
}

// rejection-free cases: element i comes from a fixed word
// kind 0: constant low; 1: low + next_uint32 (halves of word i/2); 2: low + next_uint64; 3: Lemire32 without rejection
// (threshold 0); 4: Lemire64 without rejection
__global__ void __launch_bounds__(256)
k_int_direct(int64_t n, int kind, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo, uint64_t rng_excl,
             uint64_t low, long long* __restrict__ out) {
  const uint64_t T = (uint64_t)gridDim.x * 256;
  const uint64_t t = (uint64_t)blockIdx.x * 256 + threadIdx.x;
  if (kind == 0) {
    for (uint64_t i = t; i < (uint64_t)n; i += T) out[i] = (long long)low;
    return;
  }
  const bool half = (kind == 1 || kind == 3);
  const uint64_t nwords = half ? ((uint64_t)n + 1) / 2 : (uint64_t)n;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  u128 m, p;
  lcg_advance(a, inc, t + 1, m, p);
  u128 s = make128(s_hi, s_lo) * m + p;
  u128 mt, pt;
  lcg_advance(a, inc, T, mt, pt);
  for (uint64_t k = t; k < nwords; k += T) {
    const uint64_t w = xsl_rr(s);
    if (kind == 2) out[k] = (long long)(low + w);
    else if (kind == 4) out[k] = (long long)(low + __umul64hi(w, rng_excl));
    else {
      const uint32_t x0 = (uint32_t)w, x1 = (uint32_t)(w >> 32);
      const uint64_t v0 = (kind == 1) ? x0 : (((uint64_t)x0 * (uint32_t)rng_excl) >> 32);
      const uint64_t v1 = (kind == 1) ? x1 : (((uint64_t)x1 * (uint32_t)rng_excl) >> 32);
      out[2 * k] = (long long)(low + v0);
      if (2 * k + 1 < (uint64_t)n) out[2 * k + 1] = (long long)(low + v1);
    }
    s = s * mt + pt;
  }
}

struct ScanScratch {
  long long* header;       // [0] outputs produced, [1] final state
  uint64_t* maps;
  ulonglong2* lcg;
  long long* offs;
  unsigned char* entry;
};

static size_t scan_scratch_bytes(int64_t nchunks) {
  return 256 + (size_t)nchunks * (8 + 16 + 8) + (((size_t)nchunks + 255) & ~(size_t)255);
}

static ScanScratch carve(void* scratch, int64_t nchunks) {
  ScanScratch s;
  unsigned char* p = (unsigned char*)scratch;
  s.header = (long long*)p; p += 256;
  s.lcg = (ulonglong2*)p; p += (size_t)nchunks * 16;
  s.maps = (uint64_t*)p; p += (size_t)nchunks * 8;
  s.offs = (long long*)p; p += (size_t)nchunks * 8;
  s.entry = p;
  return s;
}

// words to scan so that at least n normals come out: the ziggurat takes the first word of a draw 98.8 % of the time
static int64_t normal_words(int64_t n) { return n + n / 32 + 4096; }

// 32/64-bit Lemire: rejection probability is threshold / 2^bits; returns the samples to scan for n accepted ones
static int64_t lemire_samples(int64_t n, double p_rej) {
  const double need = (double)n / (1.0 - p_rej);
  const double sd = sqrt((double)n * p_rej) / (1.0 - p_rej);
  return (int64_t)(need + 12.0 * sd + 256.0);
}

static int64_t chunks_for(int64_t words) { return (words + ZL - 1) / ZL; }

struct IntPlan {
  int kind;                // 0..4 direct kinds, 5 = Lemire32 with rejection, 6 = Lemire64 with rejection
  uint64_t rng_excl, threshold;
  int64_t nchunks;
};

static IntPlan plan_integers(int64_t n, uint64_t rng) {
  IntPlan pl{};
  pl.rng_excl = rng + 1;
  if (rng == 0) { pl.kind = 0; return pl; }
  if (rng <= 0xFFFFFFFFull) {
    if (rng == 0xFFFFFFFFull) { pl.kind = 1; return pl; }
    const uint32_t re = (uint32_t)rng + 1u;
    const uint32_t th = (uint32_t)((0xFFFFFFFFu - (uint32_t)rng) % re);
    pl.threshold = th;
    if (th == 0) { pl.kind = 3; return pl; }
    pl.kind = 5;
    const int64_t samples = lemire_samples(n, (double)th / 4294967296.0);
    pl.nchunks = chunks_for((samples + 1) / 2);
    return pl;
  }
  if (rng == 0xFFFFFFFFFFFFFFFFull) { pl.kind = 2; return pl; }
  const uint64_t th = (0xFFFFFFFFFFFFFFFFull - rng) % pl.rng_excl;
  pl.threshold = th;
  if (th == 0) { pl.kind = 4; return pl; }
  pl.kind = 6;
  pl.nchunks = chunks_for(lemire_samples(n, (double)th / 18446744073709551616.0));
  return pl;
}

}  // namespace nb

using namespace nb;

extern "C" int64_t nb_pcg64_normal_scratch_bytes(int64_t n) {
  if (n <= 0) return 0;
  return (int64_t)scan_scratch_bytes(chunks_for(normal_words(n)));
}

extern "C" int nb_pcg64_normal(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                               double loc, double scale, int standard, double* out_d, void* scratch_d,
                               int64_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  if (n <= 0) return 0;
  const int64_t nchunks = chunks_for(normal_words(n));
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= (int64_t)scan_scratch_bytes(nchunks),
             "nb_pcg64_normal: scratch smaller than nb_pcg64_normal_scratch_bytes(n)");
  ScanScratch sc = carve(scratch_d, nchunks);
  const int grid = (int)((nchunks + 127) / 128);
  k_zig_map<<<grid, 128, 0, st>>>(nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1], sc.maps, sc.lcg);
  NB_LAUNCH_CHECK();
  k_scan_maps<<<1, SCAN_T, 0, st>>>(nchunks, sc.maps, sc.entry, sc.offs, sc.header);
  NB_LAUNCH_CHECK();
  k_zig_emit<<<grid, 128, 0, st>>>(nchunks, n, inc_hi_lo[0], inc_hi_lo[1], sc.lcg, sc.entry, sc.offs, loc, scale,
                                   standard ? 0 : 1, out_d);
  NB_LAUNCH_CHECK();
  long long produced = 0;
  NB_CHECK(cudaMemcpyAsync(&produced, sc.header, sizeof(long long), cudaMemcpyDeviceToHost, st));
  NB_CHECK(cudaStreamSynchronize(st));
  NB_REQUIRE(produced >= n, "nb_pcg64_normal: the scanned words produced fewer than n values (retry with more slack)");
  return 0;
}

extern "C" int64_t nb_pcg64_integers_scratch_bytes(int64_t n, uint64_t rng) {
  if (n <= 0) return 0;
  const IntPlan pl = plan_integers(n, rng);
  return pl.kind >= 5 ? (int64_t)scan_scratch_bytes(pl.nchunks) : 0;
}

extern "C" int nb_pcg64_integers(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                                 int64_t low, uint64_t rng, int64_t* out_d, void* scratch_d, int64_t scratch_bytes,
                                 void* stream) {
  NB_ENTER(device);
  if (n <= 0) return 0;
  const IntPlan pl = plan_integers(n, rng);
  if (pl.kind < 5) {
    const int64_t work = (pl.kind == 1 || pl.kind == 3) ? (n + 1) / 2 : n;
    const int grid = grid_for(device, work, 256, 4);
    k_int_direct<<<grid, 256, 0, st>>>(n, pl.kind, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1],
                                       pl.rng_excl, (uint64_t)low, (long long*)out_d);
    NB_LAUNCH_CHECK();
    return 0;
  }
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= (int64_t)scan_scratch_bytes(pl.nchunks),
             "nb_pcg64_integers: scratch smaller than nb_pcg64_integers_scratch_bytes(n, rng)");
  ScanScratch sc = carve(scratch_d, pl.nchunks);
  const int grid = (int)((pl.nchunks + 127) / 128);
  if (pl.kind == 5) {
    k_int_map<0><<<grid, 128, 0, st>>>(pl.nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1],
                                       pl.rng_excl, pl.threshold, sc.maps, sc.lcg);
  } else {
    k_int_map<1><<<grid, 128, 0, st>>>(pl.nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1],
                                       pl.rng_excl, pl.threshold, sc.maps, sc.lcg);
  }
  NB_LAUNCH_CHECK();
  k_scan_maps<<<1, SCAN_T, 0, st>>>(pl.nchunks, sc.maps, sc.entry, sc.offs, sc.header);
  NB_LAUNCH_CHECK();
  if (pl.kind == 5) {
    k_int_emit<0><<<grid, 128, 0, st>>>(pl.nchunks, n, inc_hi_lo[0], inc_hi_lo[1], sc.lcg, sc.offs, pl.rng_excl,
                                        pl.threshold, (uint64_t)low, (long long*)out_d);
  } else {
    k_int_emit<1><<<grid, 128, 0, st>>>(pl.nchunks, n, inc_hi_lo[0], inc_hi_lo[1], sc.lcg, sc.offs, pl.rng_excl,
                                        pl.threshold, (uint64_t)low, (long long*)out_d);
  }
  NB_LAUNCH_CHECK();
  long long produced = 0;
  NB_CHECK(cudaMemcpyAsync(&produced, sc.header, sizeof(long long), cudaMemcpyDeviceToHost, st));
  NB_CHECK(cudaStreamSynchronize(st));
  NB_REQUIRE(produced >= n, "nb_pcg64_integers: the scanned words produced fewer than n values");
  return 0;
}
