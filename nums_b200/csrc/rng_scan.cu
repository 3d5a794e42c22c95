// Block RNG draws whose word consumption varies per output -- Generator.normal (ziggurat) and Generator.integers
// (Lemire rejection) -- produced on the device in the reference's host order (SURVEY.md §8(f) rank 1;
// nums/core/kernel/numpy_kernel.py:31-32,103-110 calls Generator(PCG64(seed).jumped(j)).<rfunc>(*args)).
//
// A sequential sampler reads the 64-bit word stream left to right and emits an output every so often; which word an
// output comes from depends on every rejection before it.  Both samplers are finite-state transducers over the word
// stream, so the dependence is a prefix scan of state maps:
//   pass 1 (k_*_map)   each warp owns a chunk of ZL = 2048 consecutive words (reached by the O(log k) LCG jump) and
//                      computes, for every state the chunk could be entered in, the exit state and the number of
//                      outputs emitted -- one walk from the start state plus walks from the other entry states that
//                      last until they meet it (usually within the first 32 words);
//   pass 2 (k_scan_*)  the chunk maps are composed in stream order (three-level scan) -> true entry state and output
//                      offset per chunk;
//   pass 3 (k_*_emit)  each warp re-walks its chunk from the true entry state and writes the values.
// A warp walks a chunk 32 words per step, lane l taking word 32*step + l (its LCG strides by 32), so the values of a
// step go to consecutive addresses.  Every lane first evaluates its word as if entered in S; when the whole step
// fast-accepts (68 % of steps) that is the answer, otherwise the few lanes entered mid-draw are resolved in stream
// order by a warp-uniform walk over the non-S exits (resolve_walk).
// Element i of the result is therefore exactly the i-th value the host generator would return, independent of launch
// geometry.  Ziggurat states: S (fresh word), W (wedge test pending on the previous word), T1/T2 (tail loop, first /
// second uniform of a pair) with the sign bit of the word that entered the tail.
#include "pcg64.cuh"
#include "libm_exact.cuh"
#define NB_ZIG_QUAL static __device__
#include "zig_tables.h"

namespace nb {

constexpr int ZL = 2048;           // words per chunk (one warp, ZL / 32 steps)
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
    const double arg = __dmul_rn(__dmul_rn(-0.5, x), x);          // in [-6.7, 0]
    // the single-precision exponential (relative error < 1e-5 on this range) decides all but ~1e-4 of the tests;
    // the others are settled by the double-precision one, so the decision is always the double-precision decision
    const double approx = (double)__expf((float)arg);
    bool accept;
    if (lhs < approx * (1.0 - 1e-4)) accept = true;
    else if (lhs > approx * (1.0 + 1e-4)) accept = false;
    else accept = lhs < exp(arg);
    if (accept) { val = x; return true; }
    return false;
  }
  if (base == ZT1) { st = (st & ~3) | ZT2; return false; }
  // ZT2: the pair (prev, w)
  const double xx = __dmul_rn(-NB_ZIG_INV_R, nb_libm::log1p_glibc_fma(-u53(prev)));
  const double yy = -nb_libm::log1p_glibc_fma(-u53(w));
  if (__dadd_rn(yy, yy) > __dmul_rn(xx, xx)) {
    const double t = __dadd_rn(NB_ZIG_R, xx);
    val = (st & ZSIGN) ? -t : t;
    st = ZS;
    return true;
  }
  st = (st & ~3) | ZT1;
  return false;
}

// chunk map: byte b (b = entry base state 0..3) of .x = exit code (base | sign | inherit), 16-bit field b of .y =
// outputs emitted when entered in b
typedef ulonglong2 ChunkMap;

__device__ __forceinline__ ChunkMap pack_map(const int exits[4], const int cnt[4]) {
  uint64_t lo = 0, hi = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) { lo |= (uint64_t)(exits[b] & 0xff) << (8 * b); hi |= (uint64_t)(cnt[b] & 0xffff) << (16 * b); }
  return make_ulonglong2(lo, hi);
}
__device__ __forceinline__ int map_exit(const ChunkMap& mp, int base) { return (int)((mp.x >> (8 * base)) & 0xff); }
__device__ __forceinline__ long long map_count(const ChunkMap& mp, int base) { return (long long)((mp.y >> (16 * base)) & 0xffff); }

__device__ __forceinline__ void load_tables(uint64_t* ski, double* swi, double* sfi) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    ski[i] = zig_ki[i];
    swi[i] = __longlong_as_double((long long)zig_wi_bits[i]);
    sfi[i] = __longlong_as_double((long long)zig_fi_bits[i]);
  }
  __syncthreads();
}

// Given every lane's exit when entered in S (st_s, with not_s = ballot(st_s != S)) and the state `carry` lane 0 is
// entered in, re-evaluates the lanes that are entered mid-draw (emit / val are overwritten for those) and returns
// the state after lane 31.
__device__ __forceinline__ int resolve_walk(int carry, unsigned not_s, int st_s, uint64_t w, uint64_t prev, int lane,
                                            bool& emit, double& val, const ZigTables& T) {
  const unsigned full = 0xffffffffu;
  int cur = carry, l = 0;
  while (l < 32) {
    if ((cur & 3) == ZS) {
      const unsigned rest = not_s & (full << l);
      if (rest == 0u) break;
      const int l2 = __ffs(rest) - 1;
      cur = __shfl_sync(full, st_s, l2);
      l = l2 + 1;
    } else {
      int st2 = cur;
      if (lane == l) {
        double v2 = 0.0;
        emit = zig_step(st2, w, prev, v2, T);
        val = v2;
      }
      cur = __shfl_sync(full, st2, l);
      l = l + 1;
    }
  }
  return cur;
}

// per-lane LCG of a warp that walks chunk c: `s` gives lane's first word, (m32, p32) strides by 32 words;
// returns the chunk's start state (whose output is the word before the chunk)
__device__ __forceinline__ u128 lane_stream(u128 s_start, u128 inc, int lane, u128& s, u128& m32, u128& p32) {
  const u128 a = NB_PCG_MULT;
  u128 m, p;
  lcg_advance(a, inc, (uint64_t)lane + 1, m, p);
  lcg_advance(a, inc, 32, m32, p32);
  s = s_start * m + p;
  return s_start;
}

__global__ void __launch_bounds__(128)
k_zig_map(int64_t nchunks, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo,
          ChunkMap* __restrict__ maps, ulonglong2* __restrict__ lcg) {
  __shared__ uint64_t ski[256];
  __shared__ double swi[256], sfi[256];
  load_tables(ski, swi, sfi);
  const ZigTables T{ski, swi, sfi};
  const int lane = threadIdx.x & 31;
  const int64_t c = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (c >= nchunks) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  u128 m, p;
  lcg_advance(a, inc, (uint64_t)c * ZL, m, p);
  const u128 s_start = make128(s_hi, s_lo) * m + p;      // state whose output is word c*ZL - 1
  if (lane == 0) lcg[c] = make_ulonglong2((uint64_t)s_start, (uint64_t)(s_start >> 64));
  u128 s, m32, p32;
  lane_stream(s_start, inc, lane, s, m32, p32);
  uint64_t last_prev = xsl_rr(s_start);
  const unsigned full = 0xffffffffu;
  int carry = ZS, cnt = 0;
  int alt_carry[3] = {ZW, ZT1 | ZINHERIT, ZT2 | ZINHERIT};
  int alt_cnt[3] = {0, 0, 0};
  bool alt_live[3] = {true, true, true};
  bool any_live = true;
  for (int it = 0; it < ZL / 32; ++it) {
    const uint64_t w = xsl_rr(s);
    s = s * m32 + p32;
    uint64_t prev = __shfl_up_sync(full, w, 1);
    if (lane == 0) prev = last_prev;
    last_prev = __shfl_sync(full, w, 31);
    int st_s = ZS;
    double val_s = 0.0;
    const bool emit_s = zig_step(st_s, w, prev, val_s, T);
    const unsigned not_s = __ballot_sync(full, st_s != ZS);
    bool emit = emit_s;
    double val = val_s;
    if (not_s != 0u || carry != ZS) carry = resolve_walk(carry, not_s, st_s, w, prev, lane, emit, val, T);
    cnt += __popc(__ballot_sync(full, emit));
    if (any_live) {
      any_live = false;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (!alt_live[k]) continue;
        bool e2 = emit_s;
        double v2 = val_s;
        alt_carry[k] = resolve_walk(alt_carry[k], not_s, st_s, w, prev, lane, e2, v2, T);
        alt_cnt[k] += __popc(__ballot_sync(full, e2));
        if (alt_carry[k] == carry) {       // same state after this step: the rest of the chunk is the S-entry walk
          alt_live[k] = false;
          alt_cnt[k] -= cnt;               // becomes (emitted so far) - (S-entry walk emitted so far)
        } else {
          any_live = true;
        }
      }
    }
  }
  if (lane == 0) {
    int exits[4], counts[4];
    exits[ZS] = carry;
    counts[ZS] = cnt;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      exits[k + 1] = alt_live[k] ? alt_carry[k] : carry;
      counts[k + 1] = alt_live[k] ? alt_cnt[k] : alt_cnt[k] + cnt;
    }
    maps[c] = pack_map(exits, counts);
  }
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

// The composition runs on three levels so that no thread walks more than SCAN_PER chunks or SCAN_T partial maps:
//   k_scan_groups   one CTA per group of SCAN_T * SCAN_PER chunks: each thread composes SCAN_PER chunk maps, thread 0
//                   composes the SCAN_T partial maps -> one map per group;
//   k_scan_top      one thread walks the group maps in stream order -> entry state and offset of every group;
//   k_scan_apply    one CTA per group again: partial maps as before, thread 0 turns them into the entry of every
//                   thread's first chunk, every thread walks its chunks writing entry[c] and offs[c].
constexpr int SCAN_T = 256;
constexpr int SCAN_PER = 16;
constexpr int SCAN_GROUP = SCAN_T * SCAN_PER;

__device__ __forceinline__ void range_identity(RangeMap& r) {
#pragma unroll
  for (int b = 0; b < 4; ++b) { r.ex[b] = (b >= ZT1) ? (b | ZINHERIT) : b; r.cnt[b] = 0; }
}

__device__ __forceinline__ void range_push(RangeMap& r, const ChunkMap& mp) {
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const int cur = r.ex[b];
    const int base = cur & 3;
    r.cnt[b] += map_count(mp, base);
    r.ex[b] = apply_exit(cur, map_exit(mp, base));
  }
}

// r then g (both range maps)
__device__ __forceinline__ void range_then(RangeMap& r, const RangeMap& g) {
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const int cur = r.ex[b];
    const int base = cur & 3;
    r.cnt[b] += g.cnt[base];
    r.ex[b] = apply_exit(cur, g.ex[base]);
  }
}

__device__ __forceinline__ void thread_range(int64_t nchunks, const ChunkMap* __restrict__ maps, int64_t& lo,
                                             int64_t& hi, RangeMap& r) {
  lo = min(nchunks, (int64_t)blockIdx.x * SCAN_GROUP + (int64_t)threadIdx.x * SCAN_PER);
  hi = min(nchunks, lo + SCAN_PER);
  range_identity(r);
  for (int64_t c = lo; c < hi; ++c) range_push(r, maps[c]);
}

__global__ void __launch_bounds__(SCAN_T)
k_scan_groups(int64_t nchunks, const ChunkMap* __restrict__ maps, RangeMap* __restrict__ group_maps) {
  __shared__ RangeMap rm[SCAN_T];
  int64_t lo, hi;
  RangeMap r;
  thread_range(nchunks, maps, lo, hi, r);
  rm[threadIdx.x] = r;
  __syncthreads();
  if (threadIdx.x == 0) {
    RangeMap g;
    range_identity(g);
    for (int i = 0; i < SCAN_T; ++i) range_then(g, rm[i]);
    group_maps[blockIdx.x] = g;
  }
}

__global__ void __launch_bounds__(SCAN_T)
k_scan_top(int ngroups, const RangeMap* __restrict__ group_maps, int* __restrict__ group_state,
           long long* __restrict__ group_off, long long* __restrict__ header) {
  __shared__ RangeMap tile[SCAN_T];
  __shared__ int cur_s;
  __shared__ long long off_s;
  if (threadIdx.x == 0) { cur_s = ZS; off_s = 0; }
  for (int g0 = 0; g0 < ngroups; g0 += SCAN_T) {
    const int g = g0 + threadIdx.x;
    if (g < ngroups) tile[threadIdx.x] = group_maps[g];
    __syncthreads();
    if (threadIdx.x == 0) {
      int cur = cur_s;
      long long off = off_s;
      const int m = min(SCAN_T, ngroups - g0);
      for (int i = 0; i < m; ++i) {
        group_state[g0 + i] = cur;
        group_off[g0 + i] = off;
        const int base = cur & 3;
        off += tile[i].cnt[base];
        cur = apply_exit(cur, tile[i].ex[base]) & 7;
      }
      cur_s = cur;
      off_s = off;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { header[0] = off_s; header[1] = cur_s; }
}

__global__ void __launch_bounds__(SCAN_T)
k_scan_apply(int64_t nchunks, const ChunkMap* __restrict__ maps, const int* __restrict__ group_state,
             const long long* __restrict__ group_off, unsigned char* __restrict__ entry,
             long long* __restrict__ offs) {
  __shared__ RangeMap rm[SCAN_T];
  __shared__ int ent_state[SCAN_T];
  __shared__ long long ent_off[SCAN_T];
  int64_t lo, hi;
  RangeMap r;
  thread_range(nchunks, maps, lo, hi, r);
  rm[threadIdx.x] = r;
  __syncthreads();
  if (threadIdx.x == 0) {
    int cur = group_state[blockIdx.x];
    long long off = group_off[blockIdx.x];
    for (int i = 0; i < SCAN_T; ++i) {
      ent_state[i] = cur;
      ent_off[i] = off;
      const int base = cur & 3;
      off += rm[i].cnt[base];
      cur = apply_exit(cur, rm[i].ex[base]) & 7;
    }
  }
  __syncthreads();
  int cur = ent_state[threadIdx.x];
  long long off = ent_off[threadIdx.x];
  for (int64_t c = lo; c < hi; ++c) {
    const ChunkMap mp = maps[c];
    entry[c] = (unsigned char)cur;
    offs[c] = off;
    const int base = cur & 3;
    off += map_count(mp, base);
    cur = apply_exit(cur, map_exit(mp, base)) & 7;
  }
}

// Pass 3: the warp of chunk c walks it from its true entry state and writes the values at their offsets.
__global__ void __launch_bounds__(128)
k_zig_emit(int64_t nchunks, int64_t n, uint64_t i_hi, uint64_t i_lo, const ulonglong2* __restrict__ lcg,
           const unsigned char* __restrict__ entry, const long long* __restrict__ offs,
           double loc, double scale, int affine, double* __restrict__ out) {
  __shared__ uint64_t ski[256];
  __shared__ double swi[256], sfi[256];
  load_tables(ski, swi, sfi);
  const ZigTables T{ski, swi, sfi};
  const int lane = threadIdx.x & 31;
  const int64_t c = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (c >= nchunks) return;
  long long off = offs[c];
  if (off >= n) return;
  int carry = entry[c];
  const ulonglong2 ls = lcg[c];
  const u128 s_start = make128(ls.y, ls.x);
  u128 s, m32, p32;
  lane_stream(s_start, make128(i_hi, i_lo), lane, s, m32, p32);
  uint64_t last_prev = xsl_rr(s_start);           // the word before this chunk
  const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
  for (int it = 0; it < ZL / 32; ++it) {
    const uint64_t w = xsl_rr(s);
    s = s * m32 + p32;
    uint64_t prev = __shfl_up_sync(full, w, 1);
    if (lane == 0) prev = last_prev;
    last_prev = __shfl_sync(full, w, 31);
    int st_s = ZS;
    double val = 0.0;
    bool emit = zig_step(st_s, w, prev, val, T);
    const unsigned not_s = __ballot_sync(full, st_s != ZS);
    if (not_s != 0u || carry != ZS) carry = resolve_walk(carry, not_s, st_s, w, prev, lane, emit, val, T);
    const unsigned em = __ballot_sync(full, emit);
    const long long at = off + __popc(em & lt);
    if (emit && at < n) out[at] = affine ? __dadd_rn(loc, __dmul_rn(scale, val)) : val;
    off += __popc(em);
    if (off >= n) break;
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
          uint64_t threshold, ChunkMap* __restrict__ maps, ulonglong2* __restrict__ lcg) {
  const int lane = threadIdx.x & 31;
  const int64_t c = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (c >= nchunks) return;
  const u128 a = NB_PCG_MULT, inc = make128(i_hi, i_lo);
  u128 m, p;
  lcg_advance(a, inc, (uint64_t)c * ZL, m, p);
  const u128 s_start = make128(s_hi, s_lo) * m + p;
  if (lane == 0) lcg[c] = make_ulonglong2((uint64_t)s_start, (uint64_t)(s_start >> 64));
  u128 s, m32, p32;
  lane_stream(s_start, inc, lane, s, m32, p32);
  int n = 0;
#pragma unroll 4
  for (int it = 0; it < ZL / 32; ++it) {
    const uint64_t w = xsl_rr(s);
    s = s * m32 + p32;
    uint64_t v;
    if (MODE == 0) {
      n += lemire32((uint32_t)w, (uint32_t)rng_excl, (uint32_t)threshold, v) ? 1 : 0;
      n += lemire32((uint32_t)(w >> 32), (uint32_t)rng_excl, (uint32_t)threshold, v) ? 1 : 0;
    } else {
      n += lemire64(w, rng_excl, threshold, v) ? 1 : 0;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if (lane == 0) {
    int exits[4] = {ZS, ZS, ZS, ZS}, cnt[4] = {n, n, n, n};
    maps[c] = pack_map(exits, cnt);
  }
}

template <int MODE>
__global__ void __launch_bounds__(128)
k_int_emit(int64_t nchunks, int64_t n, uint64_t i_hi, uint64_t i_lo, const ulonglong2* __restrict__ lcg,
           const long long* __restrict__ offs, uint64_t rng_excl, uint64_t threshold, uint64_t low,
           long long* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t c = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (c >= nchunks) return;
  long long off = offs[c];
  if (off >= n) return;
  const ulonglong2 ls = lcg[c];
  u128 s, m32, p32;
  lane_stream(make128(ls.y, ls.x), make128(i_hi, i_lo), lane, s, m32, p32);
  const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
  for (int it = 0; it < ZL / 32; ++it) {
    const uint64_t w = xsl_rr(s);
    s = s * m32 + p32;
    if (MODE == 0) {
      uint64_t v0, v1;
      const bool a0 = lemire32((uint32_t)w, (uint32_t)rng_excl, (uint32_t)threshold, v0);
      const bool a1 = lemire32((uint32_t)(w >> 32), (uint32_t)rng_excl, (uint32_t)threshold, v1);
      const unsigned b0 = __ballot_sync(full, a0), b1 = __ballot_sync(full, a1);
      const long long at = off + __popc(b0 & lt) + __popc(b1 & lt);
      if (a0 && at < n) out[at] = (long long)(low + v0);
      if (a1 && at + (a0 ? 1 : 0) < n) out[at + (a0 ? 1 : 0)] = (long long)(low + v1);
      off += __popc(b0) + __popc(b1);
    } else {
      uint64_t v;
      const bool ok = lemire64(w, rng_excl, threshold, v);
      const unsigned b = __ballot_sync(full, ok);
      const long long at = off + __popc(b & lt);
      if (ok && at < n) out[at] = (long long)(low + v);
      off += __popc(b);
    }
    if (off >= n) break;
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
  ChunkMap* maps;
  ulonglong2* lcg;
  long long* offs;
  unsigned char* entry;
  RangeMap* group_maps;
  long long* group_off;
  int* group_state;
  int ngroups;
};

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int groups_for(int64_t nchunks) { return (int)((nchunks + SCAN_GROUP - 1) / SCAN_GROUP); }

static size_t scan_scratch_bytes(int64_t nchunks) {
  const size_t g = (size_t)groups_for(nchunks);
  return 256 + (size_t)nchunks * (16 + 16 + 8) + up256((size_t)nchunks) + up256(g * sizeof(RangeMap)) +
         up256(g * 8) + up256(g * 4);
}

static ScanScratch carve(void* scratch, int64_t nchunks) {
  ScanScratch s;
  unsigned char* p = (unsigned char*)scratch;
  s.ngroups = groups_for(nchunks);
  s.header = (long long*)p; p += 256;
  s.lcg = (ulonglong2*)p; p += (size_t)nchunks * 16;
  s.maps = (ChunkMap*)p; p += (size_t)nchunks * 16;
  s.offs = (long long*)p; p += (size_t)nchunks * 8;
  s.entry = p; p += up256((size_t)nchunks);
  s.group_maps = (RangeMap*)p; p += up256((size_t)s.ngroups * sizeof(RangeMap));
  s.group_off = (long long*)p; p += up256((size_t)s.ngroups * 8);
  s.group_state = (int*)p;
  return s;
}

// maps -> entry state and output offset of every chunk, total in header[0]
static int run_scan(const ScanScratch& sc, int64_t nchunks, cudaStream_t st) {
  k_scan_groups<<<sc.ngroups, SCAN_T, 0, st>>>(nchunks, sc.maps, sc.group_maps);
  NB_LAUNCH_CHECK();
  k_scan_top<<<1, SCAN_T, 0, st>>>(sc.ngroups, sc.group_maps, sc.group_state, sc.group_off, sc.header);
  NB_LAUNCH_CHECK();
  k_scan_apply<<<sc.ngroups, SCAN_T, 0, st>>>(nchunks, sc.maps, sc.group_state, sc.group_off, sc.entry, sc.offs);
  NB_LAUNCH_CHECK();
  return 0;
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
  const int grid = (int)((nchunks + 3) / 4);
  k_zig_map<<<grid, 128, 0, st>>>(nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1], sc.maps, sc.lcg);
  NB_LAUNCH_CHECK();
  if (int rc = run_scan(sc, nchunks, st)) return rc;
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
  const int grid = (int)((pl.nchunks + 3) / 4);
  if (pl.kind == 5) {
    k_int_map<0><<<grid, 128, 0, st>>>(pl.nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1],
                                       pl.rng_excl, pl.threshold, sc.maps, sc.lcg);
  } else {
    k_int_map<1><<<grid, 128, 0, st>>>(pl.nchunks, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1],
                                       pl.rng_excl, pl.threshold, sc.maps, sc.lcg);
  }
  NB_LAUNCH_CHECK();
  if (int rc = run_scan(sc, pl.nchunks, st)) return rc;
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
