// Generator.permutation(n) on the device, in the host generator's order (SURVEY.md §8(f) rank 1;
// nums/core/kernel/numpy_kernel.py:112-114 -> Generator(PCG64(seed).jumped(j)).permutation(size)).
//
// NumPy shuffles arange(n) with Fisher-Yates from the top: for i = n-1 .. 1, j = random_interval(i), swap(a[i], a[j]),
// and random_interval draws 32-bit half-words (low half of a PCG64 word first) masked to the smallest 2^b - 1 >= i
// until the value is <= i.  Both halves of that are sequential on the host and are restated here as parallel passes.
//
// (A) Which sample does draw i use?  Within a phase (all i that share a mask) a sample with value v is accepted iff
//     v <= i_hi - t, t = accepts so far in the phase.  Samples with v <= i_lo are accepted whatever t is, samples with
//     v > i_hi never; for the conditional ones in between, with a = unconditional accepts before them (a prefix sum),
//     the rule reads c <= theta := i_hi - v - a, c = conditional accepts before.  c is bracketed, lo <= c <= hi, starting
//     from 0 <= c <= (rank among the conditional samples); every sample whose theta lies outside [lo, hi) is decided,
//     prefix sums of the decided-accept / undecided flags give new brackets, and the brackets collapse in a handful of
//     rounds (measured: ~4 per phase).  Exact, no speculation: a bracket is always a proof.
// (B) What does the sequence of swaps do?  Position i is final after step i.  With W(i) = content of position i just
//     before step i: W(i) = W(p) for p = the smallest step > i that targeted position i (the last one to write it), or i
//     if none; the value step i takes from its target j is W(q) for q = the smallest step > i with the same target, or j
//     if none (j == i: W(i)).  Per-target step lists are built with atomic exchanges (expected length ln(n / p)), the
//     W chains are resolved by pointer jumping (depth O(log n)).
#include "pcg64.cuh"

namespace nb {

constexpr int PS_T = 256;
constexpr int PS_PER = 8;
constexpr int PS_BLOCK = PS_T * PS_PER;        // elements per scan block

// ---- exclusive prefix sum of 0/1 bytes -> int32 -------------------------------------------------------------------
__device__ __forceinline__ int block_excl_scan(int v, int* warp_sums, int& block_total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < PS_T / 32 ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < PS_T / 32) warp_sums[lane] = w;              // inclusive over warps
  }
  __syncthreads();
  block_total = warp_sums[PS_T / 32 - 1];
  const int before = warp == 0 ? 0 : warp_sums[warp - 1];
  __syncthreads();
  return before + inc - v;
}

__global__ void __launch_bounds__(PS_T)
k_scan_block_sums(const unsigned char* __restrict__ f, int64_t n, int* __restrict__ bsum) {
  __shared__ int ws[PS_T / 32];
  const int64_t base = (int64_t)blockIdx.x * PS_BLOCK + (int64_t)threadIdx.x * PS_PER;
  int s = 0;
#pragma unroll
  for (int u = 0; u < PS_PER; ++u) s += (base + u < n) ? f[base + u] : 0;
  int total;
  block_excl_scan(s, ws, total);
  if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024)
k_scan_offsets(int nblocks, const int* __restrict__ bsum, int* __restrict__ boff, int* __restrict__ total) {
  __shared__ int part[1024];
  const int t = threadIdx.x;
  const int q = (nblocks + 1023) / 1024;
  const int lo = min(nblocks, t * q), hi = min(nblocks, lo + q);
  int s = 0;
  for (int b = lo; b < hi; ++b) s += bsum[b];
  part[t] = s;
  __syncthreads();
  if (t == 0) {
    int run = 0;
    for (int i = 0; i < 1024; ++i) { const int v = part[i]; part[i] = run; run += v; }
    *total = run;
  }
  __syncthreads();
  int run = part[t];
  for (int b = lo; b < hi; ++b) { boff[b] = run; run += bsum[b]; }
}

__global__ void __launch_bounds__(PS_T)
k_scan_apply_u8(const unsigned char* __restrict__ f, int64_t n, const int* __restrict__ boff, int* __restrict__ out) {
  __shared__ int ws[PS_T / 32];
  const int64_t base = (int64_t)blockIdx.x * PS_BLOCK + (int64_t)threadIdx.x * PS_PER;
  int v[PS_PER], s = 0;
#pragma unroll
  for (int u = 0; u < PS_PER; ++u) { v[u] = (base + u < n) ? f[base + u] : 0; s += v[u]; }
  int total;
  int run = block_excl_scan(s, ws, total) + boff[blockIdx.x];
#pragma unroll
  for (int u = 0; u < PS_PER; ++u) {
    if (base + u < n) out[base + u] = run;
    run += v[u];
  }
}

struct ScanTmp { int* bsum; int* boff; int* total; };

// out[i] = number of set flags before i; *tmp.total = number of set flags
static int scan_flags(const unsigned char* f, int64_t n, int* out, const ScanTmp& tmp, cudaStream_t st) {
  const int nb = (int)((n + PS_BLOCK - 1) / PS_BLOCK);
  k_scan_block_sums<<<nb, PS_T, 0, st>>>(f, n, tmp.bsum);
  NB_LAUNCH_CHECK();
  k_scan_offsets<<<1, 1024, 0, st>>>(nb, tmp.bsum, tmp.boff, tmp.total);
  NB_LAUNCH_CHECK();
  k_scan_apply_u8<<<nb, PS_T, 0, st>>>(f, n, tmp.boff, out);
  NB_LAUNCH_CHECK();
  return 0;
}

// ---- (A) samples of a phase --------------------------------------------------------------------------------------
// sample s (global index) = half (s & 1) of word s / 2; word w is the output of LCG step w + 1
__global__ void __launch_bounds__(256)
k_perm_classify(int64_t s0, int64_t wn, uint64_t s_hi, uint64_t s_lo, uint64_t i_hi64, uint64_t i_lo64,
                unsigned mask, unsigned v_lo, unsigned v_hi, unsigned* __restrict__ v,
                unsigned char* __restrict__ fa, unsigned char* __restrict__ fc) {
  const u128 a = NB_PCG_MULT, inc = make128(i_hi64, i_lo64);
  const int64_t w_first = s0 >> 1, w_last = (s0 + wn - 1) >> 1;
  const uint64_t T = (uint64_t)gridDim.x * 256;
  const uint64_t t = (uint64_t)blockIdx.x * 256 + threadIdx.x;
  if ((int64_t)(w_first + t) > w_last) return;
  u128 m, p, mt, pt;
  lcg_advance(a, inc, (uint64_t)w_first + t + 1, m, p);
  lcg_advance(a, inc, T, mt, pt);
  u128 s = make128(s_hi, s_lo) * m + p;
  for (int64_t w = w_first + (int64_t)t; w <= w_last; w += (int64_t)T) {
    const uint64_t word = xsl_rr(s);
    s = s * mt + pt;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int64_t idx = 2 * w + h - s0;
      if (idx < 0 || idx >= wn) continue;
      const unsigned val = (unsigned)(h ? (word >> 32) : word) & mask;
      v[idx] = val;
      fa[idx] = val <= v_lo ? 1 : 0;                       // accepted whatever the running index is
      fc[idx] = (val > v_lo && val <= v_hi) ? 1 : 0;       // depends on the running index
    }
  }
}

// conditional samples, compacted in stream order: theta = i_hi - v - (unconditional accepts before)
__global__ void __launch_bounds__(256)
k_perm_compact(int64_t wn, const unsigned* __restrict__ v, const unsigned char* __restrict__ fc,
               const int* __restrict__ pa, const int* __restrict__ pc, long long i_hi,
               long long* __restrict__ theta) {
  const int64_t s = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (s < wn && fc[s]) theta[pc[s]] = i_hi - (long long)v[s] - (long long)pa[s];
}

// one bracket round: sa / su = prefix sums of last round's decided-accept / undecided flags (first round: 0 / rank)
__global__ void __launch_bounds__(256)
k_perm_round(int nc, const long long* __restrict__ theta, const int* __restrict__ sa, const int* __restrict__ su,
             int first, unsigned char* __restrict__ facc, unsigned char* __restrict__ func, int* __restrict__ undecided) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  int unc = 0;
  if (j < nc) {
    const long long lo = first ? 0 : sa[j];
    const long long hi = first ? j : (long long)sa[j] + su[j];
    const long long th = theta[j];
    const bool acc = hi <= th, rej = lo > th;
    facc[j] = acc ? 1 : 0;
    unc = (!acc && !rej) ? 1 : 0;
    func[j] = (unsigned char)unc;
  }
  const unsigned any = __ballot_sync(0xffffffffu, unc);
  if ((threadIdx.x & 31) == 0 && any) atomicAdd(undecided, __popc(any));
}

// accepted = unconditional | (conditional & decided accept)
__global__ void __launch_bounds__(256)
k_perm_accepted(int64_t wn, const unsigned char* __restrict__ fa, const unsigned char* __restrict__ fc,
                const int* __restrict__ pc, const unsigned char* __restrict__ facc, unsigned char* __restrict__ acc) {
  const int64_t s = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (s < wn) acc[s] = (fa[s] || (fc[s] && facc[pc[s]])) ? 1 : 0;
}

// the first `count` accepted samples are the phase's draws: jst[i] = v for i = i_hi - rank; s_end = last one used
__global__ void __launch_bounds__(256)
k_perm_take(int64_t wn, const unsigned* __restrict__ v, const unsigned char* __restrict__ acc, const int* __restrict__ pt,
            long long i_hi, int count, int* __restrict__ jst, long long* __restrict__ s_end) {
  const int64_t s = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (s < wn && acc[s]) {
    const int r = pt[s];
    if (r < count) {
      jst[i_hi - r] = (int)v[s];
      if (r == count - 1) *s_end = s;
    }
  }
}

// ---- (B) the swaps ----------------------------------------------------------------------------------------------
// per-target lists of the steps that move something (j != i): next[i] = previous head of list j
__global__ void __launch_bounds__(256)
k_perm_lists(int n, const int* __restrict__ jst, int* __restrict__ head, int* __restrict__ next) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= 1 && i < n) {
    const int j = jst[i];
    if (j != i) next[i] = atomicExch(&head[j], i);
  }
}

// root[i] = smallest step that targeted position i (the last to write it before step i), or i itself;
// q[i] = smallest step > i with the same target as step i, or -1
__global__ void __launch_bounds__(256)
k_perm_links(int n, const int* __restrict__ jst, const int* __restrict__ head, const int* __restrict__ next,
             int* __restrict__ root, int* __restrict__ q) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  int best = 0x7fffffff;
  for (int x = head[i]; x >= 0; x = next[x]) best = min(best, x);
  root[i] = best == 0x7fffffff ? i : best;
  int qq = -1;
  if (i >= 1) {
    const int j = jst[i];
    if (j != i) {
      int b2 = 0x7fffffff;
      for (int x = head[j]; x >= 0; x = next[x]) if (x > i) b2 = min(b2, x);
      qq = b2 == 0x7fffffff ? -1 : b2;
    }
  }
  q[i] = qq;
}

// pointer jumping towards the end of every "who wrote this position" chain (chains climb to larger steps)
__global__ void __launch_bounds__(256)
k_perm_jump(int n, int* __restrict__ root, int* __restrict__ changed) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  int ch = 0;
  if (i < n) {
    const int r = root[i];
    const int rr = root[r];
    if (rr != r) { root[i] = rr; ch = 1; }
  }
  const unsigned any = __ballot_sync(0xffffffffu, ch);
  if ((threadIdx.x & 31) == 0 && any) atomicOr(changed, 1);
}

__global__ void __launch_bounds__(256)
k_perm_final(int n, const int* __restrict__ jst, const int* __restrict__ head_root0, const int* __restrict__ root,
             const int* __restrict__ q, const int* __restrict__ first_step, long long* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  if (i == 0) {
    const int p = first_step[0];                 // smallest step that targeted position 0 (or 0)
    out[0] = p == 0 ? 0 : root[p];
    return;
  }
  const int j = jst[i];
  if (j == i) out[i] = root[i];
  else out[i] = q[i] >= 0 ? root[q[i]] : j;
}

static inline size_t pm_up(size_t v) { return (v + 255) & ~(size_t)255; }

// samples to scan for `count` draws of a phase: every sample is accepted with probability >= 1/2
static int64_t phase_window(int64_t count) { return 2 * count + (int64_t)(10.0 * sqrt(2.0 * (double)count)) + 64; }

struct PermLayout {
  size_t jst, head, next, root, q, first, v, fa, fc, acc, pa, pc, pt, theta, sa, su, facc, func, bsum, boff, small, total;
};

static PermLayout perm_layout(int64_t n) {
  PermLayout L{};
  const int64_t top_count = n;                       // upper bound of a phase's draws
  const int64_t wn = phase_window(top_count);
  const int64_t nb = (wn + PS_BLOCK - 1) / PS_BLOCK + 1;
  size_t cur = 0;
  auto take = [&](size_t bytes) { size_t o = cur; cur += pm_up(bytes); return o; };
  L.small = take(256);
  L.jst = take((size_t)n * 4);
  // stage B
  L.head = take((size_t)n * 4);
  L.next = take((size_t)n * 4);
  L.root = take((size_t)n * 4);
  L.q = take((size_t)n * 4);
  L.first = take(256);
  // stage A (its buffers are dead before stage B starts, but the sum is small against the arrays it feeds)
  L.v = take((size_t)wn * 4);
  L.fa = take((size_t)wn);
  L.fc = take((size_t)wn);
  L.acc = take((size_t)wn);
  L.pa = take((size_t)wn * 4);
  L.pc = take((size_t)wn * 4);
  L.pt = take((size_t)wn * 4);
  L.theta = take((size_t)wn * 8);
  L.sa = take((size_t)wn * 4);
  L.su = take((size_t)wn * 4);
  L.facc = take((size_t)wn);
  L.func = take((size_t)wn);
  L.bsum = take((size_t)nb * 4);
  L.boff = take((size_t)nb * 4);
  L.total = cur;
  return L;
}

}  // namespace nb

using namespace nb;

extern "C" int64_t nb_pcg64_permutation_scratch_bytes(int64_t n) {
  if (n <= 1) return 256;
  return (int64_t)perm_layout(n).total;
}

extern "C" int nb_pcg64_permutation(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                                    int64_t* out_d, void* scratch_d, int64_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  if (n <= 0) return 0;
  NB_REQUIRE(n <= 0x7fffffffll, "nb_pcg64_permutation: n=%lld exceeds the 32-bit sample path", (long long)n);
  if (n == 1) { NB_CHECK(cudaMemsetAsync(out_d, 0, sizeof(int64_t), st)); return 0; }
  const PermLayout L = perm_layout(n);
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= (int64_t)L.total, "nb_pcg64_permutation: scratch too small");
  char* sp = (char*)scratch_d;
  int* jst = (int*)(sp + L.jst);
  int* head = (int*)(sp + L.head);
  int* next = (int*)(sp + L.next);
  int* root = (int*)(sp + L.root);
  int* q = (int*)(sp + L.q);
  unsigned* v = (unsigned*)(sp + L.v);
  unsigned char* fa = (unsigned char*)(sp + L.fa);
  unsigned char* fc = (unsigned char*)(sp + L.fc);
  unsigned char* acc = (unsigned char*)(sp + L.acc);
  int* pa = (int*)(sp + L.pa);
  int* pc = (int*)(sp + L.pc);
  int* pt = (int*)(sp + L.pt);
  long long* theta = (long long*)(sp + L.theta);
  int* sa = (int*)(sp + L.sa);
  int* su = (int*)(sp + L.su);
  unsigned char* facc = (unsigned char*)(sp + L.facc);
  unsigned char* func = (unsigned char*)(sp + L.func);
  // small: [0] scan total (int), [8] undecided (int), [16] s_end (long long), [24] changed (int)
  int* d_total = (int*)(sp + L.small);
  int* d_und = (int*)(sp + L.small + 8);
  long long* d_send = (long long*)(sp + L.small + 16);
  int* d_changed = (int*)(sp + L.small + 24);
  ScanTmp tmp{(int*)(sp + L.bsum), (int*)(sp + L.boff), d_total};

  // ---- (A) one phase per mask, from the top
  int64_t s0 = 0;
  int64_t i_hi = n - 1;
  while (i_hi >= 1) {
    int b = 63 - __builtin_clzll((unsigned long long)i_hi);           // 2^b <= i_hi < 2^(b+1)
    const int64_t i_lo = (int64_t)1 << b;
    const unsigned mask = (unsigned)(((uint64_t)1 << (b + 1)) - 1);
    const int64_t count = i_hi - i_lo + 1;
    int64_t wn = phase_window(count);
    for (;;) {
      const int64_t words = ((s0 + wn - 1) >> 1) - (s0 >> 1) + 1;
      const int cgrid = grid_for(device, words, 256, 4);
      k_perm_classify<<<cgrid, 256, 0, st>>>(s0, wn, state_hi_lo[0], state_hi_lo[1], inc_hi_lo[0], inc_hi_lo[1], mask,
                                             (unsigned)i_lo, (unsigned)i_hi, v, fa, fc);
      NB_LAUNCH_CHECK();
      if (int rc = scan_flags(fa, wn, pa, tmp, st)) return rc;
      if (int rc = scan_flags(fc, wn, pc, tmp, st)) return rc;
      int nc = 0;
      NB_CHECK(cudaMemcpyAsync(&nc, d_total, sizeof(int), cudaMemcpyDeviceToHost, st));
      NB_CHECK(cudaStreamSynchronize(st));
      const int sgrid = (int)((wn + 255) / 256);
      if (nc > 0) {
        k_perm_compact<<<sgrid, 256, 0, st>>>(wn, v, fc, pa, pc, (long long)i_hi, theta);
        NB_LAUNCH_CHECK();
        const int jgrid = (nc + 255) / 256;
        int first = 1;
        for (int round = 0; round < 4096; ++round) {
          NB_CHECK(cudaMemsetAsync(d_und, 0, sizeof(int), st));
          k_perm_round<<<jgrid, 256, 0, st>>>(nc, theta, sa, su, first, facc, func, d_und);
          NB_LAUNCH_CHECK();
          int und = 0;
          NB_CHECK(cudaMemcpyAsync(&und, d_und, sizeof(int), cudaMemcpyDeviceToHost, st));
          NB_CHECK(cudaStreamSynchronize(st));
          if (und == 0) break;
          NB_REQUIRE(round < 4095, "nb_pcg64_permutation: bracket rounds did not converge");
          if (int rc = scan_flags(facc, nc, sa, tmp, st)) return rc;
          if (int rc = scan_flags(func, nc, su, tmp, st)) return rc;
          first = 0;
        }
      }
      k_perm_accepted<<<sgrid, 256, 0, st>>>(wn, fa, fc, pc, facc, acc);
      NB_LAUNCH_CHECK();
      if (int rc = scan_flags(acc, wn, pt, tmp, st)) return rc;
      int got = 0;
      NB_CHECK(cudaMemcpyAsync(&got, d_total, sizeof(int), cudaMemcpyDeviceToHost, st));
      NB_CHECK(cudaStreamSynchronize(st));
      if (got >= count) break;
      wn *= 2;                                                          // (practically never) too few accepts: widen
      NB_REQUIRE(wn <= phase_window(n), "nb_pcg64_permutation: sample window exceeded the scratch plan");
    }
    const int sgrid = (int)((wn + 255) / 256);
    k_perm_take<<<sgrid, 256, 0, st>>>(wn, v, acc, pt, (long long)i_hi, (int)count, jst, d_send);
    NB_LAUNCH_CHECK();
    long long s_end = 0;
    NB_CHECK(cudaMemcpyAsync(&s_end, d_send, sizeof(long long), cudaMemcpyDeviceToHost, st));
    NB_CHECK(cudaStreamSynchronize(st));
    s0 += s_end + 1;
    i_hi = i_lo - 1;
  }

  // ---- (B) the composition of the swaps
  const int ni = (int)n;
  const int ngrid = (ni + 255) / 256;
  NB_CHECK(cudaMemsetAsync(head, 0xff, (size_t)n * 4, st));
  k_perm_lists<<<ngrid, 256, 0, st>>>(ni, jst, head, next);
  NB_LAUNCH_CHECK();
  k_perm_links<<<ngrid, 256, 0, st>>>(ni, jst, head, next, root, q);
  NB_LAUNCH_CHECK();
  // root[0] before jumping = smallest step that targeted position 0 (or 0): keep it, position 0 has no step of its own
  int* first_step = (int*)(sp + L.first);
  NB_CHECK(cudaMemcpyAsync(first_step, root, sizeof(int), cudaMemcpyDeviceToDevice, st));
  for (int round = 0; round < 64; ++round) {
    NB_CHECK(cudaMemsetAsync(d_changed, 0, sizeof(int), st));
    k_perm_jump<<<ngrid, 256, 0, st>>>(ni, root, d_changed);
    NB_LAUNCH_CHECK();
    int ch = 0;
    NB_CHECK(cudaMemcpyAsync(&ch, d_changed, sizeof(int), cudaMemcpyDeviceToHost, st));
    NB_CHECK(cudaStreamSynchronize(st));
    if (!ch) break;
  }
  k_perm_final<<<ngrid, 256, 0, st>>>(ni, jst, head, root, q, first_step, (long long*)out_d);
  NB_LAUNCH_CHECK();
  return 0;
}
