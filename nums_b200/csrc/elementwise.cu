// Elementwise block kernels: KernelCls.bop (ufunc branch), map_uop, where (3-arg), astype,
// new_block fills, strided block copies.  Reference: nums/core/kernel/numpy_kernel.py:93-155,
// 316-347, 414-431.  All of these are HBM-bound streaming kernels: grid-stride loops sized in
// multiples of the SM count, 4 independent 8-byte loads in flight per thread and operand.
#include "common.cuh"
#include <type_traits>

namespace nb {

// ------------------------------------------------------------------------------------------------
// NumPy scalar semantics
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t i_add(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }
__device__ __forceinline__ int64_t i_sub(int64_t a, int64_t b) { return (int64_t)((uint64_t)a - (uint64_t)b); }
__device__ __forceinline__ int64_t i_mul(int64_t a, int64_t b) { return (int64_t)((uint64_t)a * (uint64_t)b); }

__device__ __forceinline__ int64_t i_floordiv(int64_t a, int64_t b) {
  if (b == 0) return 0;
  if (a == INT64_MIN && b == -1) return INT64_MIN;
  int64_t q = a / b;
  if (((a > 0) != (b > 0)) && q * b != a) q--;
  return q;
}
__device__ __forceinline__ int64_t i_mod(int64_t a, int64_t b) {
  if (b == 0) return 0;
  if (b == -1) return 0;
  int64_t r = a % b;
  if (r != 0 && ((r < 0) != (b < 0))) r += b;
  return r;
}
__device__ __forceinline__ int64_t i_fmod(int64_t a, int64_t b) {
  if (b == 0) return 0;
  if (b == -1) return 0;
  return a % b;
}
__device__ __forceinline__ int64_t i_pow(int64_t a, int64_t b) {
  if (b < 0) return 0;  // the host raises ValueError before launching, as NumPy does
  uint64_t base = (uint64_t)a, r = 1;
  uint64_t e = (uint64_t)b;
  while (e) {
    if (e & 1) r *= base;
    base *= base;
    e >>= 1;
  }
  return (int64_t)r;
}
__device__ __forceinline__ int64_t i_gcd(int64_t a, int64_t b) {
  uint64_t x = a < 0 ? (uint64_t)0 - (uint64_t)a : (uint64_t)a;
  uint64_t y = b < 0 ? (uint64_t)0 - (uint64_t)b : (uint64_t)b;
  while (y) { uint64_t t = x % y; x = y; y = t; }
  return (int64_t)x;
}
__device__ __forceinline__ int64_t i_lcm(int64_t a, int64_t b) {
  int64_t g = i_gcd(a, b);
  if (g == 0) return 0;
  int64_t r = i_mul(a / g, b);
  return r < 0 ? i_sub(0, r) : r;
}
__device__ __forceinline__ double f_mod(double a, double b) {
  double m = fmod(a, b);
  if (b == 0.0) return m;
  if (m != 0.0) {
    if ((b < 0) != (m < 0)) m += b;
  } else {
    m = copysign(0.0, b);
  }
  return m;
}
__device__ __forceinline__ double f_floordiv(double a, double b) {
  if (b == 0.0) return a / b;
  double m = fmod(a, b);
  double div = (a - m) / b;
  if (m != 0.0 && ((b < 0) != (m < 0))) div -= 1.0;
  double fd;
  if (div != 0.0) {
    fd = floor(div);
    if (div - fd > 0.5) fd += 1.0;
  } else {
    fd = copysign(0.0, a / b);
  }
  return fd;
}
__device__ __forceinline__ double f_logaddexp(double a, double b) {
  if (a == b) return a + 0.693147180559945309417232121458176568;
  double t = a - b;
  if (t > 0) return a + log1p(exp(-t));
  if (t <= 0) return b + log1p(exp(t));
  return t;  // NaN
}
__device__ __forceinline__ double f_logaddexp2(double a, double b) {
  const double LOG2E = 1.442695040888963407359924681001892137;
  if (a == b) return a + 1.0;
  double t = a - b;
  if (t > 0) return a + LOG2E * log1p(exp2(-t));
  if (t <= 0) return b + LOG2E * log1p(exp2(t));
  return t;
}
__device__ __forceinline__ double f_heaviside(double x, double h) {
  if (isnan(x)) return x;
  if (x == 0.0) return h;
  return x < 0.0 ? 0.0 : 1.0;
}

// Result type of a binary op in compute type C.
template <int OP, typename C>
struct BopResult { using type = C; };
#define NB_BOOL_RESULT(OPC) \
  template <typename C> struct BopResult<OPC, C> { using type = bool; };
NB_BOOL_RESULT(NB_LT) NB_BOOL_RESULT(NB_LE) NB_BOOL_RESULT(NB_GT) NB_BOOL_RESULT(NB_GE)
NB_BOOL_RESULT(NB_EQ) NB_BOOL_RESULT(NB_NE)
NB_BOOL_RESULT(NB_LOGICAL_AND) NB_BOOL_RESULT(NB_LOGICAL_OR) NB_BOOL_RESULT(NB_LOGICAL_XOR)
#undef NB_BOOL_RESULT

template <int OP>
__device__ __forceinline__ typename BopResult<OP, double>::type bop_apply(double a, double b) {
  if constexpr (OP == NB_ADD) return a + b;
  else if constexpr (OP == NB_SUB) return a - b;
  else if constexpr (OP == NB_MUL) return a * b;
  else if constexpr (OP == NB_TRUEDIV) return a / b;
  else if constexpr (OP == NB_FLOORDIV) return f_floordiv(a, b);
  else if constexpr (OP == NB_MOD) return f_mod(a, b);
  else if constexpr (OP == NB_FMOD) return fmod(a, b);
  else if constexpr (OP == NB_POW || OP == NB_FLOAT_POW) return pow(a, b);
  else if constexpr (OP == NB_MAXIMUM) return (a >= b || isnan(a)) ? a : b;
  else if constexpr (OP == NB_MINIMUM) return (a <= b || isnan(a)) ? a : b;
  else if constexpr (OP == NB_FMAX) return fmax(a, b);
  else if constexpr (OP == NB_FMIN) return fmin(a, b);
  else if constexpr (OP == NB_LT) return a < b;
  else if constexpr (OP == NB_LE) return a <= b;
  else if constexpr (OP == NB_GT) return a > b;
  else if constexpr (OP == NB_GE) return a >= b;
  else if constexpr (OP == NB_EQ) return a == b;
  else if constexpr (OP == NB_NE) return a != b;
  else if constexpr (OP == NB_LOGICAL_AND) return (a != 0.0) && (b != 0.0);
  else if constexpr (OP == NB_LOGICAL_OR) return (a != 0.0) || (b != 0.0);
  else if constexpr (OP == NB_LOGICAL_XOR) return (a != 0.0) != (b != 0.0);
  else if constexpr (OP == NB_ARCTAN2) return atan2(a, b);
  else if constexpr (OP == NB_HYPOT) return hypot(a, b);
  else if constexpr (OP == NB_COPYSIGN) return copysign(a, b);
  else if constexpr (OP == NB_NEXTAFTER) return nextafter(a, b);
  else if constexpr (OP == NB_LDEXP) {
    double e = b > 2147483647.0 ? 2147483647.0 : (b < -2147483648.0 ? -2147483648.0 : b);
    return ldexp(a, (int)e);
  }
  else if constexpr (OP == NB_HEAVISIDE) return f_heaviside(a, b);
  else if constexpr (OP == NB_LOGADDEXP) return f_logaddexp(a, b);
  else if constexpr (OP == NB_LOGADDEXP2) return f_logaddexp2(a, b);
  else if constexpr (OP == NB_XLOGY) return (a == 0.0 && !isnan(b)) ? 0.0 : a * log(b);
  else return a;
}

template <int OP>
__device__ __forceinline__ typename BopResult<OP, int64_t>::type bop_apply(int64_t a, int64_t b) {
  if constexpr (OP == NB_ADD) return i_add(a, b);
  else if constexpr (OP == NB_SUB) return i_sub(a, b);
  else if constexpr (OP == NB_MUL) return i_mul(a, b);
  else if constexpr (OP == NB_FLOORDIV) return i_floordiv(a, b);
  else if constexpr (OP == NB_MOD) return i_mod(a, b);
  else if constexpr (OP == NB_FMOD) return i_fmod(a, b);
  else if constexpr (OP == NB_POW) return i_pow(a, b);
  else if constexpr (OP == NB_MAXIMUM || OP == NB_FMAX) return a > b ? a : b;
  else if constexpr (OP == NB_MINIMUM || OP == NB_FMIN) return a < b ? a : b;
  else if constexpr (OP == NB_LT) return a < b;
  else if constexpr (OP == NB_LE) return a <= b;
  else if constexpr (OP == NB_GT) return a > b;
  else if constexpr (OP == NB_GE) return a >= b;
  else if constexpr (OP == NB_EQ) return a == b;
  else if constexpr (OP == NB_NE) return a != b;
  else if constexpr (OP == NB_LOGICAL_AND) return (a != 0) && (b != 0);
  else if constexpr (OP == NB_LOGICAL_OR) return (a != 0) || (b != 0);
  else if constexpr (OP == NB_LOGICAL_XOR) return (a != 0) != (b != 0);
  else if constexpr (OP == NB_BIT_AND) return a & b;
  else if constexpr (OP == NB_BIT_OR) return a | b;
  else if constexpr (OP == NB_BIT_XOR) return a ^ b;
  else if constexpr (OP == NB_LSHIFT) return ((uint64_t)b < 64u) ? (int64_t)((uint64_t)a << b) : (int64_t)0;
  else if constexpr (OP == NB_RSHIFT) return ((uint64_t)b < 64u) ? (a >> b) : (a < 0 ? (int64_t)-1 : (int64_t)0);
  else if constexpr (OP == NB_GCD) return i_gcd(a, b);
  else if constexpr (OP == NB_LCM) return i_lcm(a, b);
  else return a;
}

template <int OP>
__device__ __forceinline__ bool bop_apply(bool a, bool b) {
  if constexpr (OP == NB_ADD || OP == NB_MAXIMUM || OP == NB_FMAX || OP == NB_LOGICAL_OR || OP == NB_BIT_OR) return a || b;
  else if constexpr (OP == NB_MUL || OP == NB_MINIMUM || OP == NB_FMIN || OP == NB_LOGICAL_AND || OP == NB_BIT_AND) return a && b;
  else if constexpr (OP == NB_LOGICAL_XOR || OP == NB_BIT_XOR || OP == NB_NE) return a != b;
  else if constexpr (OP == NB_EQ) return a == b;
  else if constexpr (OP == NB_LT) return !a && b;
  else if constexpr (OP == NB_LE) return !a || b;
  else if constexpr (OP == NB_GT) return a && !b;
  else if constexpr (OP == NB_GE) return a || !b;
  else return a;
}

// ------------------------------------------------------------------------------------------------
// binary kernels
// ------------------------------------------------------------------------------------------------
constexpr int EW_THREADS = 256;
constexpr int EW_UNROLL = 4;

// Flat: each operand is either contiguous with n elements or a single broadcast element.
template <int OP, typename C>
__global__ void __launch_bounds__(EW_THREADS)
k_bop_flat(int64_t n, const void* __restrict__ a, int adt, int a_scalar,
           const void* __restrict__ b, int bdt, int b_scalar, void* __restrict__ out, int odt, int pack) {
  using R = typename BopResult<OP, C>::type;
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
  C as = C(), bs = C();
  if (a_scalar) as = load_as<C>(a, adt, 0);
  if (b_scalar) bs = load_as<C>(b, bdt, 0);
  // Comparisons of 8-byte operands (x == c on 10^8 int64): a warp takes 256 consecutive elements with coalesced
  // 8-byte loads, turns the eight result ballots into bytes and every lane stores 8 bools at once (one-byte stores
  // held this at 74 % of the HBM rate).
  if constexpr (std::is_same<R, bool>::value && sizeof(C) == 8) {
    if (pack) {
      const C* pa = (const C*)a;
      const C* pb = (const C*)b;
      const int lane = threadIdx.x & 31;
      const int64_t warps = (int64_t)gridDim.x * (EW_THREADS / 32);
      const int64_t full_groups = n / 256;
      for (int64_t gi = (int64_t)blockIdx.x * (EW_THREADS / 32) + (threadIdx.x >> 5); gi < full_groups; gi += warps) {
        const int64_t e0 = gi * 256;
        C av[8], bv[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          av[u] = a_scalar ? as : pa[e0 + 32 * u + lane];
          bv[u] = b_scalar ? bs : pb[e0 + 32 * u + lane];
        }
        unsigned mine = 0;
#pragma unroll
        for (int u = 0; u < 8; u++) {
          const unsigned m = __ballot_sync(0xffffffffu, bop_apply<OP>(av[u], bv[u]));
          if ((lane >> 2) == u) mine = m;
        }
        const unsigned bits = (mine >> (8 * (lane & 3))) & 0xffu;       // results of elements 8 lane .. 8 lane + 7
        unsigned long long packed = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) packed |= (unsigned long long)((bits >> k) & 1u) << (8 * k);
        *reinterpret_cast<unsigned long long*>((unsigned char*)out + e0 + 8 * lane) = packed;
      }
      // tail (< 256 elements) by the first CTA
      if (blockIdx.x == 0) {
        for (int64_t i = full_groups * 256 + threadIdx.x; i < n; i += EW_THREADS)
          store_as<R>(out, odt, i, bop_apply<OP>(a_scalar ? as : pa[i], b_scalar ? bs : pb[i]));
      }
      return;
    }
  }
  for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n; base += stride) {
    C av[EW_UNROLL], bv[EW_UNROLL];
#pragma unroll
    for (int u = 0; u < EW_UNROLL; u++) {
      int64_t i = base + (int64_t)u * EW_THREADS;
      av[u] = as; bv[u] = bs;
      if (i < n) {
        if (!a_scalar) av[u] = load_as<C>(a, adt, i);
        if (!b_scalar) bv[u] = load_as<C>(b, bdt, i);
      }
    }
#pragma unroll
    for (int u = 0; u < EW_UNROLL; u++) {
      int64_t i = base + (int64_t)u * EW_THREADS;
      if (i < n) store_as<R>(out, odt, i, bop_apply<OP>(av[u], bv[u]));
    }
  }
}

struct Strides4 { int64_t shape[4]; int64_t sa[4]; int64_t sb[4]; };

// General broadcasting through element strides (ndim padded to 4 with leading 1s).
template <int OP, typename C>
__global__ void __launch_bounds__(EW_THREADS)
k_bop_strided(int64_t n, Strides4 s, const void* __restrict__ a, int adt,
              const void* __restrict__ b, int bdt, void* __restrict__ out, int odt) {
  using R = typename BopResult<OP, C>::type;
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
    int64_t r = i;
    int64_t i3 = r % s.shape[3]; r /= s.shape[3];
    int64_t i2 = r % s.shape[2]; r /= s.shape[2];
    int64_t i1 = r % s.shape[1]; r /= s.shape[1];
    int64_t i0 = r;
    int64_t ia = i0 * s.sa[0] + i1 * s.sa[1] + i2 * s.sa[2] + i3 * s.sa[3];
    int64_t ib = i0 * s.sb[0] + i1 * s.sb[1] + i2 * s.sb[2] + i3 * s.sb[3];
    store_as<R>(out, odt, i, bop_apply<OP>(load_as<C>(a, adt, ia), load_as<C>(b, bdt, ib)));
  }
}

// Row-broadcast fast path for the GLM-shaped (n,1) (op) (n,d) and (n,d) (op) (d,) cases:
// 2-D output [rows, cols]; operand strides (rs, cs) each 0 or contiguous.
template <int OP, typename C>
__global__ void __launch_bounds__(EW_THREADS)
k_bop_2d(int64_t rows, int64_t cols, const void* __restrict__ a, int adt, int64_t ars, int64_t acs,
         const void* __restrict__ b, int bdt, int64_t brs, int64_t bcs, void* __restrict__ out, int odt, int vec) {
  using R = typename BopResult<OP, C>::type;
  // float64 (op) float64 -> float64 with an even column count and 16-byte aligned rows: two adjacent columns per
  // thread (16-byte loads / stores on the contiguous operands, one division per pair), four pairs in flight.
  // This is the s * X product of glms.py:495-496 in the unfused path.
  if constexpr (std::is_same<C, double>::value && std::is_same<R, double>::value) {
    if (vec) {
      const int64_t hc = cols >> 1, pairs = rows * hc;
      const double* pa = (const double*)a;
      const double* pb = (const double*)b;
      double* po = (double*)out;
      const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
      const bool small = hc < (1u << 31) && pairs < (1ll << 31);
      const unsigned uhc = (unsigned)(small ? hc : 1);
      for (int64_t l0 = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; l0 < pairs; l0 += stride * 4) {
        double2 av[4], bv[4];
        int64_t off[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int64_t l = l0 + (int64_t)u * stride;
          off[u] = -1;
          if (l < pairs) {
            int64_t r, c2;
            if (small) { const unsigned q = (unsigned)l / uhc; r = q; c2 = (unsigned)l - q * uhc; }
            else { r = l / hc; c2 = l - r * hc; }
            off[u] = 2 * l;
            if (acs == 0) { const double t = pa[r * ars]; av[u] = make_double2(t, t); }
            else av[u] = *reinterpret_cast<const double2*>(pa + r * ars + 2 * c2);
            if (bcs == 0) { const double t = pb[r * brs]; bv[u] = make_double2(t, t); }
            else bv[u] = *reinterpret_cast<const double2*>(pb + r * brs + 2 * c2);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (off[u] >= 0) {
            double2 o;
            o.x = bop_apply<OP>(av[u].x, bv[u].x);
            o.y = bop_apply<OP>(av[u].y, bv[u].y);
            *reinterpret_cast<double2*>(po + off[u]) = o;
          }
        }
      }
      return;
    }
  }
  // Each CTA walks bands of whole rows; inside a band the (row, col) split is a 32-bit division and four
  // independent element loads are in flight per thread.
  const int64_t band = cols >= (1 << 20) ? 1 : ((int64_t)EW_THREADS * EW_UNROLL * 4 + cols - 1) / cols;   // rows per band
  const int64_t nbands = (rows + band - 1) / band;
  const unsigned ucols = (unsigned)(cols < (1u << 31) ? cols : 1);
  for (int64_t bb = blockIdx.x; bb < nbands; bb += gridDim.x) {
    const int64_t r0 = bb * band;
    const int64_t nr = (r0 + band < rows) ? band : rows - r0;
    const int64_t cnt = nr * cols;
    const int64_t obase = r0 * cols;
    for (int64_t l0 = threadIdx.x; l0 < cnt; l0 += (int64_t)EW_THREADS * EW_UNROLL) {
      C av[EW_UNROLL], bv[EW_UNROLL];
#pragma unroll
      for (int u = 0; u < EW_UNROLL; u++) {
        int64_t l = l0 + (int64_t)u * EW_THREADS;
        av[u] = C(); bv[u] = C();
        if (l < cnt) {
          int64_t r, c;
          if (cols < (1u << 31) && cnt < (1u << 31)) { unsigned q = (unsigned)l / ucols; r = q; c = (unsigned)l - q * ucols; }
          else { r = l / cols; c = l - r * cols; }
          r += r0;
          av[u] = load_as<C>(a, adt, r * ars + c * acs);
          bv[u] = load_as<C>(b, bdt, r * brs + c * bcs);
        }
      }
#pragma unroll
      for (int u = 0; u < EW_UNROLL; u++) {
        int64_t l = l0 + (int64_t)u * EW_THREADS;
        if (l < cnt) store_as<R>(out, odt, obase + l, bop_apply<OP>(av[u], bv[u]));
      }
    }
  }
}

template <int OP, typename C>
static int launch_bop(int device, int ndim, const int64_t* shape, const void* a, int adt, const int64_t* sa,
                      const void* b, int bdt, const int64_t* sb, void* out, int odt, cudaStream_t st) {
  int64_t n = 1;
  for (int i = 0; i < ndim; i++) n *= shape[i];
  if (n == 0) return 0;
  // classify operands: contiguous-with-output or all-zero strides
  bool a_contig = true, b_contig = true, a_scalar = true, b_scalar = true;
  int64_t expect = 1;
  for (int i = ndim - 1; i >= 0; i--) {
    if (shape[i] != 1) {
      if (sa[i] != expect) a_contig = false;
      if (sb[i] != expect) b_contig = false;
      if (sa[i] != 0) a_scalar = false;
      if (sb[i] != 0) b_scalar = false;
    }
    expect *= shape[i];
  }
  if ((a_contig || a_scalar) && (b_contig || b_scalar)) {
    int grid = grid_for(device, n, EW_THREADS, EW_UNROLL);
    const int ccode = std::is_same<C, double>::value ? NB_F64 : (std::is_same<C, int64_t>::value ? NB_I64 : NB_BOOL);
    const int pack = (odt == NB_BOOL && adt == ccode && bdt == ccode && sizeof(C) == 8 && n >= 4096 &&
                      (((uintptr_t)out) & 7) == 0) ? 1 : 0;
    k_bop_flat<OP, C><<<grid, EW_THREADS, 0, st>>>(n, a, adt, a_scalar && !a_contig ? 1 : 0, b, bdt,
                                                  b_scalar && !b_contig ? 1 : 0, out, odt, pack);
    NB_LAUNCH_CHECK();
    return 0;
  }
  Strides4 s;
  for (int i = 0; i < 4; i++) { s.shape[i] = 1; s.sa[i] = 0; s.sb[i] = 0; }
  for (int i = 0; i < ndim; i++) {
    s.shape[4 - ndim + i] = shape[i];
    s.sa[4 - ndim + i] = shape[i] == 1 ? 0 : sa[i];
    s.sb[4 - ndim + i] = shape[i] == 1 ? 0 : sb[i];
  }
  int grid = grid_for(device, n, EW_THREADS, s.shape[0] == 1 && s.shape[1] == 1 ? EW_UNROLL * 4 : 1);
  if (s.shape[0] == 1 && s.shape[1] == 1) {
    // vector path: all float64, even column count, every operand either column-broadcast or contiguous along the
    // columns with 16-byte aligned rows
    auto side_ok = [&](const void* p, int64_t rs, int64_t cs) {
      if (cs == 0) return true;
      return cs == 1 && (rs & 1) == 0 && (((uintptr_t)p) & 15) == 0;
    };
    const int vec = (adt == NB_F64 && bdt == NB_F64 && odt == NB_F64 && (s.shape[3] & 1) == 0 &&
                     (((uintptr_t)out) & 15) == 0 && side_ok(a, s.sa[2], s.sa[3]) && side_ok(b, s.sb[2], s.sb[3])) ? 1 : 0;
    k_bop_2d<OP, C><<<grid, EW_THREADS, 0, st>>>(s.shape[2], s.shape[3], a, adt, s.sa[2], s.sa[3], b, bdt,
                                                s.sb[2], s.sb[3], out, odt, vec);
  } else {
    k_bop_strided<OP, C><<<grid, EW_THREADS, 0, st>>>(n, s, a, adt, b, bdt, out, odt);
  }
  NB_LAUNCH_CHECK();
  return 0;
}

#define NB_BOP_CASE(OPC, CT) \
  case OPC: return launch_bop<OPC, CT>(device, ndim, shape, a, adt, sa, b, bdt, sb, out, odt, st);

static int dispatch_bop(int device, int op, int cdt, int ndim, const int64_t* shape, const void* a, int adt,
                        const int64_t* sa, const void* b, int bdt, const int64_t* sb, void* out, int odt,
                        cudaStream_t st) {
  if (cdt == NB_F64) {
    switch (op) {
      NB_BOP_CASE(NB_ADD, double) NB_BOP_CASE(NB_SUB, double) NB_BOP_CASE(NB_MUL, double)
      NB_BOP_CASE(NB_TRUEDIV, double) NB_BOP_CASE(NB_FLOORDIV, double) NB_BOP_CASE(NB_MOD, double)
      NB_BOP_CASE(NB_FMOD, double) NB_BOP_CASE(NB_POW, double) NB_BOP_CASE(NB_FLOAT_POW, double)
      NB_BOP_CASE(NB_MAXIMUM, double) NB_BOP_CASE(NB_MINIMUM, double) NB_BOP_CASE(NB_FMAX, double)
      NB_BOP_CASE(NB_FMIN, double) NB_BOP_CASE(NB_LT, double) NB_BOP_CASE(NB_LE, double)
      NB_BOP_CASE(NB_GT, double) NB_BOP_CASE(NB_GE, double) NB_BOP_CASE(NB_EQ, double)
      NB_BOP_CASE(NB_NE, double) NB_BOP_CASE(NB_LOGICAL_AND, double) NB_BOP_CASE(NB_LOGICAL_OR, double)
      NB_BOP_CASE(NB_LOGICAL_XOR, double) NB_BOP_CASE(NB_ARCTAN2, double) NB_BOP_CASE(NB_HYPOT, double)
      NB_BOP_CASE(NB_COPYSIGN, double) NB_BOP_CASE(NB_NEXTAFTER, double) NB_BOP_CASE(NB_LDEXP, double)
      NB_BOP_CASE(NB_HEAVISIDE, double) NB_BOP_CASE(NB_LOGADDEXP, double) NB_BOP_CASE(NB_LOGADDEXP2, double)
      NB_BOP_CASE(NB_XLOGY, double)
      default: break;
    }
  } else if (cdt == NB_I64) {
    switch (op) {
      NB_BOP_CASE(NB_ADD, int64_t) NB_BOP_CASE(NB_SUB, int64_t) NB_BOP_CASE(NB_MUL, int64_t)
      NB_BOP_CASE(NB_FLOORDIV, int64_t) NB_BOP_CASE(NB_MOD, int64_t) NB_BOP_CASE(NB_FMOD, int64_t)
      NB_BOP_CASE(NB_POW, int64_t) NB_BOP_CASE(NB_MAXIMUM, int64_t) NB_BOP_CASE(NB_MINIMUM, int64_t)
      NB_BOP_CASE(NB_FMAX, int64_t) NB_BOP_CASE(NB_FMIN, int64_t) NB_BOP_CASE(NB_LT, int64_t)
      NB_BOP_CASE(NB_LE, int64_t) NB_BOP_CASE(NB_GT, int64_t) NB_BOP_CASE(NB_GE, int64_t)
      NB_BOP_CASE(NB_EQ, int64_t) NB_BOP_CASE(NB_NE, int64_t) NB_BOP_CASE(NB_LOGICAL_AND, int64_t)
      NB_BOP_CASE(NB_LOGICAL_OR, int64_t) NB_BOP_CASE(NB_LOGICAL_XOR, int64_t) NB_BOP_CASE(NB_BIT_AND, int64_t)
      NB_BOP_CASE(NB_BIT_OR, int64_t) NB_BOP_CASE(NB_BIT_XOR, int64_t) NB_BOP_CASE(NB_LSHIFT, int64_t)
      NB_BOP_CASE(NB_RSHIFT, int64_t) NB_BOP_CASE(NB_GCD, int64_t) NB_BOP_CASE(NB_LCM, int64_t)
      default: break;
    }
  } else if (cdt == NB_BOOL) {
    switch (op) {
      NB_BOP_CASE(NB_ADD, bool) NB_BOP_CASE(NB_MUL, bool) NB_BOP_CASE(NB_MAXIMUM, bool)
      NB_BOP_CASE(NB_MINIMUM, bool) NB_BOP_CASE(NB_FMAX, bool) NB_BOP_CASE(NB_FMIN, bool)
      NB_BOP_CASE(NB_LT, bool) NB_BOP_CASE(NB_LE, bool) NB_BOP_CASE(NB_GT, bool) NB_BOP_CASE(NB_GE, bool)
      NB_BOP_CASE(NB_EQ, bool) NB_BOP_CASE(NB_NE, bool) NB_BOP_CASE(NB_LOGICAL_AND, bool)
      NB_BOP_CASE(NB_LOGICAL_OR, bool) NB_BOP_CASE(NB_LOGICAL_XOR, bool) NB_BOP_CASE(NB_BIT_AND, bool)
      NB_BOP_CASE(NB_BIT_OR, bool) NB_BOP_CASE(NB_BIT_XOR, bool)
      default: break;
    }
  }
  set_error("nb_bop: op %d is not defined for compute dtype %d", op, cdt);
  return 2;
}

// ------------------------------------------------------------------------------------------------
// unary
// ------------------------------------------------------------------------------------------------
template <int OP, typename C>
struct UopResult { using type = C; };
#define NB_UBOOL(OPC) \
  template <typename C> struct UopResult<OPC, C> { using type = bool; };
NB_UBOOL(NB_ISNAN) NB_UBOOL(NB_ISINF) NB_UBOOL(NB_ISFINITE) NB_UBOOL(NB_SIGNBIT) NB_UBOOL(NB_LOGICAL_NOT)
#undef NB_UBOOL

template <int OP>
__device__ __forceinline__ typename UopResult<OP, double>::type uop_apply(double x) {
  if constexpr (OP == NB_ABS) return fabs(x);
  else if constexpr (OP == NB_NEGATIVE) return -x;
  else if constexpr (OP == NB_POSITIVE) return x;
  else if constexpr (OP == NB_SIGN) return isnan(x) ? x : (x > 0.0 ? 1.0 : (x < 0.0 ? -1.0 : 0.0));
  else if constexpr (OP == NB_SQUARE) return x * x;
  else if constexpr (OP == NB_RECIPROCAL) return 1.0 / x;
  else if constexpr (OP == NB_SQRT) return sqrt(x);
  else if constexpr (OP == NB_CBRT) return cbrt(x);
  else if constexpr (OP == NB_EXP) return exp(x);
  else if constexpr (OP == NB_EXP2) return exp2(x);
  else if constexpr (OP == NB_EXPM1) return expm1(x);
  else if constexpr (OP == NB_LOG) return log(x);
  else if constexpr (OP == NB_LOG2) return log2(x);
  else if constexpr (OP == NB_LOG10) return log10(x);
  else if constexpr (OP == NB_LOG1P) return log1p(x);
  else if constexpr (OP == NB_SIN) return sin(x);
  else if constexpr (OP == NB_COS) return cos(x);
  else if constexpr (OP == NB_TAN) return tan(x);
  else if constexpr (OP == NB_ARCSIN) return asin(x);
  else if constexpr (OP == NB_ARCCOS) return acos(x);
  else if constexpr (OP == NB_ARCTAN) return atan(x);
  else if constexpr (OP == NB_SINH) return sinh(x);
  else if constexpr (OP == NB_COSH) return cosh(x);
  else if constexpr (OP == NB_TANH) return tanh(x);
  else if constexpr (OP == NB_ARCSINH) return asinh(x);
  else if constexpr (OP == NB_ARCCOSH) return acosh(x);
  else if constexpr (OP == NB_ARCTANH) return atanh(x);
  else if constexpr (OP == NB_CEIL) return ceil(x);
  else if constexpr (OP == NB_FLOOR) return floor(x);
  else if constexpr (OP == NB_TRUNC) return trunc(x);
  else if constexpr (OP == NB_RINT) return rint(x);
  else if constexpr (OP == NB_DEG2RAD) return x * (3.141592653589793238462643383279502884 / 180.0);
  else if constexpr (OP == NB_RAD2DEG) return x * (180.0 / 3.141592653589793238462643383279502884);
  else if constexpr (OP == NB_ISNAN) return isnan(x);
  else if constexpr (OP == NB_ISINF) return isinf(x);
  else if constexpr (OP == NB_ISFINITE) return isfinite(x);
  else if constexpr (OP == NB_SIGNBIT) return signbit(x);
  else if constexpr (OP == NB_LOGICAL_NOT) return x == 0.0;
  else if constexpr (OP == NB_SPACING) {
    if (isinf(x)) return nan("");
    if (isnan(x)) return x;
    return nextafter(x, copysign(INFINITY, x)) - x;
  }
  else return x;
}

template <int OP>
__device__ __forceinline__ typename UopResult<OP, int64_t>::type uop_apply(int64_t x) {
  if constexpr (OP == NB_ABS) return x < 0 ? i_sub(0, x) : x;
  else if constexpr (OP == NB_NEGATIVE) return i_sub(0, x);
  else if constexpr (OP == NB_POSITIVE) return x;
  else if constexpr (OP == NB_SIGN) return x > 0 ? 1 : (x < 0 ? -1 : 0);
  else if constexpr (OP == NB_SQUARE) return i_mul(x, x);
  else if constexpr (OP == NB_RECIPROCAL) return x == 0 ? INT64_MIN : (x == 1 ? 1 : (x == -1 ? -1 : 0));
  else if constexpr (OP == NB_CEIL || OP == NB_FLOOR || OP == NB_TRUNC || OP == NB_RINT) return x;
  else if constexpr (OP == NB_ISNAN || OP == NB_ISINF || OP == NB_SIGNBIT) return OP == NB_SIGNBIT ? (x < 0) : false;
  else if constexpr (OP == NB_ISFINITE) return true;
  else if constexpr (OP == NB_LOGICAL_NOT) return x == 0;
  else if constexpr (OP == NB_INVERT) return ~x;
  else return x;
}

template <int OP>
__device__ __forceinline__ bool uop_apply(bool x) {
  if constexpr (OP == NB_LOGICAL_NOT || OP == NB_INVERT) return !x;
  else if constexpr (OP == NB_ISFINITE) return true;
  else if constexpr (OP == NB_ISNAN || OP == NB_ISINF || OP == NB_SIGNBIT) return false;
  else return x;  // abs/positive/sign/square/ceil/floor/trunc/rint keep a bool unchanged
}

template <int OP, typename C>
__global__ void __launch_bounds__(EW_THREADS)
k_uop(int64_t n, const void* __restrict__ a, int adt, void* __restrict__ out, int odt) {
  using R = typename UopResult<OP, C>::type;
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
  for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n; base += stride) {
    C v[EW_UNROLL];
#pragma unroll
    for (int u = 0; u < EW_UNROLL; u++) {
      int64_t i = base + (int64_t)u * EW_THREADS;
      v[u] = C();
      if (i < n) v[u] = load_as<C>(a, adt, i);
    }
#pragma unroll
    for (int u = 0; u < EW_UNROLL; u++) {
      int64_t i = base + (int64_t)u * EW_THREADS;
      if (i < n) store_as<R>(out, odt, i, uop_apply<OP>(v[u]));
    }
  }
}

template <int OP, typename C>
static int launch_uop(int device, int64_t n, const void* a, int adt, void* out, int odt, cudaStream_t st) {
  if (n == 0) return 0;
  int grid = grid_for(device, n, EW_THREADS, EW_UNROLL);
  k_uop<OP, C><<<grid, EW_THREADS, 0, st>>>(n, a, adt, out, odt);
  NB_LAUNCH_CHECK();
  return 0;
}

#define NB_UOP_CASE(OPC, CT) case OPC: return launch_uop<OPC, CT>(device, n, a, adt, out, odt, st);

static int dispatch_uop(int device, int op, int cdt, int64_t n, const void* a, int adt, void* out, int odt,
                        cudaStream_t st) {
  if (cdt == NB_F64) {
    switch (op) {
      NB_UOP_CASE(NB_ABS, double) NB_UOP_CASE(NB_NEGATIVE, double) NB_UOP_CASE(NB_POSITIVE, double)
      NB_UOP_CASE(NB_SIGN, double) NB_UOP_CASE(NB_SQUARE, double) NB_UOP_CASE(NB_RECIPROCAL, double)
      NB_UOP_CASE(NB_SQRT, double) NB_UOP_CASE(NB_CBRT, double) NB_UOP_CASE(NB_EXP, double)
      NB_UOP_CASE(NB_EXP2, double) NB_UOP_CASE(NB_EXPM1, double) NB_UOP_CASE(NB_LOG, double)
      NB_UOP_CASE(NB_LOG2, double) NB_UOP_CASE(NB_LOG10, double) NB_UOP_CASE(NB_LOG1P, double)
      NB_UOP_CASE(NB_SIN, double) NB_UOP_CASE(NB_COS, double) NB_UOP_CASE(NB_TAN, double)
      NB_UOP_CASE(NB_ARCSIN, double) NB_UOP_CASE(NB_ARCCOS, double) NB_UOP_CASE(NB_ARCTAN, double)
      NB_UOP_CASE(NB_SINH, double) NB_UOP_CASE(NB_COSH, double) NB_UOP_CASE(NB_TANH, double)
      NB_UOP_CASE(NB_ARCSINH, double) NB_UOP_CASE(NB_ARCCOSH, double) NB_UOP_CASE(NB_ARCTANH, double)
      NB_UOP_CASE(NB_CEIL, double) NB_UOP_CASE(NB_FLOOR, double) NB_UOP_CASE(NB_TRUNC, double)
      NB_UOP_CASE(NB_RINT, double) NB_UOP_CASE(NB_DEG2RAD, double) NB_UOP_CASE(NB_RAD2DEG, double)
      NB_UOP_CASE(NB_ISNAN, double) NB_UOP_CASE(NB_ISINF, double) NB_UOP_CASE(NB_ISFINITE, double)
      NB_UOP_CASE(NB_SIGNBIT, double) NB_UOP_CASE(NB_LOGICAL_NOT, double) NB_UOP_CASE(NB_SPACING, double)
      default: break;
    }
  } else if (cdt == NB_I64) {
    switch (op) {
      NB_UOP_CASE(NB_ABS, int64_t) NB_UOP_CASE(NB_NEGATIVE, int64_t) NB_UOP_CASE(NB_POSITIVE, int64_t)
      NB_UOP_CASE(NB_SIGN, int64_t) NB_UOP_CASE(NB_SQUARE, int64_t) NB_UOP_CASE(NB_RECIPROCAL, int64_t)
      NB_UOP_CASE(NB_CEIL, int64_t) NB_UOP_CASE(NB_FLOOR, int64_t) NB_UOP_CASE(NB_TRUNC, int64_t)
      NB_UOP_CASE(NB_RINT, int64_t) NB_UOP_CASE(NB_ISNAN, int64_t) NB_UOP_CASE(NB_ISINF, int64_t)
      NB_UOP_CASE(NB_ISFINITE, int64_t) NB_UOP_CASE(NB_SIGNBIT, int64_t) NB_UOP_CASE(NB_LOGICAL_NOT, int64_t)
      NB_UOP_CASE(NB_INVERT, int64_t)
      default: break;
    }
  } else if (cdt == NB_BOOL) {
    switch (op) {
      NB_UOP_CASE(NB_ABS, bool) NB_UOP_CASE(NB_POSITIVE, bool) NB_UOP_CASE(NB_SQUARE, bool)
      NB_UOP_CASE(NB_CEIL, bool) NB_UOP_CASE(NB_FLOOR, bool) NB_UOP_CASE(NB_TRUNC, bool)
      NB_UOP_CASE(NB_RINT, bool) NB_UOP_CASE(NB_ISNAN, bool) NB_UOP_CASE(NB_ISINF, bool)
      NB_UOP_CASE(NB_ISFINITE, bool) NB_UOP_CASE(NB_SIGNBIT, bool) NB_UOP_CASE(NB_LOGICAL_NOT, bool)
      NB_UOP_CASE(NB_INVERT, bool)
      default: break;
    }
  }
  set_error("nb_uop: op %d is not defined for compute dtype %d", op, cdt);
  return 2;
}

// ------------------------------------------------------------------------------------------------
// where (3-arg), astype, fill, eye, strided copy
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(EW_THREADS)
k_where3(int64_t n, const void* __restrict__ cond, int cdt, const T* __restrict__ x, const T* __restrict__ y,
         T* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = load_as<bool>(cond, cdt, i) ? x[i] : y[i];
}

template <typename C>
__global__ void __launch_bounds__(EW_THREADS)
k_astype(int64_t n, const void* __restrict__ a, int adt, void* __restrict__ out, int odt) {
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
    store_as<C>(out, odt, i, load_as<C>(a, adt, i));
}

__global__ void __launch_bounds__(EW_THREADS)
k_fill(int64_t n, void* __restrict__ out, int odt, double fv, int64_t iv) {
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
    if (odt == NB_F64 || odt == NB_F32) store_as<double>(out, odt, i, fv);
    else store_as<int64_t>(out, odt, i, iv);
  }
}

__global__ void __launch_bounds__(EW_THREADS)
k_eye(int64_t rows, int64_t cols, void* __restrict__ out, int odt) {
  const int64_t n = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
    int64_t r = i / cols, c = i - r * cols;
    store_as<int64_t>(out, odt, i, r == c ? 1 : 0);
  }
}

struct CopyDesc { int64_t shape[4]; int64_t ss[4]; int64_t ds[4]; };

template <typename T>
__global__ void __launch_bounds__(EW_THREADS)
k_copy_strided(int64_t n, CopyDesc d, const T* __restrict__ src, T* __restrict__ dst) {
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
    int64_t r = i;
    int64_t i3 = r % d.shape[3]; r /= d.shape[3];
    int64_t i2 = r % d.shape[2]; r /= d.shape[2];
    int64_t i1 = r % d.shape[1]; r /= d.shape[1];
    int64_t i0 = r;
    dst[i0 * d.ds[0] + i1 * d.ds[1] + i2 * d.ds[2] + i3 * d.ds[3]] =
        src[i0 * d.ss[0] + i1 * d.ss[1] + i2 * d.ss[2] + i3 * d.ss[3]];
  }
}

// 2-D transpose-style copies go through a padded shared-memory tile so both sides stay coalesced.
template <typename T>
__global__ void __launch_bounds__(256)
k_transpose_tile(int64_t rows, int64_t cols, const T* __restrict__ src, int64_t s_rs, T* __restrict__ dst,
                 int64_t d_cs) {
  // src[r*s_rs + c] (c contiguous)  ->  dst[c*d_cs + r] (r contiguous)
  __shared__ T tile[32][33];
  const int64_t tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int64_t t = blockIdx.x; t < tiles_c * tiles_r; t += gridDim.x) {
    int64_t tr = t / tiles_c, tc = t - tr * tiles_c;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      int64_t r = tr * 32 + ty + j, c = tc * 32 + tx;
      if (r < rows && c < cols) tile[ty + j][tx] = src[r * s_rs + c];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      int64_t c = tc * 32 + ty + j, r = tr * 32 + tx;
      if (r < rows && c < cols) dst[c * d_cs + r] = tile[tx][ty + j];
    }
    __syncthreads();
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_bop(int device, int op, int ndim, const int64_t* shape, const void* a_d, int adt, const int64_t* sa,
           const void* b_d, int bdt, const int64_t* sb, void* out_d, int odt, int cdt, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(ndim >= 0 && ndim <= 4, "nb_bop: ndim %d > 4 is not supported", ndim);
  NB_REQUIRE(op >= 0 && op < NB_BOP_COUNT, "nb_bop: unknown op %d", op);
  return dispatch_bop(device, op, cdt, ndim, shape, a_d, adt, sa, b_d, bdt, sb, out_d, odt, st);
}

int nb_uop(int device, int op, int64_t n, const void* a_d, int adt, void* out_d, int odt, int cdt, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(op >= 0 && op < NB_UOP_COUNT, "nb_uop: unknown op %d", op);
  return dispatch_uop(device, op, cdt, n, a_d, adt, out_d, odt, st);
}

int nb_where3(int device, int64_t n, const void* cond_d, int cond_dt, const void* x_d, const void* y_d,
              void* out_d, int xdt, void* stream) {
  NB_ENTER(device);
  if (n == 0) return 0;
  int grid = grid_for(device, n, EW_THREADS, 1);
  if (item_size(xdt) == 8)
    k_where3<uint64_t><<<grid, EW_THREADS, 0, st>>>(n, cond_d, cond_dt, (const uint64_t*)x_d, (const uint64_t*)y_d, (uint64_t*)out_d);
  else if (item_size(xdt) == 4)
    k_where3<uint32_t><<<grid, EW_THREADS, 0, st>>>(n, cond_d, cond_dt, (const uint32_t*)x_d, (const uint32_t*)y_d, (uint32_t*)out_d);
  else
    k_where3<uint8_t><<<grid, EW_THREADS, 0, st>>>(n, cond_d, cond_dt, (const uint8_t*)x_d, (const uint8_t*)y_d, (uint8_t*)out_d);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_astype(int device, int64_t n, const void* a_d, int adt, void* out_d, int odt, void* stream) {
  NB_ENTER(device);
  if (n == 0) return 0;
  int grid = grid_for(device, n, EW_THREADS, 1);
  if (odt == NB_BOOL) {
    k_astype<bool><<<grid, EW_THREADS, 0, st>>>(n, a_d, adt, out_d, odt);
  } else if (adt == NB_F64 || adt == NB_F32) {
    k_astype<double><<<grid, EW_THREADS, 0, st>>>(n, a_d, adt, out_d, odt);
  } else {
    k_astype<int64_t><<<grid, EW_THREADS, 0, st>>>(n, a_d, adt, out_d, odt);
  }
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_fill(int device, int64_t n, void* out_d, int odt, double fvalue, int64_t ivalue, void* stream) {
  NB_ENTER(device);
  if (n == 0) return 0;
  int grid = grid_for(device, n, EW_THREADS, 1);
  k_fill<<<grid, EW_THREADS, 0, st>>>(n, out_d, odt, fvalue, ivalue);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_eye(int device, int64_t rows, int64_t cols, void* out_d, int odt, void* stream) {
  NB_ENTER(device);
  if (rows * cols == 0) return 0;
  int grid = grid_for(device, rows * cols, EW_THREADS, 1);
  k_eye<<<grid, EW_THREADS, 0, st>>>(rows, cols, out_d, odt);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_copy_strided(int device, int ndim, const int64_t* shape, int itemsize, const void* src_d,
                    const int64_t* sstrides, void* dst_d, const int64_t* dstrides, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(ndim >= 0 && ndim <= 4, "nb_copy_strided: ndim %d > 4 is not supported", ndim);
  NB_REQUIRE(itemsize == 8 || itemsize == 4 || itemsize == 1, "nb_copy_strided: itemsize %d", itemsize);
  CopyDesc d;
  for (int i = 0; i < 4; i++) { d.shape[i] = 1; d.ss[i] = 0; d.ds[i] = 0; }
  int64_t n = 1;
  for (int i = 0; i < ndim; i++) {
    d.shape[4 - ndim + i] = shape[i];
    d.ss[4 - ndim + i] = sstrides[i];
    d.ds[4 - ndim + i] = dstrides[i];
    n *= shape[i];
  }
  if (n == 0) return 0;
  // 2-D transposing copy (src contiguous along cols, dst contiguous along rows): tiled path.
  if (itemsize == 8 && ndim == 2 && sstrides[1] == 1 && dstrides[0] == 1 && shape[0] >= 32 && shape[1] >= 32) {
    int64_t tiles = ((shape[0] + 31) / 32) * ((shape[1] + 31) / 32);
    int grid = (int)(tiles < (int64_t)dev_info(device).sms * 8 ? tiles : (int64_t)dev_info(device).sms * 8);
    k_transpose_tile<uint64_t><<<grid, 256, 0, st>>>(shape[0], shape[1], (const uint64_t*)src_d, sstrides[0],
                                                     (uint64_t*)dst_d, dstrides[1]);
    NB_LAUNCH_CHECK();
    return 0;
  }
  int grid = grid_for(device, n, EW_THREADS, 1);
  if (itemsize == 8)
    k_copy_strided<uint64_t><<<grid, EW_THREADS, 0, st>>>(n, d, (const uint64_t*)src_d, (uint64_t*)dst_d);
  else if (itemsize == 4)
    k_copy_strided<uint32_t><<<grid, EW_THREADS, 0, st>>>(n, d, (const uint32_t*)src_d, (uint32_t*)dst_d);
  else
    k_copy_strided<uint8_t><<<grid, EW_THREADS, 0, st>>>(n, d, (const uint8_t*)src_d, (uint8_t*)dst_d);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
