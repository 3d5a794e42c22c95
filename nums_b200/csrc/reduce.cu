// Reductions, arg-reductions and stream compaction.
// Reference: nums/core/kernel/numpy_kernel.py — reduce_axis :294-313, sum_reduce :291-292,
// bop_reduce :433-453, arg_op :492-508, where (1-arg) :320-330.
// All HBM-bound.  Every reduction is a fixed-order two-stage tree (per-thread strided partials ->
// warp shuffles -> shared memory -> ordered second pass): no floating-point atomics, so results
// are run-to-run identical.
#include "common.cuh"
#include <mutex>

namespace nb {

constexpr int RED_THREADS = 256;
constexpr int64_t RED_TARGET_BLOCKS = 592;  // 4 x 148 SMs; fixed so that the tree shape is device independent

// ------------------------------------------------------------------------------------------------
// combine functors
// ------------------------------------------------------------------------------------------------
template <int RED, typename T> struct Red;

template <typename T> struct Red<NB_RSUM, T> {
  static __device__ __forceinline__ T identity() { return T(0); }
  static __device__ __forceinline__ T combine(T a, T b) { return a + b; }
  static __device__ __forceinline__ T pre(T v) { return v; }
};
template <> struct Red<NB_RSUM, int64_t> {
  static __device__ __forceinline__ int64_t identity() { return 0; }
  static __device__ __forceinline__ int64_t combine(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }
  static __device__ __forceinline__ int64_t pre(int64_t v) { return v; }
};
template <typename T> struct Red<NB_RPROD, T> {
  static __device__ __forceinline__ T identity() { return T(1); }
  static __device__ __forceinline__ T combine(T a, T b) { return a * b; }
  static __device__ __forceinline__ T pre(T v) { return v; }
};
template <> struct Red<NB_RPROD, int64_t> {
  static __device__ __forceinline__ int64_t identity() { return 1; }
  static __device__ __forceinline__ int64_t combine(int64_t a, int64_t b) { return (int64_t)((uint64_t)a * (uint64_t)b); }
  static __device__ __forceinline__ int64_t pre(int64_t v) { return v; }
};
template <> struct Red<NB_RMIN, double> {
  static __device__ __forceinline__ double identity() { return INFINITY; }
  static __device__ __forceinline__ double combine(double a, double b) { return (a <= b || isnan(a)) ? a : b; }
  static __device__ __forceinline__ double pre(double v) { return v; }
};
template <> struct Red<NB_RMAX, double> {
  static __device__ __forceinline__ double identity() { return -INFINITY; }
  static __device__ __forceinline__ double combine(double a, double b) { return (a >= b || isnan(a)) ? a : b; }
  static __device__ __forceinline__ double pre(double v) { return v; }
};
template <> struct Red<NB_RMIN, int64_t> {
  static __device__ __forceinline__ int64_t identity() { return INT64_MAX; }
  static __device__ __forceinline__ int64_t combine(int64_t a, int64_t b) { return a < b ? a : b; }
  static __device__ __forceinline__ int64_t pre(int64_t v) { return v; }
};
template <> struct Red<NB_RMAX, int64_t> {
  static __device__ __forceinline__ int64_t identity() { return INT64_MIN; }
  static __device__ __forceinline__ int64_t combine(int64_t a, int64_t b) { return a > b ? a : b; }
  static __device__ __forceinline__ int64_t pre(int64_t v) { return v; }
};
template <> struct Red<NB_RMIN, bool> {
  static __device__ __forceinline__ bool identity() { return true; }
  static __device__ __forceinline__ bool combine(bool a, bool b) { return a && b; }
  static __device__ __forceinline__ bool pre(bool v) { return v; }
};
template <> struct Red<NB_RMAX, bool> {
  static __device__ __forceinline__ bool identity() { return false; }
  static __device__ __forceinline__ bool combine(bool a, bool b) { return a || b; }
  static __device__ __forceinline__ bool pre(bool v) { return v; }
};
template <> struct Red<NB_RALL, bool> {
  static __device__ __forceinline__ bool identity() { return true; }
  static __device__ __forceinline__ bool combine(bool a, bool b) { return a && b; }
  static __device__ __forceinline__ bool pre(bool v) { return v; }
};
template <> struct Red<NB_RANY, bool> {
  static __device__ __forceinline__ bool identity() { return false; }
  static __device__ __forceinline__ bool combine(bool a, bool b) { return a || b; }
  static __device__ __forceinline__ bool pre(bool v) { return v; }
};
template <> struct Red<NB_RNANSUM, double> {
  static __device__ __forceinline__ double identity() { return 0.0; }
  static __device__ __forceinline__ double combine(double a, double b) { return a + b; }
  static __device__ __forceinline__ double pre(double v) { return isnan(v) ? 0.0 : v; }
};
template <> struct Red<NB_RNANMIN, double> {
  static __device__ __forceinline__ double identity() { return nan(""); }
  static __device__ __forceinline__ double combine(double a, double b) { return fmin(a, b); }
  static __device__ __forceinline__ double pre(double v) { return v; }
};
template <> struct Red<NB_RNANMAX, double> {
  static __device__ __forceinline__ double identity() { return nan(""); }
  static __device__ __forceinline__ double combine(double a, double b) { return fmax(a, b); }
  static __device__ __forceinline__ double pre(double v) { return v; }
};

template <typename T> struct ShflT { using type = T; };
template <> struct ShflT<bool> { using type = int; };

template <typename R, typename T>
__device__ __forceinline__ T warp_red(T v) {
  using S = typename ShflT<T>::type;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    T other = (T)__shfl_down_sync(0xffffffffu, (S)v, o);
    v = R::combine(v, other);
  }
  return v;  // valid in lane 0
}

// Block reduction, result valid in thread 0.  Order: lanes (tree), then warps in index order.
template <typename R, typename T>
__device__ __forceinline__ T block_red(T v, T* sm /* >= 8 */) {
  v = warp_red<R, T>(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  T r = R::identity();
  if (threadIdx.x == 0) {
    r = sm[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) r = R::combine(r, sm[w]);
  }
  return r;
}

// ------------------------------------------------------------------------------------------------
// Kernel A: inner == 1.  Block (s, o) reduces a[o, s*chunk : (s+1)*chunk] -> part[o*S + s].
// ------------------------------------------------------------------------------------------------
template <int RED, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_rows(const void* __restrict__ a, int adt, int64_t outer, int64_t len, int64_t S, int64_t chunk,
              void* __restrict__ out, int odt) {
  using R = Red<RED, T>;
  __shared__ T sm[8];
  const int64_t s = blockIdx.x;
  for (int64_t o = blockIdx.y; o < outer; o += gridDim.y) {
    const int64_t lo = s * chunk, hi = (lo + chunk < len) ? lo + chunk : len;
    const int64_t base = o * len;
    T acc[4];
#pragma unroll
    for (int u = 0; u < 4; u++) acc[u] = R::identity();
    int64_t i = lo + threadIdx.x;
    for (; i + 3 * RED_THREADS < hi; i += 4 * RED_THREADS) {
      T v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) v[u] = load_as<T>(a, adt, base + i + u * RED_THREADS);
#pragma unroll
      for (int u = 0; u < 4; u++) acc[u] = R::combine(acc[u], R::pre(v[u]));
    }
    for (; i < hi; i += RED_THREADS) acc[0] = R::combine(acc[0], R::pre(load_as<T>(a, adt, base + i)));
    T v = R::combine(R::combine(acc[0], acc[1]), R::combine(acc[2], acc[3]));
    v = block_red<R, T>(v, sm);
    if (threadIdx.x == 0) store_as<T>(out, odt, o * S + s, v);
  }
}

// Kernel A2: inner == 1, short rows: LPR lanes per row (a power of two <= 32), 32/LPR rows per warp, so a
// warp always covers >= 256 contiguous bytes per load instruction.
template <int RED, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_rows_warp(const void* __restrict__ a, int adt, int64_t outer, int64_t len, int lpr, void* __restrict__ out,
                   int odt) {
  using R = Red<RED, T>;
  using S = typename ShflT<T>::type;
  const int lane = threadIdx.x & 31;
  const int rpw = 32 / lpr;                        // rows per warp
  const int sub = lane / lpr, sl = lane % lpr;
  const int64_t warps = (int64_t)gridDim.x * (RED_THREADS / 32);
  for (int64_t o0 = ((int64_t)blockIdx.x * (RED_THREADS / 32) + (threadIdx.x >> 5)) * rpw; o0 < outer; o0 += warps * rpw) {
    const int64_t o = o0 + sub;
    T acc0 = R::identity(), acc1 = R::identity();
    if (o < outer) {
      const int64_t base = o * len;
      int64_t i = sl;
      for (; i + lpr < len; i += 2 * lpr) {
        T v0 = load_as<T>(a, adt, base + i), v1 = load_as<T>(a, adt, base + i + lpr);
        acc0 = R::combine(acc0, R::pre(v0));
        acc1 = R::combine(acc1, R::pre(v1));
      }
      if (i < len) acc0 = R::combine(acc0, R::pre(load_as<T>(a, adt, base + i)));
    }
    T acc = R::combine(acc0, acc1);
    for (int off = lpr >> 1; off > 0; off >>= 1) {
      T other = (T)__shfl_down_sync(0xffffffffu, (S)acc, off);
      acc = R::combine(acc, other);
    }
    if (sl == 0 && o < outer) store_as<T>(out, odt, o, acc);
  }
}

// Kernel A3: inner == 1, very short rows (len <= 16): one thread per row, sequential (NumPy order).
template <int RED, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_rows_thread(const void* __restrict__ a, int adt, int64_t outer, int64_t len, void* __restrict__ out, int odt) {
  using R = Red<RED, T>;
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  for (int64_t o = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; o < outer; o += stride) {
    T acc = R::identity();
    for (int64_t i = 0; i < len; i++) acc = R::combine(acc, R::pre(load_as<T>(a, adt, o * len + i)));
    store_as<T>(out, odt, o, acc);
  }
}

// ------------------------------------------------------------------------------------------------
// Kernel B: inner > 1.  out[o, s, i] = red over k in split s of a[o, k, i].
// Thread t of the first T_active = (256/tile)*tile threads owns column i = t % tile and walks
// k = k0 + t / tile with step 256/tile, so every warp reads contiguous memory.
// ------------------------------------------------------------------------------------------------
template <int RED, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_cols(const void* __restrict__ a, int adt, int64_t outer, int64_t len, int64_t inner, int64_t S,
              int64_t chunk, int tile, void* __restrict__ out, int odt) {
  using R = Red<RED, T>;
  __shared__ T sm[RED_THREADS];
  const int kpar = RED_THREADS / tile;       // rows walked in parallel
  const int t = threadIdx.x;
  const bool active = t < kpar * tile;
  const int ti = t % tile, tk = t / tile;
  const int64_t col0 = (int64_t)blockIdx.x * tile;
  const int64_t s = blockIdx.y;
  for (int64_t o = blockIdx.z; o < outer; o += gridDim.z) {
    const int64_t lo = s * chunk, hi = (lo + chunk < len) ? lo + chunk : len;
    T acc0 = R::identity(), acc1 = R::identity();
    const int64_t col = col0 + ti;
    if (active && col < inner) {
      const int64_t base = o * len * inner + col;
      int64_t k = lo + tk;
      // eight rows in flight per thread: the walk is a pure stream, latency is hidden by the loads outstanding
      T acc[8];
#pragma unroll
      for (int u = 0; u < 8; u++) acc[u] = R::identity();
      for (; k + 7 * (int64_t)kpar < hi; k += 8 * kpar) {
        T v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = load_as<T>(a, adt, base + (k + (int64_t)u * kpar) * inner);
#pragma unroll
        for (int u = 0; u < 8; u++) acc[u] = R::combine(acc[u], R::pre(v[u]));
      }
      for (; k < hi; k += kpar) acc[0] = R::combine(acc[0], R::pre(load_as<T>(a, adt, base + k * inner)));
      acc0 = R::combine(R::combine(acc[0], acc[2]), R::combine(acc[4], acc[6]));
      acc1 = R::combine(R::combine(acc[1], acc[3]), R::combine(acc[5], acc[7]));
    }
    __syncthreads();
    sm[t] = R::combine(acc0, acc1);
    __syncthreads();
    if (t < tile && col < inner) {
      T r = sm[t];
      for (int j = 1; j < kpar; j++) r = R::combine(r, sm[t + j * tile]);
      store_as<T>(out, odt, (o * S + s) * inner + col, r);
    }
  }
}

struct RedPlan { int64_t S; int64_t chunk; int tile; int64_t col_blocks; };

static RedPlan plan_reduce(int64_t outer, int64_t len, int64_t inner) {
  RedPlan p;
  if (inner == 1) {
    p.tile = 1; p.col_blocks = 1;
    int64_t want = RED_TARGET_BLOCKS / (outer < RED_TARGET_BLOCKS ? (outer > 0 ? outer : 1) : RED_TARGET_BLOCKS);
    if (want < 1) want = 1;
    int64_t max_s = (len + 4095) / 4096;
    p.S = want < max_s ? want : max_s;
    if (p.S < 1) p.S = 1;
  } else {
    p.tile = inner >= RED_THREADS ? RED_THREADS : (int)inner;
    p.col_blocks = (inner + p.tile - 1) / p.tile;
    int64_t cells = outer * p.col_blocks;
    int64_t want = cells >= RED_TARGET_BLOCKS ? 1 : RED_TARGET_BLOCKS / (cells > 0 ? cells : 1);
    int64_t kpar = RED_THREADS / p.tile;
    int64_t max_s = (len + 16 * kpar - 1) / (16 * kpar);
    p.S = want < max_s ? want : max_s;
    if (p.S < 1) p.S = 1;
  }
  p.chunk = (len + p.S - 1) / p.S;
  if (p.chunk < 1) p.chunk = 1;
  p.S = (len + p.chunk - 1) / p.chunk;
  if (p.S < 1) p.S = 1;
  return p;
}

template <int RED, typename T>
static int run_reduce(int device, int64_t outer, int64_t len, int64_t inner, const void* a, int adt, void* out,
                      int odt, int acc_dt, void* scratch, size_t scratch_bytes, cudaStream_t st,
                      bool second_stage = false) {
  const int sms = dev_info(device).sms;
  if (inner == 1 && len <= 16 && outer >= 1024) {
    int grid = grid_for(device, outer, RED_THREADS, 1);
    k_reduce_rows_thread<RED, T><<<grid, RED_THREADS, 0, st>>>(a, adt, outer, len, out, odt);
    NB_LAUNCH_CHECK();
    return 0;
  }
  if (inner == 1 && len <= 2048 && outer >= 64) {
    int lpr = 32;
    while (lpr > 1 && (int64_t)(lpr / 2) * 8 >= len) lpr >>= 1;      // ~8 elements per lane
    int64_t blocks = (outer + 8 * (32 / lpr) - 1) / (8 * (32 / lpr));
    int grid = (int)(blocks < (int64_t)sms * 8 ? blocks : (int64_t)sms * 8);
    k_reduce_rows_warp<RED, T><<<grid, RED_THREADS, 0, st>>>(a, adt, outer, len, lpr, out, odt);
    NB_LAUNCH_CHECK();
    return 0;
  }
  RedPlan p = plan_reduce(outer, len, inner);
  if (second_stage) { p.S = 1; p.chunk = len; }   // partials of stage one are reduced in one go
  void* dst = out;
  int dst_dt = odt;
  if (p.S > 1) {
    size_t need = (size_t)outer * p.S * inner * 8;
    NB_REQUIRE(scratch != nullptr && scratch_bytes >= need, "nb_reduce: scratch too small (%zu < %zu)",
               scratch_bytes, need);
    dst = scratch;
    dst_dt = acc_dt == NB_BOOL ? NB_I64 : acc_dt;  // partials are stored 8 bytes wide
  }
  if (inner == 1) {
    dim3 grid((unsigned)p.S, (unsigned)(outer < 65535 ? outer : 65535));
    k_reduce_rows<RED, T><<<grid, RED_THREADS, 0, st>>>(a, adt, outer, len, p.S, p.chunk, dst, dst_dt);
  } else {
    dim3 grid((unsigned)p.col_blocks, (unsigned)p.S, (unsigned)(outer < 65535 ? outer : 65535));
    k_reduce_cols<RED, T><<<grid, RED_THREADS, 0, st>>>(a, adt, outer, len, inner, p.S, p.chunk, p.tile, dst, dst_dt);
  }
  NB_LAUNCH_CHECK();
  if (p.S > 1) {
    // second stage: [outer, S, inner] -> [outer, inner]; S <= 592 so no further split is needed
    return run_reduce<RED == NB_RNANSUM ? NB_RSUM : RED, T>(device, outer, p.S, inner, scratch, dst_dt, out, odt,
                                                           acc_dt, nullptr, 0, st, true);
  }
  return 0;
}

#define NB_RED_CASE(REDC, T) \
  case REDC: return run_reduce<REDC, T>(device, outer, len, inner, a, adt, out, odt, acc_dt, scratch, scratch_bytes, st);

static int dispatch_reduce(int device, int red, int64_t outer, int64_t len, int64_t inner, const void* a, int adt,
                           void* out, int odt, int acc_dt, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  if (acc_dt == NB_F64) {
    switch (red) {
      NB_RED_CASE(NB_RSUM, double) NB_RED_CASE(NB_RPROD, double) NB_RED_CASE(NB_RMIN, double)
      NB_RED_CASE(NB_RMAX, double) NB_RED_CASE(NB_RNANSUM, double) NB_RED_CASE(NB_RNANMIN, double)
      NB_RED_CASE(NB_RNANMAX, double)
      default: break;
    }
  } else if (acc_dt == NB_I64) {
    switch (red) {
      NB_RED_CASE(NB_RSUM, int64_t) NB_RED_CASE(NB_RPROD, int64_t) NB_RED_CASE(NB_RMIN, int64_t)
      NB_RED_CASE(NB_RMAX, int64_t)
      case NB_RNANSUM: return run_reduce<NB_RSUM, int64_t>(device, outer, len, inner, a, adt, out, odt, acc_dt, scratch, scratch_bytes, st);
      case NB_RNANMIN: return run_reduce<NB_RMIN, int64_t>(device, outer, len, inner, a, adt, out, odt, acc_dt, scratch, scratch_bytes, st);
      case NB_RNANMAX: return run_reduce<NB_RMAX, int64_t>(device, outer, len, inner, a, adt, out, odt, acc_dt, scratch, scratch_bytes, st);
      default: break;
    }
  } else if (acc_dt == NB_BOOL) {
    switch (red) {
      NB_RED_CASE(NB_RALL, bool) NB_RED_CASE(NB_RANY, bool) NB_RED_CASE(NB_RMIN, bool) NB_RED_CASE(NB_RMAX, bool)
      default: break;
    }
  }
  set_error("nb_reduce: reduction %d is not defined for accumulator dtype %d", red, acc_dt);
  return 2;
}

// ------------------------------------------------------------------------------------------------
// bop_reduce / sum_reduce
// ------------------------------------------------------------------------------------------------
template <int RED, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_pair_reduce(int64_t n, const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out) {
  using R = Red<RED, T>;
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = R::combine(R::pre(a[i]), R::pre(b[i]));
}

template <int RED, typename T>
static int run_pair(int device, int64_t n, const void* a, const void* b, void* out, cudaStream_t st) {
  if (n == 0) return 0;
  int grid = grid_for(device, n, RED_THREADS, 1);
  k_pair_reduce<RED, T><<<grid, RED_THREADS, 0, st>>>(n, (const T*)a, (const T*)b, (T*)out);
  NB_LAUNCH_CHECK();
  return 0;
}

constexpr int SUMRED_MAX = 16;
struct PtrPack { const void* p[SUMRED_MAX]; };

template <typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_sum_reduce(int k, int64_t n, PtrPack pk, T* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; i < n; i += stride) {
    T acc = ((const T*)pk.p[0])[i];
    for (int j = 1; j < k; j++) acc = Red<NB_RSUM, T>::combine(acc, ((const T*)pk.p[j])[i]);
    out[i] = acc;
  }
}

// ------------------------------------------------------------------------------------------------
// arg_op
// ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ bool is_nan_t(T) { return false; }
template <> __device__ __forceinline__ bool is_nan_t<double>(double v) { return isnan(v); }

// true if (va, ia) should replace (vb, ib) under NumPy's argmax/argmin rule:
// first NaN wins; otherwise the strictly better value; ties -> lower index.
template <bool IS_MAX, typename T>
__device__ __forceinline__ bool arg_better(T va, int64_t ia, T vb, int64_t ib) {
  if (ib < 0) return ia >= 0;
  if (ia < 0) return false;
  const bool na = is_nan_t(va), nb_ = is_nan_t(vb);
  if (na || nb_) {
    if (na && nb_) return ia < ib;
    return na;
  }
  if (IS_MAX ? (va > vb) : (va < vb)) return true;
  if (va == vb) return ia < ib;
  return false;
}

template <bool IS_MAX, typename T>
__device__ __forceinline__ void arg_block_reduce(T& v, int64_t& idx, T* smv, int64_t* smi) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    T ov = __shfl_down_sync(0xffffffffu, v, o);
    int64_t oi = __shfl_down_sync(0xffffffffu, idx, o);
    if (arg_better<IS_MAX, T>(ov, oi, v, idx)) { v = ov; idx = oi; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) { smv[warp] = v; smi[warp] = idx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); w++)
      if (arg_better<IS_MAX, T>(smv[w], smi[w], v, idx)) { v = smv[w]; idx = smi[w]; }
  }
}

template <bool IS_MAX, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_argop_stage1(int64_t n, const T* __restrict__ a, int64_t chunk, T* __restrict__ pv, int64_t* __restrict__ pi) {
  __shared__ T smv[8];
  __shared__ int64_t smi[8];
  const int64_t lo = (int64_t)blockIdx.x * chunk, hi = (lo + chunk < n) ? lo + chunk : n;
  T best = T(0);
  int64_t bi = -1;
  int64_t i = lo + threadIdx.x;
  // eight loads in flight, two independent (value, index) chains merged with the same ordering rule
  T best2 = T(0);
  int64_t bi2 = -1;
  for (; i + 7 * RED_THREADS < hi; i += 8 * RED_THREADS) {
    T v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) v[u] = a[i + u * RED_THREADS];
#pragma unroll
    for (int u = 0; u < 8; u += 2) {
      if (arg_better<IS_MAX, T>(v[u], i + u * RED_THREADS, best, bi)) { best = v[u]; bi = i + u * RED_THREADS; }
      if (arg_better<IS_MAX, T>(v[u + 1], i + (u + 1) * RED_THREADS, best2, bi2)) { best2 = v[u + 1]; bi2 = i + (u + 1) * RED_THREADS; }
    }
  }
  if (arg_better<IS_MAX, T>(best2, bi2, best, bi)) { best = best2; bi = bi2; }
  for (; i < hi; i += RED_THREADS) {
    T v = a[i];
    if (arg_better<IS_MAX, T>(v, i, best, bi)) { best = v; bi = i; }
  }
  arg_block_reduce<IS_MAX, T>(best, bi, smv, smi);
  if (threadIdx.x == 0) { pv[blockIdx.x] = best; pi[blockIdx.x] = bi; }
}

template <bool IS_MAX, typename T>
__global__ void __launch_bounds__(RED_THREADS)
k_argop_stage2(int64_t nparts, const T* __restrict__ pv, const int64_t* __restrict__ pi, int64_t start,
               const int64_t* __restrict__ other_idx, const T* __restrict__ other_val, int64_t* __restrict__ out_idx,
               T* __restrict__ out_val) {
  __shared__ T smv[8];
  __shared__ int64_t smi[8];
  T best = T(0);
  int64_t bi = -1;
  for (int64_t i = threadIdx.x; i < nparts; i += RED_THREADS)
    if (arg_better<IS_MAX, T>(pv[i], pi[i], best, bi)) { best = pv[i]; bi = pi[i]; }
  arg_block_reduce<IS_MAX, T>(best, bi, smv, smi);
  if (threadIdx.x == 0) {
    int64_t ridx = start + bi;
    T rval = best;
    if (other_val != nullptr) {
      // numpy_kernel.py:498/504: keep the earlier optimum only if strictly better (plain comparison,
      // false when either side is NaN) -> a later block wins ties.
      T ov = *other_val;
      if (IS_MAX ? (ov > best) : (ov < best)) { ridx = *other_idx; rval = ov; }
    }
    *out_idx = ridx;
    *out_val = rval;
  }
}

static int64_t argop_parts(int64_t n) {
  int64_t parts = (n + 4095) / 4096;
  if (parts > RED_TARGET_BLOCKS) parts = RED_TARGET_BLOCKS;
  if (parts < 1) parts = 1;
  return parts;
}

template <bool IS_MAX, typename T>
static int run_argop(int64_t n, const void* a, int64_t start, const int64_t* oi, const void* ov, int64_t* out_idx,
                     void* out_val, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  int64_t parts = argop_parts(n);
  size_t need = (size_t)parts * 16;
  NB_REQUIRE(scratch != nullptr && scratch_bytes >= need, "nb_argop: scratch too small (%zu < %zu)", scratch_bytes, need);
  int64_t chunk = (n + parts - 1) / parts;
  parts = (n + chunk - 1) / chunk;
  T* pv = (T*)scratch;
  int64_t* pi = (int64_t*)((char*)scratch + (size_t)argop_parts(n) * 8);
  k_argop_stage1<IS_MAX, T><<<(unsigned)parts, RED_THREADS, 0, st>>>(n, (const T*)a, chunk, pv, pi);
  NB_LAUNCH_CHECK();
  k_argop_stage2<IS_MAX, T><<<1, RED_THREADS, 0, st>>>(parts, pv, pi, start, oi, (const T*)ov, out_idx, (T*)out_val);
  NB_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// where (1-arg): ordered stream compaction
// ------------------------------------------------------------------------------------------------
constexpr int WH_STEPS = 16;                          // 32-element rows per warp per chunk
constexpr int64_t WH_CHUNK = 8 * 32 * WH_STEPS;       // 4096 elements per chunk

__device__ __forceinline__ void where_masks(const void* cond, int cdt, int64_t n, int64_t wbase, int lane,
                                            unsigned (&m)[WH_STEPS]) {
  if (cdt == NB_BOOL && wbase + 32 * WH_STEPS <= n && ((((uintptr_t)cond) & 15) == 0)) {
    // 1-byte conditions: one 16-byte load per lane covers the warp's 512 elements; the per-step ballots are
    // rebuilt from the lanes' 16-bit masks (step j = lanes 2j and 2j+1)
    const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(cond) + wbase + lane * 16);
    const unsigned w[4] = {v.x, v.y, v.z, v.w};
    unsigned mk = 0;
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
      for (int b = 0; b < 4; b++)
        if ((w[q] >> (8 * b)) & 0xffu) mk |= 1u << (4 * q + b);
#pragma unroll
    for (int j = 0; j < WH_STEPS; j++)
      m[j] = __shfl_sync(0xffffffffu, mk, 2 * j) | (__shfl_sync(0xffffffffu, mk, 2 * j + 1) << 16);
    return;
  }
  bool f[WH_STEPS];
#pragma unroll
  for (int j = 0; j < WH_STEPS; j++) {
    int64_t i = wbase + j * 32 + lane;
    f[j] = (i < n) ? load_as<bool>(cond, cdt, i) : false;
  }
#pragma unroll
  for (int j = 0; j < WH_STEPS; j++) m[j] = __ballot_sync(0xffffffffu, f[j]);
}

__global__ void __launch_bounds__(256)
k_where_count(int64_t n, const void* __restrict__ cond, int cdt, int64_t nchunks, int64_t* __restrict__ counts) {
  __shared__ int wc[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    unsigned m[WH_STEPS];
    where_masks(cond, cdt, n, c * WH_CHUNK + (int64_t)warp * 32 * WH_STEPS, lane, m);
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < WH_STEPS; j++) cnt += __popc(m[j]);
    __syncthreads();
    if (lane == 0) wc[warp] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 8; w++) t += wc[w];
      counts[c] = t;
    }
  }
}

// In-place exclusive scan of counts[0..m) by one block; writes the total.
__global__ void __launch_bounds__(1024)
k_scan_counts(int64_t m, int64_t* __restrict__ counts, int64_t* __restrict__ total) {
  __shared__ int64_t sums[1024];
  const int t = threadIdx.x;
  const int64_t per = (m + 1023) / 1024;
  const int64_t lo = t * per, hi = (lo + per < m) ? lo + per : m;
  int64_t s = 0;
  for (int64_t i = lo; i < hi; i++) s += counts[i];
  sums[t] = s;
  __syncthreads();
  // Hillis-Steele inclusive scan over 1024 partial sums
  for (int off = 1; off < 1024; off <<= 1) {
    int64_t v = (t >= off) ? sums[t - off] : 0;
    __syncthreads();
    sums[t] += v;
    __syncthreads();
  }
  int64_t run = (t == 0) ? 0 : sums[t - 1];
  for (int64_t i = lo; i < hi; i++) {
    int64_t c = counts[i];
    counts[i] = run;
    run += c;
  }
  if (t == 1023) *total = sums[1023];
}

struct WhereOut { int64_t* out[4]; int64_t shape[4]; int64_t off[4]; int ndim; };

__global__ void __launch_bounds__(256)
k_where_fill(int64_t n, const void* __restrict__ cond, int cdt, int64_t nchunks,
             const int64_t* __restrict__ offsets, WhereOut wo) {
  __shared__ int wc[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    unsigned m[WH_STEPS];
    const int64_t wbase = c * WH_CHUNK + (int64_t)warp * 32 * WH_STEPS;
    where_masks(cond, cdt, n, wbase, lane, m);
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < WH_STEPS; j++) cnt += __popc(m[j]);
    __syncthreads();
    if (lane == 0) wc[warp] = cnt;
    __syncthreads();
    int64_t pos = offsets[c];
    for (int w = 0; w < warp; w++) pos += wc[w];
#pragma unroll
    for (int j = 0; j < WH_STEPS; j++) {
      if (m[j] & (1u << lane)) {
        int64_t p = pos + __popc(m[j] & lt);
        int64_t flat = wbase + j * 32 + lane;
        if (wo.ndim == 1) {
          wo.out[0][p] = flat + wo.off[0];
        } else {
          int64_t r = flat;
          for (int ax = wo.ndim - 1; ax >= 0; ax--) {
            int64_t q = r / wo.shape[ax];
            wo.out[ax][p] = (r - q * wo.shape[ax]) + wo.off[ax];
            r = q;
          }
        }
      }
      pos += __popc(m[j]);
    }
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_reduce_scratch_bytes(int64_t outer, int64_t len, int64_t inner) {
  if (outer <= 0 || len <= 0 || inner <= 0) return 0;
  RedPlan p = plan_reduce(outer, len, inner);
  return p.S > 1 ? (size_t)outer * p.S * inner * 8 : 0;
}

int nb_reduce(int device, int red, int64_t outer, int64_t len, int64_t inner, const void* a_d, int adt,
              void* out_d, int odt, int acc_dt, void* scratch_d, size_t scratch_bytes, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(outer >= 0 && len >= 1 && inner >= 1, "nb_reduce: empty reduction axis (len=%lld)", (long long)len);
  if (outer == 0) return 0;
  return dispatch_reduce(device, red, outer, len, inner, a_d, adt, out_d, odt, acc_dt, scratch_d, scratch_bytes, st);
}

int nb_bop_reduce(int device, int red, int64_t n, const void* a_d, const void* b_d, void* out_d, int dt,
                  void* stream) {
  NB_ENTER(device);
#define NB_PAIR_CASE(REDC, T) case REDC: return run_pair<REDC, T>(device, n, a_d, b_d, out_d, st);
  if (dt == NB_F64) {
    switch (red) {
      NB_PAIR_CASE(NB_RSUM, double) NB_PAIR_CASE(NB_RPROD, double) NB_PAIR_CASE(NB_RMIN, double)
      NB_PAIR_CASE(NB_RMAX, double) NB_PAIR_CASE(NB_RNANSUM, double) NB_PAIR_CASE(NB_RNANMIN, double)
      NB_PAIR_CASE(NB_RNANMAX, double)
      default: break;
    }
  } else if (dt == NB_I64) {
    switch (red) {
      NB_PAIR_CASE(NB_RSUM, int64_t) NB_PAIR_CASE(NB_RPROD, int64_t) NB_PAIR_CASE(NB_RMIN, int64_t)
      NB_PAIR_CASE(NB_RMAX, int64_t)
      case NB_RNANSUM: return run_pair<NB_RSUM, int64_t>(device, n, a_d, b_d, out_d, st);
      case NB_RNANMIN: return run_pair<NB_RMIN, int64_t>(device, n, a_d, b_d, out_d, st);
      case NB_RNANMAX: return run_pair<NB_RMAX, int64_t>(device, n, a_d, b_d, out_d, st);
      default: break;
    }
  } else if (dt == NB_BOOL) {
    switch (red) {
      NB_PAIR_CASE(NB_RALL, bool) NB_PAIR_CASE(NB_RANY, bool) NB_PAIR_CASE(NB_RMIN, bool) NB_PAIR_CASE(NB_RMAX, bool)
      case NB_RSUM: return run_pair<NB_RANY, bool>(device, n, a_d, b_d, out_d, st);   // bool + bool
      case NB_RPROD: return run_pair<NB_RALL, bool>(device, n, a_d, b_d, out_d, st);  // bool * bool
      default: break;
    }
  }
#undef NB_PAIR_CASE
  set_error("nb_bop_reduce: reduction %d is not defined for dtype %d", red, dt);
  return 2;
}

int nb_sum_reduce(int device, int k, int64_t n, const void* const* ptrs, void* out_d, int dt, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(k >= 1, "nb_sum_reduce: need at least one array");
  NB_REQUIRE(dt == NB_F64 || dt == NB_I64, "nb_sum_reduce: dtype %d not supported", dt);
  if (n == 0) return 0;
  int grid = grid_for(device, n, RED_THREADS, 1);
  // np.add.reduce(arrs) accumulates left to right; chain launches of up to 16 operands keep that order.
  int done = 0;
  while (done < k) {
    PtrPack pk;
    int cnt = 0;
    if (done > 0) pk.p[cnt++] = out_d;
    while (cnt < SUMRED_MAX && done < k) pk.p[cnt++] = ptrs[done++];
    if (dt == NB_F64) k_sum_reduce<double><<<grid, RED_THREADS, 0, st>>>(cnt, n, pk, (double*)out_d);
    else k_sum_reduce<int64_t><<<grid, RED_THREADS, 0, st>>>(cnt, n, pk, (int64_t*)out_d);
    NB_LAUNCH_CHECK();
  }
  return 0;
}

size_t nb_argop_scratch_bytes(int64_t n) { return (size_t)argop_parts(n) * 16; }

int nb_argop(int device, int is_max, int64_t n, const void* a_d, int adt, int64_t start, const int64_t* other_idx_d,
             const void* other_val_d, int64_t* out_idx_d, void* out_val_d, void* scratch_d, size_t scratch_bytes,
             void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(n >= 1, "nb_argop: attempt to get arg%s of an empty sequence", is_max ? "max" : "min");
  NB_REQUIRE(adt == NB_F64 || adt == NB_I64, "nb_argop: dtype %d not supported", adt);
  if (adt == NB_F64) {
    if (is_max) return run_argop<true, double>(n, a_d, start, other_idx_d, other_val_d, out_idx_d, out_val_d, scratch_d, scratch_bytes, st);
    return run_argop<false, double>(n, a_d, start, other_idx_d, other_val_d, out_idx_d, out_val_d, scratch_d, scratch_bytes, st);
  }
  if (is_max) return run_argop<true, int64_t>(n, a_d, start, other_idx_d, other_val_d, out_idx_d, out_val_d, scratch_d, scratch_bytes, st);
  return run_argop<false, int64_t>(n, a_d, start, other_idx_d, other_val_d, out_idx_d, out_val_d, scratch_d, scratch_bytes, st);
}

int64_t nb_where_chunks(int64_t n) { return n <= 0 ? 1 : (n + WH_CHUNK - 1) / WH_CHUNK; }

int nb_where_count(int device, int64_t n, const void* cond_d, int cond_dt, int64_t* chunk_counts_d, int64_t* total_d,
                   void* stream) {
  NB_ENTER(device);
  int64_t nchunks = nb_where_chunks(n);
  int64_t cap = (int64_t)dev_info(device).sms * 8;
  int grid = (int)(nchunks < cap ? nchunks : cap);
  k_where_count<<<grid, 256, 0, st>>>(n, cond_d, cond_dt, nchunks, chunk_counts_d);
  NB_LAUNCH_CHECK();
  k_scan_counts<<<1, 1024, 0, st>>>(nchunks, chunk_counts_d, total_d);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_where_fill(int device, int64_t n, const void* cond_d, int cond_dt, int ndim, const int64_t* shape,
                  const int64_t* offsets, const int64_t* chunk_counts_d, int64_t* const* idx_out_d, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(ndim >= 1 && ndim <= 4, "nb_where_fill: ndim %d not in 1..4", ndim);
  WhereOut wo;
  wo.ndim = ndim;
  for (int i = 0; i < 4; i++) { wo.out[i] = nullptr; wo.shape[i] = 1; wo.off[i] = 0; }
  for (int i = 0; i < ndim; i++) { wo.out[i] = idx_out_d[i]; wo.shape[i] = shape[i]; wo.off[i] = offsets[i]; }
  int64_t nchunks = nb_where_chunks(n);
  int64_t cap = (int64_t)dev_info(device).sms * 8;
  int grid = (int)(nchunks < cap ? nchunks : cap);
  k_where_fill<<<grid, 256, 0, st>>>(n, cond_d, cond_dt, nchunks, chunk_counts_d, wo);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// nb_reduce_static: the reduction for callers that cannot allocate device memory
// ------------------------------------------------------------------------------------------------
// Some users of this library only ever hold the input and output pointers somebody else gave them:
// a cgo/JNI/ctypes host that wraps buffers owned by another runtime's memory pool, or a plug-in running
// inside a service whose HBM budget is fixed at start-up.  They have nowhere to get the split-reduction
// scratch from, so this variant keeps it inside the library: a fixed array in device memory that is part
// of the loaded module (every GPU that loads libnumsb200.so carries these NB_REDUCE_STATIC_BYTES until
// the process exits).
//
// `g_reduce_static` below is a host-side handle in host code, not an address a kernel can use; the
// real device address is looked up once per device with cudaGetSymbolAddress while that device is
// current, and cached.
static __device__ __align__(256) unsigned char g_reduce_static[NB_REDUCE_STATIC_BYTES];

namespace nb {
static void* g_static_addr[64];
static bool g_static_ready[64];
static std::mutex g_static_mu;

// Resolves the device address of g_reduce_static on `device` (which must be the current device).
static int static_scratch_for(int device, void** out) {
  NB_REQUIRE(device >= 0 && device < 64, "nb_reduce_static: device %d out of range", device);
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ready[device]) {
    void* p = nullptr;
    NB_CHECK(cudaGetSymbolAddress(&p, g_reduce_static));
    NB_REQUIRE(p != nullptr, "nb_reduce_static: cudaGetSymbolAddress returned a null address");
    g_static_addr[device] = p;
    g_static_ready[device] = true;
  }
  *out = g_static_addr[device];
  return 0;
}
}  // namespace nb

extern "C" {

// Same contract and same kernels as nb_reduce, minus the scratch arguments.
//  * The first call on each device performs the address look-up described above.  Make that first call
//    before any stream capture (it is an ordinary synchronous runtime query, not stream work).
//  * A reduction whose split needs more than NB_REDUCE_STATIC_BYTES (1 MiB) is refused with an error;
//    use nb_reduce with caller-provided scratch for those (nb_reduce_scratch_bytes tells the size).
//  * All calls on a device share the one array and it keeps whatever the previous call left there.
//    Nothing in it has to start at zero: stage one overwrites every partial that stage two reads.
//    Because it is shared, two calls whose kernels can overlap on the GPU (different streams, or
//    different host threads on different streams) race on it; issue them on one stream or serialise them.
//  * No cudaMalloc / cudaFree / allocating library routine is called on this path.
int nb_reduce_static(int device, int red, int64_t outer, int64_t len, int64_t inner, const void* a_d, int adt,
                     void* out_d, int odt, int acc_dt, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(outer >= 0 && len >= 1 && inner >= 1, "nb_reduce_static: empty reduction axis (len=%lld)", (long long)len);
  if (outer == 0) return 0;
  const size_t need = nb_reduce_scratch_bytes(outer, len, inner);
  NB_REQUIRE(need <= (size_t)NB_REDUCE_STATIC_BYTES,
             "nb_reduce_static: this reduction needs %zu bytes of scratch but the built-in buffer holds %u; "
             "call nb_reduce with your own scratch instead", need, (unsigned)NB_REDUCE_STATIC_BYTES);
  void* scratch = nullptr;
  int rc = static_scratch_for(device, &scratch);
  if (rc) return rc;
  return dispatch_reduce(device, red, outer, len, inner, a_d, adt, out_d, odt, acc_dt, scratch,
                         NB_REDUCE_STATIC_BYTES, st);
}

}  // extern "C"
