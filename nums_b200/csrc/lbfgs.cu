// The p-sized algebra of L-BFGS (nums/models/lbfgs.py) as three single-CTA kernels on a device-resident workspace.
//
// The reference runs the two-loop recursion (lbfgs.py:86-100), the memory update (:55-62) and the Armijo test
// (:37-52) as block-array operations: a dot product, a scale and an axpy per memory slot, each its own kernel call, and
// a host comparison per trial.  With p <= a few thousand every one of those calls is pure launch latency — ~120 launches
// for the recursion at m = 10 against one pass over X.  Here the memory (S, Y, rho) is a ring in HBM and
//   nb_lbfgs_direction  p = -H g by the two-loop recursion, and slope = g.p           (one launch)
//   nb_lbfgs_trial      theta + alpha p                                               (one launch)
//   nb_lbfgs_judge      Armijo test of a trial; on acceptance the pair (s, y) = (theta' - theta, g' - g) enters the
//                       ring with rho = 1 / (s.y + 1e-30) and gamma = s.y / (y.y + 1e-30), and max|g'| is formed
//                       for the convergence test; the verdict is two doubles for ONE read-back  (one launch)
// Dot products are summed in a fixed order (thread-strided partial sums, then a tree), so results do not depend on timing.
#include "common.cuh"

namespace nb {

constexpr int LB_THREADS = 512;

struct LbfgsHeader {          // start of the workspace
  int32_t count;              // pairs stored (<= m)
  int32_t head;               // slot of the newest pair
  int32_t pad[2];
  double gamma;               // scaling of the initial inverse Hessian (1 before the first pair)
};
// layout: header (64 B) | rho[m] | alpha[m] | q[p] | S[m][p] | Y[m][p]
__host__ __device__ inline size_t lb_off_rho() { return 64; }
__host__ __device__ inline size_t lb_off_alpha(int m) { return lb_off_rho() + (size_t)m * 8; }
__host__ __device__ inline size_t lb_off_q(int m) { return lb_off_alpha(m) + (size_t)m * 8; }
__host__ __device__ inline size_t lb_off_S(int m, int64_t p) { return ((lb_off_q(m) + (size_t)p * 8) + 255) & ~(size_t)255; }
__host__ __device__ inline size_t lb_off_Y(int m, int64_t p) { return lb_off_S(m, p) + (size_t)m * p * 8; }

__device__ __forceinline__ double lb_block_sum(double v, double* sm) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();                       // sm may still be read from the previous reduction
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  double s = 0.0;
#pragma unroll
  for (int w = 0; w < LB_THREADS / 32; w++) s += sm[w];
  return s;
}
__device__ __forceinline__ double lb_dot(const double* a, const double* b, int64_t p, double* sm) {
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) acc = fma(a[i], b[i], acc);
  return lb_block_sum(acc, sm);
}

__global__ void __launch_bounds__(LB_THREADS)
k_lbfgs_direction(int64_t p, int m, char* ws, const double* __restrict__ g, double* __restrict__ dir, double* __restrict__ slope) {
  __shared__ double sm[LB_THREADS / 32];
  const LbfgsHeader* h = (const LbfgsHeader*)ws;
  const double* rho = (const double*)(ws + lb_off_rho());
  double* alpha = (double*)(ws + lb_off_alpha(m));
  double* q = (double*)(ws + lb_off_q(m));
  const double* S = (const double*)(ws + lb_off_S(m, p));
  const double* Y = (const double*)(ws + lb_off_Y(m, p));
  const int count = h->count, head = h->head;
  for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) q[i] = g[i];
  __syncthreads();
  // newest to oldest: alpha_i = rho_i s_i.q ; q -= alpha_i y_i            (lbfgs.py:89-95)
  for (int k = 0; k < count; k++) {
    const int slot = (head - k + m) % m;
    const double a = rho[slot] * lb_dot(S + (size_t)slot * p, q, p, sm);
    if (threadIdx.x == 0) alpha[slot] = a;
    for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) q[i] = fma(-a, Y[(size_t)slot * p + i], q[i]);
    __syncthreads();
  }
  // r = (gamma I) q                                                       (:78-84, :96)
  const double gamma = count > 0 ? h->gamma : 1.0;
  for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) q[i] *= gamma;
  __syncthreads();
  // oldest to newest: beta = rho_i y_i.r ; r += s_i (alpha_i - beta)      (:97-99)
  for (int k = count - 1; k >= 0; k--) {
    const int slot = (head - k + m) % m;
    const double b = rho[slot] * lb_dot(Y + (size_t)slot * p, q, p, sm);
    const double c = alpha[slot] - b;
    for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) q[i] = fma(S[(size_t)slot * p + i], c, q[i]);
    __syncthreads();
  }
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) { const double v = -q[i]; dir[i] = v; acc = fma(g[i], v, acc); }
  const double s = lb_block_sum(acc, sm);
  if (threadIdx.x == 0) slope[0] = s;
}

__global__ void __launch_bounds__(256)
k_lbfgs_trial(int64_t p, const double* __restrict__ theta, const double* __restrict__ dir, double alpha, double* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < p; i += (int64_t)gridDim.x * 256) out[i] = fma(alpha, dir[i], theta[i]);
}

// verdict[0] = 1 accepted / 0 rejected; verdict[1] = max|g_trial| (when accepted and remembered)
__global__ void __launch_bounds__(LB_THREADS)
k_lbfgs_judge(int64_t p, int m, char* ws, const double* __restrict__ theta, const double* __restrict__ g,
              const double* __restrict__ f_val, const double* __restrict__ theta_t, const double* __restrict__ g_t,
              const double* __restrict__ f_t, const double* __restrict__ slope, double alpha, double c, int force,
              int remember, double* __restrict__ verdict) {
  __shared__ double sm[LB_THREADS / 32];
  const double fn = f_t[0];
  // lbfgs.py:45-48: shrink while f is NaN or above the Armijo line
  const bool reject = (fn != fn) || (fn > f_val[0] + c * alpha * slope[0]);
  if (reject && !force) { if (threadIdx.x == 0) { verdict[0] = 0.0; verdict[1] = 0.0; } return; }
  double gm = 0.0;
  if (remember) {
    LbfgsHeader* h = (LbfgsHeader*)ws;
    double* rho = (double*)(ws + lb_off_rho());
    double* S = (double*)(ws + lb_off_S(m, p));
    double* Y = (double*)(ws + lb_off_Y(m, p));
    const int slot = h->count == 0 ? 0 : (h->head + 1) % m;
    double sy = 0.0, yy = 0.0;
    bool nan_seen = false;
    for (int64_t i = threadIdx.x; i < p; i += LB_THREADS) {
      const double s = theta_t[i] - theta[i], y = g_t[i] - g[i];
      S[(size_t)slot * p + i] = s; Y[(size_t)slot * p + i] = y;
      sy = fma(s, y, sy); yy = fma(y, y, yy);
      const double a = fabs(g_t[i]);
      if (a != a) nan_seen = true;
      gm = fmax(gm, a);
    }
    sy = lb_block_sum(sy, sm);
    yy = lb_block_sum(yy, sm);
    // max|g| with NumPy's rule (a NaN propagates)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_xor_sync(0xffffffffu, gm, o));
    const int any_nan = __syncthreads_or(nan_seen ? 1 : 0);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = gm;
    __syncthreads();
    if (threadIdx.x == 0) {
      double mx = 0.0;
      for (int w = 0; w < LB_THREADS / 32; w++) mx = fmax(mx, sm[w]);
      gm = any_nan ? __longlong_as_double(0x7ff8000000000000ll) : mx;
      rho[slot] = 1.0 / (sy + 1e-30);                    // lbfgs.py:60-62
      h->gamma = sy / (yy + 1e-30);
      h->head = slot;
      if (h->count < m) h->count += 1;
    }
  }
  if (threadIdx.x == 0) { verdict[0] = 1.0; verdict[1] = gm; }
}

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_lbfgs_ws_bytes(int64_t p, int m) { return lb_off_Y(m, p) + (size_t)m * p * 8; }

int nb_lbfgs_reset(int device, int64_t p, int m, void* ws_d, size_t ws_bytes, void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(p >= 1 && m >= 1 && ws_d != nullptr && ws_bytes >= nb_lbfgs_ws_bytes(p, m), "nb_lbfgs_reset: workspace too small");
  NB_CHECK(cudaMemsetAsync(ws_d, 0, lb_off_q(m), st));
  return 0;
}

int nb_lbfgs_direction(int device, int64_t p, int m, void* ws_d, const double* g_d, double* dir_d, double* slope_d,
                       void* stream) {
  NB_ENTER(device);
  NB_REQUIRE(p >= 1 && m >= 1, "nb_lbfgs_direction: bad p=%lld m=%d", (long long)p, m);
  k_lbfgs_direction<<<1, LB_THREADS, 0, st>>>(p, m, (char*)ws_d, g_d, dir_d, slope_d);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_lbfgs_trial(int device, int64_t p, const double* theta_d, const double* dir_d, double alpha, double* out_d,
                   void* stream) {
  NB_ENTER(device);
  const int64_t blocks = (p + 255) / 256;
  k_lbfgs_trial<<<(unsigned)(blocks < 64 ? blocks : 64), 256, 0, st>>>(p, theta_d, dir_d, alpha, out_d);
  NB_LAUNCH_CHECK();
  return 0;
}

int nb_lbfgs_judge(int device, int64_t p, int m, void* ws_d, const double* theta_d, const double* g_d,
                   const double* f_d, const double* theta_t_d, const double* g_t_d, const double* f_t_d,
                   const double* slope_d, double alpha, double c, int force, int remember, double* verdict_d,
                   void* stream) {
  NB_ENTER(device);
  k_lbfgs_judge<<<1, LB_THREADS, 0, st>>>(p, m, (char*)ws_d, theta_d, g_d, f_d, theta_t_d, g_t_d, f_t_d, slope_d, alpha, c,
                                          force, remember, verdict_d);
  NB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
