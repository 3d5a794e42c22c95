// Link functions of the GLM families, shared by glm.cu and glm_fused.cu.
#pragma once
#include "common.cuh"

namespace nb {

struct LinkOut { double resid, wgt, obj; };

// the same sum by a whole warp (lanes stride over the partials, fixed xor tree): for the short statistics vectors,
// where one thread per statistic would leave a single CTA walking all the partials
__device__ __forceinline__ double sum_partials_warp(const double* __restrict__ p, size_t stride, int count) {
  const int lane = threadIdx.x & 31;
  double a0 = 0.0, a1 = 0.0;
  int c = lane;
  for (; c + 32 < count; c += 64) { a0 += p[(size_t)c * stride]; a1 += p[(size_t)(c + 32) * stride]; }
  if (c < count) a0 += p[(size_t)c * stride];
  double s = a0 + a1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}

// sum_{c < count} p[c * stride] with eight loads in flight and a fixed association (the partials of a pass are summed
// by one thread per statistic: a serial chain of dependent loads made the finish kernels cost 40-90 us)
__device__ __forceinline__ double sum_partials(const double* __restrict__ p, size_t stride, int count) {
  double acc[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  int c = 0;
  for (; c + 8 <= count; c += 8) {
    double v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) v[u] = p[(size_t)(c + u) * stride];
#pragma unroll
    for (int u = 0; u < 8; u++) acc[u] += v[u];
  }
  for (; c < count; c++) acc[c & 7] += p[(size_t)c * stride];
  return ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
}

template <int FAMILY>
__device__ __forceinline__ LinkOut glm_link(double eta, double y) {
  LinkOut o;
  if constexpr (FAMILY == NB_GLM_LOGISTIC || FAMILY == NB_GLM_LOGISTIC_IRLS) {
    // glms.py:457-458 link_inv, :470 objective, :485 gradient residual, :495 hessian weights (:692 for irls)
    double mu = 1.0 / (1.0 + exp(-eta));
    o.resid = mu - y;
    o.wgt = mu * (1.0 - mu) + (FAMILY == NB_GLM_LOGISTIC_IRLS ? 1e-16 : 0.0);
    o.obj = -(y * log(mu) + (1.0 - y) * log(1.0 - mu));
  } else if constexpr (FAMILY == NB_GLM_POISSON) {
    // glms.py:516-517, :531, :544, :548
    double mu = exp(eta);
    o.resid = mu - y;
    o.wgt = mu;
    o.obj = mu - y * eta;
  } else {
    // glms.py:283-284, :296, :311, :318
    o.resid = eta - y;
    o.wgt = 1.0;
    double e = y - eta;
    o.obj = e * e;
  }
  return o;
}

template <int FAMILY>
__device__ __forceinline__ double glm_mu(double eta) {
  if constexpr (FAMILY == NB_GLM_LOGISTIC || FAMILY == NB_GLM_LOGISTIC_IRLS) return 1.0 / (1.0 + exp(-eta));
  else if constexpr (FAMILY == NB_GLM_POISSON) return exp(eta);
  else return eta;
}

__device__ __forceinline__ double load_y(const void* y, int ydt, int64_t i) {
  return ydt == NB_F64 ? ((const double*)y)[i] : (double)((const int64_t*)y)[i];
}

// Newton needs only the residual and the weight: skips the objective's two logarithms.
__device__ __forceinline__ LinkOut glm_link_noobj_rt(int family, double eta, double y) {
  LinkOut o;
  o.obj = 0.0;
  if (family == NB_GLM_LOGISTIC || family == NB_GLM_LOGISTIC_IRLS) {
    double mu = 1.0 / (1.0 + exp(-eta));
    o.resid = mu - y;
    o.wgt = mu * (1.0 - mu) + (family == NB_GLM_LOGISTIC_IRLS ? 1e-16 : 0.0);
  } else if (family == NB_GLM_POISSON) {
    double mu = exp(eta);
    o.resid = mu - y;
    o.wgt = mu;
  } else {
    o.resid = eta - y;
    o.wgt = 1.0;
  }
  return o;
}

__device__ __forceinline__ LinkOut glm_link_rt(int family, double eta, double y) {
  if (family == NB_GLM_LOGISTIC) return glm_link<NB_GLM_LOGISTIC>(eta, y);
  if (family == NB_GLM_LOGISTIC_IRLS) {       // irls(): s = mu * (1 - mu) + 1e-16 (glms.py:692)
    LinkOut o = glm_link<NB_GLM_LOGISTIC>(eta, y);
    o.wgt += 1e-16;
    return o;
  }
  if (family == NB_GLM_POISSON) return glm_link<NB_GLM_POISSON>(eta, y);
  return glm_link<NB_GLM_LINEAR>(eta, y);
}


}  // namespace nb
