// Internal interface of the FP64 contraction kernels (matmul.cu), shared with glm.cu.
#pragma once
#include "common.cuh"

namespace nb {

// C[m,n] = sum_k opA[m,k] * w[k] * opB[k,n]   (w may be null).
// a_rs/a_cs: strides of opA along m / k; b_rs/b_cs: strides of opB along k / n.
// Enqueues on `st`; scratch as reported by gemm_scratch_bytes.
int gemm_f64(int device, int64_t m, int64_t n, int64_t k, const double* a, int64_t a_rs, int64_t a_cs,
             const double* b, int64_t b_rs, int64_t b_cs, const double* w, double* c, int64_t ldc,
             void* scratch, size_t scratch_bytes, cudaStream_t st);

size_t gemm_scratch_bytes(int device, int64_t m, int64_t n, int64_t k);

}  // namespace nb
