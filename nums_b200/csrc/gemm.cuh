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

// warp-specialised TMA kernel (matmul_ws.cu)
struct WsPlan {
  bool ok;
  int tiles_m, tiles_n, ntiles, splits, sym;
  int64_t kchunk;
  size_t scratch;        // != 0: partial tiles, to be summed by k_gemm_reduce<128,128>
};
WsPlan ws_plan(int device, int64_t m, int64_t n, int64_t k, bool sym, bool k_rows);
bool gemm_ws_aligned(const double* a, int64_t lda, const double* b, int64_t ldb, const double* w, const double* c,
                     int64_t ldc);
int gemm_ws_f64(const WsPlan& plan, int64_t m, int64_t n, int64_t k, const double* a, int64_t lda, bool a_mc,
                const double* b, int64_t ldb, bool b_nc, const double* w, double* c, int64_t ldc, double* part,
                cudaStream_t st);

}  // namespace nb
