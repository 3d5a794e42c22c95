/*
 * nums_b200.h — C ABI of libnumsb200.so, the sm_100a block-compute library behind
 * the NumS `Backend` + kernel-module plug-in (nums_b200.cuda_backend / cuda_kernel).
 *
 * Conventions
 *   - Every pointer named *_d / documented "device" is a raw device pointer on `device`
 *     (e.g. torch.Tensor.data_ptr()).  No torch / C++ types cross this boundary.
 *   - Every entry point selects `device`, enqueues its work on `stream`
 *     (a cudaStream_t passed as void*; NULL = legacy default stream) and returns
 *     without synchronising unless stated.  Return value: 0 = ok, non-zero = error;
 *     nb_last_error() returns the thread-local message.
 *   - Arrays are C-contiguous unless strides are passed.  Strides are in ELEMENTS.
 *   - Kernels that need scratch take (scratch, scratch_bytes) from the caller; the
 *     matching *_scratch_bytes() query tells how much.  The library never allocates
 *     device memory.
 *
 * Each declaration cites the reference (nums-project/nums v0.2.8) interface it replaces.
 */
#ifndef NUMS_B200_H
#define NUMS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* element types of a block (nums/core/kernel/numpy_kernel.py works on np.ndarray of these) */
typedef enum {
  NB_F64 = 0,
  NB_I64 = 1,
  NB_BOOL = 2, /* 1 byte, 0/1 */
  NB_F32 = 3,  /* astype only */
  NB_I32 = 4   /* astype only */
} nb_dtype;

/* binary ufuncs reachable through KernelCls.bop (numpy_kernel.py:414-431, settings.py:50-63) */
typedef enum {
  NB_ADD = 0, NB_SUB, NB_MUL, NB_TRUEDIV, NB_FLOORDIV, NB_MOD, NB_FMOD, NB_POW, NB_FLOAT_POW,
  NB_MAXIMUM, NB_MINIMUM, NB_FMAX, NB_FMIN,
  NB_LT, NB_LE, NB_GT, NB_GE, NB_EQ, NB_NE,
  NB_LOGICAL_AND, NB_LOGICAL_OR, NB_LOGICAL_XOR,
  NB_BIT_AND, NB_BIT_OR, NB_BIT_XOR, NB_LSHIFT, NB_RSHIFT,
  NB_ARCTAN2, NB_HYPOT, NB_COPYSIGN, NB_NEXTAFTER, NB_LDEXP, NB_HEAVISIDE,
  NB_LOGADDEXP, NB_LOGADDEXP2, NB_GCD, NB_LCM, NB_XLOGY,
  NB_BOP_COUNT
} nb_bop_code;

/* unary ufuncs reachable through KernelCls.map_uop (numpy_kernel.py:316-318) */
typedef enum {
  NB_ABS = 0, NB_NEGATIVE, NB_POSITIVE, NB_SIGN, NB_SQUARE, NB_RECIPROCAL, NB_SQRT, NB_CBRT,
  NB_EXP, NB_EXP2, NB_EXPM1, NB_LOG, NB_LOG2, NB_LOG10, NB_LOG1P,
  NB_SIN, NB_COS, NB_TAN, NB_ARCSIN, NB_ARCCOS, NB_ARCTAN,
  NB_SINH, NB_COSH, NB_TANH, NB_ARCSINH, NB_ARCCOSH, NB_ARCTANH,
  NB_CEIL, NB_FLOOR, NB_TRUNC, NB_RINT, NB_DEG2RAD, NB_RAD2DEG,
  NB_ISNAN, NB_ISINF, NB_ISFINITE, NB_SIGNBIT, NB_LOGICAL_NOT, NB_INVERT, NB_SPACING,
  NB_UOP_COUNT
} nb_uop_code;

/* reductions reachable through KernelCls.reduce_axis / bop_reduce (numpy_kernel.py:294-313,433-453) */
typedef enum {
  NB_RSUM = 0, NB_RPROD, NB_RMIN, NB_RMAX, NB_RALL, NB_RANY, NB_RNANSUM, NB_RNANMIN, NB_RNANMAX,
  NB_RED_COUNT
} nb_red_code;

/* GLM families of nums/models/glms.py */
typedef enum {
  NB_GLM_LINEAR = 0,   /* LinearRegressionBase  glms.py:275-327 */
  NB_GLM_LOGISTIC = 1, /* LogisticRegression    glms.py:430-512 */
  NB_GLM_POISSON = 2,  /* PoissonRegression     glms.py:515-556 */
  NB_GLM_LOGISTIC_IRLS = 3 /* logistic with the IRLS weights mu(1-mu) + 1e-16, glms.py:679-700 */
} nb_glm_family;

/* ---- library ------------------------------------------------------------------------- */

/* Thread-local message of the last failing call ("" if none). */
const char* nb_last_error(void);
/* ABI version (bumped on any signature change). */
int nb_version(void);
/* Number of visible CUDA devices; <0 on error.  Replaces Backend.devices() discovery
 * (nums/core/backends/serial.py:44, ray.py:55-125). */
int nb_device_count(void);
/* Fills sm_count / compute capability of `device`. */
int nb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor);

/* ---- elementwise: KernelCls.bop (non-tensordot branch), numpy_kernel.py:426-431 -------- */

/* out[idx] = op(a[idx], b[idx]) with NumPy broadcasting expressed as strides.
 * ndim <= 4; shape[] is the broadcast output shape; sa[]/sb[] are element strides of the
 * operands in that shape (0 on broadcast axes; deferred transposes a1_T/a2_T are just swapped
 * strides).  `cdt` is the type the ufunc loop computes in (NumPy's common type: NB_F64, NB_I64
 * or NB_BOOL); `odt` the dtype NumPy returns (decided on the host exactly like
 * nums/core/array/utils.py:40-51).  Integer semantics follow NumPy, not C (floor division,
 * sign of mod, x//0 == 0, shifts >= 64, wrapping power). */
int nb_bop(int device, int op, int ndim, const int64_t* shape,
           const void* a_d, int adt, const int64_t* sa,
           const void* b_d, int bdt, const int64_t* sb,
           void* out_d, int odt, int cdt, void* stream);

/* ---- elementwise: KernelCls.map_uop, numpy_kernel.py:316-318 --------------------------- */

/* out[i] = op(a[i]) over n contiguous elements; cdt/odt as in nb_bop. */
int nb_uop(int device, int op, int64_t n, const void* a_d, int adt,
           void* out_d, int odt, int cdt, void* stream);

/* ---- KernelCls.where, 3-argument form, numpy_kernel.py:331-334 ------------------------- */
int nb_where3(int device, int64_t n, const void* cond_d, int cond_dt,
              const void* x_d, const void* y_d, void* out_d, int xdt, void* stream);

/* ---- KernelCls.where, 1-argument form (np.where(arr) + block offsets), :320-330 -------- */

/* Pass 1: counts non-zeros.  chunk_counts_d: int64[nb_where_chunks(n)] scratch,
 * total_d: int64[1].  The host reads *total_d (the one sync the reference also has,
 * application.py:634) to size the outputs. */
int64_t nb_where_chunks(int64_t n);
int nb_where_count(int device, int64_t n, const void* cond_d, int cond_dt,
                   int64_t* chunk_counts_d, int64_t* total_d, void* stream);
/* Pass 2: writes, in C order, for every non-zero element its index along each of `ndim`
 * axes plus offsets[axis] into idx_out_d[axis] (int64[total] each; idx_out_d itself, shape and
 * offsets are HOST arrays of length ndim <= 4).
 * chunk_counts_d must be the array produced by nb_where_count for the same input. */
int nb_where_fill(int device, int64_t n, const void* cond_d, int cond_dt,
                  int ndim, const int64_t* shape, const int64_t* offsets,
                  const int64_t* chunk_counts_d, int64_t* const* idx_out_d, void* stream);

/* ---- KernelCls.astype :339, new_block :93, create_block/update_block :116-155,
 *      transpose :343, reshape :510 (views are host-side; these are the copies) ---------- */
int nb_astype(int device, int64_t n, const void* a_d, int adt, void* out_d, int odt, void* stream);
int nb_fill(int device, int64_t n, void* out_d, int odt, double fvalue, int64_t ivalue, void* stream);
/* eye(rows, cols) block with 1 on the diagonal (new_block "eye", :97-99) */
int nb_eye(int device, int64_t rows, int64_t cols, void* out_d, int odt, void* stream);
/* Strided N-D (<=4) copy of itemsize-byte elements: dst[idx*dstrides] = src[idx*sstrides]. */
int nb_copy_strided(int device, int ndim, const int64_t* shape, int itemsize,
                    const void* src_d, const int64_t* sstrides,
                    void* dst_d, const int64_t* dstrides, void* stream);

/* ---- KernelCls.random_block for rfunc "random" (numpy_kernel.py:103-110, block_rng :31-32) ----------- */

/* n uniform doubles in [0,1), bit-identical to Generator(PCG64(seed).jumped(j)).random(n): the host passes
 * that bit generator's 128-bit state and increment as {hi, lo} pairs (seeding/jumping stay in NumPy);
 * the device advances each thread to its position with the O(log k) LCG jump. */
int nb_pcg64_random(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                    double* out_d, void* stream);

/* rfunc "normal" / "standard_normal" (random.py:158-195 -> numpy_kernel.py:103-110): n doubles in the order
 * Generator(PCG64(seed).jumped(j)).normal(loc, scale, n) returns them.  NumPy's ziggurat consumes a
 * data-dependent number of words per value; the device resolves that with a prefix scan of per-chunk state
 * maps (csrc/rng_scan.cu), so element i is the host generator's i-th value: identical bits wherever the
 * value is a table product (99.97 % of draws), within 1 ulp of libm's log1p on the tail draws.
 * standard != 0 returns the values unscaled (standard_normal).  The call reads back one 8-byte count and
 * therefore synchronises `stream`; it fails if the scanned words yielded fewer than n values. */
int64_t nb_pcg64_normal_scratch_bytes(int64_t n);
int nb_pcg64_normal(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                    double loc, double scale, int standard, double* out_d,
                    void* scratch_d, int64_t scratch_bytes, void* stream);

/* rfunc "integers" with dtype int64 (random.py:158-195): out[i] = low + (i-th bounded draw over [0, rng]),
 * rng = high - low - 1 (or high - low with endpoint=True), bit-identical to Generator.integers: Lemire's
 * multiply-shift on 32-bit half-words when rng < 2^32 (low half first), on 64-bit words otherwise, with its
 * rejections placed by the same prefix scan.  Synchronises `stream` when the range needs rejection. */
int64_t nb_pcg64_integers_scratch_bytes(int64_t n, uint64_t rng);
int nb_pcg64_integers(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                      int64_t low, uint64_t rng, int64_t* out_d,
                      void* scratch_d, int64_t scratch_bytes, void* stream);

/* KernelCls.permutation (numpy_kernel.py:112-114): out = Generator(PCG64(seed).jumped(j)).permutation(n), bit-identical,
 * n < 2^31.  NumPy's Fisher-Yates with masked-rejection draws is restated as parallel passes (csrc/perm.cu): the
 * sample each draw uses is found phase by phase with bracketed prefix sums, the composition of the n - 1 swaps by
 * per-target step lists and pointer jumping.  Synchronises `stream` several times per phase (convergence tests). */
int64_t nb_pcg64_permutation_scratch_bytes(int64_t n);
int nb_pcg64_permutation(int device, int64_t n, const uint64_t* state_hi_lo, const uint64_t* inc_hi_lo,
                         int64_t* out_d, void* scratch_d, int64_t scratch_bytes, void* stream);

/* ---- CSV ingest: the text of one read_csv_block call (nums/core/filesystem.py:227-283) parsed in HBM ------------
 * The host uploads the lines of the block's byte range as raw text ending in '\n'.  nb_csv_count_rows counts
 * them (one 8-byte read-back: the host sizes the output); nb_csv_parse trims "\r\n" from both ends of every
 * line, splits it at `delimiter` into exactly `cols` fields and converts each like Python's float() (odt = NB_F64:
 * correctly rounded, bit-identical to float(text); whitespace, sign, underscores, inf / nan) or int()
 * (odt = NB_I64).  error_host = {failing line + 1 or 0, code, column}: 1 bad field, 2 wrong field count,
 * 3 more than 19 significant digits with a rounding the device cannot decide, 4 hexadecimal float, 5 out of range.
 * Both calls use the same scratch (nb_csv_scratch_bytes(nbytes, rows)) and synchronise `stream`. */
int64_t nb_csv_scratch_bytes(int64_t nbytes, int64_t max_rows);
int nb_csv_count_rows(int device, const unsigned char* text_d, int64_t nbytes, void* scratch_d, int64_t scratch_bytes,
                      int64_t* rows_host, void* stream);
int nb_csv_parse(int device, const unsigned char* text_d, int64_t nbytes, int64_t rows, int cols, int delimiter,
                 int odt, void* out_d, void* scratch_d, int64_t scratch_bytes, int64_t* error_host, void* stream);

/* ---- reductions: KernelCls.reduce_axis :294-313, sum_reduce :291, bop_reduce :433-453 -- */

/* Views the input as [outer, len, inner] (C order) and reduces the middle axis:
 * out[o, i] = red_{k} a[o, k, i].  axis=None is outer=inner=1.  `acc_dt` is the accumulator
 * type (NB_F64 for the reference's matmul(ones, arr) path on integer input, quirk Q1).
 * Deterministic: fixed two-stage order, no atomics. */
size_t nb_reduce_scratch_bytes(int64_t outer, int64_t len, int64_t inner);
int nb_reduce(int device, int red, int64_t outer, int64_t len, int64_t inner,
              const void* a_d, int adt, void* out_d, int odt, int acc_dt,
              void* scratch_d, size_t scratch_bytes, void* stream);
/* nb_reduce for callers that have no way to allocate device memory: a foreign-runtime host (cgo, JNI,
 * another process's memory pool) that only holds the input and output buffers it was lent, or a plug-in
 * inside a service whose HBM budget was fixed at start-up.  No counterpart in the reference: NumPy
 * allocates its own temporaries.  The split scratch lives in a built-in device array of
 * NB_REDUCE_STATIC_BYTES (1 MiB) that every GPU loading the library holds until exit; reductions whose
 * split needs more are refused with an error (use nb_reduce).  The first call on a device resolves that
 * array's address and should come before any stream capture.  All calls share the array: it keeps the
 * previous call's partials (every partial that is read is overwritten first), and calls that may
 * overlap on the GPU race. */
#define NB_REDUCE_STATIC_BYTES (1u << 20)
int nb_reduce_static(int device, int red, int64_t outer, int64_t len, int64_t inner,
                     const void* a_d, int adt, void* out_d, int odt, int acc_dt, void* stream);
/* out = a (red) b elementwise over n elements of one dtype (pairwise tree combine). */
int nb_bop_reduce(int device, int red, int64_t n, const void* a_d, const void* b_d,
                  void* out_d, int dt, void* stream);
/* out = sum of k equal-shape arrays, accumulated left to right like np.add.reduce(arrs).
 * ptrs: HOST array of k device pointers. */
int nb_sum_reduce(int device, int k, int64_t n, const void* const* ptrs, void* out_d, int dt,
                  void* stream);

/* ---- KernelCls.arg_op, numpy_kernel.py:492-508 ----------------------------------------- */

/* (start + argmax/argmin(a), optimum) of a flat array with NumPy's NaN rule (first NaN wins),
 * then the reference's chain rule: the earlier pair (other_idx_d, other_val_d; NULL = none) is
 * kept only if STRICTLY better, so a later block wins ties (quirk Q2). */
size_t nb_argop_scratch_bytes(int64_t n);
int nb_argop(int device, int is_max, int64_t n, const void* a_d, int adt, int64_t start,
             const int64_t* other_idx_d, const void* other_val_d,
             int64_t* out_idx_d, void* out_val_d,
             void* scratch_d, size_t scratch_bytes, void* stream);

/* ---- KernelCls.bop "tensordot" branch, numpy_kernel.py:419-425 (np.matmul, <=2-D) ------ */

/* C[m,n] = op(A)[m,k] . op(B)[k,n], float64, FP64 tensor-core (DMMA) tiles.
 * a_rs/a_cs: element strides of op(A) along m and k (one of them must be 1);
 * b_rs/b_cs: strides of op(B) along k and n (one of them must be 1).
 * n == 1 / m == 1 cover mat.vec, vec.mat and dot.  `w_d` (nullable, length k) scales the
 * contraction index: C = op(A) . diag(w) . op(B)  — the Hessian X^T.diag(s).X of
 * glms.py:491-499 without materialising s*X.  Deterministic split-K through `scratch`. */
size_t nb_matmul_scratch_bytes(int device, int64_t m, int64_t n, int64_t k);
int nb_matmul_f64(int device, int64_t m, int64_t n, int64_t k,
                  const double* a_d, int64_t a_rs, int64_t a_cs,
                  const double* b_d, int64_t b_rs, int64_t b_cs,
                  const double* w_d, double* c_d, int64_t ldc,
                  void* scratch_d, size_t scratch_bytes, void* stream);
/* Same contraction for int64 blocks (NumPy uses non-BLAS loops; wrapping arithmetic). */
int nb_matmul_i64(int device, int64_t m, int64_t n, int64_t k,
                  const int64_t* a_d, int64_t a_rs, int64_t a_cs,
                  const int64_t* b_d, int64_t b_rs, int64_t b_cs,
                  int64_t* c_d, int64_t ldc, void* stream);

/* ---- KernelCls.inv, numpy_kernel.py:471-472 (np.linalg.inv) ----------------------------- */

/* out = inverse of the p x p row-major matrix a (partial pivoting).  info_d: int32[1] set to
 * 0, or k>0 if pivot k was exactly zero (singular, NumPy raises LinAlgError).
 * scratch: nb_inv_scratch_bytes(p). */
size_t nb_inv_scratch_bytes(int64_t p);
int nb_inv_f64(int device, int64_t p, const double* a_d, double* out_d, int32_t* info_d,
               void* scratch_d, size_t scratch_bytes, void* stream);

/* ---- KernelCls.cholesky / qr / svd (numpy_kernel.py:455-469): the p x p ends of the tall-skinny routines of
 *      nums/core/linalg.py:55-243 (indirect_tsr, direct_tsqr, svd, pca); p <= 1024 --------------------------- */

/* l = lower Cholesky factor of the symmetric positive definite a (np.linalg.cholesky, numpy_kernel.py:463-464;
 * only the lower triangle of a is read).  *info_d = 0, or j+1 when the leading minor of order j+1 is not positive. */
int nb_cholesky_f64(int device, int64_t p, const double* a_d, double* l_d, int32_t* info_d, void* stream);

/* Signs of diag(R) that np.linalg.qr (LAPACK dgeqrf, numpy_kernel.py:455-461) gives for a matrix whose
 * positive-diagonal QR factor has the top p x p block qtop (row-major, leading dimension ldq): signs_d[j] = +-1,
 * R = diag(signs) R+, Q = Q+ diag(signs).  `square` != 0 when the factored matrix is p x p (LAPACK's last
 * reflector is then the identity). */
int64_t nb_qr_signs_scratch_bytes(int64_t p);
int nb_qr_signs_f64(int device, int64_t p, const double* qtop_d, int64_t ldq, int square, double* signs_d,
                    void* scratch_d, int64_t scratch_bytes, void* stream);

/* a = u diag(s) vt for a square p x p matrix (np.linalg.svd, numpy_kernel.py:466-469), singular values in
 * descending order, by one-sided Jacobi.  The signs of paired columns of u / rows of vt are not LAPACK's (any
 * consistent choice is a valid SVD).  *info_d = sweeps used, or -1 if 60 sweeps did not converge. */
int64_t nb_svd_scratch_bytes(int64_t p);
int nb_svd_f64(int device, int64_t p, const double* a_d, double* u_d, double* s_d, double* vt_d,
               int32_t* info_d, void* scratch_d, int64_t scratch_bytes, void* stream);

/* ---- the GLM inner loop: glms.py:457-499 (Logistic), :275-327 (Linear), :515-556 (Poisson)
 *      as called from newton() :660-676 and LBFGS lbfgs.py:102-155 ------------------------ */

/* One pass over a row block X[n, d] (row stride ldx) and y[n] (NB_F64 or NB_I64) with
 * beta[p], p = d + 1, the intercept last (GLM.fit appends a ones column, glms.py:124-135 —
 * never materialised here).  Computes for this block
 *     eta = X.beta[:d] + beta[d];  mu = link_inv(eta)
 *     obj  = family objective summed over rows            (stats_d[p*p + p])
 *     g    = [X,1]^T (mu - y)                              (stats_d[p*p .. p*p+p))
 *     H    = [X,1]^T diag(w) [X,1], w = mu(1-mu) | 1 | mu  (stats_d[0 .. p*p), row-major)
 * into ONE packed buffer stats_d of nb_glm_stats_len(d) doubles, so a single all-reduce sums
 * the row shards.  want_hess == 0 skips H (L-BFGS / gradient passes) and leaves that part of
 * stats_d untouched.  want_hess == NB_GLM_HESS_NO_OBJECTIVE (3) computes g and H but not the
 * objective, whose slot is set to 0: newton() (glms.py:660-676) never evaluates it, and its two
 * logarithms per row are most of the link's cost on narrow blocks (d <= 7 always computes it).
 * Penalty terms are not included (they are added once, after the
 * all-reduce, by nb_glm_newton_update or the host).  Deterministic. */
#define NB_GLM_HESS_NO_OBJECTIVE 3
int64_t nb_glm_stats_len(int64_t d);
size_t nb_glm_scratch_bytes(int device, int64_t n, int64_t d, int want_hess);
int nb_glm_stats(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx,
                 const void* y_d, int ydt, const double* beta_d, int want_hess,
                 double* stats_d, void* scratch_d, size_t scratch_bytes, void* stream);

/* The same statistics with the d x d part of the Hessian, X^T diag(w) X (glms.py:491-499: the
 * `(X.T * s) @ X` of LogisticRegression.hessian, the Gram matrix of LinearRegression), formed on the
 * INTEGER tensor path (tcgen05.mma kind::i8, int32 accumulators in tensor memory) instead of FP64 DMMA:
 * x sqrt(w) is rounded to a 39-bit fixed-point number per column scale and split into five int8 digit
 * planes whose pairwise products are exact; relative error <= ~1e-11 of the column scale (see
 * csrc/glm_ozaki.cu).  Requested with want_hess | NB_GLM_INT8_HESSIAN (also in nb_glm_scratch_bytes: the
 * planes need 5 bytes per element of X) and the output of nb_glm_prepare for this X block; applies to
 * logistic / linear families, 128 < d <= 256, d % 4 == 0, ldx % 4 == 0, 32-byte aligned X — otherwise
 * the call takes the FP64 path of nb_glm_stats.  A value outside the prepared column scale (prep of
 * another block, Inf/NaN in X) turns H into NaN instead of a wrong number.
 *
 * nb_glm_prepare: per-column maxima of |X| -> power-of-two scales (one pass over X; X blocks are immutable,
 * numpy_kernel.py:131-135, so once per block).  prep_d: nb_glm_prep_bytes(d) bytes of device memory.  Its first
 * 258 int32 are readable by the host: colexp[256] (max|x[:,i]| < 2^colexp[i]), bad (Inf/NaN seen), range_log2 =
 * ceil(log2(max_i max|x[:,i]| / rms(x[:,i]))).  The fixed-point error of H, relative to sqrt(H_ii H_jj), is about
 * 2^-38 * 2^range_log2 (less by ~sqrt(n) for rounding errors of random sign); a caller that needs 1e-10 should
 * take the FP64 path when range_log2 > 5 (heavy-tailed columns, single outliers). */
#define NB_GLM_INT8_HESSIAN 4
/* measurement only, together with NB_GLM_INT8_HESSIAN: run just the row pass (link, g, u, digit planes) or just the
 * integer products + finish on the planes the previous call left in scratch_d */
/* with NB_GLM_INT8_HESSIAN: leave out the two least significant of the eleven plane products (the level-5 terms).  Their
 * relative contribution to H is ~ 2^(2 r - 38.6) / sqrt(n) (r = range_log2 of nb_glm_prepare, n rows): the caller sets this
 * when that is far below its tolerance (nums_b200.fused: <= 1e-11), which takes a quarter off the operand traffic */
#define NB_GLM_INT8_NO_LEVEL5 32
/* with NB_GLM_INT8_HESSIAN: FOUR digit planes (x sqrt(w) rounded to 31 bits per column scale) and eight plane products.
 * The rounding error of H grows to ~ 2^(r - 31) / sqrt(n) relative to sqrt(H_ii H_jj) (random sign over n rows): for
 * blocks of millions of rows that is still below 1e-11 (nums_b200.fused sets the flag only then), and the pass writes and
 * reads a fifth fewer plane bytes and runs 8 instead of 9 products on fully used accumulators */
#define NB_GLM_INT8_FOUR_PLANES 64
#define NB_GLM_INT8_ROWS_ONLY 8
#define NB_GLM_INT8_PRODUCTS_ONLY 16
size_t nb_glm_prep_bytes(int64_t d);
int nb_glm_prepare(int device, const double* x_d, int64_t n, int64_t d, int64_t ldx, void* prep_d,
                   size_t prep_bytes, void* stream);
int nb_glm_stats_prepared(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx,
                          const void* y_d, int ydt, const double* beta_d, int want_hess, const void* prep_d,
                          double* stats_d, void* scratch_d, size_t scratch_bytes, void* stream);

/* The Newton update of glms.py:660-676 on the (all-reduced) packed stats:
 *     g += l2vec * beta;  H += diag(l2vec)           (grad_penalty/hessian_penalty :254-272)
 *     beta_out = beta - inv(H) @ g                    (explicit inverse, linalg.py:246-261)
 *     gmax_d[0] = max |g|                             (convergence test :674)
 * l2vec_d may be NULL (no penalty).  info_d as in nb_inv_f64.  stats_d is CONSUMED: the penalty is added to its g and
 * to the diagonal of its H in place.  scratch: nb_inv_scratch_bytes(p)
 * + 2*p*p*sizeof(double). */
size_t nb_glm_newton_scratch_bytes(int64_t p);
int nb_glm_newton_update(int device, int64_t p, double* stats_d, const double* l2vec_d,
                         const double* beta_d, double* beta_out_d, double* gmax_d,
                         int32_t* info_d, void* scratch_d, size_t scratch_bytes, void* stream);

/* newton() of glms.py:660-676 as ONE call: per iteration nb_glm_stats_prepared on every row block, the FIFO pairwise sum
 * of the partial statistics (BlockArray.tree_reduce, blockarray.py:643-689), nb_glm_newton_update, and the read-back of
 * the 16-byte verdict that the `max|g| <= tol` test needs (the stream is synchronised once per iteration; nothing else
 * crosses to the host).  All blocks live on `device`; x_blocks / y_blocks / rows / ldx / preps are HOST arrays of device
 * pointers and sizes (preps may be NULL, or hold NULL for blocks that take the FP64 Hessian).  beta_d (p doubles) is
 * updated in place.  iters_out = iterations run, gmax_out = max|g| of the last one, info_out != 0 if the last update
 * hit a singular or non-finite Hessian (beta_d then holds the iterate before it).  Stops after max_iter iterations or when
 * max|g| <= tol; tol < 0 runs exactly max_iter iterations. */
size_t nb_glm_newton_fit_scratch_bytes(int device, int nblocks, int64_t max_rows, int64_t d, int want_hess_flags);
int nb_glm_newton_fit(int device, int family, int nblocks, const double* const* x_blocks, const int64_t* rows,
                      int64_t d, const int64_t* ldx, const void* const* y_blocks, int ydt,
                      const void* const* preps, int want_hess_flags, const double* l2vec_d, double* beta_d,
                      double tol, int max_iter, int32_t* iters_out, double* gmax_out, int32_t* info_out,
                      void* scratch_d, size_t scratch_bytes, void* stream);

/* L-BFGS (nums/models/lbfgs.py) on a device-resident workspace: the memory of m (s, y) pairs is a ring in HBM, and
 * the p-sized algebra the reference runs as dozens of block-array calls per iteration is three launches.
 *   nb_lbfgs_direction  dir = -H g by the two-loop recursion (LBFGS.get_p :86-100 with H = gamma I, get_H :78-84),
 *                       slope_d[0] = g . dir
 *   nb_lbfgs_trial      out = theta + alpha dir                                     (the line-search trial point :47,:51)
 *   nb_lbfgs_judge      BackTrackingLineSearch's test (:45-48): rejected iff f_t is NaN or f_t > f + c alpha slope.
 *                       verdict_d[0] = 1 / 0.  On acceptance (or `force`, the min_alpha exit :50) with `remember`, the
 *                       pair (theta_t - theta, g_t - g) enters the ring with rho = 1 / (s.y + 1e-30) and
 *                       gamma = s.y / (y.y + 1e-30) (LBFGSMemory :55-62) and verdict_d[1] = max |g_t| (converged() :157).
 * The workspace (nb_lbfgs_ws_bytes) is owned by the caller; nb_lbfgs_reset empties the ring.  All vectors are
 * contiguous float64 of length p on `device`; f / slope / verdict are device scalars, so one 16-byte read-back per
 * trial is the only host synchronisation. */
size_t nb_lbfgs_ws_bytes(int64_t p, int m);
int nb_lbfgs_reset(int device, int64_t p, int m, void* ws_d, size_t ws_bytes, void* stream);
int nb_lbfgs_direction(int device, int64_t p, int m, void* ws_d, const double* g_d, double* dir_d,
                       double* slope_d, void* stream);
int nb_lbfgs_trial(int device, int64_t p, const double* theta_d, const double* dir_d, double alpha,
                   double* out_d, void* stream);
int nb_lbfgs_judge(int device, int64_t p, int m, void* ws_d, const double* theta_d, const double* g_d,
                   const double* f_d, const double* theta_t_d, const double* g_t_d, const double* f_t_d,
                   const double* slope_d, double alpha, double c, int force, int remember,
                   double* verdict_d, void* stream);

/* mu = link_inv(X.beta[:d] + beta[d]) for predict / forward (glms.py:194-197). */
int nb_glm_predict(int device, int family, const double* x_d, int64_t n, int64_t d, int64_t ldx,
                   const double* beta_d, double* mu_d, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NUMS_B200_H */
