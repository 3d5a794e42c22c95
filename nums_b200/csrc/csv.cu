// CSV ingest on the device: the byte range of a file that nums/core/filesystem.py:227-283
// (read_csv_block) would hand to Python's float() / np.int64() line by line is uploaded as raw text and parsed in HBM.
//
//   k_csv_count   newlines per 2048-byte chunk
//   k_csv_scan    exclusive scan of the chunk counts (one CTA) -> rows
//   k_csv_lines   start offset of every line
//   k_csv_parse   one thread per line: trim "\r\n", split at the delimiter, convert every field
//
// float fields follow Python's float() grammar (surrounding whitespace, sign, digits with single underscores, optional
// fraction and exponent, inf / infinity / nan) and are CORRECTLY ROUNDED: the decimal significand (up to 19 digits in a
// 64-bit integer) and exponent go through the Eisel-Lemire algorithm with the 128-bit powers of five of pow5_table.h,
// so every value equals float(text) bit for bit.  Significands longer than 19 digits are truncated and converted twice
// (w and w + 1); if the two roundings differ the field is reported as ambiguous instead of guessed.  Hexadecimal
// floats ("0x", float.fromhex in the reference) are reported as unsupported.  int64 fields follow int(): whitespace,
// sign, digits with underscores, range-checked.
#include "common.cuh"
#include "pow5_table.h"

namespace nb {

constexpr int CSV_CHUNK = 2048;
constexpr int CSV_T = 256;

enum CsvError { CSV_OK = 0, CSV_BAD_FIELD = 1, CSV_RAGGED = 2, CSV_AMBIGUOUS = 3, CSV_HEX = 4, CSV_RANGE = 5 };

__global__ void __launch_bounds__(CSV_T)
k_csv_count(const unsigned char* __restrict__ text, int64_t n, int64_t nchunks, int* __restrict__ counts) {
  __shared__ int sm[CSV_T / 32];
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int64_t lo = c * CSV_CHUNK;
    int cnt = 0;
    // 8 bytes per thread: CSV_CHUNK = CSV_T * 8; text + lo is 8-byte aligned (cudaMalloc base, chunk multiple of 8)
    const int64_t p = lo + (int64_t)threadIdx.x * 8;
    if (p + 8 <= n) {
      const unsigned long long v = *reinterpret_cast<const unsigned long long*>(text + p);
#pragma unroll
      for (int b = 0; b < 8; ++b) cnt += (((v >> (8 * b)) & 0xff) == '\n') ? 1 : 0;
    } else {
      for (int64_t q = p; q < n; ++q) cnt += text[q] == '\n' ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < CSV_T / 32; ++w) t += sm[w];
      counts[c] = t;
    }
  }
}

// offs[c] = newlines before chunk c; header[0] = total
__global__ void __launch_bounds__(1024)
k_csv_scan(int64_t nchunks, const int* __restrict__ counts, long long* __restrict__ offs, long long* __restrict__ header) {
  __shared__ long long part[1024];
  const int t = threadIdx.x;
  const int64_t q = (nchunks + 1023) / 1024;
  const int64_t lo = min(nchunks, (int64_t)t * q), hi = min(nchunks, lo + q);
  long long s = 0;
  for (int64_t c = lo; c < hi; ++c) s += counts[c];
  part[t] = s;
  __syncthreads();
  if (t == 0) {
    long long run = 0;
    for (int i = 0; i < 1024; ++i) { const long long v = part[i]; part[i] = run; run += v; }
    header[0] = run;
  }
  __syncthreads();
  long long run = part[t];
  for (int64_t c = lo; c < hi; ++c) { offs[c] = run; run += counts[c]; }
}

// line_start[r] = offset of the first byte of line r (line 0 starts at 0); line_start[rows] = n
__global__ void __launch_bounds__(32)
k_csv_lines(const unsigned char* __restrict__ text, int64_t n, int64_t nchunks, const long long* __restrict__ offs,
            long long* __restrict__ line_start, int64_t rows) {
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    // the block is one warp: it walks the chunk 32 bytes at a time; ballot gives the rank of every newline
    const int lane = threadIdx.x;
    long long base = offs[c];
    const int64_t lo = c * CSV_CHUNK, hi = min(n, lo + CSV_CHUNK);
    for (int64_t p = lo; p < hi; p += 32) {
      const int64_t q = p + lane;
      const bool nl = q < hi && text[q] == '\n';
      const unsigned m = __ballot_sync(0xffffffffu, nl);
      if (nl) {
        const long long r = base + __popc(m & ((1u << lane) - 1u)) + 1;
        if (r <= rows) line_start[r] = q + 1;
      }
      base += __popc(m);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) line_start[0] = 0;
}

// ---- decimal -> binary64 (Eisel-Lemire) ---------------------------------------------------------------------
struct Am { uint64_t mantissa; int power2; };

__device__ __forceinline__ Am el_compute(int64_t q, uint64_t w) {
  Am a;
  if (w == 0 || q < NB_POW5_LO) { a.mantissa = 0; a.power2 = 0; return a; }
  if (q > NB_POW5_HI) { a.mantissa = 0; a.power2 = 0x7FF; return a; }
  const int lz = __clzll((long long)w);
  w <<= lz;
  const int index = 2 * (int)(q - NB_POW5_LO);
  uint64_t hi = __umul64hi(w, pow5_128[index]);
  uint64_t lo = w * pow5_128[index];
  const uint64_t precision_mask = 0xFFFFFFFFFFFFFFFFull >> 55;
  if ((hi & precision_mask) == precision_mask) {
    const uint64_t second_hi = __umul64hi(w, pow5_128[index + 1]);
    lo += second_hi;
    if (second_hi > lo) hi++;
  }
  const int upperbit = (int)(hi >> 63);
  const int shift = upperbit + 64 - 52 - 3;
  a.mantissa = hi >> shift;
  a.power2 = (int)((((152170 + 65536) * q) >> 16) + 63) + upperbit - lz + 1023;
  if (a.power2 <= 0) {                       // subnormal
    if (-a.power2 + 1 >= 64) { a.mantissa = 0; a.power2 = 0; return a; }
    a.mantissa >>= -a.power2 + 1;
    a.mantissa += (a.mantissa & 1);
    a.mantissa >>= 1;
    a.power2 = (a.mantissa < (1ull << 52)) ? 0 : 1;
    return a;
  }
  if (lo <= 1 && q >= -4 && q <= 23 && (a.mantissa & 3) == 1) {
    if ((a.mantissa << shift) == hi) a.mantissa &= ~1ull;      // exactly halfway: round to even
  }
  a.mantissa += (a.mantissa & 1);
  a.mantissa >>= 1;
  if (a.mantissa >= (2ull << 52)) { a.mantissa = 1ull << 52; a.power2++; }
  a.mantissa &= ~(1ull << 52);
  if (a.power2 >= 0x7FF) { a.power2 = 0x7FF; a.mantissa = 0; }
  return a;
}

__device__ __forceinline__ bool is_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }
__device__ __forceinline__ bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
__device__ __forceinline__ unsigned char lower(unsigned char c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

__device__ __forceinline__ bool match_word(const unsigned char* p, const unsigned char* e, const char* wd, int len) {
  if (e - p != len) return false;
  for (int i = 0; i < len; ++i) if (lower(p[i]) != (unsigned char)wd[i]) return false;
  return true;
}

// float(text[p, e)); returns a CsvError
__device__ int parse_float(const unsigned char* p, const unsigned char* e, double* out) {
  while (p < e && is_space(*p)) ++p;
  while (e > p && is_space(e[-1])) --e;
  if (p == e) return CSV_BAD_FIELD;
  for (const unsigned char* s = p; s + 1 < e; ++s)
    if (s[0] == '0' && lower(s[1]) == 'x') return CSV_HEX;
  bool neg = false;
  if (*p == '+' || *p == '-') { neg = (*p == '-'); ++p; }
  if (p == e) return CSV_BAD_FIELD;
  if (!is_digit(*p) && *p != '.') {
    double v;
    if (match_word(p, e, "inf", 3) || match_word(p, e, "infinity", 8)) v = __longlong_as_double(0x7FF0000000000000ll);
    else if (match_word(p, e, "nan", 3)) v = __longlong_as_double(0x7FF8000000000000ll);
    else return CSV_BAD_FIELD;
    *out = neg ? -v : v;
    return CSV_OK;
  }
  uint64_t w = 0;
  int sig = 0;                 // significant digits taken into w (<= 19)
  int64_t exp10 = 0;
  bool truncated = false, any_digit = false, seen_nonzero = false;
  // integer part
  bool prev_digit = false;
  while (p < e) {
    const unsigned char c = *p;
    if (is_digit(c)) {
      any_digit = true; prev_digit = true;
      if (c != '0') seen_nonzero = true;
      if (seen_nonzero) {
        if (sig < 19) { w = w * 10 + (c - '0'); ++sig; }
        else { ++exp10; if (c != '0') truncated = true; }
      }
      ++p;
    } else if (c == '_') {
      if (!prev_digit || p + 1 >= e || !is_digit(p[1])) return CSV_BAD_FIELD;
      prev_digit = false; ++p;
    } else break;
  }
  if (p < e && *p == '.') {
    ++p;
    prev_digit = false;
    while (p < e) {
      const unsigned char c = *p;
      if (is_digit(c)) {
        any_digit = true; prev_digit = true;
        if (c != '0') seen_nonzero = true;
        if (seen_nonzero) {
          if (sig < 19) { w = w * 10 + (c - '0'); ++sig; --exp10; }
          else if (c != '0') truncated = true;
        } else {
          --exp10;               // leading zeros of a pure fraction only move the exponent
        }
        ++p;
      } else if (c == '_') {
        if (!prev_digit || p + 1 >= e || !is_digit(p[1])) return CSV_BAD_FIELD;
        prev_digit = false; ++p;
      } else break;
    }
  }
  if (!any_digit) return CSV_BAD_FIELD;
  if (p < e && lower(*p) == 'e') {
    ++p;
    bool eneg = false;
    if (p < e && (*p == '+' || *p == '-')) { eneg = (*p == '-'); ++p; }
    if (p == e || !is_digit(*p)) return CSV_BAD_FIELD;
    int64_t ev = 0;
    prev_digit = false;
    while (p < e) {
      const unsigned char c = *p;
      if (is_digit(c)) { prev_digit = true; if (ev < 100000) ev = ev * 10 + (c - '0'); ++p; }
      else if (c == '_') {
        if (!prev_digit || p + 1 >= e || !is_digit(p[1])) return CSV_BAD_FIELD;
        prev_digit = false; ++p;
      } else break;
    }
    exp10 += eneg ? -ev : ev;
  }
  if (p != e) return CSV_BAD_FIELD;
  Am a = el_compute(exp10, w);
  if (truncated) {
    const Am b = el_compute(exp10, w + 1);
    if (a.mantissa != b.mantissa || a.power2 != b.power2) return CSV_AMBIGUOUS;
  }
  uint64_t bits = a.mantissa | ((uint64_t)a.power2 << 52);
  if (neg) bits |= 0x8000000000000000ull;
  *out = __longlong_as_double((long long)bits);
  return CSV_OK;
}

// int(text[p, e)) into int64; returns a CsvError
__device__ int parse_int(const unsigned char* p, const unsigned char* e, long long* out) {
  while (p < e && is_space(*p)) ++p;
  while (e > p && is_space(e[-1])) --e;
  if (p == e) return CSV_BAD_FIELD;
  bool neg = false;
  if (*p == '+' || *p == '-') { neg = (*p == '-'); ++p; }
  if (p == e || !is_digit(*p)) return CSV_BAD_FIELD;
  unsigned long long v = 0;
  bool prev_digit = false, over = false;
  while (p < e) {
    const unsigned char c = *p;
    if (is_digit(c)) {
      prev_digit = true;
      if (v > (0xFFFFFFFFFFFFFFFFull - (c - '0')) / 10) over = true; else v = v * 10 + (c - '0');
      ++p;
    } else if (c == '_') {
      if (!prev_digit || p + 1 >= e || !is_digit(p[1])) return CSV_BAD_FIELD;
      prev_digit = false; ++p;
    } else return CSV_BAD_FIELD;
  }
  if (over || v > (neg ? 0x8000000000000000ull : 0x7FFFFFFFFFFFFFFFull)) return CSV_RANGE;
  *out = neg ? (long long)(0ull - v) : (long long)v;
  return CSV_OK;
}

// err[0] = smallest failing line + 1 (0 = none), err[1] = its error code, err[2] = its column
__global__ void __launch_bounds__(CSV_T)
k_csv_parse(const unsigned char* __restrict__ text, const long long* __restrict__ line_start, int64_t rows, int cols,
            unsigned char delim, int odt, void* __restrict__ out, unsigned long long* __restrict__ err) {
  const int64_t stride = (int64_t)gridDim.x * CSV_T;
  for (int64_t r = (int64_t)blockIdx.x * CSV_T + threadIdx.x; r < rows; r += stride) {
    const unsigned char* p = text + line_start[r];
    const unsigned char* e = text + line_start[r + 1] - 1;          // the line's '\n'
    while (e > p && (e[-1] == '\r' || e[-1] == '\n')) --e;           // str.strip("\r\n")
    while (p < e && (*p == '\r' || *p == '\n')) ++p;
    int code = CSV_OK, col = 0;
    const unsigned char* f = p;
    for (;;) {
      const unsigned char* g = f;
      while (g < e && *g != delim) ++g;
      if (col >= cols) { code = CSV_RAGGED; break; }
      if (odt == NB_F64) {
        double v = 0.0;
        code = parse_float(f, g, &v);
        if (code == CSV_OK) ((double*)out)[r * cols + col] = v;
      } else {
        long long v = 0;
        code = parse_int(f, g, &v);
        if (code == CSV_OK) ((long long*)out)[r * cols + col] = v;
      }
      if (code != CSV_OK) break;
      ++col;
      if (g >= e) break;
      f = g + 1;
    }
    if (code == CSV_OK && col != cols) { code = CSV_RAGGED; }
    if (code != CSV_OK) {
      // keep the first failing line: (line + 1) in the high bits orders the atomicMin
      const unsigned long long key = ((unsigned long long)(r + 1) << 16) | ((unsigned long long)(col & 0xfff) << 4) |
                                     (unsigned long long)code;
      atomicMin(err, key);
    }
  }
}

static inline size_t csv_up(size_t v) { return (v + 255) & ~(size_t)255; }
static inline int64_t csv_chunks(int64_t n) { return (n + CSV_CHUNK - 1) / CSV_CHUNK; }

}  // namespace nb

using namespace nb;

// scratch: [header 256 B | counts int32 per chunk | offsets int64 per chunk | line starts int64 x (max_rows + 1)]
extern "C" int64_t nb_csv_scratch_bytes(int64_t nbytes, int64_t max_rows) {
  if (nbytes <= 0) return 256;
  const int64_t ch = csv_chunks(nbytes);
  return (int64_t)(256 + csv_up((size_t)ch * 4) + csv_up((size_t)ch * 8) + csv_up((size_t)(max_rows + 1) * 8));
}

extern "C" int nb_csv_count_rows(int device, const unsigned char* text_d, int64_t nbytes, void* scratch_d,
                                 int64_t scratch_bytes, int64_t* rows_host, void* stream) {
  NB_ENTER(device);
  *rows_host = 0;
  if (nbytes <= 0) return 0;
  NB_REQUIRE((((uintptr_t)text_d) & 7) == 0, "nb_csv_count_rows: text must be 8-byte aligned");
  const int64_t ch = csv_chunks(nbytes);
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_csv_scratch_bytes(nbytes, 0), "nb_csv_count_rows: scratch too small");
  char* sp = (char*)scratch_d;
  long long* header = (long long*)sp;
  int* counts = (int*)(sp + 256);
  long long* offs = (long long*)(sp + 256 + csv_up((size_t)ch * 4));
  const int grid = (int)(ch < (int64_t)dev_info(device).sms * 16 ? ch : (int64_t)dev_info(device).sms * 16);
  k_csv_count<<<grid, CSV_T, 0, st>>>(text_d, nbytes, ch, counts);
  NB_LAUNCH_CHECK();
  k_csv_scan<<<1, 1024, 0, st>>>(ch, counts, offs, header);
  NB_LAUNCH_CHECK();
  long long rows = 0;
  NB_CHECK(cudaMemcpyAsync(&rows, header, sizeof(long long), cudaMemcpyDeviceToHost, st));
  NB_CHECK(cudaStreamSynchronize(st));
  *rows_host = rows;
  return 0;
}

extern "C" int nb_csv_parse(int device, const unsigned char* text_d, int64_t nbytes, int64_t rows, int cols, int delimiter,
                            int odt, void* out_d, void* scratch_d, int64_t scratch_bytes, int64_t* error_host,
                            void* stream) {
  NB_ENTER(device);
  error_host[0] = 0; error_host[1] = 0; error_host[2] = 0;
  if (rows <= 0) return 0;
  NB_REQUIRE(odt == NB_F64 || odt == NB_I64, "nb_csv_parse: output dtype %d not supported", odt);
  NB_REQUIRE(cols >= 1, "nb_csv_parse: cols < 1");
  NB_REQUIRE(scratch_d != nullptr && scratch_bytes >= nb_csv_scratch_bytes(nbytes, rows), "nb_csv_parse: scratch too small");
  const int64_t ch = csv_chunks(nbytes);
  char* sp = (char*)scratch_d;
  unsigned long long* err = (unsigned long long*)(sp + 64);
  long long* offs = (long long*)(sp + 256 + csv_up((size_t)ch * 4));
  long long* line_start = (long long*)(sp + 256 + csv_up((size_t)ch * 4) + csv_up((size_t)ch * 8));
  const unsigned long long none = ~0ull;
  NB_CHECK(cudaMemcpyAsync(err, &none, sizeof(none), cudaMemcpyHostToDevice, st));
  const int grid = (int)(ch < (int64_t)dev_info(device).sms * 16 ? ch : (int64_t)dev_info(device).sms * 16);
  k_csv_lines<<<grid, 32, 0, st>>>(text_d, nbytes, ch, offs, line_start, rows);
  NB_LAUNCH_CHECK();
  const int64_t pb = (rows + CSV_T - 1) / CSV_T;
  const int pgrid = (int)(pb < (int64_t)dev_info(device).sms * 8 ? pb : (int64_t)dev_info(device).sms * 8);
  k_csv_parse<<<pgrid, CSV_T, 0, st>>>(text_d, line_start, rows, cols, (unsigned char)delimiter, odt, out_d, err);
  NB_LAUNCH_CHECK();
  unsigned long long key = 0;
  NB_CHECK(cudaMemcpyAsync(&key, err, sizeof(key), cudaMemcpyDeviceToHost, st));
  NB_CHECK(cudaStreamSynchronize(st));
  if (key != none) {
    error_host[0] = (int64_t)(key >> 16);            // failing line + 1
    error_host[1] = (int64_t)(key & 0xf);            // CsvError
    error_host[2] = (int64_t)((key >> 4) & 0xfff);   // column
  }
  return 0;
}
