// Shared helpers for libnumsb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <math.h>
#include <stdlib.h>
#include "../../include/nums_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libnumsb200 is written for sm_100a (B200) only"
#endif

namespace nb {

void set_error(const char* fmt, ...);

// Per-device properties, cached.
struct DevInfo { int sms; int max_smem_optin; };
const DevInfo& dev_info(int device);

#define NB_CHECK(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      nb::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

#define NB_REQUIRE(cond, ...)              \
  do {                                     \
    if (!(cond)) {                         \
      nb::set_error(__VA_ARGS__);          \
      return 2;                            \
    }                                      \
  } while (0)

#define NB_LAUNCH_CHECK() NB_CHECK(cudaGetLastError())

// A debug switch read from the environment once per process.
#define NB_ENV_FLAG(name) ([]() -> bool { static const bool v = getenv(name) != nullptr; return v; }())

// Launch set-up that belongs to (kernel, device), not to every call: the dynamic shared memory opt-in and the
// occupancy query cost microseconds each and sit in the per-iteration launch path of the GLM kernels.  One static
// instance per call site (i.e. per kernel instantiation).
constexpr int NB_MAX_DEVICES = 16;
struct KernelSetup {
  bool done[NB_MAX_DEVICES] = {};
  int occ[NB_MAX_DEVICES] = {};
  template <typename K>
  int get(int device, K kern, int threads, size_t smem, int* out_occ) {
    const bool cached = device >= 0 && device < NB_MAX_DEVICES && done[device];
    if (!cached) {
      NB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int o = 1;
      NB_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, threads, smem));
      if (o < 1) o = 1;
      if (device >= 0 && device < NB_MAX_DEVICES) { occ[device] = o; done[device] = true; }
      *out_occ = o;
      return 0;
    }
    *out_occ = occ[device];
    return 0;
  }
};

static inline int item_size(int dt) {
  switch (dt) {
    case NB_F64: case NB_I64: return 8;
    case NB_BOOL: return 1;
    case NB_F32: case NB_I32: return 4;
  }
  return 0;
}

// Grid for a grid-stride elementwise kernel: a multiple of the SM count.
static inline int grid_for(int device, int64_t work_items, int threads, int per_thread, int max_waves = 8) {
  int sms = dev_info(device).sms;
  int64_t blocks = (work_items + (int64_t)threads * per_thread - 1) / ((int64_t)threads * per_thread);
  int64_t cap = (int64_t)sms * max_waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ---- typed loads with conversion to a compute type -------------------------------------------
template <typename C>
__device__ __forceinline__ C load_as(const void* p, int dt, int64_t i) {
  switch (dt) {
    case NB_F64: return (C)(((const double*)p)[i]);
    case NB_I64: return (C)(((const int64_t*)p)[i]);
    case NB_BOOL: return (C)(((const uint8_t*)p)[i] != 0);
    case NB_F32: return (C)(((const float*)p)[i]);
    default: return (C)(((const int32_t*)p)[i]);
  }
}
template <>
__device__ __forceinline__ bool load_as<bool>(const void* p, int dt, int64_t i) {
  switch (dt) {
    case NB_F64: return ((const double*)p)[i] != 0.0;   // NaN is truthy, as in NumPy
    case NB_I64: return ((const int64_t*)p)[i] != 0;
    case NB_BOOL: return ((const uint8_t*)p)[i] != 0;
    case NB_F32: return ((const float*)p)[i] != 0.0f;
    default: return ((const int32_t*)p)[i] != 0;
  }
}

template <typename R>
__device__ __forceinline__ void store_as(void* p, int dt, int64_t i, R v) {
  switch (dt) {
    case NB_F64: ((double*)p)[i] = (double)v; break;
    case NB_I64: ((int64_t*)p)[i] = (int64_t)v; break;
    case NB_BOOL: ((uint8_t*)p)[i] = (uint8_t)(v != (R)0); break;
    case NB_F32: ((float*)p)[i] = (float)v; break;
    default: ((int32_t*)p)[i] = (int32_t)v; break;
  }
}

// ---- warp / block reductions -------------------------------------------------------------------
template <typename T, typename Op>
__device__ __forceinline__ T warp_reduce(T v, Op op) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ double shfl_xor_d(double v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int device) {
    target = device;
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != device && cudaSetDevice(device) != cudaSuccess) ok = false;
  }
  int target = -1;
  ~DeviceGuard() {
    if (ok && prev >= 0 && prev != target) cudaSetDevice(prev);
  }
};

#define NB_ENTER(device)                                                        \
  nb::DeviceGuard _guard(device);                                               \
  if (!_guard.ok) { nb::set_error("cannot select CUDA device %d", device); return 3; } \
  cudaStream_t st = (cudaStream_t)stream; (void)st

}  // namespace nb
