// Library-level entry points: error reporting, device discovery.
#include "common.cuh"
#include <mutex>
#include <string.h>

namespace nb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static DevInfo g_info[64];
static bool g_info_ok[64];
static std::mutex g_info_mu;

const DevInfo& dev_info(int device) {
  std::lock_guard<std::mutex> lk(g_info_mu);
  if (device < 0 || device >= 64) device = 0;
  if (!g_info_ok[device]) {
    int sms = 148, smem = 227 * 1024;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    g_info[device].sms = sms > 0 ? sms : 148;
    g_info[device].max_smem_optin = smem;
    g_info_ok[device] = true;
  }
  return g_info[device];
}

}  // namespace nb

extern "C" {

const char* nb_last_error(void) { return nb::g_err; }

int nb_version(void) { return 1; }

int nb_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    nb::set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return -1;
  }
  return n;
}

int nb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor) {
  int a = 0, b = 0, c = 0;
  NB_CHECK(cudaDeviceGetAttribute(&a, cudaDevAttrMultiProcessorCount, device));
  NB_CHECK(cudaDeviceGetAttribute(&b, cudaDevAttrComputeCapabilityMajor, device));
  NB_CHECK(cudaDeviceGetAttribute(&c, cudaDevAttrComputeCapabilityMinor, device));
  if (sm_count) *sm_count = a;
  if (cc_major) *cc_major = b;
  if (cc_minor) *cc_minor = c;
  return 0;
}

}  // extern "C"
