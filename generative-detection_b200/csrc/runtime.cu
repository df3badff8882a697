// Error bookkeeping, device queries, status strings.
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

namespace odvae {

static thread_local char g_last_error[512] = "";

void set_last_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}

static std::atomic<unsigned long long> g_launches{0};
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }

int check_launch(const char* where) {
  g_launches.fetch_add(1, std::memory_order_relaxed);  // called exactly once after every kernel launch
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_cuda_error(e, where);
    return ODVAE_ERR_CUDA;
  }
  return ODVAE_OK;
}

int num_sms() {
  static int sms[64];
  static bool have[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!have[dev]) {
    int v = 148;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) v = 148;
    sms[dev] = v;
    have[dev] = true;
  }
  return sms[dev];
}

}  // namespace odvae

extern "C" {

const char* odvae_status_string(int status) {
  switch (status) {
    case ODVAE_OK: return "ok";
    case ODVAE_ERR_SHAPE: return "unsupported or inconsistent shape/argument";
    case ODVAE_ERR_ARCH: return "device is not sm_100 (B200)";
    case ODVAE_ERR_WORKSPACE: return "workspace too small";
    case ODVAE_ERR_CUDA: return "CUDA call failed";
    default: return "unknown status";
  }
}

const char* odvae_last_cuda_error(void) { return odvae::g_last_error; }

int odvae_device_supported(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

int odvae_version(void) { return 100; }

int64_t odvae_launch_count(void) { return static_cast<int64_t>(odvae::launch_count()); }

}  // extern "C"
