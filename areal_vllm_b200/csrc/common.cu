// Library-wide host state: last-error string, version, cached device properties.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

#include <mutex>

namespace b200rl {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return kNumSMs;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kNumSMs;
    cached[dev] = n;
  }
  return cached[dev];
}

// cudaFuncAttributeMaxDynamicSharedMemorySize, once per (kernel, device). Callers on different host
// threads are serialised by a mutex: the table is tiny and the lock is uncontended in practice
// (every rank drives its GPU from one thread), but an FFI host may call from several.
int ensure_dynamic_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static const void* seen_fn[128];
  static int seen_dev[128];
  static int n_seen = 0;
  int dev = 0;
  B200RL_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n_seen; i++)
    if (seen_fn[i] == kernel && seen_dev[i] == dev) return 0;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_last_error("cudaFuncSetAttribute(smem=%d): %s", bytes, cudaGetErrorString(e));
    return (int)e;
  }
  if (n_seen < 128) {  // a full table only means the attribute is set again on later calls
    seen_fn[n_seen] = kernel;
    seen_dev[n_seen] = dev;
    n_seen++;
  }
  return 0;
}

}  // namespace b200rl

extern "C" {
const char* b200rl_last_error(void) { return b200rl::g_err; }
int b200rl_version(void) { return 100; }
}
