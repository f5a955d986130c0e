// Library-wide host state: last-error string, version, cached device properties.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

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

}  // namespace b200rl

extern "C" {
const char* b200rl_last_error(void) { return b200rl::g_err; }
int b200rl_version(void) { return 100; }
}
