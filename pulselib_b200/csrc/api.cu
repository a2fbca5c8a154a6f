// libpulse_envs: library-wide entry points (version, last error) -- include/pulse_envs.h.
#include "common.cuh"
#include <stdarg.h>
#include <stdio.h>

#include <atomic>

static thread_local char g_err[512] = "";

void pulse_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int pulse_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    pulse_set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return PULSE_OK;
}

int pulse_sm_count(void) {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  cache[dev].store(n, std::memory_order_relaxed);
  return n;
}

extern "C" {
int pulse_version(void) { return 200; }
const char* pulse_last_error(void) { return g_err; }

// Fills out[0..3] with one block of the chance tape (host side; used by callers that want to
// materialise the same draws the kernels consume, e.g. to inject them into another engine).
int pulse_tape_block(uint64_t seed, uint64_t env_id, uint32_t stream, uint32_t counter, uint32_t* out) {
  PULSE_REQUIRE(out != nullptr, "out is null");
  PulseU4 r = pulse_tape(seed, env_id, stream, counter);
  out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
  return PULSE_OK;
}
}
