// Library-wide state: last-error string, version, launch counter.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace dla {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...);
std::atomic<long long> g_launches{0};
static std::atomic<int> g_device{-1};

// Binds the process's device to the calling thread for THIS library's runtime instance.  Autograd
// worker threads have no current context for driver-API calls (cuTensorMapEncodeTiled returned
// CUDA_ERROR_INVALID_CONTEXT there), and would default to device 0 on multi-GPU boxes.
int ensure_device() {
  static thread_local int bound = -2;
  int want = g_device.load(std::memory_order_relaxed);
  if (want < 0) {
    if (bound == -2) {
      cudaError_t e = cudaFree(0);
      if (e != cudaSuccess) { set_error("cudaFree(0) failed: %s", cudaGetErrorString(e)); return (int)e; }
      bound = -1;
    }
    return 0;
  }
  if (bound != want) {
    cudaError_t e = cudaSetDevice(want);
    if (e == cudaSuccess) e = cudaFree(0);
    if (e != cudaSuccess) { set_error("cudaSetDevice(%d) failed: %s", want, cudaGetErrorString(e)); return (int)e; }
    bound = want;
  }
  return 0;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace dla

extern "C" {
const char* dla_last_error(void) { return dla::g_err; }
int dla_version(void) { return 100; }
int dla_set_device(int device) {
  dla::g_device.store(device);
  return dla::ensure_device();
}
long long dla_launch_count(void) { return dla::g_launches.load(); }
}
