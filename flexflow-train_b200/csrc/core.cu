// Error plumbing and device queries shared by every entry point.
#include "common.cuh"

#include <cstdarg>
#include <mutex>

namespace ffk {

namespace {
thread_local char g_err[1024] = "";
}

void set_error(char const *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, char const *what, char const *file, int line) {
  set_error("CUDA failure: %s (%d) in %s at %s:%d", cudaGetErrorString(e), (int)e, what, file,
            line);
  return FFK_ERR_CUDA;
}

::ffk_timer *&gemm_timer_slot() {
  thread_local ::ffk_timer *slot = nullptr;
  return slot;
}

DeviceInfo const &device_info() {
  // one process per GPU: the device current at first use is the device of the process
  static DeviceInfo info = [] {
    DeviceInfo d{kSMsB200, 227 * 1024};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) {
      cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
      cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    if (d.sm_count <= 0) d.sm_count = kSMsB200;
    return d;
  }();
  return info;
}

}  // namespace ffk

extern "C" {

char const *ffk_last_error(void) { return ffk::g_err; }

int ffk_abi_version(void) { return 1; }

int ffk_device_sm_count(void) {
  int dev = 0, n = 0;
  FFK_CUDA(cudaGetDevice(&dev));
  FFK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  return n;
}

}  // extern "C"
