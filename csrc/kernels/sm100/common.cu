// Error plumbing and library identity for libsmpb200_kernels.so.
#include "common.cuh"

#include <cstdarg>

namespace smpk {

char* last_error_buf() {
  static thread_local char buf[1024] = {0};
  return buf;
}

int set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 1024, fmt, ap);
  va_end(ap);
  return 1;
}

}  // namespace smpk

SMPK_API const char* smpk_last_error() { return smpk::last_error_buf(); }

SMPK_API int smpk_version() { return 100; }

// Reports the compute capability of the current device (major*10+minor), or -1 without a device.
SMPK_API int smpk_device_cc() {
  int dev = 0, major = 0, minor = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return -1;
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  return major * 10 + minor;
}
