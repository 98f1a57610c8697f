// Error reporting, device query and ABI version for libhpe_b200.so.
#include "hpe_common.cuh"
#include <string.h>

namespace hpe {

char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int num_sms() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n; cached_dev = dev;
  }
  return cached;
}

}  // namespace hpe

extern "C" int hpe_abi_version(void) { return 1; }

extern "C" const char* hpe_last_error(void) { return hpe::last_error_buf(); }

extern "C" int hpe_check_device(void) {
  int dev = 0, major = 0;
  HPE_CUDA_OK(cudaGetDevice(&dev));
  HPE_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) return hpe::fail(HPE_EARCH, "device %d has compute capability %d.x; libhpe_b200 needs sm_100a", dev, major);
  return HPE_OK;
}
