// runtime.cu -- error reporting, device checks, version.
#include "common.cuh"
#include <mutex>

namespace b2d {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace b2d

extern "C" int b2d_version(void) { return 100; }  // 0.1.0

extern "C" const char* b2d_last_error(void) { return b2d::err_buf(); }

extern "C" int b2d_check_device(void) {
  int dev = 0, major = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return b2d::set_err(B2D_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return b2d::set_err(B2D_ERR_CUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
  if (major != 10)
    return b2d::set_err(B2D_ERR_ARCH, "libb200diff is built for sm_100a only; device %d is sm_%d*", dev, major);
  return B2D_OK;
}
