// Library-wide state: last-error string, device check, shutdown.
#include "common.cuh"
#include <string.h>

namespace cvpce {

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

int require_sm100() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_res = CVPCE_ERR_NO_DEVICE;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_err(CVPCE_ERR_NO_DEVICE, "no CUDA device (%s); cvpce_b200 has no CPU fallback", cudaGetErrorString(e));
  }
  if (dev == cached_dev) {
    if (cached_res != CVPCE_OK) set_err(cached_res, "current device %d is not sm_100; cvpce_b200 has no CPU fallback", dev);
    return cached_res;
  }
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_err(CVPCE_ERR_NO_DEVICE, "cudaDeviceGetAttribute failed: %s", cudaGetErrorString(e));
  }
  cached_dev = dev;
  cached_res = (major == 10) ? CVPCE_OK : CVPCE_ERR_NO_DEVICE;
  if (cached_res != CVPCE_OK)
    set_err(cached_res, "current device %d has compute capability %d.x, need 10.x (B200); no CPU fallback", dev, major);
  return cached_res;
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

void tma_cache_clear();  // sim_topk_sm100.cu

}  // namespace cvpce

extern "C" {

const char* cvpce_last_error(void) { return cvpce::err_buf(); }
int cvpce_version(void) { return CVPCE_B200_VERSION; }
int cvpce_device_check(void) { return cvpce::require_sm100(); }
int cvpce_shutdown(void) {
  cvpce::tma_cache_clear();
  return CVPCE_OK;
}

}  // extern "C"
