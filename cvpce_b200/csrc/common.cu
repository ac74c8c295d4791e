// Library-wide state: last-error string, device check, shutdown.
#include "common.cuh"
#include <string.h>
#include <atomic>
#include <map>
#include <mutex>
#include <utility>

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

// ---- tuning table: explicit, process-wide development overrides (cvpce_tuning_set); never read from the environment
struct TuningEntry {
  const char* name;
  std::atomic<int> value;
};
static TuningEntry g_tuning[] = {
    {"crop_kernel", {0}},       // 0: automatic (column-marching kernel for antialias); 3: direct kernel
    {"crop_band_rows", {0}},    // 0: automatic; else output rows per CTA of the antialias kernels (power of two, 8..256; column-marching kernel)
    {"sim_split", {0}},         // tcgen05 search: forced gallery split count (0: cost model)
    {"sim_plan", {0}},          // tcgen05 search: 0 automatic; 1: (query block, split) grid; 2: linear item ranges whenever legal
    {"sim_keys", {0}},          // tcgen05 search: 0 automatic (key lists when a list lives for <= 16 tiles); 1: never
    {"mac_cluster", {0}},       // 0: automatic; 2 / 4 / 8: cluster size of the dense channels-last MAC cluster kernel
    {"mac_kernel", {0}},        // 0: automatic (balanced streaming kernel when a workspace is given); 1: always the cluster kernel
    {"mac_ctas_per_sm", {0}},   // 0: automatic (2); CTAs per SM of the balanced streaming MAC kernel
    {"rmac_kernel", {0}},       // region lists / per-region L2: 0 automatic (staged kernel from B >= SMs/2); 1: always the cluster kernel; 2: rmac_staged_kernel whenever it applies; 4: whenever they apply, the warp-specialised kernel where it covers the maps
    {"rmac_ws", {0}},           // automatic choice between the two staged kernels: 0 warp-specialised where it covers the maps; 1: never
    {"rmac_stages", {0}},       // 0: automatic (as many 16 KB ring stages as fit, <= 8); else an upper limit (>= 2)
};

int tuning(const char* name) {
  for (auto& e : g_tuning)
    if (strcmp(e.name, name) == 0) return e.value.load(std::memory_order_relaxed);
  return 0;
}

int ensure_dynamic_smem(const void* func, size_t bytes) {
  // cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device, per-function attribute.  Setting it on every launch costs
  // ~10 us of host time in front of the launch (measured: it showed up inside the event-timed spans of the short kernels), so
  // the largest size opted into is remembered per (device, function) under a mutex -- never per process or per thread, which
  // a host thread that moves to a second GPU would find stale.
  if (bytes <= 32 * 1024) return CVPCE_OK;  // the 48 KB default limit also counts the kernel's static shared memory
  int dev = 0;
  CVPCE_CUDA_OK(cudaGetDevice(&dev));
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, size_t> opted;
  std::lock_guard<std::mutex> lock(mu);
  size_t& have = opted[std::make_pair(dev, func)];
  if (bytes > have) {
    CVPCE_CUDA_OK(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    have = bytes;
  }
  return CVPCE_OK;
}

}  // namespace cvpce

extern "C" {

const char* cvpce_last_error(void) { return cvpce::err_buf(); }
int cvpce_version(void) { return CVPCE_B200_VERSION; }
int cvpce_device_check(void) { return cvpce::require_sm100(); }
int cvpce_shutdown(void) {
  cvpce::tma_cache_clear();
  return CVPCE_OK;
}
int cvpce_tuning_set(const char* name, int value) {
  if (name)
    for (auto& e : cvpce::g_tuning)
      if (strcmp(e.name, name) == 0) {
        e.value.store(value, std::memory_order_relaxed);
        return CVPCE_OK;
      }
  return cvpce::set_err(CVPCE_ERR_INVALID_ARG, "cvpce_tuning_set: unknown knob '%s'", name ? name : "(null)");
}
int cvpce_tuning_get(const char* name) { return name ? cvpce::tuning(name) : 0; }

}  // extern "C"
