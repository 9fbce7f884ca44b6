// Version, error reporting and device-property cache of libvtrace_b200.so.
#include <mutex>
#include <unordered_map>

#include "common.h"

namespace vtr {

static thread_local char g_error[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t err, const char* what) {
  snprintf(g_error, sizeof(g_error), "CUDA error %d (%s) in %s", static_cast<int>(err),
           cudaGetErrorString(err), what);
  // Clear the sticky "last error" so the next call does not see it again.
  cudaGetLastError();
  return static_cast<int>(err);
}

int device_info(DeviceInfo* out) {
  constexpr int kMaxDevices = 64;
  static DeviceInfo cache[kMaxDevices];
  static bool valid[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  VTR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(VTR_ERR_UNSUPPORTED, "device index %d", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!valid[dev]) {
    DeviceInfo d;
    d.device = dev;
    VTR_CUDA_CHECK(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
    VTR_CUDA_CHECK(
        cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    VTR_CUDA_CHECK(
        cudaDeviceGetAttribute(&d.smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
    cache[dev] = d;
    valid[dev] = true;
  }
  *out = cache[dev];
  return VTR_OK;
}

// cudaFuncSetAttribute costs microseconds of host time on every call; the
// opt-in only ever has to grow, so remember what each kernel already has on
// each device.
int allow_dynamic_smem(const void* kernel, size_t bytes) {
  static std::unordered_map<unsigned long long, size_t> granted;
  static std::mutex mu;
  int dev = 0;
  VTR_CUDA_CHECK(cudaGetDevice(&dev));
  const unsigned long long key =
      (static_cast<unsigned long long>(reinterpret_cast<uintptr_t>(kernel)) << 8) ^
      static_cast<unsigned>(dev & 0xff);
  std::lock_guard<std::mutex> lock(mu);
  auto it = granted.find(key);
  if (it != granted.end() && it->second >= bytes) return VTR_OK;
  VTR_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(bytes)));
  granted[key] = bytes;
  return VTR_OK;
}

}  // namespace vtr

extern "C" int vtr_version(void) { return VTR_VERSION; }

extern "C" const char* vtr_last_error(void) { return vtr::g_error; }
