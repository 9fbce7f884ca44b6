// Host-side helpers shared by the translation units of libvtrace_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vtrace_b200.h"

namespace vtr {

// Records a thread-local message and returns `code` (see api.cu).
int fail(int code, const char* fmt, ...);

// Maps a failed CUDA runtime call to the positive-return convention.
int cuda_fail(cudaError_t err, const char* what);

inline bool aligned(const void* p, size_t a) {
  return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0;
}

struct DeviceInfo {
  int device;
  int sm_count;
  int max_smem_optin;  // bytes of dynamic shared memory a block may opt in to
  int smem_per_sm;
};

// Queries (and caches per device) the properties the launch heuristics use.
int device_info(DeviceInfo* out);

// Raises a kernel's dynamic shared memory opt-in to `bytes` on the current
// device if it is not already that high (cached: no driver call once granted).
int allow_dynamic_smem(const void* kernel, size_t bytes);

constexpr float kLog2e = 1.4426950408889634f;

}  // namespace vtr

#define VTR_CUDA_CHECK(expr)                                   \
  do {                                                         \
    cudaError_t err__ = (expr);                                \
    if (err__ != cudaSuccess) return ::vtr::cuda_fail(err__, #expr); \
  } while (0)
