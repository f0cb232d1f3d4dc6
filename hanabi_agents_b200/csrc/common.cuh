// Shared host-side helpers for the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/hanabi_b200.h"

namespace hb {

// Thread-local last-error message (hb_last_error).
void set_error(const char* fmt, ...);

inline int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("%s failed at %s:%d: %s", what, file, line, cudaGetErrorString(e));
  return HB_ERR_CUDA;
}

#define HB_CUDA(call)                                                   \
  do {                                                                  \
    cudaError_t e_ = (call);                                            \
    if (e_ != cudaSuccess) return hb::cuda_fail(e_, #call, __FILE__, __LINE__); \
  } while (0)

#define HB_REQUIRE(cond, ...)       \
  do {                              \
    if (!(cond)) {                  \
      hb::set_error(__VA_ARGS__);   \
      return HB_ERR_INVALID;        \
    }                               \
  } while (0)

// RAII device switch for entry points that own a device id.
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    want = dev;
  }
  ~DeviceGuard() {
    if (prev >= 0 && prev != want) cudaSetDevice(prev);
  }
  int want = -1;
};

inline int num_sms(int device) {
  int n = 148;
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
  return n;
}

}  // namespace hb
