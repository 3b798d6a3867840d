// common.h -- error plumbing shared by the translation units of libsimubox_b200.so
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/sbx.h"

namespace sbx {

char* last_error_buf();  // thread-local, 512 bytes

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define SBX_CUDA(call)                                                                    \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return ::sbx::fail(SBX_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,          \
                         cudaGetErrorString(e_));                                         \
  } while (0)

// RAII device switch
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (dev != prev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

int require_sm100(int device);  // SBX_OK or SBX_ENODEVICE

}  // namespace sbx
