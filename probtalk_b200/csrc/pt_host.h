// probtalk_b200 -- host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

namespace pt {

// Thread-local last-error text behind pt_last_error().
char* error_buffer();
constexpr int kErrorBufferSize = 512;

inline int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), kErrorBufferSize, fmt, ap);
  va_end(ap);
  return 1;
}

// SM count of the current device, cached (cudaDeviceGetAttribute costs microseconds per call)
inline int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  int& c = cached[dev & 63];
  if (c == 0) {
    int v = 148;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    c = v;
  }
  return c;
}

}  // namespace pt

#define PT_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) return pt::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                           __FILE__, __LINE__);                                    \
  } while (0)
