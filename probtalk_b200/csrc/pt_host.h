// probtalk_b200 -- host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <atomic>

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

// Per-device bookkeeping for set-up that CUDA keeps per device (cudaFuncSetAttribute opt-ins, symbol addresses):
// one slot per device ordinal, indexed by the CURRENT device, so a process that drives several GPUs configures each.
// value() is the largest size already configured on the current device (0 = never).
struct PerDevice {
  std::atomic<size_t> v[64];
  PerDevice() { for (auto& x : v) x.store(0, std::memory_order_relaxed); }
  std::atomic<size_t>& slot() {
    int dev = 0;
    cudaGetDevice(&dev);
    return v[dev & 63];
  }
};

}  // namespace pt

// opt a kernel in to `bytes` of dynamic shared memory once per device (and again if a later call needs more)
#define PT_SMEM_OPT_IN(kernel, bytes)                                                                          \
  do {                                                                                                         \
    static pt::PerDevice _pd;                                                                                  \
    std::atomic<size_t>& _s = _pd.slot();                                                                      \
    if ((size_t)(bytes) > _s.load(std::memory_order_acquire)) {                                                \
      PT_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));        \
      _s.store((size_t)(bytes), std::memory_order_release);                                                    \
    }                                                                                                          \
  } while (0)

#define PT_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) return pt::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                           __FILE__, __LINE__);                                    \
  } while (0)
