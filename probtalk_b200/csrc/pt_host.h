// probtalk_b200 -- host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
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

// PT_PDL=0 turns programmatic dependent launch off (A/B measurements); read once
inline bool pdl_enabled() {
  static const bool on = !(getenv("PT_PDL") && atoi(getenv("PT_PDL")) == 0);
  return on;
}

// Launch `kernel` so that it may overlap the tail of its predecessor in the stream (programmatic dependent launch;
// the kernel calls pt::pdl_wait() before it touches dependent memory), optionally as clusters of `cluster` CTAs.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

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
