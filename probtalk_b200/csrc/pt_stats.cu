// probtalk_b200 -- feature statistics for the Frechet gesture distance, sm_100a.
//
// mean and unbiased covariance (np.mean / np.cov(rowvar=False), evaluation/FGD.py:172-176) of N feature vectors of
// dimension D in fp64, deterministic: the rows are cut into a fixed number of chunks, every chunk is reduced by one CTA
// in a fixed order, and the per-chunk partials are combined in chunk order.  No atomics, so two runs give the same bits.
// The per-chunk partials live in a caller-provided workspace (pt_feature_stats_workspace_bytes).
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kStatChunks = 16;      // row chunks = CTAs of the two partial kernels
constexpr int kStatMaxD = 128;
constexpr int kStatThreads = 256;

__device__ __forceinline__ void chunk_range(int n, int chunk, int& lo, int& hi) {
  const int per = (n + kStatChunks - 1) / kStatChunks;
  lo = chunk * per;
  hi = min(n, lo + per);
}

// partial column sums: ws_sum[chunk][D]
__global__ void __launch_bounds__(kStatThreads) stat_sum_kernel(const float* __restrict__ x, int n, int d,
                                                                double* __restrict__ ws_sum) {
  __shared__ double part[kStatThreads];
  int lo, hi;
  chunk_range(n, blockIdx.x, lo, hi);
  // thread = (row lane r, column c): kStatThreads / d row lanes walk the chunk's rows with a fixed stride
  const int lanes = kStatThreads / d, c = threadIdx.x % d, r = threadIdx.x / d;
  double s = 0.0;
  if (r < lanes)
    for (int i = lo + r; i < hi; i += lanes) s += (double)x[(size_t)i * d + c];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x < d) {
    double t = 0.0;
    for (int k = 0; k < lanes; ++k) t += part[k * d + threadIdx.x];   // fixed order
    ws_sum[(size_t)blockIdx.x * d + threadIdx.x] = t;
  }
}

// partial centred second moments: ws_cov[chunk][D*D]; CTA 0 also publishes the mean
__global__ void __launch_bounds__(kStatThreads) stat_cov_kernel(const float* __restrict__ x, int n, int d,
                                                                const double* __restrict__ ws_sum,
                                                                double* __restrict__ ws_cov, double* __restrict__ mean_out) {
  __shared__ double mean_s[kStatMaxD];
  __shared__ double row_s[8][kStatMaxD];               // 8 centred rows per step
  for (int c = threadIdx.x; c < d; c += kStatThreads) {
    double t = 0.0;
    for (int k = 0; k < kStatChunks; ++k) t += ws_sum[(size_t)k * d + c];   // chunk order, same in every CTA
    t /= (double)n;
    mean_s[c] = t;
    if (blockIdx.x == 0) mean_out[c] = t;
  }
  __syncthreads();
  int lo, hi;
  chunk_range(n, blockIdx.x, lo, hi);
  // thread owns the (i, j) entries  e = threadIdx.x + k * kStatThreads  of the D x D matrix
  const int per_thread = (d * d + kStatThreads - 1) / kStatThreads;      // <= 64 for D = 128
  double acc[kStatMaxD * kStatMaxD / kStatThreads];
#pragma unroll
  for (int k = 0; k < kStatMaxD * kStatMaxD / kStatThreads; ++k) acc[k] = 0.0;
  for (int i0 = lo; i0 < hi; i0 += 8) {
    __syncthreads();
    for (int e = threadIdx.x; e < 8 * d; e += kStatThreads) {
      const int rr = e / d, c = e - rr * d, i = i0 + rr;
      row_s[rr][c] = (i < hi) ? (double)x[(size_t)i * d + c] - mean_s[c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kStatMaxD * kStatMaxD / kStatThreads; ++k) {
      if (k < per_thread) {
        const int e = threadIdx.x + k * kStatThreads;
        if (e < d * d) {
          const int a = e / d, b = e - a * d;
          double t = acc[k];
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) t = fma(row_s[rr][a], row_s[rr][b], t);
          acc[k] = t;
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kStatMaxD * kStatMaxD / kStatThreads; ++k) {
    const int e = threadIdx.x + k * kStatThreads;
    if (k < per_thread && e < d * d) ws_cov[(size_t)blockIdx.x * d * d + e] = acc[k];
  }
}

// cov = (sum of the chunk partials, chunk order) / (n - 1)
__global__ void stat_final_kernel(const double* __restrict__ ws_cov, int n, int d, double* __restrict__ cov_out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= d * d) return;
  double t = 0.0;
  for (int k = 0; k < kStatChunks; ++k) t += ws_cov[(size_t)k * d * d + e];
  cov_out[e] = t / (double)(n - 1);
}

}  // namespace

namespace pt {
// shared by the entry points of this file: validates and enqueues the three kernels on `ws`
int feature_stats_launch(const float* feats, int N, int D, double* mean_out, double* cov_out, double* ws,
                         cudaStream_t st, const char* who) {
  if (!feats || !mean_out || !cov_out || !ws) return pt::fail("%s: null pointer", who);
  if (N < 2) return pt::fail("%s: need at least 2 samples (got %d)", who, N);
  if (D < 1 || D > kStatMaxD) return pt::fail("%s: feature dimension must be 1..%d (got %d)", who, kStatMaxD, D);
  double* ws_sum = ws;
  double* ws_cov = ws + (size_t)kStatChunks * D;
  stat_sum_kernel<<<kStatChunks, kStatThreads, 0, st>>>(feats, N, D, ws_sum);
  stat_cov_kernel<<<kStatChunks, kStatThreads, 0, st>>>(feats, N, D, ws_sum, ws_cov, mean_out);
  stat_final_kernel<<<(D * D + 255) / 256, 256, 0, st>>>(ws_cov, N, D, cov_out);
  PT_CUDA(cudaGetLastError());
  return 0;
}
size_t feature_stats_ws_doubles(int D) { return (size_t)kStatChunks * ((size_t)D + (size_t)D * D); }
}  // namespace pt

extern "C" long long pt_feature_stats_workspace_bytes(int D) {
  if (D < 1 || D > kStatMaxD) return -1;
  return (long long)(pt::feature_stats_ws_doubles(D) * sizeof(double));
}

extern "C" int pt_feature_stats(const float* feats, int N, int D, double* mean_out, double* cov_out, void* workspace,
                                long long workspace_bytes, void* stream) {
  if (D >= 1 && D <= kStatMaxD && workspace_bytes < pt_feature_stats_workspace_bytes(D))
    return pt::fail("pt_feature_stats: workspace of %lld bytes is too small (need %lld)", workspace_bytes,
                    pt_feature_stats_workspace_bytes(D));
  if (reinterpret_cast<uintptr_t>(workspace) & 7) return pt::fail("pt_feature_stats: workspace must be 8-byte aligned");
  return pt::feature_stats_launch(feats, N, D, mean_out, cov_out, static_cast<double*>(workspace), (cudaStream_t)stream,
                                  "pt_feature_stats");
}

// ---------------------------------------------------------------------------------------------
// The same statistics for callers that must not allocate device memory.
//
// pt_feature_stats wants a workspace, and whoever provides one has to get it from somewhere.  Some callers cannot: a
// section of a program in which allocation calls are forbidden, a real-time loop that promised its host not to touch the
// allocator after start-up, a binding from a runtime whose buffers are all accounted for up front.  For them the partial
// sums go into a fixed array that is part of the library image itself, so the call needs nothing but its inputs and
// outputs (and it never calls an allocating routine, torch's caching allocator included).
//
//   * size: kStatChunks * (kStatMaxD + kStatMaxD^2) doubles = 16 * 16512 * 8 B = 2 113 536 bytes per device, enough for
//     the largest feature dimension pt_feature_stats accepts (128); a larger D is refused with an error, never truncated.
//   * nothing in it has to start at zero: every launch first overwrites each partial it later reads.
//   * it is ONE array per device, shared by every caller in the process: two calls that may run at the same time (different
//     streams, different host threads) race on it and corrupt each other's result.  Callers serialise such calls
//     themselves (same stream, or an event between them); calls on one stream are ordered by the stream.
//   * the address is looked up with cudaGetSymbolAddress the first time a device is used and remembered.  That lookup is
//     a runtime call that may initialise the module on the device: do this first call per device during set-up, before
//     any stream capture begins (no statement is made here about what happens inside a capture).
static __device__ __align__(256) double g_stat_scratch[kStatChunks * (kStatMaxD + kStatMaxD * kStatMaxD)];

extern "C" int pt_feature_stats_static(const float* feats, int N, int D, double* mean_out, double* cov_out, void* stream) {
  if (D > kStatMaxD)
    return pt::fail("pt_feature_stats_static: feature dimension %d exceeds the built-in scratch (at most %d; %zu bytes)", D,
                    kStatMaxD, sizeof(double) * kStatChunks * (kStatMaxD + kStatMaxD * kStatMaxD));
  // one cached address per device ordinal; atomics because two host threads may make the first call together (both
  // would look up the same address, so the duplicate store is harmless)
  static std::atomic<double*> scratch_of_device[64];
  int dev = 0;
  PT_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return pt::fail("pt_feature_stats_static: device index %d out of range", dev);
  double* ws = scratch_of_device[dev].load(std::memory_order_acquire);
  if (!ws) {
    cudaError_t rc = cudaGetSymbolAddress(reinterpret_cast<void**>(&ws), g_stat_scratch);
    if (rc != cudaSuccess || !ws)
      return pt::fail("pt_feature_stats_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(rc));
    scratch_of_device[dev].store(ws, std::memory_order_release);
  }
  // the array belongs to the CURRENT device: feats / outputs / stream must live there too
  {
    cudaPointerAttributes pa;
    const cudaError_t qrc = cudaPointerGetAttributes(&pa, feats);
    if (qrc != cudaSuccess) (void)cudaGetLastError();        // not a CUDA pointer: the launch below reports it
    else if (pa.type == cudaMemoryTypeDevice && pa.device != dev)
      return pt::fail("pt_feature_stats_static: feats live on device %d but the current device is %d (make the owning "
                      "device current before the call)", pa.device, dev);
  }
  return pt::feature_stats_launch(feats, N, D, mean_out, cov_out, ws, (cudaStream_t)stream, "pt_feature_stats_static");
}
