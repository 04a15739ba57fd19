// probtalk_b200 -- motion metrics of the evaluator: landmark velocity difference and diversity.
//
//   pt_lvd        evaluation/metrics.py:79-94 -> Batch_LVD :26-77, the branch the evaluator takes
//                 (scripts/test_holisticbody.py:120,130: symmetrical = False, weight = False):
//                     lvd = mean_b  sum_t ( sum_j | |pr[b,t+1,j] - pr[b,t,j]|_2 - |gt[t+1,j] - gt[t,j]|_2 | ) / (T - 1)
//   pt_diversity  evaluation/metrics.py:96-110: mean over sample pairs i < j of mean |x_i - x_j|.
// Both are tiny reductions: one CTA per unit (sample / pair) sums in fp64 in a fixed order and writes the unit's value,
// a one-CTA kernel averages the units in index order.  Deterministic; every buffer belongs to the caller.
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kThreads = 256;

// fixed-order block sum: per-thread partials -> shared -> thread 0 adds them in thread order
__device__ __forceinline__ double block_sum_ordered(double v, double* sh) {
  sh[threadIdx.x] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < kThreads; ++i) t += sh[i];
  __syncthreads();
  return t;                        // valid in thread 0
}

__global__ void __launch_bounds__(kThreads) lvd_kernel(const float* __restrict__ gt, const float* __restrict__ pr, int T,
                                                       int J, int D, double* __restrict__ per_sample) {
  __shared__ double sh[kThreads];
  const int b = blockIdx.x;
  const float* p = pr + (size_t)b * T * J * D;
  const long long n = (long long)(T - 1) * J;
  double acc = 0.0;
  for (long long e = threadIdx.x; e < n; e += kThreads) {      // e = (t, j), a fixed stride per thread
    const long long o = e * D;                                 // (t * J + j) * D ; the next frame is J * D further
    float sg = 0.f, sp = 0.f;
    for (int d = 0; d < D; ++d) {
      const float dg = gt[o + (long long)J * D + d] - gt[o + d];
      const float dp = p[o + (long long)J * D + d] - p[o + d];
      sg = fmaf(dg, dg, sg);
      sp = fmaf(dp, dp, sp);
    }
    acc += (double)fabsf(sqrtf(sp) - sqrtf(sg));
  }
  const double t = block_sum_ordered(acc, sh);
  if (threadIdx.x == 0) per_sample[b] = t / (double)(T - 1);
}

__global__ void __launch_bounds__(kThreads) pair_l1_kernel(const float* __restrict__ x, int B, long long L,
                                                           double* __restrict__ per_pair) {
  __shared__ double sh[kThreads];
  // pair index -> (i, j), i < j, in the reference's loop order (i outer, j inner)
  int pidx = blockIdx.x, i = 0;
  while (pidx >= B - 1 - i) { pidx -= B - 1 - i; ++i; }
  const int j = i + 1 + pidx;
  const float* a = x + (size_t)i * L;
  const float* b = x + (size_t)j * L;
  double acc = 0.0;
  for (long long e = threadIdx.x; e < L; e += kThreads) acc += (double)fabsf(a[e] - b[e]);
  const double t = block_sum_ordered(acc, sh);
  if (threadIdx.x == 0) per_pair[blockIdx.x] = t / (double)L;
}

__global__ void mean_ordered_kernel(const double* __restrict__ v, int n, double* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < n; ++i) t += v[i];
    out[0] = t / (double)n;
  }
}

}  // namespace

extern "C" int pt_lvd(const float* gt, const float* pr, int B, int T, int J, int D, double* per_sample, double* mean_out,
                      void* stream) {
  if (!gt || !pr || !per_sample || !mean_out) return pt::fail("pt_lvd: null pointer");
  if (B < 1 || T < 2 || J < 1 || D < 1) return pt::fail("pt_lvd: need B >= 1, T >= 2, J >= 1, D >= 1 (got %d, %d, %d, %d)", B, T, J, D);
  cudaStream_t st = (cudaStream_t)stream;
  lvd_kernel<<<B, kThreads, 0, st>>>(gt, pr, T, J, D, per_sample);
  mean_ordered_kernel<<<1, 32, 0, st>>>(per_sample, B, mean_out);
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_diversity(const float* kps, int B, long long L, double* per_pair, double* mean_out, void* stream) {
  if (!kps || !per_pair || !mean_out) return pt::fail("pt_diversity: null pointer");
  if (B < 2 || L < 1) return pt::fail("pt_diversity: need at least 2 samples of at least 1 element (got %d, %lld)", B, L);
  if (B > 4096) return pt::fail("pt_diversity: at most 4096 samples (got %d)", B);
  const int pairs = B * (B - 1) / 2;
  cudaStream_t st = (cudaStream_t)stream;
  pair_l1_kernel<<<pairs, kThreads, 0, st>>>(kps, B, L, per_pair);
  mean_ordered_kernel<<<1, 32, 0, st>>>(per_pair, pairs, mean_out);
  PT_CUDA(cudaGetLastError());
  return 0;
}
