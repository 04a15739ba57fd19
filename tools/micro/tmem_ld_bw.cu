// Micro-benchmark: tcgen05.ld (32x32b.x32) throughput per SM with 4 or 8 reading warps.  nvcc -arch=sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../probtalk_b200/csrc/pt_common.cuh"

__global__ void __launch_bounds__(384, 1) k(int warps_reading, int iters, long long* cycles, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) { pt::tmem_alloc(&slot, 512); pt::tmem_relinquish(); }
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  const uint32_t tmem = slot;
  float acc = 0.f;
  __syncthreads();
  long long t0 = clock64();
  if (warp >= 4 && warp < 4 + warps_reading) {
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    for (int it = 0; it < iters; ++it) {
      uint32_t v[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) pt::tmem_ld32(tmem + lane_off + ((it & 3) * 128 + 32 * c) % 512, v[c]);
      pt::tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 4; ++c) acc += __uint_as_float(v[c][0]) + __uint_as_float(v[c][31]);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 128) cycles[blockIdx.x] = t1 - t0;
  __syncthreads();
  if (acc == 123.456f) sink[0] = acc;
  pt::tc_fence_before();
  __syncthreads();
  if (warp == 0) pt::tmem_dealloc(tmem, 512);
}

int main() {
  long long* d; float* s;
  cudaMalloc(&d, 8 * 256); cudaMalloc(&s, 4);
  for (int w : {1, 4, 8}) {
    const int iters = 2000;
    k<<<148, 384>>>(w, iters, d, s);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d, 8 * 148, cudaMemcpyDeviceToHost);
    const double bytes = (double)w * iters * 4 * 32 * 32 * 4;
    printf("warps=%d cycles=%lld bytes/clk/SM=%.1f err=%s\n", w, h[0], bytes / (double)h[0], cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
