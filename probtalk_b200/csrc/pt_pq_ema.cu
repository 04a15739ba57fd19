// probtalk_b200 -- training-side update of the product quantizer: EMA codebook update and dead-code
// reset of ProductQuantization.forward in training mode, nets/spg/vqvae_modules.py:495-534 (share_code = False branch,
// the configuration of config/vq.json), with ExponentialMovingAverage.forward :271-276.
//
// Per group g, given the codes idx[., g] of this batch's rows x[., g, :]:
//     counts[k]   = #{rows with idx = k}                      dw[k] = sum of those rows
//     cs_hidden  -= (cs_hidden - counts) (1 - decay)          cs = cs_hidden / (1 - decay^step)
//     n = sum cs                                              cs_n = (cs + eps) / (n + K eps) n
//     dw_hidden  -= (dw_hidden - dw) (1 - decay)              w   = (dw_hidden / (1 - decay^step)) / cs_n
//     usage[k] = cs_n[k] >= 1                                 rand[k] = x[k, g, :]   (the `_tile` branch with rows >= K)
//     dw_hidden = usage ? dw_hidden : rand                    cs_hidden = usage ? cs_hidden : 1
//     codebook  = usage ? w : rand
// and the commitment term  sum |x - c_idx|^2  of the same pass (F.mse_loss(x, quantized), :537-538).
//
// Two kernels, no atomics, fixed summation order (bit-reproducible): one CTA per (group, code) walks the group's index
// column and sums its rows in fp64 (the reference's fp32 one-hot matmul is the same sum with a library-chosen order), one
// CTA per group applies the EMA recurrences in fp32 exactly as the reference's tensor expressions do.
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kEmaThreads = 512;

struct EmaArgs {
  const float* z; long long zs_b, zs_w, zs_c;       // x[b, g*D + d, w] at z[b*zs_b + (g*D+d)*zs_c + w*zs_w]
  const long long* idx;                             // [B*W, G]
  int B, W, G, K, D;
  float* dw;                                        // scratch [G, K, D]
  float* counts;                                    // scratch [G, K]
  double* sqerr;                                    // scratch [G, K]: sum |x - c_k|^2 over the rows of code k
  const float* codebooks;                           // [G, K, D] (before the update)
};

// one CTA per (g, k): thread = (row part p, dimension d); part p sums rows p, p + parts, ... in row order
__global__ void __launch_bounds__(kEmaThreads) ema_scatter_kernel(EmaArgs a) {
  extern __shared__ double part_s[];                 // [parts][D] sums, then [parts][D] squared errors
  const int g = blockIdx.x / a.K, k = blockIdx.x % a.K;
  const int D = a.D, parts = kEmaThreads / D;
  const int d = threadIdx.x % D, p = threadIdx.x / D;
  const long long rows = (long long)a.B * a.W;
  const float c = a.codebooks[((size_t)g * a.K + k) * D + d];
  double s = 0.0, e = 0.0;
  int n = 0;
  if (p < parts) {
    for (long long r = p; r < rows; r += parts) {
      if (a.idx[r * a.G + g] == k) {
        const long long b = r / a.W, w = r - b * a.W;
        const float x = a.z[b * a.zs_b + w * a.zs_w + (long long)(g * D + d) * a.zs_c];
        s += (double)x;
        const float df = x - c;
        e += (double)df * (double)df;
        ++n;
      }
    }
    part_s[p * D + d] = s;
    part_s[(parts + p) * D + d] = e;
  }
  __shared__ int n_s[kEmaThreads];
  n_s[threadIdx.x] = (d == 0 && p < parts) ? n : 0;
  __syncthreads();
  if (threadIdx.x < D) {
    double t = 0.0;
    for (int q = 0; q < parts; ++q) t += part_s[q * D + threadIdx.x];          // fixed order
    a.dw[((size_t)g * a.K + k) * D + threadIdx.x] = (float)t;
  }
  if (threadIdx.x == 0) {
    int cnt = 0;
    for (int q = 0; q < parts; ++q) cnt += n_s[q * D];
    a.counts[(size_t)g * a.K + k] = (float)cnt;
    double t = 0.0;
    for (int q = 0; q < parts; ++q)
      for (int dd = 0; dd < D; ++dd) t += part_s[(parts + q) * D + dd];
    a.sqerr[(size_t)g * a.K + k] = t;
  }
}

struct EmaFinalArgs {
  const float* z; long long zs_b, zs_w, zs_c;
  int B, W, G, K, D;
  const float* dw; const float* counts; const double* sqerr;
  float* codebooks; float* ema_dw; float* ema_cs;   // state, updated in place
  float one_minus_decay, bias_corr, eps;            // (1 - decay), (1 - decay^step), epsilon
  double* loss_out;                                 // [1]: sum over groups of sum |x - c|^2 / (rows * G * D)  (mse), or null
  int* dead_out;                                    // [G]: codes re-seeded this step, or null
};

// one CTA per group
__global__ void __launch_bounds__(kEmaThreads) ema_final_kernel(EmaFinalArgs a) {
  extern __shared__ float cs_s[];                    // [K] normalised cluster sizes, then usage flags
  __shared__ float n_s;
  __shared__ int dead_s;
  const int g = blockIdx.x, K = a.K, D = a.D;
  float* cs_h = a.ema_cs + (size_t)g * K;
  for (int k = threadIdx.x; k < K; k += kEmaThreads) {
    float h = cs_h[k];
    h -= (h - a.counts[(size_t)g * K + k]) * a.one_minus_decay;              // hidden.sub_((hidden - value) * (1 - decay))
    cs_h[k] = h;
    cs_s[k] = h / a.bias_corr;                                               // average = hidden / (1 - decay^counter)
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float n = 0.f;
    for (int k = 0; k < K; ++k) n += cs_s[k];                                // torch.sum (fp32); fixed order here
    n_s = n;
    dead_s = 0;
  }
  __syncthreads();
  const float n = n_s;
  for (int k = threadIdx.x; k < K; k += kEmaThreads)
    cs_s[k] = (cs_s[k] + a.eps) / (n + (float)K * a.eps) * n;
  __syncthreads();
  for (int e = threadIdx.x; e < K * D; e += kEmaThreads) {
    const int k = e / D, d = e - k * D;
    const size_t o = ((size_t)g * K + k) * D + d;
    float h = a.ema_dw[o];
    h -= (h - a.dw[o]) * a.one_minus_decay;
    const float w = (h / a.bias_corr) / cs_s[k];
    // code_rand = _tile(flat_x[:, g])[:K] = the first K rows of this batch (rows >= K: no noise branch)
    const long long b = k / a.W, ww = k - b * a.W;
    const float rnd = a.z[b * a.zs_b + ww * a.zs_w + (long long)(g * D + d) * a.zs_c];
    const bool used = cs_s[k] >= 1.0f;
    a.ema_dw[o] = used ? h : rnd;                                            // usage * hidden + (1 - usage) * code_rand
    a.codebooks[o] = used ? w : rnd;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += kEmaThreads) {
    const bool used = cs_s[k] >= 1.0f;
    if (!used) { cs_h[k] = 1.0f; atomicAdd(&dead_s, 1); }                    // hidden * usage + (1 - usage)
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (a.dead_out) a.dead_out[g] = dead_s;
    if (a.loss_out && g == 0) {
      double t = 0.0;
      for (int i = 0; i < a.G * K; ++i) t += a.sqerr[i];                     // fixed order
      a.loss_out[0] = t / ((double)a.B * a.W * a.G * D);
    }
  }
}

}  // namespace

extern "C" long long pt_pq_ema_workspace_bytes(int G, int K, int D) {
  if (G < 1 || K < 1 || D < 1) return -1;
  return (long long)G * K * ((long long)D * 4 + 4 + 8) + 64;
}

extern "C" int pt_pq_ema_update(const float* z, long long zs_b, long long zs_c, long long zs_w, int B, int W, int G, int K,
                                int D, const long long* idx, float* codebooks, float* ema_dw, float* ema_cluster_size,
                                double decay, double epsilon, int step, double* mse_out, int* dead_out, void* workspace,
                                long long workspace_bytes, void* stream) {
  if (!z || !idx || !codebooks || !ema_dw || !ema_cluster_size || !workspace) return pt::fail("pt_pq_ema_update: null pointer");
  if (G < 1 || K < 1 || D < 1 || kEmaThreads % D) return pt::fail("pt_pq_ema_update: D must divide %d (got G=%d K=%d D=%d)", kEmaThreads, G, K, D);
  if (step < 1) return pt::fail("pt_pq_ema_update: step counts from 1 (the reference increments its counter before use)");
  const long long rows = (long long)B * W;
  if (rows < K)
    return pt::fail("pt_pq_ema_update: %lld rows < %d codes: the reference re-seeds dead codes from noisy repeats of the batch "
                    "(torch.randn_like, vqvae_modules.py:570-579), which is not reproduced", rows, K);
  if (workspace_bytes < pt_pq_ema_workspace_bytes(G, K, D)) return pt::fail("pt_pq_ema_update: workspace too small");
  if (reinterpret_cast<uintptr_t>(workspace) & 7) return pt::fail("pt_pq_ema_update: workspace must be 8-byte aligned");
  double* sqerr = static_cast<double*>(workspace);
  float* dw = reinterpret_cast<float*>(sqerr + (size_t)G * K);
  float* counts = dw + (size_t)G * K * D;
  cudaStream_t st = (cudaStream_t)stream;
  EmaArgs a{z, zs_b, zs_w, zs_c, idx, B, W, G, K, D, dw, counts, sqerr, codebooks};
  const int parts = kEmaThreads / D;
  ema_scatter_kernel<<<G * K, kEmaThreads, sizeof(double) * 2 * parts * D, st>>>(a);
  PT_CUDA(cudaGetLastError());
  EmaFinalArgs f{z, zs_b, zs_w, zs_c, B, W, G, K, D, dw, counts, sqerr, codebooks, ema_dw, ema_cluster_size,
                 (float)(1.0 - decay), (float)(1.0 - pow(decay, (double)step)), (float)epsilon, mse_out, dead_out};
  if ((size_t)K * sizeof(float) > 48 * 1024) return pt::fail("pt_pq_ema_update: K = %d exceeds the 12288 codes one CTA can stage", K);
  ema_final_kernel<<<G, kEmaThreads, sizeof(float) * K, st>>>(f);
  PT_CUDA(cudaGetLastError());
  return 0;
}
