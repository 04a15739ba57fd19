// probtalk_b200 -- wav2vec2 front end and resampling: the audio features the predictor / refiner are
// conditioned on (scripts/demo.py:105-111: Wav2Vec2Model(**processor(audio)).last_hidden_state, then
// data_utils/utils.py:14-16 linear_interpolation to 30 fps).
//
// pt_w2v_frontend = Wav2Vec2FeatureExtractor's zero-mean / unit-variance normalisation + the first feature-encoder layer
// (Conv1d(1, 512, k = 10, s = 5, no bias) -> GroupNorm(512 groups = per channel over time) -> GELU) in ONE pass over the
// waveform.  The layer is linear in the waveform up to the GroupNorm, so the per-channel mean / variance over the ~192 k
// output frames of a minute of audio follow from ten window means m_j and the 10 x 10 window correlation R_jk of the
// waveform:  mean_c = sum_j w_cj m_j,  E[y_c^2] = sum_jk w_cj w_ck R_jk.  A first kernel accumulates those 67 sums
// (fp32 per thread over a few windows, fp64 across threads and chunks, fixed order), a second folds normalisation and
// GroupNorm into per-utterance effective weights W'[c][j] and biases b'[c], and the third writes
//     out[b, t, c] = GELU( sum_j W'[b][c][j] x[b, 5 t + j] + b'[b][c] )
// as bf16 rows: the 196 MB-per-minute activation is written once and never re-read for its statistics.
// pt_w2v_interp: last_hidden_state rows [B, T, C] -> [B, C, t_out] fp32 channel-first, F.interpolate(mode = 'linear').
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kC0 = 512, kK0 = 10, kS0 = 5;
constexpr int kChunks = 64;                       // CTAs per utterance in the statistics pass
constexpr int kNS = 2 + kK0 + kK0 * (kK0 + 1) / 2;  // S1, S2, m_j (10), R_jk (55, j <= k) = 67
constexpr int kStatThreads = 256;

__global__ void __launch_bounds__(kStatThreads) w2v_stats_kernel(const float* __restrict__ wave, long long L, long long T0,
                                                                 double* __restrict__ partial) {
  __shared__ double red[kStatThreads / 32][kNS];
  const int b = blockIdx.y, c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* x = wave + (long long)b * L;
  float acc[kNS];
#pragma unroll
  for (int i = 0; i < kNS; ++i) acc[i] = 0.f;
  // normalisation sums over this chunk's samples
  const long long s0 = L * c / kChunks, s1 = L * (c + 1) / kChunks;
  for (long long i = s0 + tid; i < s1; i += kStatThreads) { const float v = x[i]; acc[0] += v; acc[1] = fmaf(v, v, acc[1]); }
  // window sums over this chunk's output frames
  const long long t0 = T0 * c / kChunks, t1 = T0 * (c + 1) / kChunks;
  for (long long t = t0 + tid; t < t1; t += kStatThreads) {
    float v[kK0];
#pragma unroll
    for (int j = 0; j < kK0; ++j) v[j] = x[kS0 * t + j];
    int o = 2 + kK0;
#pragma unroll
    for (int j = 0; j < kK0; ++j) {
      acc[2 + j] += v[j];
#pragma unroll
      for (int k = j; k < kK0; ++k) { acc[o] = fmaf(v[j], v[k], acc[o]); ++o; }
    }
  }
  // warp sums in fp64 (xor tree: fixed order), then the 8 warps in warp order
#pragma unroll 1
  for (int i = 0; i < kNS; ++i) {
    double d = (double)acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (lane == 0) red[warp][i] = d;
  }
  __syncthreads();
  if (tid < kNS) {
    double d = 0.0;
    for (int w = 0; w < kStatThreads / 32; ++w) d += red[w][tid];
    partial[((size_t)b * kChunks + c) * kNS + tid] = d;
  }
}

// one CTA per utterance, thread = channel: effective weights / biases of normalisation + conv0 + GroupNorm
__global__ void __launch_bounds__(kC0) w2v_fold_kernel(const double* __restrict__ partial, long long L, long long T0,
                                                       const float* __restrict__ w, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, float* __restrict__ weff,
                                                       float* __restrict__ beff) {
  __shared__ double S[kNS];
  const int b = blockIdx.x, c = threadIdx.x;
  if (c < kNS) {
    double d = 0.0;
    for (int k = 0; k < kChunks; ++k) d += partial[((size_t)b * kChunks + k) * kNS + c];     // chunk order
    S[c] = d;
  }
  __syncthreads();
  const double mu = S[0] / (double)L;
  const double var = S[1] / (double)L - mu * mu;                       // population variance (numpy .var())
  const double r = 1.0 / sqrt(var + 1e-7);                             // Wav2Vec2FeatureExtractor.zero_mean_unit_var_norm
  double wc[kK0], mean = 0.0, e2 = 0.0, wsum = 0.0;
  for (int j = 0; j < kK0; ++j) { wc[j] = (double)w[c * kK0 + j]; wsum += wc[j]; }
  int o = 2 + kK0;
  for (int j = 0; j < kK0; ++j) {
    const double mj = S[2 + j] / (double)T0;
    mean += wc[j] * (mj - mu) * r;
    for (int k = j; k < kK0; ++k) {
      const double mk = S[2 + k] / (double)T0;
      const double rjk = r * r * (S[o] / (double)T0 - mu * (mj + mk) + mu * mu);   // correlation of the normalised waveform
      e2 += (j == k ? 1.0 : 2.0) * wc[j] * wc[k] * rjk;
      ++o;
    }
  }
  const double vc = fmax(e2 - mean * mean, 0.0);                       // biased variance over time (GroupNorm)
  const double sc = (double)gamma[c] / sqrt(vc + 1e-5);
  for (int j = 0; j < kK0; ++j) weff[((size_t)b * kC0 + c) * kK0 + j] = (float)(wc[j] * r * sc);
  beff[(size_t)b * kC0 + c] = (float)((double)beta[c] - sc * mean - sc * r * mu * wsum);
}

// packed fp32x2 arithmetic (sm_100): both lanes are independent outputs (two adjacent channels)
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 pack2(float a, float b) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f2 x, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(x.v)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// erf GELU of two values, 0.5 y (1 + erf(y / sqrt 2)), with erf from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, far below
// the bf16 rounding of the output): for z = |y| / sqrt 2,  erf z = 1 - t (a1 + t (a2 + t (a3 + t (a4 + t a5)))) exp(-z^2),
// t = 1 / (1 + 0.3275911 z);  GELU = 0.5 y + 0.5 |y| - 0.5 |y| poly exp(-z^2).  Two MUFU ops + ~12 packed / scalar FMA-pipe
// instructions per pair instead of two erff calls (~25 instructions each): this kernel is bound by instruction issue.
__device__ __forceinline__ void gelu2(float y0, float y1, float& g0, float& g1) {
  const float z0 = fabsf(y0) * 0.70710678118654752f, z1 = fabsf(y1) * 0.70710678118654752f;
  const float t0 = rcp_approx(fmaf(0.3275911f, z0, 1.f)), t1 = rcp_approx(fmaf(0.3275911f, z1, 1.f));
  const float e0 = ex2_approx(z0 * z0 * -1.4426950408889634f), e1 = ex2_approx(z1 * z1 * -1.4426950408889634f);
  const f2 t = pack2(t0, t1);
  f2 p = fma2(t, pack2(1.061405429f, 1.061405429f), pack2(-1.453152027f, -1.453152027f));
  p = fma2(p, t, pack2(1.421413741f, 1.421413741f));
  p = fma2(p, t, pack2(-0.284496736f, -0.284496736f));
  p = fma2(p, t, pack2(0.254829592f, 0.254829592f));
  p = mul2(mul2(p, t), pack2(e0, e1));                                     // poly(t) exp(-z^2) = 1 - erf z
  const f2 h = pack2(0.5f * fabsf(y0), 0.5f * fabsf(y1));
  const f2 lin = fma2(pack2(0.5f, 0.5f), pack2(y0, y1), h);                // 0.5 y + 0.5 |y|
  const f2 nh = pack2(-0.5f * fabsf(y0), -0.5f * fabsf(y1));
  unpack2(fma2(nh, p, lin), g0, g1);
}

constexpr int kTT = 64;                            // output frames per CTA of the conv pass
__global__ void __launch_bounds__(256) w2v_conv0_kernel(const float* __restrict__ wave, long long L, long long T0,
                                                        const float* __restrict__ weff, const float* __restrict__ beff,
                                                        __nv_bfloat16* __restrict__ out, long long rows_per_sample) {
  __shared__ float xs[kS0 * kTT + kK0];
  const int b = blockIdx.y, tid = threadIdx.x;
  const long long tb = (long long)blockIdx.x * kTT;
  const float* x = wave + (long long)b * L;
  for (int i = tid; i < kS0 * kTT + kK0 - kS0; i += 256) {
    const long long s = kS0 * tb + i;
    xs[i] = s < L ? x[s] : 0.f;
  }
  const int c = tid * 2;                           // two adjacent channels: one 4-byte store per frame, 128 B per warp
  f2 w[kK0];                                       // (W'[c][j], W'[c + 1][j]) pairs: one packed FMA per tap for both channels
#pragma unroll
  for (int j = 0; j < kK0; ++j) w[j] = pack2(weff[((size_t)b * kC0 + c) * kK0 + j], weff[((size_t)b * kC0 + c + 1) * kK0 + j]);
  const f2 bias = pack2(beff[(size_t)b * kC0 + c], beff[(size_t)b * kC0 + c + 1]);
  __syncthreads();
  const int nt = (int)min((long long)kTT, T0 - tb);
  __nv_bfloat16* orow = out + ((size_t)b * rows_per_sample + tb) * kC0 + c;
#pragma unroll 2
  for (int t = 0; t < nt; ++t) {
    f2 y = bias;
#pragma unroll
    for (int j = 0; j < kK0; ++j) { const float v = xs[kS0 * t + j]; y = fma2(w[j], pack2(v, v), y); }
    float y0, y1, g0, g1;
    unpack2(y, y0, y1);
    gelu2(y0, y1, g0, g1);
    *reinterpret_cast<uint32_t*>(orow + (size_t)t * kC0) = pt::pack_bf16x2(g0, g1);
  }
}

// rows [B, Tin_pitch, C] (bf16 or fp32) -> [B, C, Tout] fp32, PyTorch's linear rule (align_corners = False)
template <typename InT>
__global__ void __launch_bounds__(256) w2v_interp_kernel(const InT* __restrict__ in, long long in_rows_per_sample, int Tin, int C,
                                                         float* __restrict__ out, int Tout) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, t0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // (32, 8)
  const float scale = (float)Tin / (float)Tout;
  for (int i = ty; i < 32; i += 8) {                            // rows = output frames, lanes = channels (coalesced reads)
    const int t = t0 + i, c = c0 + tx;
    float v = 0.f;
    if (t < Tout && c < C) {
      float src = scale * ((float)t + 0.5f) - 0.5f;
      if (src < 0.f) src = 0.f;
      const int i0 = min((int)src, Tin - 1), i1 = i0 + (i0 < Tin - 1 ? 1 : 0);
      const float l1 = src - (float)i0, l0 = 1.f - l1;
      const InT* p = in + ((size_t)b * in_rows_per_sample) * C + c;
      v = l0 * (float)p[(size_t)i0 * C] + l1 * (float)p[(size_t)i1 * C];
    }
    tile[i][tx] = v;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {                            // rows = channels, lanes = frames (coalesced writes)
    const int c = c0 + i, t = t0 + tx;
    if (c < C && t < Tout) out[((size_t)b * C + c) * Tout + t] = tile[tx][i];
  }
}

}  // namespace

extern "C" long long pt_w2v_frontend_workspace_bytes(int B) {
  if (B < 1) return -1;
  return (long long)B * ((long long)kChunks * kNS * 8 + (long long)kC0 * kK0 * 4 + (long long)kC0 * 4) + 64;
}

extern "C" int pt_w2v_frontend(const float* wave, int B, long long L, const float* conv0_w, const float* gn_gamma,
                               const float* gn_beta, void* out_bf16, long long out_rows_per_sample, void* workspace,
                               long long workspace_bytes, void* stream) {
  if (B <= 0) return 0;
  if (!wave || !conv0_w || !gn_gamma || !gn_beta || !out_bf16 || !workspace) return pt::fail("pt_w2v_frontend: null pointer");
  if (L < kK0) return pt::fail("pt_w2v_frontend: need at least %d samples (got %lld)", kK0, L);
  const long long T0 = (L - kK0) / kS0 + 1;
  if (out_rows_per_sample < T0) return pt::fail("pt_w2v_frontend: output holds %lld rows per utterance, %lld needed", out_rows_per_sample, T0);
  if (workspace_bytes < pt_w2v_frontend_workspace_bytes(B)) return pt::fail("pt_w2v_frontend: workspace too small");
  if ((reinterpret_cast<uintptr_t>(workspace) & 7) || (reinterpret_cast<uintptr_t>(out_bf16) & 3))
    return pt::fail("pt_w2v_frontend: workspace must be 8-byte aligned, output 4-byte aligned");
  if (B > 65535) return pt::fail("pt_w2v_frontend: at most 65535 utterances per call");
  double* partial = static_cast<double*>(workspace);
  float* weff = reinterpret_cast<float*>(partial + (size_t)B * kChunks * kNS);
  float* beff = weff + (size_t)B * kC0 * kK0;
  cudaStream_t st = (cudaStream_t)stream;
  w2v_stats_kernel<<<dim3(kChunks, B), kStatThreads, 0, st>>>(wave, L, T0, partial);
  w2v_fold_kernel<<<B, kC0, 0, st>>>(partial, L, T0, conv0_w, gn_gamma, gn_beta, weff, beff);
  w2v_conv0_kernel<<<dim3((unsigned)((T0 + kTT - 1) / kTT), B), 256, 0, st>>>(wave, L, T0, weff, beff, (__nv_bfloat16*)out_bf16,
                                                                              out_rows_per_sample);
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_w2v_interp(const void* in, int in_f32, long long in_rows_per_sample, int B, int Tin, int C, float* out,
                             int Tout, void* stream) {
  if (B <= 0 || Tout <= 0) return 0;
  if (!in || !out) return pt::fail("pt_w2v_interp: null pointer");
  if (Tin < 1 || C < 1 || in_rows_per_sample < Tin) return pt::fail("pt_w2v_interp: bad shape");
  if (B > 65535) return pt::fail("pt_w2v_interp: at most 65535 utterances per call");
  dim3 grid((Tout + 31) / 32, (C + 31) / 32, B);
  cudaStream_t st = (cudaStream_t)stream;
  if (in_f32) w2v_interp_kernel<float><<<grid, 256, 0, st>>>((const float*)in, in_rows_per_sample, Tin, C, out, Tout);
  else w2v_interp_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, in_rows_per_sample, Tin, C, out, Tout);
  PT_CUDA(cudaGetLastError());
  return 0;
}
