// probtalk_b200 -- HBM-bound layout / normalisation / sampler kernels (sm_100a).
//
// These are the non-GEMM ops of the path; each is a single pass over its operands with coalesced,
// vectorised accesses.  Reference functions replaced are cited at every entry point.
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kDModel = 512;

// PyTorch's linear (align_corners=False) source index rule, fp32 like the reference's F.interpolate
__device__ __forceinline__ float interp_linear(const float* __restrict__ row, int t, int Tin, int Tout) {
  if (Tin == Tout) return row[t];
  const float scale = (float)Tin / (float)Tout;
  float src = scale * ((float)t + 0.5f) - 0.5f;
  if (src < 0.f) src = 0.f;
  const int i0 = (int)src;
  const int i1 = i0 + (i0 < Tin - 1 ? 1 : 0);
  const float l1 = src - (float)i0, l0 = 1.f - l1;
  return l0 * row[i0] + l1 * row[i1];
}

// ---------------------------------------------------------------------------------------------
// pack: channel-first fp32 [B, C, T]  ->  channels-last bf16 rows [B, Tp, Cp] (halo H), via a 32x32 smem transpose
//   mode 0: src                      mode 1: src * mask            (predictor.py:425  input_poses = gt * mask)
//   mode 2: src*mask + interp(pred)*(1-mask)                        (refiner.py:526-529 blank)
struct PackArgs {
  const float* src; const float* mask; const float* pred;
  __nv_bfloat16* dst;
  int B, C, T, Tp, H, Cp, c_off, mode, Tpred;
  __nv_bfloat16* dst_lo;        // split arithmetic: the lo plane (same layout as dst) or null
};

__global__ void pack_rows_kernel(PackArgs a) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, t0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;   // (32, 8)
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, t = t0 + tx;
    float v = 0.f;
    if (c < a.C && t < a.T) {
      v = a.src[((size_t)b * a.C + c) * a.T + t];
      if (a.mode >= 1) {
        const float m = a.mask[(size_t)b * a.T + t];
        v *= m;
        if (a.mode == 2) v += interp_linear(a.pred + ((size_t)b * a.C + c) * a.Tpred, t, a.Tpred, a.T) * (1.f - m);
      }
    }
    tile[i][tx] = v;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, c = c0 + tx;
    if (t < a.T && c < a.C) {
      const size_t o = ((size_t)b * a.Tp + a.H + t) * a.Cp + a.c_off + c;
      const __nv_bfloat16 h = __float2bfloat16_rn(tile[tx][i]);
      a.dst[o] = h;
      if (a.dst_lo) a.dst_lo[o] = __float2bfloat16_rn(tile[tx][i] - __bfloat162float(h));
    }
  }
}

// Same operation, one block per (sample, 32-channel slab) over the whole time axis: every channel row is read as one
// contiguous run (T floats), the slab sits in shared memory, and the store side writes 64-byte channel runs per time
// step (8 lanes x 4 bf16).  ~22 independent loads per thread keep enough bytes in flight to approach HBM speed.
constexpr int kPackMaxT = 384;
template <int kIters>                                  // time steps per lane: T <= 32 * kIters
__global__ void __launch_bounds__(256, 5) pack_rows_slab_kernel(PackArgs a) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  extern __shared__ float slab[];                      // [32][T + 1]
  __shared__ float mask_s[kPackMaxT];
  const int b = blockIdx.y, c0 = blockIdx.x * 32;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int ld = a.T + 1;
  if (a.mode >= 1)
    for (int t = tid; t < a.T; t += 256) mask_s[t] = a.mask[(size_t)b * a.T + t];
  __syncthreads();
  for (int i = w; i < 32; i += 8) {                    // warp w: channels c0 + w, c0 + w + 8, ...
    const int c = c0 + i;
    if (c < a.C) {
      const float* src = a.src + ((size_t)b * a.C + c) * a.T;
      const float* prd = (a.mode == 2) ? a.pred + ((size_t)b * a.C + c) * a.Tpred : nullptr;
      // all of this lane's loads of the row are issued before the first use
      float v[kIters], pv[kIters];
#pragma unroll
      for (int k = 0; k < kIters; ++k) {
        const int t = lane + 32 * k;
        v[k] = (t < a.T) ? __ldg(src + t) : 0.f;
      }
      if (a.mode == 2) {
        if (a.Tpred == a.T) {
#pragma unroll
          for (int k = 0; k < kIters; ++k) {
            const int t = lane + 32 * k;
            pv[k] = (t < a.T) ? __ldg(prd + t) : 0.f;
          }
        } else {
#pragma unroll
          for (int k = 0; k < kIters; ++k) {
            const int t = lane + 32 * k;
            pv[k] = (t < a.T) ? interp_linear(prd, t, a.Tpred, a.T) : 0.f;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kIters; ++k) {
        const int t = lane + 32 * k;
        if (t < a.T) {
          float x = v[k];
          if (a.mode >= 1) {
            const float m = mask_s[t];
            x *= m;
            if (a.mode == 2) x += pv[k] * (1.f - m);
          }
          slab[i * ld + t] = x;
        }
      }
    } else {
      for (int t = lane; t < a.T; t += 32) slab[i * ld + t] = 0.f;
    }
  }
  __syncthreads();
  const int cg = (tid & 7) * 4;                        // 4 consecutive channels of the slab
  for (int t = tid >> 3; t < a.T; t += 32) {
    const size_t off = ((size_t)b * a.Tp + a.H + t) * a.Cp + a.c_off + c0 + cg;
    __nv_bfloat16* d = a.dst + off;
    const float v0 = slab[(cg + 0) * ld + t], v1 = slab[(cg + 1) * ld + t];
    const float v2 = slab[(cg + 2) * ld + t], v3 = slab[(cg + 3) * ld + t];
    uint2 hi, lo;
    pt::split_bf16x2(v0, v1, hi.x, lo.x);
    pt::split_bf16x2(v2, v3, hi.y, lo.y);
    if (c0 + cg + 3 < a.C && ((reinterpret_cast<uintptr_t>(d) & 7) == 0)) {
      *reinterpret_cast<uint2*>(d) = hi;
      if (a.dst_lo) *reinterpret_cast<uint2*>(a.dst_lo + off) = lo;
    } else {
      const uint32_t h4[2] = {hi.x, hi.y}, l4[2] = {lo.x, lo.y};
      for (int j = 0; j < 4; ++j) {
        if (c0 + cg + j < a.C) {
          reinterpret_cast<uint16_t*>(d)[j] = (uint16_t)(h4[j >> 1] >> (16 * (j & 1)));
          if (a.dst_lo) reinterpret_cast<uint16_t*>(a.dst_lo + off)[j] = (uint16_t)(l4[j >> 1] >> (16 * (j & 1)));
        }
      }
    }
  }
}

// mask channels: mode 0 -> one channel = mask (predictornet.py:85 cat([motion, mask]));
//                mode 1 -> two channels one_hot(mask.long(), 2) (RTransformer.py:218): ch0 = (mask==0), ch1 = (mask==1)
__global__ void pack_mask_kernel(const float* __restrict__ mask, __nv_bfloat16* __restrict__ dst, int B, int T, int Tp,
                                 int H, int Cp, int c_off, int mode) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T) return;
  const int b = i / T, t = i % T;
  const float m = mask[i];
  __nv_bfloat16* p = dst + ((size_t)b * Tp + H + t) * Cp + c_off;
  if (mode == 0) {
    p[0] = __float2bfloat16_rn(m);
  } else {
    const long long k = (long long)m;
    p[0] = __float2bfloat16_rn(k == 0 ? 1.f : 0.f);
    p[1] = __float2bfloat16_rn(k == 1 ? 1.f : 0.f);
  }
}

// out[b,c,t] = known[b,c,t] * mask^(1 or 2) + interp(pred)[b,c,t] * (1 - mask)        (all channel-first fp32)
//   predictor.py:445-446: pred = interpolate(pred, T); out = (gt*mask) * mask + pred * (1 - mask)   -> mask_pow = 2
//   refiner.py:553:       out = gt * mask + pred * (1 - mask)                                       -> mask_pow = 1
__global__ void blend_kernel(const float* __restrict__ known, const float* __restrict__ mask,
                             const float* __restrict__ pred, float* __restrict__ out, int B, int C, int T, int Tpred,
                             int mask_pow) {
  const long long n = (long long)B * C * T;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i % T);
    const long long bc = i / T;
    const int b = (int)(bc / C);
    const float m = mask[(size_t)b * T + t];
    float k = known[i] * m;
    if (mask_pow == 2) k *= m;
    out[i] = k + interp_linear(pred + bc * Tpred, t, Tpred, T) * (1.f - m);
  }
}

// resampling path (pred has Tpred != T frames): one warp per (sample, channel) row, lanes stride the time axis, so the
// row bases cost one division per warp and the two interpolation taps of neighbouring lanes share cache lines
__global__ void __launch_bounds__(256) blend_rows_kernel(const float* __restrict__ known, const float* __restrict__ mask,
                                                         const float* __restrict__ pred, float* __restrict__ out,
                                                         unsigned rows, unsigned C, int T, int Tpred, int mask_pow) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  const unsigned row = blockIdx.x * 8u + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const unsigned b = row / C;
  const float* k = known + (size_t)row * T;
  const float* m = mask + (size_t)b * T;
  const float* p = pred + (size_t)row * Tpred;
  float* o = out + (size_t)row * T;
  for (int t0 = 0; t0 < T; t0 += 128) {
    float kv[4], mv[4], pv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + lane + 32 * j;
      kv[j] = (t < T) ? __ldg(k + t) : 0.f;
      mv[j] = (t < T) ? __ldg(m + t) : 0.f;
      pv[j] = (t < T) ? interp_linear(p, t, Tpred, T) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + lane + 32 * j;
      if (t < T) {
        float x = kv[j] * mv[j];
        if (mask_pow == 2) x *= mv[j];
        o[t] = x + pv[j] * (1.f - mv[j]);
      }
    }
  }
}

// fast path: no resampling (Tpred == T), T a multiple of 4, 16-byte aligned rows: one float4 per thread, 32-bit indexing
__global__ void blend_vec4_kernel(const float4* __restrict__ known, const float* __restrict__ mask,
                                  const float4* __restrict__ pred, float4* __restrict__ out, unsigned n4, unsigned T4,
                                  unsigned C, int mask_pow) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
    const unsigned bc = i / T4, t4 = i - bc * T4, b = bc / C;
    const float4 m = __ldg(reinterpret_cast<const float4*>(mask) + (size_t)b * T4 + t4);
    const float4 k = __ldg(known + i), p = __ldg(pred + i);
    float4 o;
    if (mask_pow == 2) {
      o.x = k.x * m.x * m.x + p.x * (1.f - m.x); o.y = k.y * m.y * m.y + p.y * (1.f - m.y);
      o.z = k.z * m.z * m.z + p.z * (1.f - m.z); o.w = k.w * m.w * m.w + p.w * (1.f - m.w);
    } else {
      o.x = k.x * m.x + p.x * (1.f - m.x); o.y = k.y * m.y + p.y * (1.f - m.y);
      o.z = k.z * m.z + p.z * (1.f - m.z); o.w = k.w * m.w + p.w * (1.f - m.w);
    }
    out[i] = o;
  }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm over 512 channels, one warp per row (16 values per lane), fp32 two-pass statistics.
template <typename InT>
__device__ __forceinline__ void load_row16(const InT* row, int lane, float (&v)[16]) {
  if constexpr (sizeof(InT) == 4) {
    const float4* p = reinterpret_cast<const float4*>(row);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float4 x = __ldg(p + lane + 32 * i);
      v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
    }
  } else {
    const uint2* p = reinterpret_cast<const uint2*>(row);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint2 x = __ldg(p + lane + 32 * i);
      __nv_bfloat162 h0 = *reinterpret_cast<__nv_bfloat162*>(&x.x), h1 = *reinterpret_cast<__nv_bfloat162*>(&x.y);
      v[4 * i] = __bfloat162float(h0.x); v[4 * i + 1] = __bfloat162float(h0.y);
      v[4 * i + 2] = __bfloat162float(h1.x); v[4 * i + 3] = __bfloat162float(h1.y);
    }
  }
}
// lane's channels: c = 4*(lane + 32*i) + j
__device__ __forceinline__ void ln_stats(const float (&v)[16], float& mean, float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  mean = s * (1.f / kDModel);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { const float d = v[i] - mean; q = fmaf(d, d, q); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  rstd = rsqrtf(q * (1.f / kDModel) + 1e-5f);
}

// nn.LayerNorm (norm1/2/3 of DeepSupervisionLayer, PTransformer.py:329-331): fp32 in -> fp32 and/or bf16 out
template <typename InT>
__global__ void ln_kernel(const InT* __restrict__ in, const __nv_bfloat16* __restrict__ res,
                          const __nv_bfloat16* __restrict__ res_lo,
                          const float* __restrict__ gamma, const float* __restrict__ beta,
                          float* __restrict__ out32, __nv_bfloat16* __restrict__ out16, __nv_bfloat16* __restrict__ out16_lo,
                          int M) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  float v[16];
  load_row16(in + (size_t)row * kDModel, lane, v);
  if (res) {                                   // x + sublayer(x): the residual add of the post-norm layer
    float r[16];
    load_row16(res + (size_t)row * kDModel, lane, r);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += r[i];
    if (res_lo) {                              // split residual: the lo plane carries the next 8 mantissa bits
      load_row16(res_lo + (size_t)row * kDModel, lane, r);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] += r[i];
    }
  }
  float mean, rstd;
  ln_stats(v, mean, rstd);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = 4 * (lane + 32 * i);
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c)), bb = __ldg(reinterpret_cast<const float4*>(beta + c));
    const float y0 = (v[4 * i] - mean) * rstd * g.x + bb.x, y1 = (v[4 * i + 1] - mean) * rstd * g.y + bb.y;
    const float y2 = (v[4 * i + 2] - mean) * rstd * g.z + bb.z, y3 = (v[4 * i + 3] - mean) * rstd * g.w + bb.w;
    if (out32) *reinterpret_cast<float4*>(out32 + (size_t)row * kDModel + c) = make_float4(y0, y1, y2, y3);
    if (out16) {
      uint2 hi, lo;
      pt::split_bf16x2(y0, y1, hi.x, lo.x);
      pt::split_bf16x2(y2, y3, hi.y, lo.y);
      *reinterpret_cast<uint2*>(out16 + (size_t)row * kDModel + c) = hi;
      if (out16_lo) *reinterpret_cast<uint2*>(out16_lo + (size_t)row * kDModel + c) = lo;
    }
  }
}

// nn.LayerNorm over C = 128 * kVec channels of bf16 rows (+ bf16 residual), one warp per row: the 768-wide norms of the
// wav2vec2 encoder (feature projection, encoder.layer_norm, layer_norm / final_layer_norm of every layer)
template <int kVec>
__global__ void ln_rows_kernel(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ res,
                               const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                               __nv_bfloat16* __restrict__ out, int M) {
  constexpr int C = 128 * kVec;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  float v[4 * kVec];
  const uint2* p = reinterpret_cast<const uint2*>(in + (size_t)row * C);
  const uint2* pr = res ? reinterpret_cast<const uint2*>(res + (size_t)row * C) : nullptr;
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const uint2 x = __ldg(p + lane + 32 * i);
    const __nv_bfloat162 h0 = *reinterpret_cast<const __nv_bfloat162*>(&x.x), h1 = *reinterpret_cast<const __nv_bfloat162*>(&x.y);
    v[4 * i] = __bfloat162float(h0.x); v[4 * i + 1] = __bfloat162float(h0.y);
    v[4 * i + 2] = __bfloat162float(h1.x); v[4 * i + 3] = __bfloat162float(h1.y);
    if (pr) {
      const uint2 y = __ldg(pr + lane + 32 * i);
      const __nv_bfloat162 g0 = *reinterpret_cast<const __nv_bfloat162*>(&y.x), g1 = *reinterpret_cast<const __nv_bfloat162*>(&y.y);
      v[4 * i] += __bfloat162float(g0.x); v[4 * i + 1] += __bfloat162float(g0.y);
      v[4 * i + 2] += __bfloat162float(g1.x); v[4 * i + 3] += __bfloat162float(g1.y);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4 * kVec; ++i) s += v[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s * (1.f / C);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 4 * kVec; ++i) { const float d = v[i] - mean; q = fmaf(d, d, q); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q * (1.f / C) + eps);
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int c = 4 * (lane + 32 * i);
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c)), bb = __ldg(reinterpret_cast<const float4*>(beta + c));
    const float y0 = (v[4 * i] - mean) * rstd * g.x + bb.x, y1 = (v[4 * i + 1] - mean) * rstd * g.y + bb.y;
    const float y2 = (v[4 * i + 2] - mean) * rstd * g.z + bb.z, y3 = (v[4 * i + 3] - mean) * rstd * g.w + bb.w;
    *reinterpret_cast<uint2*>(out + (size_t)row * C + c) = make_uint2(pt::pack_bf16x2(y0, y1), pt::pack_bf16x2(y2, y3));
  }
}

// StylizationBlock_V3 'sb' branch up to the output Linear (PTransformer.py:134-139):
//   out = SiLU( LN(h) * (1 + scale[label]) + shift[label] ),  mod = [n_classes, 1024] = [scale | shift]
template <typename InT>
__global__ void ln_mod_silu_kernel(const InT* __restrict__ in, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, const float* __restrict__ mod,
                                   const int* __restrict__ labels, int rows_per_sample,
                                   __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo, int M) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  float v[16];
  load_row16(in + (size_t)row * kDModel, lane, v);
  float mean, rstd;
  ln_stats(v, mean, rstd);
  const float* m = mod + (size_t)__ldg(labels + row / rows_per_sample) * (2 * kDModel);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = 4 * (lane + 32 * i);
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c)), bb = __ldg(reinterpret_cast<const float4*>(beta + c));
    const float4 sc = __ldg(reinterpret_cast<const float4*>(m + c)), sh = __ldg(reinterpret_cast<const float4*>(m + kDModel + c));
    float y[4] = {(v[4 * i] - mean) * rstd * g.x + bb.x, (v[4 * i + 1] - mean) * rstd * g.y + bb.y,
                  (v[4 * i + 2] - mean) * rstd * g.z + bb.z, (v[4 * i + 3] - mean) * rstd * g.w + bb.w};
    const float s4[4] = {sc.x, sc.y, sc.z, sc.w}, h4[4] = {sh.x, sh.y, sh.z, sh.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float u = y[j] * (1.f + s4[j]) + h4[j];
      y[j] = u / (1.f + __expf(-u));
    }
    uint2 hi, lo;
    pt::split_bf16x2(y[0], y[1], hi.x, lo.x);
    pt::split_bf16x2(y[2], y[3], hi.y, lo.y);
    *reinterpret_cast<uint2*>(out + (size_t)row * kDModel + c) = hi;
    if (out_lo) *reinterpret_cast<uint2*>(out_lo + (size_t)row * kDModel + c) = lo;
  }
}

// tok_emb lookup + positional encodings (PTransformer.py:659-679): x = float32( emb[tok] + pe64[n] ), n = row % N
__global__ void embed_tokens_kernel(const int* __restrict__ tokens, const float* __restrict__ emb,
                                    const double* __restrict__ pe64, float* __restrict__ out32,
                                    __nv_bfloat16* __restrict__ out16, __nv_bfloat16* __restrict__ out16_lo, int M, int N,
                                    int n_emb, int* bad) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  int tok = __ldg(tokens + row);
  if (tok < 0 || tok >= n_emb) { if (lane == 0 && bad) atomicAdd(bad, 1); tok = 0; }
  const float* e = emb + (size_t)tok * kDModel;
  const double* p = pe64 + (size_t)(row % N) * kDModel;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = 4 * (lane + 32 * i);
    const float4 x = __ldg(reinterpret_cast<const float4*>(e + c));
    const double2 p0 = __ldg(reinterpret_cast<const double2*>(p + c)), p1 = __ldg(reinterpret_cast<const double2*>(p + c + 2));
    const float y0 = (float)((double)x.x + p0.x), y1 = (float)((double)x.y + p0.y);
    const float y2 = (float)((double)x.z + p1.x), y3 = (float)((double)x.w + p1.y);
    if (out32) *reinterpret_cast<float4*>(out32 + (size_t)row * kDModel + c) = make_float4(y0, y1, y2, y3);
    if (out16) {
      uint2 hi, lo;
      pt::split_bf16x2(y0, y1, hi.x, lo.x);
      pt::split_bf16x2(y2, y3, hi.y, lo.y);
      *reinterpret_cast<uint2*>(out16 + (size_t)row * kDModel + c) = hi;
      if (out16_lo) *reinterpret_cast<uint2*>(out16_lo + (size_t)row * kDModel + c) = lo;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// One MaskGIT refinement step (PTransformer.sample_good :603-638 + mask_by_random_topk :696-704), one CTA per
// sample.  The two random draws are inputs: u (uniform, categorical by inverse CDF in index order) and gumbel.
constexpr int kMaxTokens = 1024;
constexpr int kCodes = 128;

__global__ void __launch_bounds__(128) maskgit_step_kernel(const float* __restrict__ logits, const int* __restrict__ cur,
                                                            const float* __restrict__ u, const float* __restrict__ gumbel,
                                                            int* __restrict__ next, int N, int mask_id, int base_len,
                                                            float temperature) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  __shared__ float conf[kMaxTokens];
  __shared__ int samp[kMaxTokens];
  __shared__ float pbuf[4][kCodes];
  __shared__ int unknown_count;
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) unknown_count = 0;
  __syncthreads();
  for (int n = warp; n < N; n += 4) {
    const size_t tok = (size_t)b * N + n;
    const float4 l = __ldg(reinterpret_cast<const float4*>(logits + tok * kCodes) + lane);
    float mx = fmaxf(fmaxf(l.x, l.y), fmaxf(l.z, l.w));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float e0 = expf(l.x - mx), e1 = expf(l.y - mx), e2 = expf(l.z - mx), e3 = expf(l.w - mx);
    pbuf[warp][4 * lane] = e0; pbuf[warp][4 * lane + 1] = e1; pbuf[warp][4 * lane + 2] = e2; pbuf[warp][4 * lane + 3] = e3;
    __syncwarp();
    const int c = __ldg(cur + tok);
    const bool unknown = (c == mask_id);
    int id = c;
    float total = 0.f;
    if (lane == 0) {
      // sequential fp32 CDF in index order (the oracle's rule): first k with cdf[k] > u * total
      float cdf = 0.f;
      for (int k = 0; k < kCodes; ++k) cdf += pbuf[warp][k];
      total = cdf;
      if (unknown) {
        const float thr = __ldg(u + tok) * total;
        cdf = 0.f;
        int k = 0;
        for (; k < kCodes - 1; ++k) { cdf += pbuf[warp][k]; if (cdf > thr) break; }
        id = k;
      }
    }
    id = __shfl_sync(0xffffffffu, id, 0);
    total = __shfl_sync(0xffffffffu, total, 0);
    if (lane == 0) {
      samp[n] = id;
      float cf = INFINITY;                                   // known tokens: +inf confidence (:617)
      if (unknown) {
        cf = logf(pbuf[warp][id] / total) + temperature * __ldg(gumbel + tok);
        atomicAdd(&unknown_count, 1);
      }
      conf[n] = cf;
    }
    __syncwarp();
  }
  __syncthreads();
  int mask_len = min(unknown_count - 1, base_len);
  mask_len = max(mask_len, 0);
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const float cf = conf[n];
    int le = 0;
    for (int j = 0; j < N; ++j) le += (conf[j] <= cf) ? 1 : 0;
    // conf[n] < sorted(conf)[mask_len]  <=>  #{j : conf[j] <= conf[n]} <= mask_len
    next[(size_t)b * N + n] = (le <= mask_len) ? mask_id : samp[n];
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
extern "C" int pt_pack_rows(const float* src, const float* mask, const float* pred, int pred_T, void* dst_bf16,
                            void* dst_lo, int B, int C, int T, int Tp, int H, int Cp, int c_off, int mode, void* stream) {
  if (B <= 0 || C <= 0 || T <= 0) return 0;
  if (!src || !dst_bf16) return pt::fail("pt_pack_rows: null pointer");
  if (mode < 0 || mode > 2) return pt::fail("pt_pack_rows: bad mode %d", mode);
  if (mode >= 1 && !mask) return pt::fail("pt_pack_rows: mode %d needs a mask", mode);
  if (mode == 2 && (!pred || pred_T <= 0)) return pt::fail("pt_pack_rows: mode 2 needs pred");
  if (c_off + C > Cp || H + T > Tp) return pt::fail("pt_pack_rows: destination too small");
  PackArgs a{src, mask, pred, (__nv_bfloat16*)dst_bf16, B, C, T, Tp, H, Cp, c_off, mode, pred_T, (__nv_bfloat16*)dst_lo};
  if (T <= kPackMaxT && B <= 65535) {
    const size_t smem = (size_t)32 * (T + 1) * sizeof(float);   // <= 48 KB
    if (T <= 192) PT_CUDA(pt::launch_pdl(pack_rows_slab_kernel<6>, dim3((C + 31) / 32, B), dim3(256), smem, (cudaStream_t)stream, 1, a));
    else PT_CUDA(pt::launch_pdl(pack_rows_slab_kernel<12>, dim3((C + 31) / 32, B), dim3(256), smem, (cudaStream_t)stream, 1, a));
  } else {
    dim3 grid((T + 31) / 32, (C + 31) / 32, B), block(32, 8);
    pack_rows_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(a);
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_pack_mask(const float* mask, void* dst_bf16, int B, int T, int Tp, int H, int Cp, int c_off,
                            int onehot, void* stream) {
  if (B <= 0 || T <= 0) return 0;
  if (!mask || !dst_bf16) return pt::fail("pt_pack_mask: null pointer");
  const int n = B * T;
  PT_CUDA(pt::launch_pdl(pack_mask_kernel, dim3((n + 255) / 256), dim3(256), 0, (cudaStream_t)stream, 1, mask,
                         (__nv_bfloat16*)dst_bf16, B, T, Tp, H, Cp, c_off, onehot ? 1 : 0));
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_interp_blend(const float* known, const float* mask, const float* pred, int pred_T, float* out, int B,
                               int C, int T, int mask_pow, void* stream) {
  const long long n = (long long)B * C * T;
  if (n <= 0) return 0;
  if (!known || !mask || !pred || !out) return pt::fail("pt_interp_blend: null pointer");
  if (mask_pow != 1 && mask_pow != 2) return pt::fail("pt_interp_blend: mask_pow must be 1 or 2");
  const bool aligned = ((reinterpret_cast<uintptr_t>(known) | reinterpret_cast<uintptr_t>(mask) |
                         reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (pred_T == T && T % 4 == 0 && aligned && n / 4 < (1LL << 31)) {
    const unsigned n4 = (unsigned)(n / 4);
    const unsigned blocks = (n4 + 255) / 256;          // one float4 per thread: no loop-carried latency
    PT_CUDA(pt::launch_pdl(blend_vec4_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, 1, (const float4*)known, mask,
                           (const float4*)pred, (float4*)out, n4, (unsigned)(T / 4), (unsigned)C, mask_pow));
  } else if ((long long)B * C < (1LL << 31)) {
    const unsigned rows = (unsigned)B * (unsigned)C;
    PT_CUDA(pt::launch_pdl(blend_rows_kernel, dim3((rows + 7) / 8), dim3(256), 0, (cudaStream_t)stream, 1, known, mask, pred, out,
                           rows, (unsigned)C, T, pred_T, mask_pow));
  } else {
    const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
    blend_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(known, mask, pred, out, B, C, T, pred_T, mask_pow);
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_layernorm512(const void* in, int in_f32, const void* residual_bf16, const void* residual_lo,
                               const float* gamma, const float* beta, float* out_f32, void* out_bf16, void* out_lo, int M,
                               void* stream) {
  if (M <= 0) return 0;
  if (!in || !gamma || !beta || (!out_f32 && !out_bf16)) return pt::fail("pt_layernorm512: null pointer");
  if ((residual_lo && !residual_bf16) || (out_lo && !out_bf16)) return pt::fail("pt_layernorm512: a lo plane needs its hi plane");
  const __nv_bfloat16* res = (const __nv_bfloat16*)residual_bf16;
  const __nv_bfloat16* rlo = (const __nv_bfloat16*)residual_lo;
  if (in_f32)
    ln_kernel<float><<<(M + 7) / 8, 256, 0, (cudaStream_t)stream>>>((const float*)in, res, rlo, gamma, beta, out_f32, (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, M);
  else
    ln_kernel<__nv_bfloat16><<<(M + 7) / 8, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, res, rlo, gamma, beta, out_f32, (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, M);
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_layernorm_rows(const void* in_bf16, const void* residual_bf16, const float* gamma, const float* beta,
                                 float eps, void* out_bf16, int M, int C, void* stream) {
  if (M <= 0) return 0;
  if (!in_bf16 || !gamma || !beta || !out_bf16) return pt::fail("pt_layernorm_rows: null pointer");
  if (C < 128 || C > 1024 || (C & 127)) return pt::fail("pt_layernorm_rows: C must be a multiple of 128 in [128, 1024] (got %d)", C);
  const __nv_bfloat16* in = (const __nv_bfloat16*)in_bf16;
  const __nv_bfloat16* res = (const __nv_bfloat16*)residual_bf16;
  __nv_bfloat16* out = (__nv_bfloat16*)out_bf16;
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = (M + 7) / 8;
  switch (C / 128) {
    case 1: ln_rows_kernel<1><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 2: ln_rows_kernel<2><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 3: ln_rows_kernel<3><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 4: ln_rows_kernel<4><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 5: ln_rows_kernel<5><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 6: ln_rows_kernel<6><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    case 7: ln_rows_kernel<7><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
    default: ln_rows_kernel<8><<<blocks, 256, 0, st>>>(in, res, gamma, beta, eps, out, M); break;
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_ln_mod_silu512(const void* in, int in_f32, const float* gamma, const float* beta, const float* mod,
                                 const int* labels, int rows_per_sample, void* out_bf16, void* out_lo, int M, void* stream) {
  if (M <= 0) return 0;
  if (!in || !gamma || !beta || !mod || !labels || !out_bf16) return pt::fail("pt_ln_mod_silu512: null pointer");
  if (rows_per_sample <= 0) return pt::fail("pt_ln_mod_silu512: rows_per_sample must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  if (in_f32)
    ln_mod_silu_kernel<float><<<(M + 7) / 8, 256, 0, st>>>((const float*)in, gamma, beta, mod, labels, rows_per_sample, (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, M);
  else
    ln_mod_silu_kernel<__nv_bfloat16><<<(M + 7) / 8, 256, 0, st>>>((const __nv_bfloat16*)in, gamma, beta, mod, labels, rows_per_sample, (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, M);
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_embed_tokens512(const int* tokens, const float* emb, int n_emb, const double* pe64, int N,
                                  float* out_f32, void* out_bf16, void* out_lo, int M, int* bad_count, void* stream) {
  if (M <= 0) return 0;
  if (!tokens || !emb || !pe64 || (!out_f32 && !out_bf16)) return pt::fail("pt_embed_tokens512: null pointer");
  if (N <= 0) return pt::fail("pt_embed_tokens512: N must be positive");
  PT_CUDA(pt::launch_pdl(embed_tokens_kernel, dim3((M + 7) / 8), dim3(256), 0, (cudaStream_t)stream, 1, tokens, emb, pe64, out_f32,
                         (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, M, N, n_emb, bad_count));
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_maskgit_step(const float* logits, const int* cur_ids, const float* u, const float* gumbel,
                               int* next_ids, int B, int N, int n_codes, int mask_id, int base_mask_len,
                               float temperature, void* stream) {
  if (B <= 0 || N <= 0) return 0;
  if (!logits || !cur_ids || !u || !gumbel || !next_ids) return pt::fail("pt_maskgit_step: null pointer");
  if (n_codes != kCodes) return pt::fail("pt_maskgit_step: only %d codes are built (got %d)", kCodes, n_codes);
  if (N > kMaxTokens) return pt::fail("pt_maskgit_step: at most %d tokens per sample (got %d)", kMaxTokens, N);
  PT_CUDA(pt::launch_pdl(maskgit_step_kernel, dim3(B), dim3(128), 0, (cudaStream_t)stream, 1, logits, cur_ids, u, gumbel, next_ids,
                         N, mask_id, base_mask_len, temperature));
  PT_CUDA(cudaGetLastError());
  return 0;
}
