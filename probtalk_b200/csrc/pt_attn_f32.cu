// probtalk_b200 -- fp32 attention core for the tolerance-tight ("bf16x3") arithmetic mode (sm_100a).
//
// softmax(q k^T * scale) v  with fp32 inputs, fp32 CUDA-core arithmetic throughout (no tensor core: a bf16 P or Q
// would put 2^-9 relative error into every score, 250x the split-bf16 GEMMs around it), and a split-bf16 (hi, lo)
// output that feeds the next split GEMM.  Replaces the scaled-dot-product part of nn.MultiheadAttention as called at
// nets/inpainting/PTransformer.py:340 (self) and :350 (cross), exactly like pt_attention, at fp32 fidelity.
//
// One CTA per (sample, head): K and V of the head are staged in shared memory once (K rows padded to 65 floats so that
// lanes = keys read conflict-free), eight warps walk the queries four at a time:
//   scores   lane = key (j = lane + 32 i), the four queries' values of dimension d arrive as one broadcast LDS.128
//   softmax  fp32, expf, normalised before P.V like the reference
//   P.V      lane = output dimension (d, d + 32), the four queries' probabilities of key j as one broadcast LDS.128
// Shared-memory traffic, not FLOPs, bounds it (about 10 LDS per 24-32 FMA); it is a fidelity mode, the throughput
// path is attention_tc_kernel (pt_attn_tc.cu).
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kHeadDim = 64;
constexpr int kWarps = 8;
constexpr int kQB = 4;                       // queries per warp pass
constexpr int kKPitch = kHeadDim + 1;        // padded K row

struct AttnF32Args {
  const float* q; const float* k; const float* v;
  long long qs, ks, vs, os;                  // row strides in elements
  __nv_bfloat16* out_hi; __nv_bfloat16* out_lo;
  int heads, Nq, Nk;
  float scale;
};

template <int NI>                            // ceil(Nk / 32)
__global__ void __launch_bounds__(kWarps * 32) attention_f32_kernel(AttnF32Args a) {
  extern __shared__ __align__(16) float sm[];
  const int Nk = a.Nk, Nq = a.Nq;
  float* Ks = sm;                                          // [Nk][65]
  float* Vs = Ks + ((Nk * kKPitch + 3) & ~3);              // [Nk][64]
  float* Qs = Vs + Nk * kHeadDim;                          // [warps][64][4]
  float* Ps = Qs + kWarps * kHeadDim * kQB;                // [warps][NI * 32][4]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x / a.heads, h = blockIdx.x % a.heads;
  const float* kb = a.k + (long long)b * Nk * a.ks + h * kHeadDim;
  const float* vb = a.v + (long long)b * Nk * a.vs + h * kHeadDim;
  for (int i = tid; i < Nk * (kHeadDim / 4); i += kWarps * 32) {
    const int r = i >> 4, c = (i & 15) * 4;
    const float4 kk = __ldg(reinterpret_cast<const float4*>(kb + (long long)r * a.ks + c));
    const float4 vv = __ldg(reinterpret_cast<const float4*>(vb + (long long)r * a.vs + c));
    float* kd = Ks + r * kKPitch + c;
    kd[0] = kk.x; kd[1] = kk.y; kd[2] = kk.z; kd[3] = kk.w;
    *reinterpret_cast<float4*>(Vs + r * kHeadDim + c) = vv;
  }
  __syncthreads();
  float* qs = Qs + warp * kHeadDim * kQB;
  float* ps = Ps + warp * NI * 32 * kQB;
  const float* qb = a.q + (long long)b * Nq * a.qs + h * kHeadDim;
  for (int q0 = warp * kQB; q0 < Nq; q0 += kWarps * kQB) {
#pragma unroll
    for (int qi = 0; qi < kQB; ++qi) {
      const int qr = q0 + qi;
      float x0 = 0.f, x1 = 0.f;
      if (qr < Nq) {
        x0 = __ldg(qb + (long long)qr * a.qs + lane) * a.scale;           // q * scaling, then q k^T (torch MHA order)
        x1 = __ldg(qb + (long long)qr * a.qs + lane + 32) * a.scale;
      }
      qs[lane * kQB + qi] = x0;
      qs[(lane + 32) * kQB + qi] = x1;
    }
    __syncwarp();
    float acc[kQB][NI];
#pragma unroll
    for (int qi = 0; qi < kQB; ++qi)
#pragma unroll
      for (int i = 0; i < NI; ++i) acc[qi][i] = 0.f;
    const float* krow[NI];
#pragma unroll
    for (int i = 0; i < NI; ++i) krow[i] = Ks + min(lane + 32 * i, Nk - 1) * kKPitch;
#pragma unroll 8
    for (int d = 0; d < kHeadDim; ++d) {
      const float4 qv = *reinterpret_cast<const float4*>(qs + d * kQB);
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const float kv = krow[i][d];
        acc[0][i] = fmaf(qv.x, kv, acc[0][i]); acc[1][i] = fmaf(qv.y, kv, acc[1][i]);
        acc[2][i] = fmaf(qv.z, kv, acc[2][i]); acc[3][i] = fmaf(qv.w, kv, acc[3][i]);
      }
    }
#pragma unroll
    for (int qi = 0; qi < kQB; ++qi) {
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < NI; ++i) if (lane + 32 * i < Nk) mx = fmaxf(mx, acc[qi][i]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const float e = (lane + 32 * i < Nk) ? expf(acc[qi][i] - mx) : 0.f;
        acc[qi][i] = e;
        sum += e;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      const float inv = 1.f / sum;
#pragma unroll
      for (int i = 0; i < NI; ++i) ps[(lane + 32 * i) * kQB + qi] = acc[qi][i] * inv;
    }
    __syncwarp();
    float o0[kQB] = {0.f, 0.f, 0.f, 0.f}, o1[kQB] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int j = 0; j < Nk; ++j) {
      const float4 pv = *reinterpret_cast<const float4*>(ps + j * kQB);
      const float v0 = Vs[j * kHeadDim + lane], v1 = Vs[j * kHeadDim + lane + 32];
      o0[0] = fmaf(pv.x, v0, o0[0]); o1[0] = fmaf(pv.x, v1, o1[0]);
      o0[1] = fmaf(pv.y, v0, o0[1]); o1[1] = fmaf(pv.y, v1, o1[1]);
      o0[2] = fmaf(pv.z, v0, o0[2]); o1[2] = fmaf(pv.z, v1, o1[2]);
      o0[3] = fmaf(pv.w, v0, o0[3]); o1[3] = fmaf(pv.w, v1, o1[3]);
    }
#pragma unroll
    for (int qi = 0; qi < kQB; ++qi) {
      const int qr = q0 + qi;
      if (qr < Nq) {
        const long long off = ((long long)b * Nq + qr) * a.os + h * kHeadDim + lane;
        const __nv_bfloat16 h0 = __float2bfloat16_rn(o0[qi]), h1 = __float2bfloat16_rn(o1[qi]);
        a.out_hi[off] = h0; a.out_hi[off + 32] = h1;
        if (a.out_lo) {
          a.out_lo[off] = __float2bfloat16_rn(o0[qi] - __bfloat162float(h0));
          a.out_lo[off + 32] = __float2bfloat16_rn(o1[qi] - __bfloat162float(h1));
        }
      }
    }
    __syncwarp();
  }
}

template <int NI>
int launch_f32(const AttnF32Args& a, int B, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024) PT_SMEM_OPT_IN(attention_f32_kernel<NI>, smem);
  attention_f32_kernel<NI><<<B * a.heads, kWarps * 32, smem, st>>>(a);
  PT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" int pt_attention_f32(const float* q, long long q_stride, const float* k, long long k_stride, const float* v,
                                long long v_stride, void* out_bf16, void* out_lo, long long out_stride, int B, int heads,
                                int head_dim, int Nq, int Nk, float scale, void* stream) {
  if (B <= 0 || heads <= 0 || Nq <= 0) return 0;
  if (!q || !k || !v || !out_bf16) return pt::fail("pt_attention_f32: null pointer");
  if (head_dim != kHeadDim) return pt::fail("pt_attention_f32: head_dim must be %d (got %d)", kHeadDim, head_dim);
  if (Nk < 1 || Nk > 256) return pt::fail("pt_attention_f32: 1 <= Nk <= 256 keys are built (got %d)", Nk);
  if (((reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15) || (k_stride & 3) || (v_stride & 3))
    return pt::fail("pt_attention_f32: k / v rows must be 16-byte aligned");
  AttnF32Args a{q, k, v, q_stride, k_stride, v_stride, out_stride, (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo,
                heads, Nq, Nk, scale};
  const int ni = (Nk + 31) / 32;
  const size_t smem = sizeof(float) * (size_t)(((Nk * kKPitch + 3) & ~3) + Nk * kHeadDim + kWarps * kHeadDim * kQB +
                                               kWarps * ni * 32 * kQB);
  cudaStream_t st = (cudaStream_t)stream;
  switch (ni) {
    case 1: return launch_f32<1>(a, B, smem, st);
    case 2: return launch_f32<2>(a, B, smem, st);
    case 3: return launch_f32<3>(a, B, smem, st);
    case 4: return launch_f32<4>(a, B, smem, st);
    case 5: return launch_f32<5>(a, B, smem, st);
    case 6: return launch_f32<6>(a, B, smem, st);
    case 7: return launch_f32<7>(a, B, smem, st);
    default: return launch_f32<8>(a, B, smem, st);
  }
}
