// probtalk_b200 -- small-sequence multi-head attention (sm_100a).
//
// Replaces the scaled-dot-product core of nn.MultiheadAttention as called by DeepSupervisionLayer
// (reference nets/inpainting/PTransformer.py:340 self-attention, :350 cross-attention): 8 heads x 64,
// no masks, softmax over all keys.  Sequences on this path are tiny (88 tokens x 22 memory steps for the
// predictor, 180 x 180 for the refiner), so one CTA owns one (sample, head): K, V and Q live in shared
// memory, each warp owns 16-query blocks and walks the keys in blocks of 32 with an online softmax.
// Tensor-core work uses warp-level mma.sync m16n8k16 (bf16 in, fp32 accumulate): at <= 192 keys per head a
// 128-row tcgen05 tile would be mostly padding; attention is < 10 % of the path's FLOPs.
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"
#include <stdlib.h>

// tcgen05 implementation for up to 256 keys (pt_attn_tc.cu)
int pt_attention_tc(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                    long long v_stride, void* out, long long out_stride, int B, int heads, int Nq, int Nk, float scale,
                    cudaStream_t st);
// tcgen05 implementation with streamed keys for longer sequences (pt_attn_flash.cu)
int pt_attention_flash(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                       long long v_stride, void* out, long long out_stride, int B, int heads, int Nq, int Nk, float scale,
                       cudaStream_t st);

namespace {

constexpr int kHeadDim = 64;
constexpr int kRowPad = 72;        // bf16 elements per smem row (144 B): conflict-free fragment loads
constexpr int kWarps = 6;

__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(pt::smem_u32(p)));
}

struct AttnArgs {
  const __nv_bfloat16 *q, *k, *v;
  __nv_bfloat16* out;
  long long q_stride, k_stride, v_stride, out_stride;   // elements between consecutive rows
  int heads, Nq, Nk;
  float scale_log2e;
};

// Persistent CTAs walk (sample, head) pairs; the next pair's K, V and Q are prefetched with cp.async into the
// second half of a double-buffered smem tile while the tensor cores work on the current one.  Keys are processed
// in blocks of 32 with an online softmax (running max / sum, accumulator rescale), so the register footprint is
// independent of the sequence length.
constexpr int kKeyBlock = 32;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
  const uint32_t d = pt::smem_u32(smem_dst);
  const int sz = valid ? 16 : 0;                          // src-size 0 -> 16 bytes of zeros
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(kWarps * 32, 2) attention_kernel(AttnArgs a, int n_pairs, int nbuf) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int NKP = ((a.Nk + kKeyBlock - 1) / kKeyBlock) * kKeyBlock;
  const int NQP = (a.Nq + 15) & ~15;
  const int buf_elems = (2 * NKP + NQP) * kRowPad;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  auto stage = [&](int pair, int buf) {
    __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw) + (size_t)buf * buf_elems;
    __nv_bfloat16* Vs = Ks + NKP * kRowPad;
    __nv_bfloat16* Qs = Vs + NKP * kRowPad;
    const int b = pair / a.heads, h = pair % a.heads;
    const __nv_bfloat16* kg = a.k + (size_t)b * a.Nk * a.k_stride + h * kHeadDim;
    const __nv_bfloat16* vg = a.v + (size_t)b * a.Nk * a.v_stride + h * kHeadDim;
    const __nv_bfloat16* qg = a.q + (size_t)b * a.Nq * a.q_stride + h * kHeadDim;
    for (int i = tid; i < NKP * 8; i += kWarps * 32) {
      const int r = i >> 3, c = (i & 7) * 8;
      const bool ok = r < a.Nk;
      const size_t rr = ok ? r : 0;
      cp_async16(Ks + r * kRowPad + c, kg + rr * a.k_stride + c, ok);
      cp_async16(Vs + r * kRowPad + c, vg + rr * a.v_stride + c, ok);
    }
    for (int i = tid; i < NQP * 8; i += kWarps * 32) {
      const int r = i >> 3, c = (i & 7) * 8;
      const bool ok = r < a.Nq;
      cp_async16(Qs + r * kRowPad + c, qg + (size_t)(ok ? r : 0) * a.q_stride + c, ok);
    }
  };

  int pair = blockIdx.x;
  if (pair < n_pairs) stage(pair, 0);
  cp_async_commit();
  const int g = lane >> 2, tig = lane & 3;
  for (int it = 0; pair < n_pairs; pair += gridDim.x, ++it) {
    const int buf = (nbuf == 2) ? (it & 1) : 0;
    const int next = pair + gridDim.x;
    if (nbuf == 2) {
      if (next < n_pairs) stage(next, buf ^ 1);          // the other buffer was released by the barrier below
      cp_async_commit();
      cp_async_wait<1>();                                // this pair's group has landed (the newest may be in flight)
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const __nv_bfloat16* Ks = reinterpret_cast<const __nv_bfloat16*>(smem_raw) + (size_t)buf * buf_elems;
    const __nv_bfloat16* Vs = Ks + NKP * kRowPad;
    const __nv_bfloat16* Qs = Vs + NKP * kRowPad;
    const int b = pair / a.heads, h = pair % a.heads;

    for (int qb = warp; qb * 16 < a.Nq; qb += kWarps) {
      uint32_t qf[4][4];
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const __nv_bfloat16* q0 = Qs + (qb * 16 + g) * kRowPad + ks * 16 + tig * 2;
        qf[ks][0] = *reinterpret_cast<const uint32_t*>(q0);
        qf[ks][1] = *reinterpret_cast<const uint32_t*>(q0 + 8 * kRowPad);
        qf[ks][2] = *reinterpret_cast<const uint32_t*>(q0 + 8);
        qf[ks][3] = *reinterpret_cast<const uint32_t*>(q0 + 8 * kRowPad + 8);
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, sum0 = 0.f, sum1 = 0.f;
      float oacc[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) oacc[nt][0] = oacc[nt][1] = oacc[nt][2] = oacc[nt][3] = 0.f;

      for (int kb = 0; kb < NKP; kb += kKeyBlock) {
        // S = Q K^T for 32 keys: 4 tiles of 8 keys
        float sacc[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const __nv_bfloat16* kp = Ks + (kb + nt * 8 + g) * kRowPad + ks * 16 + tig * 2;
            mma_bf16(sacc[nt], qf[ks], *reinterpret_cast<const uint32_t*>(kp), *reinterpret_cast<const uint32_t*>(kp + 8));
          }
        }
        float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const int key = kb + nt * 8 + tig * 2;
          if (key >= a.Nk) { sacc[nt][0] = -INFINITY; sacc[nt][2] = -INFINITY; }
          if (key + 1 >= a.Nk) { sacc[nt][1] = -INFINITY; sacc[nt][3] = -INFINITY; }
          bm0 = fmaxf(bm0, fmaxf(sacc[nt][0], sacc[nt][1]));
          bm1 = fmaxf(bm1, fmaxf(sacc[nt][2], sacc[nt][3]));
        }
        bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1)); bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
        bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1)); bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
        const float nm0 = fmaxf(mx0, bm0), nm1 = fmaxf(mx1, bm1);     // finite: every block holds at least one valid key
        const float sc0 = exp2f((mx0 - nm0) * a.scale_log2e), sc1 = exp2f((mx1 - nm1) * a.scale_log2e);
        mx0 = nm0; mx1 = nm1;
        sum0 *= sc0; sum1 *= sc1;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) { oacc[nt][0] *= sc0; oacc[nt][1] *= sc0; oacc[nt][2] *= sc1; oacc[nt][3] *= sc1; }
        uint32_t pf[4][2];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const float p0 = exp2f((sacc[nt][0] - mx0) * a.scale_log2e), p1 = exp2f((sacc[nt][1] - mx0) * a.scale_log2e);
          const float p2 = exp2f((sacc[nt][2] - mx1) * a.scale_log2e), p3 = exp2f((sacc[nt][3] - mx1) * a.scale_log2e);
          sum0 += p0 + p1; sum1 += p2 + p3;
          pf[nt][0] = pt::pack_bf16x2(p0, p1);
          pf[nt][1] = pt::pack_bf16x2(p2, p3);
        }
        // O += P V for this key block (two 16-key steps)
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) {
          const uint32_t af[4] = {pf[2 * kt][0], pf[2 * kt][1], pf[2 * kt + 1][0], pf[2 * kt + 1][1]};
#pragma unroll
          for (int np = 0; np < 4; ++np) {
            uint32_t bf[4];
            // lanes 0-7: keys +0..7 @ dim np*16 ; 8-15: keys +8..15 ; 16-23: keys +0..7 @ dim np*16+8 ; 24-31: keys +8..15
            const __nv_bfloat16* vp = Vs + (kb + kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * kRowPad + np * 16 + (lane >> 4) * 8;
            ldmatrix_x4_trans(bf, vp);
            mma_bf16(oacc[2 * np], af, bf[0], bf[1]);
            mma_bf16(oacc[2 * np + 1], af, bf[2], bf[3]);
          }
        }
      }
      sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
      sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
      const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;
      const int r0 = qb * 16 + g, r1 = r0 + 8;
      __nv_bfloat16* o0 = a.out + ((size_t)b * a.Nq + r0) * a.out_stride + h * kHeadDim + tig * 2;
      __nv_bfloat16* o1 = a.out + ((size_t)b * a.Nq + r1) * a.out_stride + h * kHeadDim + tig * 2;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        if (r0 < a.Nq) *reinterpret_cast<uint32_t*>(o0 + nt * 8) = pt::pack_bf16x2(oacc[nt][0] * inv0, oacc[nt][1] * inv0);
        if (r1 < a.Nq) *reinterpret_cast<uint32_t*>(o1 + nt * 8) = pt::pack_bf16x2(oacc[nt][2] * inv1, oacc[nt][3] * inv1);
      }
    }
    __syncthreads();                                     // everyone is done with `buf` before it is refilled
    if (nbuf == 1) {                                     // single buffer (very long sequences): no overlap
      if (next < n_pairs) stage(next, 0);
      cp_async_commit();
    }
  }
  cp_async_wait<0>();
}

// ---------------------------------------------------------------------------------------------
// Long sequences (wav2vec2 encoder: ~3000 frames of a minute of audio attend to each other): K and V no longer fit in
// shared memory, so a CTA owns 128 queries of one (sample, head) and STREAMS the keys through two cp.async buffers of 64
// keys while the same online-softmax / mma.sync inner block as above consumes the other buffer.
constexpr int kLWarps = 8, kLQ = kLWarps * 16, kLTile = 64;

__global__ void __launch_bounds__(kLWarps * 32, 2) attention_long_kernel(AttnArgs a, int q_blocks) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smem_raw);            // [128][72]
  __nv_bfloat16* KV = Qs + kLQ * kRowPad;                                     // [2 buffers][K | V][64][72]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qb = blockIdx.x % q_blocks, bh = blockIdx.x / q_blocks;
  const int b = bh / a.heads, h = bh % a.heads;
  const int q0 = qb * kLQ;
  const __nv_bfloat16* kg = a.k + (size_t)b * a.Nk * a.k_stride + h * kHeadDim;
  const __nv_bfloat16* vg = a.v + (size_t)b * a.Nk * a.v_stride + h * kHeadDim;
  const __nv_bfloat16* qg = a.q + (size_t)b * a.Nq * a.q_stride + h * kHeadDim;
  auto stage_kv = [&](int tile, int buf) {
    __nv_bfloat16* Ks = KV + (size_t)buf * 2 * kLTile * kRowPad;
    __nv_bfloat16* Vs = Ks + kLTile * kRowPad;
    for (int i = tid; i < kLTile * 8; i += kLWarps * 32) {
      const int r = i >> 3, c = (i & 7) * 8, key = tile * kLTile + r;
      const bool ok = key < a.Nk;
      const size_t rr = ok ? key : 0;
      cp_async16(Ks + r * kRowPad + c, kg + rr * a.k_stride + c, ok);
      cp_async16(Vs + r * kRowPad + c, vg + rr * a.v_stride + c, ok);
    }
  };
  for (int i = tid; i < kLQ * 8; i += kLWarps * 32) {
    const int r = i >> 3, c = (i & 7) * 8;
    const bool ok = q0 + r < a.Nq;
    cp_async16(Qs + r * kRowPad + c, qg + (size_t)(ok ? q0 + r : 0) * a.q_stride + c, ok);
  }
  stage_kv(0, 0);
  cp_async_commit();
  const int n_tiles = (a.Nk + kLTile - 1) / kLTile;
  const int g = lane >> 2, tig = lane & 3;
  uint32_t qf[4][4];
  float mx0 = -INFINITY, mx1 = -INFINITY, sum0 = 0.f, sum1 = 0.f;
  float oacc[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) oacc[nt][0] = oacc[nt][1] = oacc[nt][2] = oacc[nt][3] = 0.f;
  for (int tile = 0; tile < n_tiles; ++tile) {
    const int buf = tile & 1;
    if (tile + 1 < n_tiles) { stage_kv(tile + 1, buf ^ 1); cp_async_commit(); cp_async_wait<1>(); }
    else cp_async_wait<0>();
    __syncthreads();
    if (tile == 0) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const __nv_bfloat16* qp = Qs + (warp * 16 + g) * kRowPad + ks * 16 + tig * 2;
        qf[ks][0] = *reinterpret_cast<const uint32_t*>(qp);
        qf[ks][1] = *reinterpret_cast<const uint32_t*>(qp + 8 * kRowPad);
        qf[ks][2] = *reinterpret_cast<const uint32_t*>(qp + 8);
        qf[ks][3] = *reinterpret_cast<const uint32_t*>(qp + 8 * kRowPad + 8);
      }
    }
    const __nv_bfloat16* Ks = KV + (size_t)buf * 2 * kLTile * kRowPad;
    const __nv_bfloat16* Vs = Ks + kLTile * kRowPad;
#pragma unroll
    for (int kb = 0; kb < kLTile; kb += kKeyBlock) {
      const int key0 = tile * kLTile + kb;
      if (key0 >= a.Nk) break;                                   // block-uniform
      float sacc[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const __nv_bfloat16* kp = Ks + (kb + nt * 8 + g) * kRowPad + ks * 16 + tig * 2;
          mma_bf16(sacc[nt], qf[ks], *reinterpret_cast<const uint32_t*>(kp), *reinterpret_cast<const uint32_t*>(kp + 8));
        }
      }
      float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int key = key0 + nt * 8 + tig * 2;
        if (key >= a.Nk) { sacc[nt][0] = -INFINITY; sacc[nt][2] = -INFINITY; }
        if (key + 1 >= a.Nk) { sacc[nt][1] = -INFINITY; sacc[nt][3] = -INFINITY; }
        bm0 = fmaxf(bm0, fmaxf(sacc[nt][0], sacc[nt][1]));
        bm1 = fmaxf(bm1, fmaxf(sacc[nt][2], sacc[nt][3]));
      }
      bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1)); bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
      bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1)); bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
      const float nm0 = fmaxf(mx0, bm0), nm1 = fmaxf(mx1, bm1);     // finite: a processed block holds a valid key
      const float sc0 = exp2f((mx0 - nm0) * a.scale_log2e), sc1 = exp2f((mx1 - nm1) * a.scale_log2e);
      mx0 = nm0; mx1 = nm1;
      sum0 *= sc0; sum1 *= sc1;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) { oacc[nt][0] *= sc0; oacc[nt][1] *= sc0; oacc[nt][2] *= sc1; oacc[nt][3] *= sc1; }
      uint32_t pf[4][2];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const float p0 = exp2f((sacc[nt][0] - mx0) * a.scale_log2e), p1 = exp2f((sacc[nt][1] - mx0) * a.scale_log2e);
        const float p2 = exp2f((sacc[nt][2] - mx1) * a.scale_log2e), p3 = exp2f((sacc[nt][3] - mx1) * a.scale_log2e);
        sum0 += p0 + p1; sum1 += p2 + p3;
        pf[nt][0] = pt::pack_bf16x2(p0, p1);
        pf[nt][1] = pt::pack_bf16x2(p2, p3);
      }
#pragma unroll
      for (int kt = 0; kt < 2; ++kt) {
        const uint32_t af[4] = {pf[2 * kt][0], pf[2 * kt][1], pf[2 * kt + 1][0], pf[2 * kt + 1][1]};
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t bf[4];
          const __nv_bfloat16* vp = Vs + (kb + kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * kRowPad + np * 16 + (lane >> 4) * 8;
          ldmatrix_x4_trans(bf, vp);
          mma_bf16(oacc[2 * np], af, bf[0], bf[1]);
          mma_bf16(oacc[2 * np + 1], af, bf[2], bf[3]);
        }
      }
    }
    __syncthreads();                                     // everyone is done with `buf` before it is refilled
  }
  sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
  sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
  const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;
  const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
  __nv_bfloat16* o0 = a.out + ((size_t)b * a.Nq + r0) * a.out_stride + h * kHeadDim + tig * 2;
  __nv_bfloat16* o1 = a.out + ((size_t)b * a.Nq + r1) * a.out_stride + h * kHeadDim + tig * 2;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    if (r0 < a.Nq) *reinterpret_cast<uint32_t*>(o0 + nt * 8) = pt::pack_bf16x2(oacc[nt][0] * inv0, oacc[nt][1] * inv0);
    if (r1 < a.Nq) *reinterpret_cast<uint32_t*>(o1 + nt * 8) = pt::pack_bf16x2(oacc[nt][2] * inv1, oacc[nt][3] * inv1);
  }
}

int launch_attn_long(const AttnArgs& a, int B, cudaStream_t st) {
  const size_t smem = (size_t)(kLQ + 4 * kLTile) * kRowPad * sizeof(__nv_bfloat16);      // 55 296 bytes
  PT_SMEM_OPT_IN(attention_long_kernel, smem);
  const int q_blocks = (a.Nq + kLQ - 1) / kLQ;
  const long long grid = (long long)B * a.heads * q_blocks;
  if (grid > 0x7fffffffLL) return pt::fail("pt_attention: too many (sample, head, query block) units");
  attention_long_kernel<<<(unsigned)grid, kLWarps * 32, smem, st>>>(a, q_blocks);
  PT_CUDA(cudaGetLastError());
  return 0;
}

int launch_attn(const AttnArgs& a, int B, cudaStream_t st) {
  const int NQP = (a.Nq + 15) & ~15;
  const int NKP = ((a.Nk + kKeyBlock - 1) / kKeyBlock) * kKeyBlock;
  const size_t one = (size_t)(2 * NKP + NQP) * kRowPad * sizeof(__nv_bfloat16);
  const int nbuf = (2 * one <= 226 * 1024) ? 2 : 1;
  const size_t smem = nbuf * one;
  if (smem > 226 * 1024) return launch_attn_long(a, B, st);           // K, V, Q of a head no longer fit: stream the keys
  if (smem > 48 * 1024) PT_SMEM_OPT_IN(attention_kernel, smem);
  const int sms = pt::sm_count();
  int per_sm = (int)((227 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);   // __launch_bounds__(192, 2)
  const int n_pairs = B * a.heads;
  const int grid = n_pairs < sms * per_sm ? n_pairs : sms * per_sm;
  attention_kernel<<<grid, kWarps * 32, smem, st>>>(a, n_pairs, nbuf);
  PT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" int pt_attention(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                            long long v_stride, void* out, long long out_stride, int B, int heads, int head_dim,
                            int Nq, int Nk, float scale, void* stream) {
  if (B <= 0 || Nq <= 0) return 0;
  if (!q || !k || !v || !out) return pt::fail("pt_attention: null pointer");
  if (head_dim != kHeadDim) return pt::fail("pt_attention: only head_dim 64 is built (got %d)", head_dim);
  if (Nk <= 0) return pt::fail("pt_attention: need at least one key (got %d)", Nk);
  if ((q_stride | k_stride | v_stride) & 7) return pt::fail("pt_attention: row strides must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15)
    return pt::fail("pt_attention: q/k/v must be 16-byte aligned");
  if ((out_stride & 1) || (reinterpret_cast<uintptr_t>(out) & 3)) return pt::fail("pt_attention: out must be 4-byte aligned");
  {
    // tensor-core (tcgen05) path for the sequence lengths of this model; the mma.sync kernel below covers
    // longer key sequences (and PT_ATTN_LEGACY=1 forces it, for A/B measurements)
    static const bool legacy = getenv("PT_ATTN_LEGACY") != nullptr;
    if (Nk <= 256 && !legacy)
      return pt_attention_tc(q, q_stride, k, k_stride, v, v_stride, out, out_stride, B, heads, Nq, Nk, scale,
                             (cudaStream_t)stream);
    if (!legacy)
      return pt_attention_flash(q, q_stride, k, k_stride, v, v_stride, out, out_stride, B, heads, Nq, Nk, scale,
                                (cudaStream_t)stream);
  }
  AttnArgs a;
  a.q = (const __nv_bfloat16*)q; a.k = (const __nv_bfloat16*)k; a.v = (const __nv_bfloat16*)v;
  a.out = (__nv_bfloat16*)out;
  a.q_stride = q_stride; a.k_stride = k_stride; a.v_stride = v_stride; a.out_stride = out_stride;
  a.heads = heads; a.Nq = Nq; a.Nk = Nk;
  a.scale_log2e = scale * 1.4426950408889634f;
  cudaStream_t st = (cudaStream_t)stream;
  if (Nk > 1024 || Nq > 1024) return launch_attn_long(a, B, st);      // K / V streamed through shared memory
  return launch_attn(a, B, st);
}
