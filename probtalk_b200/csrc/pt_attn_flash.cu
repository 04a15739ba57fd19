// probtalk_b200 -- multi-head attention on tcgen05 tensor cores for LONG key sequences (sm_100a).
//
// Same contract as pt_attention.  Used where the keys of a head no longer fit the single-pass kernel of pt_attn_tc.cu
// (Nk > 256): the wav2vec2 encoder of SURVEY.md 8(f)1 (scripts/demo.py:105-111 -> transformers Wav2Vec2Model: 12 heads x
// 64, ~3 000 frames per minute of audio attend to each other).
//
// One work unit = (sample, head, PAIR of 128-query tiles).  The keys stream through shared memory in blocks of 128:
//     S_X[128 x 128] = Q_X K_j^T          tcgen05.mma, Q / K straight from TMA (SWIZZLE_128B, K-major), X = tile A, B
//     P_X            = exp2(S_X c - m c)   one thread per query row, running maximum m; P goes back into TENSOR MEMORY as
//                                          packed bf16 (tcgen05.st, 64 columns) and is the A operand of the next MMA
//     PV_X[128 x 64] = P_X V_j             tcgen05.mma with A from TMEM, V used in place (smem) as an MN-major B operand
//     O_X            = O_X exp2((m_old - m_new) c) + PV_X   in the row owner's REGISTERS (no TMEM read-modify-write)
// P in shared memory (the layout of pt_attn_tc.cu) made every P V instruction read 6 KB of operands for 32 cycles of
// math: the tensor pipe waited on shared-memory bandwidth (8 MMAs took ~560 cycles to issue) and the in-order issue queue
// delayed everything behind it.  With A in TMEM a P V instruction reads 2 KB.
// Persistent CTAs, one per SM: warp 0 = TMA producer (Q pair; 4-slot rings for K and for V), warps 1 / 2 = issuers of the
// P V MMAs of tile A / B, warp 3 = issuer of the S MMAs of both tiles (each blocked on exactly the barriers of its queue);
// warps 4-7 = rows of tile A, warps 8-11 = rows of tile B.  setmaxnreg: 40 / 232 / 232 registers.
// TMEM (512 columns): S_A, S_B (128 each) | P_A, P_B (64 each) | PV_A, PV_B (64 each).
#include <cuda.h>
#include <stdlib.h>
#include "pt_common.cuh"
#include "pt_host.h"

namespace {

constexpr int kFThreads = 12 * 32;
constexpr int kFBlk = 128;                         // keys per block = query rows per tile
constexpr int kFStages = 4;
constexpr uint32_t kFTile = 128 * 128;             // [128 rows x 64] bf16, one SWIZZLE_128B box: 16 KB
constexpr uint32_t kFSmem = (2 + 2 * kFStages) * kFTile;       // Q_A Q_B | K ring | V ring

struct FlashArgs {
  __nv_bfloat16* out;
  long long out_stride;
  int heads, Nq, Nk, B;
  int pairs;                                       // ceil(ceil(Nq / 128) / 2)
  int nb;                                          // key blocks = ceil(Nk / 128)
  float scale_log2e;
  long long* trace;                                // developer hook (PT_FLASH_TRACE): clock64 stamps of CTA 0, [3 roles][16 blocks][8]
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void fma2(float& a0, float& a1, float x0, float x1, float y0, float y1, float z0, float z1) {
  unsigned long long x, y, z, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(x0), "f"(x1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(y0), "f"(y1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(z0), "f"(z1));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(x), "l"(y), "l"(z));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(d));
}
__device__ __forceinline__ void add2(float& a0, float& a1, float b0, float b1) {
  unsigned long long a, b, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(d));
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]: the A operand (M = 128 rows = lanes, 16 bf16 of K = 8 columns) comes from tensor memory
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

#define FTRACE(role, blk, ev)                                                                             \
  do {                                                                                                  \
    if (a.trace && blockIdx.x == 0 && (blk) < 16) a.trace[((role) * 16 + (blk)) * 8 + (ev)] = clock64(); \
  } while (0)

__global__ void __launch_bounds__(kFThreads, 1)
attention_flash_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                       const __grid_constant__ CUtensorMap tmap_v, const FlashArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* q_smem = smem;                                   // [2][16 KB]
  uint8_t* k_smem = q_smem + 2 * kFTile;                    // [3][16 KB]
  uint8_t* v_smem = k_smem + kFStages * kFTile;
  __shared__ uint64_t bars[2 + 4 * kFStages + 2 + 2 + 2 + 2];
  uint64_t* q_full = bars;                 // Q pair of a unit has landed
  uint64_t* q_empty = bars + 1;            // ... and the last S of the unit has read it
  uint64_t* k_full = bars + 2;             // [3]
  uint64_t* k_empty = k_full + kFStages;   // [3]
  uint64_t* v_full = k_empty + kFStages;   // [3]
  uint64_t* v_empty = v_full + kFStages;   // [3]
  uint64_t* s_full = v_empty + kFStages;   // [2]  per tile
  uint64_t* s_free = s_full + 2;           // [2]
  uint64_t* p_full = s_free + 2;           // [2]
  uint64_t* o_full = p_full + 2;           // [2]  P V of the tile's current block is complete
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = a.B * a.heads * a.pairs;
  const int nb = a.nb;

  if (warp == 0 && lane == 0) {
    pt::tma_prefetch_desc(&tmap_q); pt::tma_prefetch_desc(&tmap_k); pt::tma_prefetch_desc(&tmap_v);
    pt::mbar_init(q_full, 1); pt::mbar_init(q_empty, 1);
    for (int i = 0; i < kFStages; ++i) {
      pt::mbar_init(&k_full[i], 1); pt::mbar_init(&k_empty[i], 1); pt::mbar_init(&v_full[i], 1); pt::mbar_init(&v_empty[i], 2);
    }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&s_full[i], 1); pt::mbar_init(&s_free[i], 4); pt::mbar_init(&p_full[i], 4); }
    for (int i = 0; i < 2; ++i) pt::mbar_init(&o_full[i], 1);
    pt::fence_mbar_init();
  }
  if (warp == 1) { pt::tmem_alloc(&tmem_slot, 512); pt::tmem_relinquish(); }
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  const uint32_t tmem = tmem_slot;
  auto TS = [&](int x) { return tmem + (uint32_t)x * 128u; };                       // S accumulator of tile x
  auto TP = [&](int x) { return tmem + 256u + (uint32_t)x * 64u; };                 // P of tile x: 128 bf16 per row = 64 columns
  auto TO = [&](int x) { return tmem + 384u + (uint32_t)x * 64u; };                 // PV accumulator of tile x

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 0) {
      // ------------------------------------------------------------ TMA producer
      if (lane == 0) {
        int it = 0;
        int g = 0;                                   // blocks loaded so far (a CTA sees far fewer than 2^31)
        for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
          const int pair = u % a.pairs, bh = u / a.pairs;
          const int h = bh % a.heads, b = bh / a.heads;
          pt::mbar_wait(q_empty, ((uint32_t)it & 1u) ^ 1u);
          pt::mbar_expect_tx(q_full, 2 * kFTile);
          pt::tma_load_2d(q_smem, &tmap_q, q_full, h * 64, b * a.Nq + pair * 256);
          pt::tma_load_2d(q_smem + kFTile, &tmap_q, q_full, h * 64, b * a.Nq + pair * 256 + 128);
          for (int j = 0; j < nb; ++j, ++g) {
            const int st = g % kFStages;
            const uint32_t ph = (uint32_t)(g / kFStages) & 1u;
            pt::mbar_wait(&k_empty[st], ph ^ 1u);
            pt::mbar_expect_tx(&k_full[st], kFTile);
            pt::tma_load_2d(k_smem + st * kFTile, &tmap_k, &k_full[st], h * 64, b * a.Nk + j * kFBlk);
            pt::mbar_wait(&v_empty[st], ph ^ 1u);
            pt::mbar_expect_tx(&v_full[st], kFTile);
            pt::tma_load_2d(v_smem + st * kFTile, &tmap_v, &v_full[st], h * 64, b * a.Nk + j * kFBlk);
          }
        }
      }
    } else {
      // ------------------------------------------------------------ MMA issuers: warp 1 / 2 = P V of tile A / B, warp 3 = S
      // The MMAs here are short (32-64 cycles of tensor time), so the ISSUE path has to be fast.  A first version -- one
      // thread polling four queues with mbarrier.test_wait, descriptors rebuilt from addresses -- needed ~1000 cycles
      // per group of 4-8 MMAs (five barrier probes per poll, ~12 dependent uniform-datapath instructions per MMA) and the
      // whole CTA waited on it: tensor pipe 25 %.  Now each queue that the softmax rows wait on has its own warp, blocked
      // on exactly the barrier it needs, and descriptors are advanced by adding to their low word.
      const int n_my = (n_units > (int)blockIdx.x) ? (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
      const int total = n_my * nb;
      if (warp == 3) {
        const uint32_t idesc_s = pt::umma_idesc_bf16(128, 128);
        const uint64_t dk0 = pt::umma_desc_sw128(pt::smem_u32(k_smem));
        const uint64_t dq0 = pt::umma_desc_sw128(pt::smem_u32(q_smem));
        int unit = 0, jb = 0;
        for (int gi = 0; gi < total; ++gi) {
          const int st = gi % kFStages;
          const uint32_t ph = (uint32_t)(gi / kFStages) & 1u;
          if (jb == 0) pt::mbar_wait(q_full, (uint32_t)unit & 1u);
          pt::mbar_wait(&k_full[st], ph);
          const uint64_t dk = dk0 + (uint64_t)((uint32_t)st * (kFTile >> 4));
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            if (gi > 0) pt::mbar_wait(&s_free[x], (uint32_t)(gi - 1) & 1u);     // the rows have read S(gi - 1) of tile x
            pt::tc_fence_after();
            if (pt::elect_one()) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                pt::umma_bf16(tmem + (uint32_t)x * 128u, dq0 + (uint64_t)(x * (kFTile >> 4) + kk * 2), dk + (uint64_t)(kk * 2), idesc_s,
                              kk ? 1u : 0u);
              pt::umma_commit(&s_full[x]);
              if (x == 1) {
                pt::umma_commit(&k_empty[st]);                 // both tiles have used this K block
                if (jb == nb - 1) pt::umma_commit(q_empty);    // ... and it was the unit's last: Q may be replaced
              }
              FTRACE(2, gi, x);
            }
            __syncwarp();
          }
          if (++jb == nb) { jb = 0; ++unit; }
        }
      } else if (warp != 0) {
        const int x = warp - 1;
        const uint32_t idesc_o = pt::umma_idesc_bf16(128, 64) | (1u << 16);          // B operand (V) is MN-major
        const uint64_t dv0 = pt::umma_desc_sw128_mn(pt::smem_u32(v_smem));
        const uint32_t tp = tmem + 256u + (uint32_t)x * 64u, to = tmem + 384u + (uint32_t)x * 64u;
        for (int gi = 0; gi < total; ++gi) {
          const int st = gi % kFStages;
          const uint32_t ph = (uint32_t)(gi / kFStages) & 1u;
          pt::mbar_wait(&v_full[st], ph);
          // P of this block is in TMEM.  (The accumulator is free by construction: the rows add PV(j-1) to their registers
          // BEFORE they publish P(j).)
          pt::mbar_wait(&p_full[x], (uint32_t)gi & 1u);
          pt::tc_fence_after();
          if (pt::elect_one()) {
            const uint64_t dv = dv0 + (uint64_t)((uint32_t)st * (kFTile >> 4));
#pragma unroll
            for (int jj = 0; jj < 8; ++jj)          // 16 keys per instruction: 8 TMEM columns of P, 2 KB of V
              umma_bf16_ts(to, tp + jj * 8, dv + (uint64_t)(jj * (2048 >> 4)), idesc_o, jj ? 1u : 0u);
            pt::umma_commit(&o_full[x]);
            pt::umma_commit(&v_empty[st]);           // (two arrivals free the slot: one per tile)
            FTRACE(2, gi, 2 + x);
          }
          __syncwarp();
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ------------------------------------------------------------ softmax + output rows (thread = query row of tile x)
    const int x = (warp >= 8) ? 1 : 0;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
    const float sc = a.scale_log2e;
    const uint32_t ts = TS(x) + lane_off;
    int g = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int pair = u % a.pairs, bh = u / a.pairs;
      const int h = bh % a.heads, b = bh / a.heads;
      float m = -INFINITY, l = 0.f;
      float alpha = 0.f;                       // exp2((m_{j-1} - m_j) c) of the block whose PV is still in TMEM
      float O[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) O[i] = 0.f;
      for (int j = 0; j < nb; ++j, ++g) {
        if (lane == 0 && qd == 0) FTRACE(x, g, 0);
        pt::mbar_wait(&s_full[x], (uint32_t)g & 1u);
        if (lane == 0 && qd == 0) FTRACE(x, g, 1);
        pt::tc_fence_after();
        uint32_t v[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) pt::tmem_ld32(ts + 32 * c, v[c]);
        pt::tmem_ld_wait();
        pt::tc_fence_before();
        __syncwarp();
        if (lane == 0) pt::mbar_arrive(&s_free[x]);
        if (lane == 0 && qd == 0) FTRACE(x, g, 2);
        const int valid = a.Nk - j * kFBlk;                           // keys of this block that exist (>= 1)
        if (valid < kFBlk) {                                          // last block of a ragged key set: mask once, up front
          asm volatile("// ragged tail: keep this a branch, not 128 selects on every block" ::: "memory");
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (32 * c + i >= valid) v[c][i] = 0xff800000u;         // -inf: exp2 gives 0, the maximum ignores it
        }
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            mx0 = fmax3(mx0, __uint_as_float(v[c][i]), __uint_as_float(v[c][i + 1]));
            mx1 = fmax3(mx1, __uint_as_float(v[c][i + 2]), __uint_as_float(v[c][i + 3]));
          }
        const float m_new = fmaxf(m, fmaxf(mx0, mx1));
        const float alpha_new = ex2f((m - m_new) * sc);               // first block: exp2(-inf) = 0
        m = m_new;
        const float nm = -m_new * sc;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        uint32_t pk[2][32];                                           // the row's 128 probabilities as bf16 pairs
#pragma unroll
        for (int c = 0; c < 4; ++c) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float t0, t1;
            fma2(t0, t1, __uint_as_float(v[c][2 * i]), __uint_as_float(v[c][2 * i + 1]), sc, sc, nm, nm);
            const float p0 = ex2f(t0), p1 = ex2f(t1);
            if (i & 1) add2(s2, s3, p0, p1); else add2(s0, s1, p0, p1);
            pk[c >> 1][(c & 1) * 16 + i] = pt::pack_bf16x2(p0, p1);
          }
        }
        l = fmaf(l, alpha_new, (s0 + s1) + (s2 + s3));
        if (lane == 0 && qd == 0) FTRACE(x, g, 3);
        if (j > 0) {
          // the P V of the previous block: complete -> its P columns may be overwritten, and its result joins the row's
          // registers, O = O alpha_{j-1} + PV(j-1)  (one FMA per element)
          pt::mbar_wait(&o_full[x], (uint32_t)(g - 1) & 1u);
          pt::tc_fence_after();
          const uint32_t to = TO(x) + lane_off;
          uint32_t o[2][32];                                          // (the score registers are dead by now)
          pt::tmem_ld32(to, o[0]);
          pt::tmem_ld32(to + 32, o[1]);
          pt::tmem_ld_wait();
#pragma unroll
          for (int hc = 0; hc < 2; ++hc)
#pragma unroll
            for (int i = 0; i < 32; i += 2)
              fma2(O[32 * hc + i], O[32 * hc + i + 1], O[32 * hc + i], O[32 * hc + i + 1], alpha, alpha, __uint_as_float(o[hc][i]),
                   __uint_as_float(o[hc][i + 1]));
        }
        if (lane == 0 && qd == 0) FTRACE(x, g, 4);
        pt::tmem_st32(TP(x) + lane_off, pk[0]);
        pt::tmem_st32(TP(x) + lane_off + 32, pk[1]);
        pt::tmem_st_wait();
        pt::tc_fence_before();
        __syncwarp();
        if (lane == 0) pt::mbar_arrive(&p_full[x]);
        if (lane == 0 && qd == 0) FTRACE(x, g, 5);
        if (lane == 0 && qd == 0) FTRACE(x, g, 6);
        alpha = alpha_new;
      }
      // last block of the unit
      {
        pt::mbar_wait(&o_full[x], (uint32_t)(g - 1) & 1u);
        pt::tc_fence_after();
        const uint32_t to = TO(x) + lane_off;
        const float inv = 1.f / l;
        const float ainv = alpha * inv;
        const int qrow = pair * 256 + x * 128 + row;
        uint4* dst = reinterpret_cast<uint4*>(a.out + ((size_t)b * a.Nq + qrow) * a.out_stride + h * 64);
#pragma unroll
        for (int hc = 0; hc < 2; ++hc) {
          uint32_t o[32];
          pt::tmem_ld32(to + 32 * hc, o);
          pt::tmem_ld_wait();
          if (qrow < a.Nq) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float r[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) r[e] = fmaf(O[32 * hc + 8 * i + e], ainv, __uint_as_float(o[8 * i + e]) * inv);
              dst[4 * hc + i] = make_uint4(pt::pack_bf16x2(r[0], r[1]), pt::pack_bf16x2(r[2], r[3]), pt::pack_bf16x2(r[4], r[5]),
                                           pt::pack_bf16x2(r[6], r[7]));
            }
          }
        }
      }
    }
  }
  pt::tc_fence_before();
  __syncthreads();
  if (warp == 1) pt::tmem_dealloc(tmem, 512);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_map128(CUtensorMap* tm, const void* base, long long rows, long long cols, long long pitch) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return pt::fail("cuTensorMapEncodeTiled entry point not found");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {64u, 128u};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return pt::fail("cuTensorMapEncodeTiled failed (%d) for attention operand", (int)r);
  return 0;
}

}  // namespace

// returns 0 on success, 1 on error; called by pt_attention for Nk > 256
int pt_attention_flash(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                       long long v_stride, void* out, long long out_stride, int B, int heads, int Nq, int Nk, float scale,
                       cudaStream_t st) {
  if ((out_stride & 7) || (reinterpret_cast<uintptr_t>(out) & 15))
    return pt::fail("pt_attention: out must be 16-byte aligned with a row stride multiple of 8 elements");
  FlashArgs a;
  a.out = (__nv_bfloat16*)out; a.out_stride = out_stride;
  a.heads = heads; a.Nq = Nq; a.Nk = Nk; a.B = B;
  a.pairs = ((Nq + 127) / 128 + 1) / 2;
  a.nb = (Nk + kFBlk - 1) / kFBlk;
  a.scale_log2e = scale * 1.4426950408889634f;
  {
    const char* e = getenv("PT_FLASH_TRACE");             // developer hook: device address of 384 int64 (hex or decimal)
    a.trace = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
  }
  const long long n_units = (long long)B * heads * a.pairs;
  if (n_units > 0x7fffffffLL || (long long)B * Nq > 0x7fffffffLL || (long long)B * Nk > 0x7fffffffLL)
    return pt::fail("pt_attention: too many rows for one launch");
  CUtensorMap tq, tk, tv;
  if (make_map128(&tq, q, (long long)B * Nq, (long long)heads * 64, q_stride)) return 1;
  if (make_map128(&tk, k, (long long)B * Nk, (long long)heads * 64, k_stride)) return 1;
  if (make_map128(&tv, v, (long long)B * Nk, (long long)heads * 64, v_stride)) return 1;
  const size_t smem = 1024 + kFSmem;
  const int sms = pt::sm_count();
  const int grid = n_units < sms ? (int)n_units : sms;
  PT_SMEM_OPT_IN(attention_flash_kernel, smem);
  PT_CUDA(pt::launch_pdl(attention_flash_kernel, dim3(grid), dim3(kFThreads), smem, st, 1, tq, tk, tv, a));
  PT_CUDA(cudaGetLastError());
  return 0;
}
