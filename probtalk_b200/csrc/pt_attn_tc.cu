// probtalk_b200 -- multi-head attention on tcgen05 tensor cores (sm_100a), sequences up to 256 keys.
//
// Same contract as pt_attention (scaled-dot-product core of nn.MultiheadAttention, reference
// nets/inpainting/PTransformer.py:340,350).  One work unit = (sample, head, 128-query tile):
//     S[128 x Nk]  = Q K^T        tcgen05.mma, Q and K straight from TMA (SWIZZLE_128B, K-major)
//     P            = softmax(S)   one thread per query row, S read from TMEM, P written to smem as bf16 (K-major)
//     O[128 x 64]  = P V          tcgen05.mma, V used in place as an MN-major operand (no transpose)
// Persistent CTAs: warp 0 = TMA producer (double-buffered Q/K and V, separate barriers), warp 1 = issuer of the Q K^T MMAs,
// warp 2 = issuer of the P V MMAs, warps 4-7 = softmax (TMEM lane = query row), warps 8-11 = output (O read-out, 1/sum
// scaling, stores).  Reading S twice (maximum, then exponentials) was the busiest part of the row warps' work: setmaxnreg
// moves registers from warpgroup 0 to the softmax warps so that up to 96 scores per row stay in registers and S is read
// ONCE; warps whose rows lie beyond Nq read nothing.
// The unit chain (load -> Q K^T -> softmax -> P V -> store) is latency bound, so the stages are decoupled as far as the
// buffers allow: the two MMA queues have their own issuer warps (each blocked on exactly its barriers; the whole warp walks
// the loop and one elected lane issues), Q/K slots are released as soon as their Q K^T has run (not after P V), and V has
// its own full / empty barriers.  With two S accumulators in TMEM the loads and the Q K^T of unit u+2 are in flight while
// unit u is still in its softmax.
#include <cuda.h>
#include <stdlib.h>
#include "pt_common.cuh"
#include "pt_host.h"

namespace {

constexpr int kThreadsTc = 12 * 32;   // warpgroup 0: TMA, MMA, 2 idle; warpgroup 1: 4 softmax warps; warpgroup 2: 4 output warps
constexpr int kQRows = 128;
constexpr uint32_t kQBytes = kQRows * 128;          // 16 KB
constexpr uint32_t kAtomBytes = kQRows * 128;       // one [128 x 64] bf16 swizzle atom of P

struct AttnTcArgs {
  __nv_bfloat16* out;
  long long out_stride;
  int heads, Nq, Nk, NKP;          // NKP = keys padded to 16
  int B, m_tiles;                  // m_tiles = ceil(Nq / 128)
  int tmem_cols, o_col;            // TMEM allocation (power of two) and the column where O starts
  int qk_stages, v_stages;         // smem slots for (Q, K) pairs and for V (1 or 2 each; V is held until its P V has run)
  int kv_rows;                     // rows of the K / V TMA boxes (NKP, or Nk rounded up to 8: see pt_attention_tc)
  int q_region;                    // bytes reserved for Q in a stage (q_box rows rounded up to 1 KB; the M = 128 MMA may
                                   // read on into K, which only produces score rows that are never stored)
  int q_box;                       // rows of the Q TMA box (Nq rounded up to 8 when one tile covers all queries)
  int p_bufs;                      // P tiles in smem and O accumulators in TMEM (2 = softmax never waits for the last P V)
  int s_bufs;                      // S accumulators in TMEM (2 = the next unit's Q K^T overlaps this unit's softmax)
  float scale_log2e;
  long long* trace;                // developer hook (PT_ATTN_TRACE): clock64 stamps of CTA 0, [5 roles][16 units][4]
};

__device__ __forceinline__ float ex2(float x) {
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
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

#define ATRACE(role, unit, ev)                                                                               \
  do {                                                                                                       \
    if (a.trace && blockIdx.x == 0 && (unit) < 16) a.trace[((role) * 16 + (unit)) * 4 + (ev)] = clock64();     \
  } while (0)

template <int kNch>                                     // 1..3: keys / 32, scores stay in registers; 0: two TMEM passes
__global__ void __launch_bounds__(kThreadsTc, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const AttnTcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t kv_bytes = (uint32_t)a.kv_rows * 128u;
  const uint32_t kv_pad = (kv_bytes + 1023u) & ~1023u;
  const uint32_t q_region = (uint32_t)a.q_region;
  const uint32_t qk_bytes = q_region + kv_pad;      // smem: [qk_stages x (Q | K)] [v_stages x V] [p_bufs x P]
  uint8_t* v_smem = smem + (size_t)a.qk_stages * qk_bytes;
  const int p_atoms = (a.NKP + 63) / 64;
  uint8_t* p_smem = v_smem + (size_t)a.v_stages * kv_pad;
  __shared__ uint64_t bars[20];
  __shared__ float sums_s[2][kQRows];          // softmax row sums, handed from the softmax warps to the output warps
  uint64_t* qk_full = bars;         // [2]  Q and K of a unit have landed
  uint64_t* qk_empty = bars + 2;    // [2]  ... and its Q K^T has finished reading them
  uint64_t* s_full = bars + 4;      // [2]  (S is double-buffered in TMEM when it fits)
  uint64_t* s_free = bars + 6;      // [2]
  uint64_t* p_full = bars + 8;      // [2] placed at the end of the array
  uint64_t* o_full = bars + 9;
  uint64_t* o_free = bars + 10;
  uint64_t* sums_full = bars + 11;  // [2]
  uint64_t* v_full = bars + 13;     // [2]  V of a unit has landed
  uint64_t* v_empty = bars + 15;    // [2]  ... and its P V has finished reading it
  uint64_t* p_full1 = bars + 17; uint64_t* o_full1 = bars + 18; uint64_t* o_free1 = bars + 19;   // [1] of each
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = a.B * a.heads * a.m_tiles;

  {
    // Operand rows that no TMA box ever writes (keys beyond the box, see kv_rows) are read by the MMAs; they must hold
    // finite numbers (they meet P = 0 or masked score columns), so the whole buffer starts as zeros.
    const uint32_t total = (uint32_t)a.qk_stages * qk_bytes + (uint32_t)a.v_stages * kv_pad + (uint32_t)(a.p_bufs * p_atoms) * kAtomBytes;
    for (uint32_t i = threadIdx.x * 16u; i < total; i += kThreadsTc * 16u) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    pt::fence_async_smem();
  }
  if (warp == 0 && lane == 0) {
    pt::tma_prefetch_desc(&tmap_q); pt::tma_prefetch_desc(&tmap_k); pt::tma_prefetch_desc(&tmap_v);
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&qk_full[i], 1); pt::mbar_init(&qk_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&v_full[i], 1); pt::mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&s_full[i], 1); pt::mbar_init(&s_free[i], 4); }
    pt::mbar_init(p_full, 4); pt::mbar_init(p_full1, 4);
    pt::mbar_init(o_full, 1); pt::mbar_init(o_free, 4); pt::mbar_init(o_full1, 1); pt::mbar_init(o_free1, 4);
    pt::mbar_init(&sums_full[0], 4); pt::mbar_init(&sums_full[1], 4);
    pt::fence_mbar_init();
  }
  if (warp == 1) { pt::tmem_alloc(&tmem_slot, (uint32_t)a.tmem_cols); pt::tmem_relinquish(); }
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  pt::pdl_launch_dependents();      // set-up done: the next kernel may start its own; q / k / v / out are dependent memory
  pt::pdl_wait();
  const uint32_t tmem = tmem_slot;
  const uint32_t tmem_s = tmem, tmem_o = tmem + (uint32_t)a.o_col;
  const uint32_t p_bytes = (uint32_t)p_atoms * kAtomBytes;
  // P / O buffer of this CTA's it-th unit, the barriers that go with it, and the parity of its use of them
  auto PB = [&](int it) { return it % a.p_bufs; };
  auto PPH = [&](int it) { return (uint32_t)(it / a.p_bufs) & 1u; };
  auto P_FULL = [&](int pb) { return pb ? p_full1 : p_full; };
  auto O_FULL = [&](int pb) { return pb ? o_full1 : o_full; };
  auto O_FREE = [&](int pb) { return pb ? o_free1 : o_free; };
  const bool dbl = (a.s_bufs == 2);

  // Register re-balancing per warpgroup (2 CTAs / SM => 80 registers per thread at launch; setmaxnreg sits at the top of
  // each warpgroup's own branch so that ptxas allocates each role against its own limit):
  //   128 * (40 + 136 + 64) = 30720 = 384 * 80
  if (warp < 4) {
   asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
   if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int it = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
        const int sq_i = it % a.qk_stages, sv_i = it % a.v_stages;
        const uint32_t qph = (uint32_t)(it / a.qk_stages) & 1u, vph = (uint32_t)(it / a.v_stages) & 1u;
        const int mt = u % a.m_tiles, bh = u / a.m_tiles;
        const int h = bh % a.heads, b = bh / a.heads;
        uint8_t* sq = smem + (size_t)sq_i * qk_bytes;
        ATRACE(0, it, 0);
        pt::mbar_wait(&qk_empty[sq_i], qph ^ 1);
        ATRACE(0, it, 1);
        pt::mbar_expect_tx(&qk_full[sq_i], (uint32_t)a.q_box * 128u + kv_bytes);
        pt::tma_load_2d(sq, &tmap_q, &qk_full[sq_i], h * 64, b * a.Nq + mt * kQRows);
        pt::tma_load_2d(sq + q_region, &tmap_k, &qk_full[sq_i], h * 64, b * a.Nk);
        pt::mbar_wait(&v_empty[sv_i], vph ^ 1);
        pt::mbar_expect_tx(&v_full[sv_i], kv_bytes);
        pt::tma_load_2d(v_smem + (size_t)sv_i * kv_pad, &tmap_v, &v_full[sv_i], h * 64, b * a.Nk);
        ATRACE(0, it, 2);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer of S = Q K^T (warp 2 issues the P V MMAs)
    // The MMAs of this kernel are short (N <= 256 keys, 16-64 cycles of tensor time each), so the issue path counts.  The
    // WHOLE warp walks the loop (every lane blocks on the same barriers) and one ELECTED lane issues: under `if (lane ==
    // 0)` the compiler wraps every tcgen05.mma in an election loop with five R2UR moves and rebuilds both descriptors
    // (~20 dependent instructions, ~70 cycles per MMA), and one thread polling both queues with mbarrier.test_wait +
    // __nanosleep added up to ~1000 cycles per group of MMAs (measured on the streamed-key kernel, pt_attn_flash.cu).
    // S runs ahead of P V as far as the S accumulators and the Q/K slots allow (the barriers say when).
    {
      const uint32_t idesc_s = pt::umma_idesc_bf16(kQRows, (uint32_t)a.NKP);
      const int n_my = (n_units > (int)blockIdx.x) ? (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
      for (int is = 0; is < n_my; ++is) {
        const int st = is % a.qk_stages, sb = is % a.s_bufs;
        const uint32_t ph = (uint32_t)(is / a.qk_stages) & 1u, sph = (uint32_t)(is / a.s_bufs) & 1u;
        pt::mbar_wait(&qk_full[st], ph);
        if (lane == 0) ATRACE(1, is, 0);
        pt::mbar_wait(&s_free[sb], sph ^ 1);
        pt::tc_fence_after();
        if (pt::elect_one()) {
          ATRACE(1, is, 1);
          const uint32_t sq = pt::smem_u32(smem + (size_t)st * qk_bytes);
          const uint64_t dq = pt::umma_desc_sw128(sq), dk = pt::umma_desc_sw128(sq + q_region);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            pt::umma_bf16(tmem_s + sb * a.NKP, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_s, k ? 1u : 0u);
          pt::umma_commit(&s_full[sb]);
          pt::umma_commit(&qk_empty[st]);                // Q and K of this slot may be overwritten once S is complete
          ATRACE(1, is, 2);
        }
        __syncwarp();
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------ MMA issuer of O = P V
    {
      const uint32_t idesc_o = pt::umma_idesc_bf16(kQRows, 64) | (1u << 16);       // B operand (V) is MN-major
      const int n_my = (n_units > (int)blockIdx.x) ? (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
      const int ksteps = a.NKP / 16;
      for (int ip = 0; ip < n_my; ++ip) {
        const int st = ip % a.v_stages;
        const uint32_t ph = (uint32_t)(ip / a.v_stages) & 1u;
        const int pb = PB(ip);
        const uint32_t pph = PPH(ip);
        // V has landed, the O accumulator's previous contents have been read out of TMEM, P (bf16) is in smem
        pt::mbar_wait(&v_full[st], ph);
        if (lane == 0) ATRACE(2, ip, 0);
        pt::mbar_wait(O_FREE(pb), pph ^ 1);
        pt::mbar_wait(P_FULL(pb), pph);
        pt::tc_fence_after();
        if (pt::elect_one()) {
          ATRACE(2, ip, 1);
          uint64_t dp = pt::umma_desc_sw128(pt::smem_u32(p_smem) + pb * p_bytes);
          uint64_t dv = pt::umma_desc_sw128_mn(pt::smem_u32(v_smem + (size_t)st * kv_pad));
          for (int j = 0; j < ksteps; ++j) {
            pt::umma_bf16(tmem_o + pb * 64, dp, dv, idesc_o, j ? 1u : 0u);
            dp += ((j & 3) == 3) ? (uint64_t)((kAtomBytes >> 4) - 6) : (uint64_t)2;      // next 16 keys: +32 B, or the next 64-key atom
            dv += (uint64_t)(2048 >> 4);
          }
          pt::umma_commit(O_FULL(pb));
          pt::umma_commit(&v_empty[st]);
          ATRACE(2, ip, 2);
        }
        __syncwarp();
      }
    }
   }
  } else if (warp < 8) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 136;");
    // ------------------------------------------------------------ softmax (thread = query row)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const uint32_t up = (uint32_t)it & 1u;
      const int pb = PB(it);
      const uint32_t pph = PPH(it);
      const int sb = dbl ? (it & 1) : 0;
      const uint32_t sph = (uint32_t)(dbl ? (it >> 1) : it) & 1u;
      const uint32_t tmem_sb = tmem_s + sb * a.NKP;
      if (q == 0 && lane == 0) ATRACE(3, it, 0);
      pt::mbar_wait(&s_full[sb], sph);
      if (q == 0 && lane == 0) ATRACE(3, it, 1);
      pt::tc_fence_after();
      // a warp whose 32 rows all lie beyond the last query of this tile (Nq = 88: rows 96..127) only keeps the barrier
      // protocol going; its P rows stay stale, which is harmless because every O row depends on its own P row only
      const int rows_valid = min(kQRows, a.Nq - (u % a.m_tiles) * kQRows);
      if (q * 32 >= rows_valid) {
        if (it >= a.p_bufs) pt::mbar_wait(O_FULL(pb), pph ^ 1);
        // sums_s / sums_full are a two-deep ring: the output warps must be done with the unit that last used this O
        // buffer (they read its row sums before releasing O) before the slot is rewritten, otherwise a fast softmax could
        // complete the same barrier twice before a slow output warp has looked at it (parity aliasing -> deadlock)
        if (it >= a.p_bufs) pt::mbar_wait(O_FREE(pb), pph ^ 1);
        sums_s[up][row] = 1.f;
        __syncwarp();
        if (lane == 0) { pt::mbar_arrive(&s_free[sb]); pt::mbar_arrive(P_FULL(pb)); pt::mbar_arrive(&sums_full[up]); }
        continue;
      }
      if constexpr (kNch > 0) {
        // up to 96 keys: ONE TMEM read, the scores stay in registers for the maximum and the exponentials, and the S
        // accumulator goes back to the issuer before the arithmetic starts
        uint32_t v[kNch][32];
#pragma unroll
        for (int c = 0; c < kNch; ++c) pt::tmem_ld32(tmem_sb + lane_off + 32 * c, v[c]);
        pt::tmem_ld_wait();
        pt::tc_fence_before();
        __syncwarp();
        if (lane == 0) pt::mbar_arrive(&s_free[sb]);
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int c = 0; c < kNch; ++c) {
          {
            if (32 * c + 32 <= a.Nk) {
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                mx0 = max3(mx0, __uint_as_float(v[c][i]), __uint_as_float(v[c][i + 1]));
                mx1 = max3(mx1, __uint_as_float(v[c][i + 2]), __uint_as_float(v[c][i + 3]));
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (32 * c + i < a.Nk) mx0 = fmaxf(mx0, __uint_as_float(v[c][i]));
            }
          }
        }
        const float mx = fmaxf(mx0, mx1);
        if (it >= a.p_bufs) pt::mbar_wait(O_FULL(pb), pph ^ 1);   // the P V MMA that last read this P tile is done
        float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
        const float nmxs = -mx * a.scale_log2e, sc = a.scale_log2e;
#pragma unroll
        for (int c = 0; c < kNch; ++c) {
          {
            uint32_t pk[16];
            const bool full = (32 * c + 32 <= a.Nk);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              float t0, t1;
              fma2(t0, t1, __uint_as_float(v[c][2 * i]), __uint_as_float(v[c][2 * i + 1]), sc, sc, nmxs, nmxs);
              float p0 = ex2(t0), p1 = ex2(t1);
              if (!full) {
                if (32 * c + 2 * i >= a.Nk) p0 = 0.f;
                if (32 * c + 2 * i + 1 >= a.Nk) p1 = 0.f;
              }
              if (i & 1) add2(sum2, sum3, p0, p1); else add2(sum0, sum1, p0, p1);
              pk[i] = pt::pack_bf16x2(p0, p1);
            }
            uint8_t* atom = p_smem + pb * p_bytes + (size_t)(c >> 1) * kAtomBytes;
            const int chunk0 = (c & 1) * 4;              // 16-byte chunk index of key 32 c within the 64-key atom
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(atom + pt::sw128_offset(row, chunk0 + i)) =
                  make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          }
        }
        // sums_s / sums_full are a two-deep ring: the output warps must be done with the unit that last used this O
        // buffer (they read its row sums before releasing O) before the slot is rewritten, otherwise a fast softmax could
        // complete the same barrier twice before a slow output warp has looked at it (parity aliasing -> deadlock)
        if (it >= a.p_bufs) pt::mbar_wait(O_FREE(pb), pph ^ 1);
        sums_s[up][row] = (sum0 + sum1) + (sum2 + sum3);
        pt::fence_async_smem();                          // P stores -> visible to the tensor-core proxy
        __syncwarp();
        if (lane == 0) { pt::mbar_arrive(P_FULL(pb)); pt::mbar_arrive(&sums_full[up]); }
        if (q == 0 && lane == 0) ATRACE(3, it, 2);
        continue;
      }
      // pass 1: row maximum (only the last 32-column chunk can hold keys beyond Nk)
      float mx = -INFINITY;
      for (int c0 = 0; c0 < a.NKP; c0 += 32) {
        uint32_t v[32];
        pt::tmem_ld32(tmem_sb + lane_off + c0, v);
        pt::tmem_ld_wait();
        if (c0 + 32 <= a.Nk) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) mx = max3(mx, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < a.Nk) mx = fmaxf(mx, __uint_as_float(v[i]));
        }
      }
      // the P tile is single-buffered: the P V MMA of the previous unit must have finished reading it
      if (it >= a.p_bufs) pt::mbar_wait(O_FULL(pb), pph ^ 1);
      // pass 2: P = exp2(s * scale*log2e - max * scale*log2e) as bf16 into the K-major swizzled A tile; fp32 row sum
      float sum0 = 0.f, sum1 = 0.f;
      const float nmxs = -mx * a.scale_log2e, sc = a.scale_log2e;
      for (int c0 = 0; c0 < a.NKP; c0 += 32) {
        uint32_t v[32];
        pt::tmem_ld32(tmem_sb + lane_off + c0, v);
        pt::tmem_ld_wait();
        uint32_t pk[16];
        const bool full = (c0 + 32 <= a.Nk);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float t0, t1;
          fma2(t0, t1, __uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]), sc, sc, nmxs, nmxs);
          float p0 = ex2(t0), p1 = ex2(t1);
          if (!full) {
            if (c0 + 2 * i >= a.Nk) p0 = 0.f;
            if (c0 + 2 * i + 1 >= a.Nk) p1 = 0.f;
          }
          add2(sum0, sum1, p0, p1);
          pk[i] = pt::pack_bf16x2(p0, p1);
        }
        uint8_t* atom = p_smem + pb * p_bytes + (size_t)(c0 >> 6) * kAtomBytes;
        const int chunk0 = (c0 & 63) >> 3;               // 16-byte chunk index of key c0 within the 64-key atom
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(atom + pt::sw128_offset(row, chunk0 + i)) =
              make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      }
      const float sum = sum0 + sum1;
      // sums_s / sums_full are a two-deep ring: the output warps must be done with the unit that last used this O
      // buffer (they read its row sums before releasing O) before the slot is rewritten, otherwise a fast softmax could
      // complete the same barrier twice before a slow output warp has looked at it (parity aliasing -> deadlock)
      if (it >= a.p_bufs) pt::mbar_wait(O_FREE(pb), pph ^ 1);
      sums_s[up][row] = sum;
      pt::tc_fence_before();
      pt::fence_async_smem();                            // P stores -> visible to the tensor-core proxy
      __syncwarp();
      if (lane == 0) { pt::mbar_arrive(&s_free[sb]); pt::mbar_arrive(P_FULL(pb)); pt::mbar_arrive(&sums_full[up]); }
      if (q == 0 && lane == 0) ATRACE(3, it, 2);
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
    // ------------------------------------------------------------ output: O = P V read-out, 1/sum, stores
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const uint32_t up = (uint32_t)it & 1u;
      const int pb = PB(it);
      const uint32_t pph = PPH(it);
      const int mt = u % a.m_tiles, bh = u / a.m_tiles;
      const int h = bh % a.heads, b = bh / a.heads;
      if (q == 0 && lane == 0) ATRACE(4, it, 0);
      pt::mbar_wait(&sums_full[up], (uint32_t)(it >> 1) & 1u);
      pt::mbar_wait(O_FULL(pb), pph);
      if (q == 0 && lane == 0) ATRACE(4, it, 1);
      pt::tc_fence_after();
      if (q * 32 >= min(kQRows, a.Nq - mt * kQRows)) {   // no query rows here: nothing to read or store
        __syncwarp();
        if (lane == 0) pt::mbar_arrive(O_FREE(pb));
        continue;
      }
      const float inv = 1.f / sums_s[up][row];
      const int qrow = mt * kQRows + row;
      uint4* dst = reinterpret_cast<uint4*>(a.out + ((size_t)b * a.Nq + qrow) * a.out_stride + h * 64);
#pragma unroll 1
      for (int hc = 0; hc < 2; ++hc) {                   // 32 output columns at a time (64 registers per thread here)
        uint32_t o[32];
        pt::tmem_ld32(tmem_o + pb * 64 + lane_off + 32 * hc, o);
        pt::tmem_ld_wait();
        if (hc == 1) {
          pt::tc_fence_before();
          __syncwarp();
          if (lane == 0) pt::mbar_arrive(O_FREE(pb));
        }
        if (qrow < a.Nq) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[4 * hc + i] = make_uint4(pt::pack_bf16x2(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv),
                                         pt::pack_bf16x2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv),
                                         pt::pack_bf16x2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv),
                                         pt::pack_bf16x2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv));
        }
      }
      if (q == 0 && lane == 0) ATRACE(4, it, 2);
    }
    (void)p_atoms;
  }
  pt::tc_fence_before();
  __syncthreads();
  if (warp == 1) pt::tmem_dealloc(tmem, (uint32_t)a.tmem_cols);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_map(CUtensorMap* tm, const void* base, long long rows, long long cols, long long pitch, int box_rows) {
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
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return pt::fail("cuTensorMapEncodeTiled failed (%d) for attention operand", (int)r);
  return 0;
}

}  // namespace

// returns 0 on success, 1 on error; called by pt_attention for Nk <= 256
int pt_attention_tc(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                    long long v_stride, void* out, long long out_stride, int B, int heads, int Nq, int Nk, float scale,
                    cudaStream_t st) {
  AttnTcArgs a;
  a.out = (__nv_bfloat16*)out; a.out_stride = out_stride;
  a.heads = heads; a.Nq = Nq; a.Nk = Nk; a.NKP = (Nk + 15) & ~15;
  a.B = B; a.m_tiles = (Nq + kQRows - 1) / kQRows;
  a.scale_log2e = scale * 1.4426950408889634f;
  {
    const char* e = getenv("PT_ATTN_TRACE");              // developer hook: device address of 320 int64 (hex or decimal)
    a.trace = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
  }
  if ((out_stride & 7) || (reinterpret_cast<uintptr_t>(out) & 15))
    return pt::fail("pt_attention: out must be 16-byte aligned with a row stride multiple of 8 elements");
  CUtensorMap tq, tk, tv;
  a.q_box = (a.m_tiles == 1) ? ((Nq + 7) & ~7) : kQRows;
  if (make_map(&tq, q, (long long)B * Nq, (long long)heads * 64, q_stride, a.q_box)) return 1;
  // Configurations in order of preference.  The unit chain (load -> MMA -> softmax -> MMA -> store) is latency bound, so
  // two co-resident CTAs per SM (two softmax warp groups) beat deeper buffering inside one CTA.
  //   kv_rows: the K / V boxes carry NKP rows, or only Nk rounded up to 8 when that is what lets a second CTA fit.  In
  //   the short-box layout the K / V regions are kv_rows * 128 bytes apart and the MMAs (which always read NKP rows)
  //   run up to 1 KB into the neighbouring region: for K those are masked score columns, for V they meet P = 0, and the
  //   neighbour holds finite bf16 data (V, Q or probabilities; zeros at start), so the products vanish.
  const size_t sm_smem = 227 * 1024, ptile = (size_t)((a.NKP + 63) / 64) * kAtomBytes;
  const int kv8 = (Nk + 7) & ~7;
  a.q_region = (a.q_box * 128 + 1023) & ~1023;
  struct Cand { int per_sm, qk_stages, v_stages, s_bufs, p_bufs, kv_rows; };
  // single-pass softmax (<= 96 keys) frees S after one read, so ONE S accumulator is enough there and the TMEM / smem go
  // to a second P tile + O accumulator instead; the two-pass path keeps S for a whole softmax and prefers two S buffers.
  // V needs its second slot more than Q / K do: a V slot is only released by the P V at the END of its unit.
  const int nch_host = (a.NKP <= 96 && a.NKP % 32 == 0) ? a.NKP / 32 : 0;
  const Cand cands_1p[] = {{2, 2, 2, 1, 2, a.NKP}, {2, 1, 2, 1, 2, a.NKP}, {2, 1, 2, 1, 2, kv8}, {2, 2, 2, 1, 1, a.NKP},
                           {2, 1, 1, 1, 2, a.NKP}, {2, 1, 1, 1, 1, a.NKP}, {1, 2, 2, 1, 2, a.NKP}, {1, 2, 2, 1, 1, a.NKP},
                           {1, 1, 1, 1, 1, a.NKP}};
  const Cand cands_2p[] = {{2, 2, 2, 2, 1, a.NKP}, {2, 2, 2, 1, 1, a.NKP}, {2, 1, 1, 2, 1, a.NKP}, {2, 1, 1, 1, 1, a.NKP},
                           {2, 1, 1, 1, 1, kv8}, {1, 2, 2, 2, 1, a.NKP}, {1, 2, 2, 1, 1, a.NKP}, {1, 1, 1, 1, 1, a.NKP}};
  const Cand* cands = nch_host ? cands_1p : cands_2p;
  const int n_cands = nch_host ? (int)(sizeof(cands_1p) / sizeof(Cand)) : (int)(sizeof(cands_2p) / sizeof(Cand));
  int per_sm = 0;
  size_t smem_need = 0;
  for (int ci = 0; ci < n_cands; ++ci) {
    const Cand& c = cands[ci];
    const int cols = c.s_bufs * a.NKP + 64 * c.p_bufs;
    int tc = 32;
    while (tc < cols) tc *= 2;
    if (tc > 512 || tc * c.per_sm > 512) continue;
    if (c.kv_rows != a.NKP && (a.NKP - c.kv_rows) * 128 > 1024) continue;
    const size_t kvp = ((size_t)c.kv_rows * 128 + 1023) & ~(size_t)1023;
    const size_t need = c.qk_stages * ((size_t)a.q_region + kvp) + c.v_stages * kvp + c.p_bufs * ptile;
    // per CTA: 1 KB alignment slack + buffers + ~1.3 KB static + 1 KB reserved by the driver
    if ((1024 + need + 1280 + 1024) * c.per_sm > sm_smem + 1024) continue;
    per_sm = c.per_sm; a.qk_stages = c.qk_stages; a.v_stages = c.v_stages; a.s_bufs = c.s_bufs; a.p_bufs = c.p_bufs;
    a.kv_rows = c.kv_rows; a.tmem_cols = tc;
    smem_need = need;
    break;
  }
  if (!per_sm) return pt::fail("pt_attention: Nk=%d does not fit in shared memory", Nk);
  a.o_col = a.s_bufs * a.NKP;
  if (make_map(&tk, k, (long long)B * Nk, (long long)heads * 64, k_stride, a.kv_rows)) return 1;
  if (make_map(&tv, v, (long long)B * Nk, (long long)heads * 64, v_stride, a.kv_rows)) return 1;
  const size_t smem = 1024 + smem_need;
  const int sms = pt::sm_count();
  const int n_units = B * heads * a.m_tiles;
  const int grid_max = sms * per_sm;
  const int grid = n_units < grid_max ? n_units : grid_max;
  const int nch = nch_host;
#define PT_ATTN_LAUNCH(K)                                                                                              \
  do {                                                                                                                 \
    PT_SMEM_OPT_IN(attention_tc_kernel<K>, smem);                                                                      \
    PT_CUDA(pt::launch_pdl(attention_tc_kernel<K>, dim3(grid), dim3(kThreadsTc), smem, st, 1, tq, tk, tv, a));         \
  } while (0)
  if (nch == 1) PT_ATTN_LAUNCH(1);
  else if (nch == 2) PT_ATTN_LAUNCH(2);
  else if (nch == 3) PT_ATTN_LAUNCH(3);
  else PT_ATTN_LAUNCH(0);
#undef PT_ATTN_LAUNCH
  PT_CUDA(cudaGetLastError());
  return 0;
}
