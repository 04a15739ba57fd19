// probtalk_b200 -- multi-head attention on tcgen05 tensor cores (sm_100a), sequences up to 256 keys.
//
// Same contract as pt_attention (scaled-dot-product core of nn.MultiheadAttention, reference
// nets/inpainting/PTransformer.py:340,350).  One work unit = (sample, head, 128-query tile):
//     S[128 x Nk]  = Q K^T        tcgen05.mma, Q and K straight from TMA (SWIZZLE_128B, K-major)
//     P            = softmax(S)   one thread per query row, S read from TMEM, P written to smem as bf16 (K-major)
//     O[128 x 64]  = P V          tcgen05.mma, V used in place as an MN-major operand (no transpose)
// Persistent CTAs: warp 0 = TMA producer (double-buffered Q/K/V), warp 1 = MMA issuer, warps 2-5 = softmax (TMEM lane =
// query row), warps 6-9 = output (O read-out, 1/sum scaling, stores).  The S MMA and softmax of unit u+1 overlap the
// P V MMA and the output stage of unit u.
#include <cuda.h>
#include "pt_common.cuh"
#include "pt_host.h"

namespace {

constexpr int kThreadsTc = 10 * 32;   // TMA, MMA, 4 softmax warps, 4 output warps
constexpr int kQRows = 128;
constexpr uint32_t kQBytes = kQRows * 128;          // 16 KB
constexpr uint32_t kAtomBytes = kQRows * 128;       // one [128 x 64] bf16 swizzle atom of P

struct AttnTcArgs {
  __nv_bfloat16* out;
  long long out_stride;
  int heads, Nq, Nk, NKP;          // NKP = keys padded to 16
  int B, m_tiles;                  // m_tiles = ceil(Nq / 128)
  int tmem_cols, o_col;            // TMEM allocation (power of two) and the column where O starts
  int nstage;                      // Q/K/V smem stages (2 = prefetch the next unit)
  int s_bufs;                      // S accumulators in TMEM (2 = the next unit's Q K^T overlaps this unit's softmax)
  float scale_log2e;
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

// smem matrix descriptor, MN-major operand, SWIZZLE_128B: 64 MN-elements (128 B) per row, rows = K index;
// SBO = stride between groups of 8 K-rows (1024 B); LBO = stride between 64-element MN atoms (single atom here)
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1024 >> 4) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__global__ void __launch_bounds__(kThreadsTc, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const AttnTcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t kv_bytes = (uint32_t)a.NKP * 128u;
  const uint32_t kv_pad = (kv_bytes + 1023u) & ~1023u;
  const uint32_t stage_bytes = kQBytes + 2 * kv_pad;
  const int p_atoms = (a.NKP + 63) / 64;
  uint8_t* p_smem = smem + (size_t)a.nstage * stage_bytes;
  __shared__ uint64_t bars[13];
  __shared__ float sums_s[2][kQRows];          // softmax row sums, handed from the softmax warps to the output warps
  uint64_t* qkv_full = bars;        // [2]
  uint64_t* qkv_empty = bars + 2;   // [2]
  uint64_t* s_full = bars + 4;      // [2]  (S is double-buffered in TMEM when it fits)
  uint64_t* s_free = bars + 6;      // [2]
  uint64_t* p_full = bars + 8;
  uint64_t* o_full = bars + 9;
  uint64_t* o_free = bars + 10;
  uint64_t* sums_full = bars + 11;  // [2]
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = a.B * a.heads * a.m_tiles;

  if (warp == 0 && lane == 0) {
    pt::tma_prefetch_desc(&tmap_q); pt::tma_prefetch_desc(&tmap_k); pt::tma_prefetch_desc(&tmap_v);
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&qkv_full[i], 1); pt::mbar_init(&qkv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&s_full[i], 1); pt::mbar_init(&s_free[i], 4); }
    pt::mbar_init(p_full, 4);
    pt::mbar_init(o_full, 1); pt::mbar_init(o_free, 4);
    pt::mbar_init(&sums_full[0], 4); pt::mbar_init(&sums_full[1], 4);
    pt::fence_mbar_init();
  }
  if (warp == 1) { pt::tmem_alloc(&tmem_slot, (uint32_t)a.tmem_cols); pt::tmem_relinquish(); }
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t tmem_s = tmem, tmem_o = tmem + (uint32_t)a.o_col;
  const bool dbl = (a.s_bufs == 2);

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int it = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
        const int st = (a.nstage == 2) ? (it & 1) : 0;
        const uint32_t ph = (uint32_t)((a.nstage == 2) ? (it >> 1) : it) & 1u;
        const int mt = u % a.m_tiles, bh = u / a.m_tiles;
        const int h = bh % a.heads, b = bh / a.heads;
        pt::mbar_wait(&qkv_empty[st], ph ^ 1);
        uint8_t* sq = smem + (size_t)st * stage_bytes;
        pt::mbar_expect_tx(&qkv_full[st], kQBytes + 2 * kv_bytes);
        pt::tma_load_2d(sq, &tmap_q, &qkv_full[st], h * 64, b * a.Nq + mt * kQRows);
        pt::tma_load_2d(sq + kQBytes, &tmap_k, &qkv_full[st], h * 64, b * a.Nk);
        pt::tma_load_2d(sq + kQBytes + kv_pad, &tmap_v, &qkv_full[st], h * 64, b * a.Nk);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc_s = pt::umma_idesc_bf16(kQRows, (uint32_t)a.NKP);
      const uint32_t idesc_o = pt::umma_idesc_bf16(kQRows, 64) | (1u << 16);       // B operand (V) is MN-major
      // S(u) = Q K^T into S buffer (u & 1 when double-buffered); issued one unit AHEAD of the P V MMA so that it runs
      // while the softmax warps are still busy with the previous unit
      auto issue_s = [&](int it) {
        const int st = (a.nstage == 2) ? (it & 1) : 0;
        const uint32_t ph = (uint32_t)((a.nstage == 2) ? (it >> 1) : it) & 1u;
        const int sb = dbl ? (it & 1) : 0;
        const uint32_t sph = (uint32_t)(dbl ? (it >> 1) : it) & 1u;
        const uint32_t sq = pt::smem_u32(smem + (size_t)st * stage_bytes);
        const uint32_t sk = sq + kQBytes;
        pt::mbar_wait(&qkv_full[st], ph);
        pt::mbar_wait(&s_free[sb], sph ^ 1);             // the softmax that last used this S buffer has read it
        pt::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          pt::umma_bf16(tmem_s + sb * a.NKP, pt::umma_desc_sw128(sq + k * 32), pt::umma_desc_sw128(sk + k * 32), idesc_s, k ? 1u : 0u);
        pt::umma_commit(&s_full[sb]);
      };
      const bool ahead = dbl && a.nstage == 2;
      int it = 0;
      if (ahead && (int)blockIdx.x < n_units) issue_s(0);
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
        const int st = (a.nstage == 2) ? (it & 1) : 0;
        const uint32_t up = (uint32_t)it & 1u;
        if (ahead) { if (u + (int)gridDim.x < n_units) issue_s(it + 1); }
        else issue_s(it);
        const uint32_t sv = pt::smem_u32(smem + (size_t)st * stage_bytes) + kQBytes + kv_pad;
        pt::mbar_wait(p_full, up);                       // P (bf16) is in smem
        pt::mbar_wait(o_free, up ^ 1);                   // the previous O has been read out of TMEM
        pt::tc_fence_after();
        const uint32_t sp = pt::smem_u32(p_smem);
        const int ksteps = a.NKP / 16;
        for (int j = 0; j < ksteps; ++j)
          pt::umma_bf16(tmem_o, pt::umma_desc_sw128(sp + (j >> 2) * kAtomBytes + (j & 3) * 32),
                        umma_desc_sw128_mn(sv + j * 2048), idesc_o, j ? 1u : 0u);
        pt::umma_commit(o_full);
        pt::umma_commit(&qkv_empty[st]);                 // Q/K/V stage (and P) consumed
      }
    }
  } else if (warp < 6) {
    // ------------------------------------------------------------ softmax (thread = query row)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const uint32_t up = (uint32_t)it & 1u;
      const int sb = dbl ? (it & 1) : 0;
      const uint32_t sph = (uint32_t)(dbl ? (it >> 1) : it) & 1u;
      const uint32_t tmem_sb = tmem_s + sb * a.NKP;
      pt::mbar_wait(&s_full[sb], sph);
      pt::tc_fence_after();
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
      if (it > 0) pt::mbar_wait(o_full, up ^ 1);
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
        uint8_t* atom = p_smem + (size_t)(c0 >> 6) * kAtomBytes;
        const int chunk0 = (c0 & 63) >> 3;               // 16-byte chunk index of key c0 within the 64-key atom
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(atom + pt::sw128_offset(row, chunk0 + i)) =
              make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      }
      const float sum = sum0 + sum1;
      sums_s[up][row] = sum;
      pt::tc_fence_before();
      pt::fence_async_smem();                            // P stores -> visible to the tensor-core proxy
      __syncwarp();
      if (lane == 0) { pt::mbar_arrive(&s_free[sb]); pt::mbar_arrive(p_full); pt::mbar_arrive(&sums_full[up]); }
    }
  } else {
    // ------------------------------------------------------------ output: O = P V read-out, 1/sum, stores
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const uint32_t up = (uint32_t)it & 1u;
      const int mt = u % a.m_tiles, bh = u / a.m_tiles;
      const int h = bh % a.heads, b = bh / a.heads;
      pt::mbar_wait(&sums_full[up], (uint32_t)(it >> 1) & 1u);
      pt::mbar_wait(o_full, up);
      pt::tc_fence_after();
      const float inv = 1.f / sums_s[up][row];
      const int qrow = mt * kQRows + row;
      uint32_t o0[32], o1[32];
      pt::tmem_ld32(tmem_o + lane_off, o0);
      pt::tmem_ld32(tmem_o + lane_off + 32, o1);
      pt::tmem_ld_wait();
      pt::tc_fence_before();
      __syncwarp();
      if (lane == 0) pt::mbar_arrive(o_free);
      if (qrow < a.Nq) {
        uint4* dst = reinterpret_cast<uint4*>(a.out + ((size_t)b * a.Nq + qrow) * a.out_stride + h * 64);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          dst[i] = make_uint4(pt::pack_bf16x2(__uint_as_float(o0[8 * i]) * inv, __uint_as_float(o0[8 * i + 1]) * inv),
                              pt::pack_bf16x2(__uint_as_float(o0[8 * i + 2]) * inv, __uint_as_float(o0[8 * i + 3]) * inv),
                              pt::pack_bf16x2(__uint_as_float(o0[8 * i + 4]) * inv, __uint_as_float(o0[8 * i + 5]) * inv),
                              pt::pack_bf16x2(__uint_as_float(o0[8 * i + 6]) * inv, __uint_as_float(o0[8 * i + 7]) * inv));
#pragma unroll
        for (int i = 0; i < 4; ++i)
          dst[4 + i] = make_uint4(pt::pack_bf16x2(__uint_as_float(o1[8 * i]) * inv, __uint_as_float(o1[8 * i + 1]) * inv),
                                  pt::pack_bf16x2(__uint_as_float(o1[8 * i + 2]) * inv, __uint_as_float(o1[8 * i + 3]) * inv),
                                  pt::pack_bf16x2(__uint_as_float(o1[8 * i + 4]) * inv, __uint_as_float(o1[8 * i + 5]) * inv),
                                  pt::pack_bf16x2(__uint_as_float(o1[8 * i + 6]) * inv, __uint_as_float(o1[8 * i + 7]) * inv));
      }
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
  if ((out_stride & 7) || (reinterpret_cast<uintptr_t>(out) & 15))
    return pt::fail("pt_attention: out must be 16-byte aligned with a row stride multiple of 8 elements");
  CUtensorMap tq, tk, tv;
  if (make_map(&tq, q, (long long)B * Nq, (long long)heads * 64, q_stride, kQRows)) return 1;
  if (make_map(&tk, k, (long long)B * Nk, (long long)heads * 64, k_stride, a.NKP)) return 1;
  if (make_map(&tv, v, (long long)B * Nk, (long long)heads * 64, v_stride, a.NKP)) return 1;
  const uint32_t kv_pad = ((uint32_t)a.NKP * 128u + 1023u) & ~1023u;
  const size_t stage = kQBytes + 2 * (size_t)kv_pad, ptile = (size_t)((a.NKP + 63) / 64) * kAtomBytes;
  // TMEM: S (NKP columns) + O (64 columns), rounded to a power of two; smaller allocations let several CTAs share an SM
  a.s_bufs = (2 * a.NKP + 64 <= 512) ? 2 : 1;
  a.o_col = a.s_bufs * a.NKP;
  a.tmem_cols = 32;
  while (a.tmem_cols < a.o_col + 64) a.tmem_cols *= 2;
  const int by_tmem = 512 / a.tmem_cols;
  // the unit chain (MMA -> softmax -> MMA -> store) is latency bound, so co-resident CTAs matter more than prefetch depth
  const size_t sm_smem = 227 * 1024;
  int per_sm = 1;
  a.nstage = 2;
  for (int want = (by_tmem < 2 ? by_tmem : 2); want >= 1; --want) {
    if ((1024 + 2 * stage + ptile + 1024) * want <= sm_smem) { per_sm = want; a.nstage = 2; break; }
    if ((1024 + stage + ptile + 1024) * want <= sm_smem) { per_sm = want; a.nstage = 1; break; }
  }
  const size_t smem = 1024 + a.nstage * stage + ptile;
  if (smem > sm_smem) return pt::fail("pt_attention: Nk=%d does not fit in shared memory", Nk);
  static size_t configured = 0;
  if (smem > configured) {
    PT_CUDA(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  const int sms = pt::sm_count();
  const int n_units = B * heads * a.m_tiles;
  const int grid_max = sms * per_sm;
  attention_tc_kernel<<<n_units < grid_max ? n_units : grid_max, kThreadsTc, smem, st>>>(tq, tk, tv, a);
  PT_CUDA(cudaGetLastError());
  return 0;
}
