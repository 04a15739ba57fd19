// probtalk_b200 -- product-quantizer nearest-code search and codebook gather (sm_100a).
//
// Replaces ProductQuantization.get_code_indices + the per-group loop of .forward (eval) and
// .quantize_all of the reference (nets/spg/vqvae_modules.py:546-555, :470-491, :557-569).
//
// Search = dense contraction:  score[r, j] = |c_j|^2 - 2 * x_r . c_j   (|x_r|^2 is constant per row
// and cannot change the argmin).  One CTA owns up to 128 query rows of ONE sub-codebook:
//   1. all 8 warps stage the rows and the 128x128 codebook into shared memory as split-bf16
//      operands (x = x_hi + x_lo, c = c_hi + c_lo), K-major, SWIZZLE_128B;
//   2. one thread issues 24 tcgen05.mma (hi*hi + hi*lo + lo*hi) into a 128x128 fp32 TMEM tile;
//      dropped terms are O(2^-16), so the TMEM scores are within
//      E = 2.5e-4 * |x| * max|c| (+ fp32 rounding) of the exact ones (checked on hardware by
//      tests/test_pq_search.py through the score dump);
//   3. the TMEM tile is scanned (lane = row, tcgen05.ld) for the smallest and second smallest score;
//   4. rows whose top-2 gap is inside the error band are re-scored in fp32 with the reference's own
//      formula  (sum x^2 + sum c^2) - 2 * dot,  FMA accumulation, lowest index winning ties -- so the
//      emitted index is the fp32 answer for every row, and differs from the reference's only where
//      the reference's own summation order decides (true gap below a few ulp of |x|^2).
//
// HBM traffic: z once (fp32), codebook from L2, int64 indices out.  No scratch memory.
#include "pt_common.cuh"
#include "pt_host.h"

namespace {

constexpr int kD = 128;          // sub-vector width
constexpr int kK = 128;          // codes per sub-codebook
constexpr int kTileRows = 128;   // UMMA M
constexpr int kThreads = 256;
constexpr uint32_t kOpBytes = kTileRows * kD * 2;   // one bf16 operand tile: 32 KB (two 64-col swizzle atoms)
constexpr uint32_t kAtomBytes = kTileRows * 128;    // 16 KB
constexpr float kBand = 2.5e-4f;                    // rigorous bound is ~9.3e-5 * |x||c| on the dot, x2 for the score

struct PqSmem {
  // operands (each 1024-B aligned): A_hi, A_lo, B_hi, B_lo
  uint8_t ops[4][kOpBytes];
  float sc2[kK];           // |c_j|^2 (fp32, deterministic order)
  float sx2[kTileRows];    // |x_r|^2 (only used for the error band)
  float pmin[2][kTileRows];
  float pmin2[2][kTileRows];
  int pidx[2][kTileRows];
  unsigned long long cand[2][kTileRows];
  float rmin[kTileRows];
  float band[kTileRows];
  int ambiguous[kTileRows];
  float cmax_w[kThreads / 32];
  uint64_t mma_done;
  uint32_t tmem_base;
  int n_amb;
};

__device__ __forceinline__ void split_store(uint8_t* hi_tile, uint8_t* lo_tile, uint32_t off, const float (&v)[8]) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * i]), h1 = __float2bfloat16_rn(v[2 * i + 1]);
    float r0 = v[2 * i] - __bfloat162float(h0), r1 = v[2 * i + 1] - __bfloat162float(h1);
    __nv_bfloat162 hh; hh.x = h0; hh.y = h1;
    h[i] = *reinterpret_cast<uint32_t*>(&hh);
    l[i] = pt::pack_bf16x2(r0, r1);
  }
  *reinterpret_cast<uint4*>(hi_tile + off) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(lo_tile + off) = make_uint4(l[0], l[1], l[2], l[3]);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct PqArgs {
  const float* z;
  long long zs_b, zs_w, zs_c;   // element strides of z[b, c, w]
  int B, W, G;
  const float* codebooks;       // [G, 128, 128] fp32
  long long* idx_out;           // [B*W, G]
  float* gap_out;               // [B*W, G] or null
  float* score_dump;            // [B*W, G, 128] or null (test hook)
  int* n_rescored;              // single counter or null (test hook)
  int rows_per_tile;
  int rows_fast;                 // 1: channels-last rows, 16-byte aligned -> float4 row loader
};

template <bool kDump>
__global__ void __launch_bounds__(kThreads, 1) pq_search_kernel(PqArgs a) {
  extern __shared__ uint8_t smem_raw[];
  PqSmem& s = *reinterpret_cast<PqSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = blockIdx.y;
  const long long total_rows = (long long)a.B * a.W;
  const long long row0 = (long long)blockIdx.x * a.rows_per_tile;
  const int nrows = (int)min((long long)a.rows_per_tile, total_rows - row0);
  const float* cb = a.codebooks + (size_t)g * kK * kD;

  if (warp == 0) {
    pt::tmem_alloc(&s.tmem_base, 128);
    pt::tmem_relinquish();
  }
  if (tid == 0) {
    pt::mbar_init(&s.mma_done, 1);
    pt::fence_mbar_init();
    s.n_amb = 0;
  }
  if (tid < kTileRows) s.sx2[tid] = 0.f;
  __syncthreads();

  // ---- stage the codebook: warp per code row, lane holds k = 4*lane .. 4*lane+3 (coalesced 512 B).
  // All 16 rows of a warp are requested before any is consumed (memory-level parallelism).
  {
    constexpr int kPerWarp = kK / (kThreads / 32);
    float4 cv[kPerWarp];
#pragma unroll
    for (int i = 0; i < kPerWarp; ++i)
      cv[i] = __ldg(reinterpret_cast<const float4*>(cb + (size_t)(warp + i * (kThreads / 32)) * kD) + lane);
    float sq[kPerWarp];
#pragma unroll
    for (int i = 0; i < kPerWarp; ++i)
      sq[i] = fmaf(cv[i].w, cv[i].w, fmaf(cv[i].z, cv[i].z, fmaf(cv[i].y, cv[i].y, cv[i].x * cv[i].x)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int i = 0; i < kPerWarp; ++i) sq[i] += __shfl_xor_sync(0xffffffffu, sq[i], o);   // 16 independent chains
    }
    float wmax = 0.f;
#pragma unroll
    for (int i = 0; i < kPerWarp; ++i) {
      const int j = warp + i * (kThreads / 32);
      const float4 c = cv[i];
      wmax = fmaxf(wmax, sq[i]);
      if (lane == 0) s.sc2[j] = sq[i];
      __nv_bfloat16 h0 = __float2bfloat16_rn(c.x), h1 = __float2bfloat16_rn(c.y), h2 = __float2bfloat16_rn(c.z),
                    h3 = __float2bfloat16_rn(c.w);
      __nv_bfloat162 a0, a1; a0.x = h0; a0.y = h1; a1.x = h2; a1.y = h3;
      uint2 hi = make_uint2(*reinterpret_cast<uint32_t*>(&a0), *reinterpret_cast<uint32_t*>(&a1));
      uint2 lo = make_uint2(pt::pack_bf16x2(c.x - __bfloat162float(h0), c.y - __bfloat162float(h1)),
                            pt::pack_bf16x2(c.z - __bfloat162float(h2), c.w - __bfloat162float(h3)));
      const int k0 = lane * 4;                       // element index within the row
      const uint32_t off = (k0 >> 6) * kAtomBytes + pt::sw128_offset(j, (k0 & 63) >> 3) + (k0 & 7) * 2;
      *reinterpret_cast<uint2*>(s.ops[2] + off) = hi;
      *reinterpret_cast<uint2*>(s.ops[3] + off) = lo;
    }
    if (lane == 0) s.cmax_w[warp] = sqrtf(wmax);
  }

  // ---- stage the query rows (row index math in 32 bits: tiles never span more than 2^31 rows)
  const int W = a.W;
  const int b0 = (int)(row0 / W), w0 = (int)(row0 % W);
  if (a.rows_fast) {
    // channels-last rows: warp per row, float4 per lane; 16 rows in flight per warp
    constexpr int kPerWarp = kTileRows / (kThreads / 32);
    float4 xv[kPerWarp];
#pragma unroll
    for (int i = 0; i < kPerWarp; ++i) {
      const int r = warp + i * (kThreads / 32);
      xv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < nrows) {
        const int q = w0 + r, b = b0 + q / W, w = q % W;
        xv[i] = __ldg(reinterpret_cast<const float4*>(a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD) + lane);
      }
    }
#pragma unroll
    for (int i = 0; i < kPerWarp; ++i) {
      const int r = warp + i * (kThreads / 32);
      const float4 x = xv[i];
      float sq = warp_sum(fmaf(x.w, x.w, fmaf(x.z, x.z, fmaf(x.y, x.y, x.x * x.x))));
      if (lane == 0) s.sx2[r] = sq;
      __nv_bfloat16 h0 = __float2bfloat16_rn(x.x), h1 = __float2bfloat16_rn(x.y), h2 = __float2bfloat16_rn(x.z),
                    h3 = __float2bfloat16_rn(x.w);
      __nv_bfloat162 a0, a1; a0.x = h0; a0.y = h1; a1.x = h2; a1.y = h3;
      uint2 hi = make_uint2(*reinterpret_cast<uint32_t*>(&a0), *reinterpret_cast<uint32_t*>(&a1));
      uint2 lo = make_uint2(pt::pack_bf16x2(x.x - __bfloat162float(h0), x.y - __bfloat162float(h1)),
                            pt::pack_bf16x2(x.z - __bfloat162float(h2), x.w - __bfloat162float(h3)));
      const int k0 = lane * 4;
      const uint32_t off = (k0 >> 6) * kAtomBytes + pt::sw128_offset(r, (k0 & 63) >> 3) + (k0 & 7) * 2;
      *reinterpret_cast<uint2*>(s.ops[0] + off) = hi;
      *reinterpret_cast<uint2*>(s.ops[1] + off) = lo;
    }
  } else {
    // channel-first (the reference's [B, C, W]): lanes walk consecutive rows (w contiguous), 8 channels per
    // unit; a thread's units share its row, and 4 units (32 loads) are in flight at a time
    const int r = tid & (kTileRows - 1);
    const bool live = r < nrows;
    const int q = w0 + r, b = b0 + q / W, w = q % W;
    const float* prow = a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD * a.zs_c;
    float sq = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int chunk = (tid >> 7) + 2 * (h * 4 + u);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[u][i] = live ? __ldg(prow + (size_t)(chunk * 8 + i) * a.zs_c) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int chunk = (tid >> 7) + 2 * (h * 4 + u);
#pragma unroll
        for (int i = 0; i < 8; ++i) sq = fmaf(v[u][i], v[u][i], sq);
        const uint32_t off = (chunk >> 3) * kAtomBytes + pt::sw128_offset(r, chunk & 7);
        split_store(s.ops[0], s.ops[1], off, v[u]);
      }
    }
    if (live) atomicAdd(&s.sx2[r], sq);
  }
  pt::fence_async_smem();        // generic-proxy smem writes -> visible to the tensor-core (async) proxy
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  // ---- 24 MMAs by one thread: D = Ahi*Bhi^T + Ahi*Blo^T + Alo*Bhi^T
  if (tid == 0) {
    const uint32_t idesc = pt::umma_idesc_bf16(128, 128);
    const uint32_t a_hi = pt::smem_u32(s.ops[0]), a_lo = pt::smem_u32(s.ops[1]);
    const uint32_t b_hi = pt::smem_u32(s.ops[2]), b_lo = pt::smem_u32(s.ops[3]);
    const uint32_t pa[3] = {a_hi, a_hi, a_lo}, pb[3] = {b_hi, b_lo, b_hi};
    uint32_t acc = 0;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
#pragma unroll
      for (int k = 0; k < kD / 16; ++k) {
        const uint32_t koff = (k >> 2) * kAtomBytes + (k & 3) * 32;
        pt::umma_bf16(tmem, pt::umma_desc_sw128(pa[p] + koff), pt::umma_desc_sw128(pb[p] + koff), idesc, acc);
        acc = 1;
      }
    }
    pt::umma_commit(&s.mma_done);
  }
  pt::mbar_wait(&s.mma_done, 0);
  pt::tc_fence_after();

  // ---- scan: warp w reads TMEM lanes 32*(w%4).., columns 64*(w/4)..+63
  const int r = (warp & 3) * 32 + lane;
  const int half = warp >> 2;
  const uint32_t t_row = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + half * 64;
  // two independent (min, argmin, second-min) trackers (even / odd columns) for instruction-level parallelism
  float ma1 = INFINITY, ma2 = INFINITY, mb1 = INFINITY, mb2 = INFINITY;
  int ia = 0, ib = 1;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint32_t v[32];
    pt::tmem_ld32(t_row + c * 32, v);
    pt::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      const int j = half * 64 + c * 32 + i;
      const float sa = fmaf(-2.f, __uint_as_float(v[i]), s.sc2[j]);
      const float sb = fmaf(-2.f, __uint_as_float(v[i + 1]), s.sc2[j + 1]);
      if constexpr (kDump) {
        if (r < nrows) {
          a.score_dump[((row0 + r) * a.G + g) * kK + j] = sa;
          a.score_dump[((row0 + r) * a.G + g) * kK + j + 1] = sb;
        }
      }
      ma2 = fminf(ma2, fmaxf(sa, ma1));           // second-min update, then min (strict < keeps the first index)
      if (sa < ma1) { ma1 = sa; ia = j; }
      mb2 = fminf(mb2, fmaxf(sb, mb1));
      if (sb < mb1) { mb1 = sb; ib = j + 1; }
    }
  }
  // merge the trackers; on equal scores the smaller index wins
  float m1, m2;
  int i1;
  if (mb1 < ma1 || (mb1 == ma1 && ib < ia)) { m1 = mb1; i1 = ib; m2 = fminf(ma1, mb2); }
  else { m1 = ma1; i1 = ia; m2 = fminf(mb1, ma2); }
  s.pmin[half][r] = m1; s.pmin2[half][r] = m2; s.pidx[half][r] = i1;
  __syncthreads();
  bool amb = false;
  if (half == 0) {
    float n1 = s.pmin[1][r], n2 = s.pmin2[1][r];
    int k1 = s.pidx[1][r];
    // merge halves; the low half wins ties (first minimum, like torch.argmin)
    float best = m1, second = fminf(m2, n1);
    int bi = i1;
    if (n1 < m1) { best = n1; bi = k1; second = fminf(m1, n2); }
    float cmax = s.cmax_w[0];
#pragma unroll
    for (int i = 1; i < kThreads / 32; ++i) cmax = fmaxf(cmax, s.cmax_w[i]);
    const float band = 2.f * kBand * sqrtf(s.sx2[r]) * cmax + 4e-7f * fabsf(best);
    s.rmin[r] = best; s.band[r] = band;
    amb = (r < nrows) && (second - best <= band);
    s.ambiguous[r] = amb ? 1 : 0;
    if (amb) atomicAdd(&s.n_amb, 1);
    if (!amb && r < nrows) {
      a.idx_out[(row0 + r) * a.G + g] = bi;
      if (a.gap_out) a.gap_out[(row0 + r) * a.G + g] = second - best;
    }
  }
  __syncthreads();
  if (s.n_amb > 0) {                      // block-uniform
    // second pass: candidate mask of every code inside the band (all warps re-read their columns)
    const float lim = s.rmin[r] + s.band[r];
    unsigned long long mask = 0ull;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      pt::tmem_ld32(t_row + c * 32, v);
      pt::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float sc = fmaf(-2.f, __uint_as_float(v[i]), s.sc2[half * 64 + c * 32 + i]);
        if (sc <= lim) mask |= 1ull << (c * 32 + i);
      }
    }
    s.cand[half][r] = mask;
    __syncthreads();
    // fp32 re-score, warp per ambiguous row, lanes split k; reference formula (vqvae_modules.py:549-553)
    for (int rr = warp; rr < nrows; rr += kThreads / 32) {
      if (!s.ambiguous[rr]) continue;
      const long long row = row0 + rr;
      const int q = w0 + rr, b = b0 + q / W, w = q % W;
      const float* px = a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD * a.zs_c;
      float x[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) x[i] = __ldg(px + (size_t)(lane + 32 * i) * a.zs_c);
      float sx2 = warp_sum(fmaf(x[3], x[3], fmaf(x[2], x[2], fmaf(x[1], x[1], x[0] * x[0]))));
      float best = INFINITY, second = INFINITY;
      int bi = 0;
      for (int h = 0; h < 2; ++h) {
        unsigned long long m = s.cand[h][rr];
        while (m) {
          const int j = h * 64 + __ffsll((long long)m) - 1;
          m &= m - 1;
          const float* pc = cb + (size_t)j * kD;
          float dot = 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) dot = fmaf(x[i], __ldg(pc + lane + 32 * i), dot);
          dot = warp_sum(dot);
          const float d = (sx2 + s.sc2[j]) - 2.f * dot;
          if (d < best) { second = best; best = d; bi = j; }      // ascending j: first minimum wins ties
          else if (d < second) { second = d; }
        }
      }
      if (lane == 0) {
        a.idx_out[row * a.G + g] = bi;
        if (a.gap_out) a.gap_out[row * a.G + g] = second - best;
        if (a.n_rescored) atomicAdd(a.n_rescored, 1);
      }
    }
  }
  pt::tc_fence_before();
  __syncthreads();
  if (warp == 0) pt::tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------------------------
// gather: indices -> quantized vectors (ProductQuantization.quantize_all, vqvae_modules.py:557-564)
// out element (b, w, g*128 + d) at out + b*os_b + w*os_w + (g*128+d)*os_c ; fp32 or bf16
template <typename OutT, typename IdxT>
__global__ void pq_gather_kernel(const IdxT* __restrict__ idx, const float* __restrict__ cb, OutT* __restrict__ out,
                                 long long os_b, long long os_w, long long os_c, int B, int W, int G, int* bad) {
  const long long n = (long long)B * W * G * (kD / 4);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % (kD / 4));
    const long long e = i / (kD / 4);          // (row, g)
    const int g = (int)(e % G);
    const long long row = e / G;
    const long long b = row / W, w = row % W;
    long long j = (long long)idx[e];
    if (j < 0 || j >= kK) { if (bad) atomicAdd(bad, 1); j = 0; }
    const float4 v = __ldg(reinterpret_cast<const float4*>(cb + ((size_t)g * kK + j) * kD) + q);
    OutT* p = out + b * os_b + w * os_w + (long long)(g * kD + q * 4) * os_c;
    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if constexpr (sizeof(OutT) == 4) p[d * os_c] = vv[d];
      else p[d * os_c] = __float2bfloat16_rn(vv[d]);
    }
  }
}

int choose_rows_per_tile(long long rows, int G, int sms) {
  // fewest waves first, then the smallest tile (tile time grows with its rows)
  int best = kTileRows;
  double best_cost = 1e30;
  for (int R = 16; R <= kTileRows; R += 8) {
    const long long units = ((rows + R - 1) / R) * G;
    const long long waves = (units + sms - 1) / sms;
    const double cost = (double)waves * (40.0 + R);     // ~40 rows' worth of fixed cost per tile (codebook, MMA, scan)
    if (cost < best_cost - 1e-9) { best_cost = cost; best = R; }
  }
  return best;
}

}  // namespace

extern "C" int pt_pq_search(const float* z, long long zs_b, long long zs_c, long long zs_w, int B, int W, int G,
                            int K, int D, const float* codebooks, long long* idx_out, float* gap_out,
                            float* score_dump, int* n_rescored, void* stream) {
  if (K != kK || D != kD) return pt::fail("pt_pq_search: only K = D = 128 is built (got K=%d D=%d)", K, D);
  if (G < 1 || G > 65535) return pt::fail("pt_pq_search: bad group count %d", G);
  if (B < 0 || W < 0) return pt::fail("pt_pq_search: negative shape");
  if ((long long)B * W == 0) return 0;
  if (!z || !codebooks || !idx_out) return pt::fail("pt_pq_search: null pointer");
  if (reinterpret_cast<uintptr_t>(codebooks) & 15) return pt::fail("pt_pq_search: codebooks must be 16-byte aligned");
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  PqArgs a;
  a.z = z; a.zs_b = zs_b; a.zs_w = zs_w; a.zs_c = zs_c; a.B = B; a.W = W; a.G = G;
  a.codebooks = codebooks; a.idx_out = idx_out; a.gap_out = gap_out; a.score_dump = score_dump;
  a.n_rescored = n_rescored;
  a.rows_fast = (zs_c == 1 && !(reinterpret_cast<uintptr_t>(z) & 15) && !(zs_b & 3) && !(zs_w & 3)) ? 1 : 0;
  const long long rows = (long long)B * W;
  a.rows_per_tile = choose_rows_per_tile(rows, G, sms);
  const size_t smem = sizeof(PqSmem) + 1024;
  static bool attr_done[64] = {};
  if (!attr_done[dev & 63]) {
    PT_CUDA(cudaFuncSetAttribute(pq_search_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PT_CUDA(cudaFuncSetAttribute(pq_search_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done[dev & 63] = true;
  }
  dim3 grid((unsigned)((rows + a.rows_per_tile - 1) / a.rows_per_tile), (unsigned)G);
  if (score_dump) pq_search_kernel<true><<<grid, kThreads, smem, (cudaStream_t)stream>>>(a);
  else pq_search_kernel<false><<<grid, kThreads, smem, (cudaStream_t)stream>>>(a);
  PT_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int pt_pq_gather(const void* idx, int idx_is_i64, const float* codebooks, void* out, int out_is_bf16,
                            long long os_b, long long os_c, long long os_w, int B, int W, int G, int K, int D,
                            int* bad_count, void* stream) {
  if (K != kK || D != kD) return pt::fail("pt_pq_gather: only K = D = 128 is built (got K=%d D=%d)", K, D);
  const long long n = (long long)B * W * G * (kD / 4);
  if (n == 0) return 0;
  if (!idx || !codebooks || !out) return pt::fail("pt_pq_gather: null pointer");
  const int threads = 256;
  const int blocks = (int)min((long long)148 * 8, (n + threads - 1) / threads);
  cudaStream_t st = (cudaStream_t)stream;
  if (idx_is_i64) {
    if (out_is_bf16) pq_gather_kernel<__nv_bfloat16, long long><<<blocks, threads, 0, st>>>((const long long*)idx, codebooks, (__nv_bfloat16*)out, os_b, os_w, os_c, B, W, G, bad_count);
    else pq_gather_kernel<float, long long><<<blocks, threads, 0, st>>>((const long long*)idx, codebooks, (float*)out, os_b, os_w, os_c, B, W, G, bad_count);
  } else {
    if (out_is_bf16) pq_gather_kernel<__nv_bfloat16, int><<<blocks, threads, 0, st>>>((const int*)idx, codebooks, (__nv_bfloat16*)out, os_b, os_w, os_c, B, W, G, bad_count);
    else pq_gather_kernel<float, int><<<blocks, threads, 0, st>>>((const int*)idx, codebooks, (float*)out, os_b, os_w, os_c, B, W, G, bad_count);
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}
