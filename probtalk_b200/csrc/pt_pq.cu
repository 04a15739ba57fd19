// probtalk_b200 -- product-quantizer nearest-code search and codebook gather (sm_100a).
//
// Replaces ProductQuantization.get_code_indices + the per-group loop of .forward (eval) and
// .quantize_all of the reference (nets/spg/vqvae_modules.py:546-555, :470-491, :557-569).
//
// Search = dense contraction:  score[r, j] = |c_j|^2 - 2 * x_r . c_j   (|x_r|^2 is constant per row and
// cannot change the argmin).  Persistent, warp-specialised CTAs; a CTA is bound to ONE sub-codebook and
// walks 128-row query tiles of it:
//   prologue      the raw fp32 codebook chunk lands in the (still idle) second query-operand buffer with cp.async -- every
//                 thread's copies in flight at once --, then 8 warps split it into bf16 operands (c = c_hi + c_lo),
//                 K-major, SWIZZLE_128B, and reduce its squared norms;
//   loader warps  (8) read the next tile's fp32 rows straight from global memory (eight 512-byte rows in flight per
//                 warp), split them (x = x_hi + x_lo) and write the bf16 operand tile of the OTHER operand buffer while
//                 the previous tile is in the tensor core / being scanned;
//   MMA warp      one thread issues 24 tcgen05.mma per tile (hi*hi + hi*lo + lo*hi) into one of two 128-column
//                 TMEM accumulators; dropped terms are O(2^-16), so scores are within
//                 E = 2.5e-4 * |x| * max|c| of exact (checked on hardware by tests/test_gpu_pq.py);
//   scan warps    (4) read the TMEM tile (lane = row) for the smallest and second smallest score; rows whose
//                 top-2 gap is inside the error band are re-scored in fp32 with the reference's own formula
//                 (sum x^2 + sum c^2) - 2 * dot, FMA accumulation, lowest index winning ties -- so the emitted
//                 index is the fp32 answer for every row, and differs from the reference's only where the
//                 reference's own summation order decides (true gap below a few ulp of |x|^2).
// HBM traffic: z once (fp32), codebook from L2, int64 indices out.  No scratch memory.
#include "pt_common.cuh"
#include "pt_host.h"
#include <stdlib.h>

namespace {

constexpr int kK = 128;          // codes per staged codebook chunk (UMMA N); larger codebooks are split into chunks
constexpr int kTileRows = 128;   // UMMA M
constexpr int kLoadWarps = 8, kScanWarps = 4;
constexpr int kThreads = (kLoadWarps + 1 + kScanWarps) * 32;   // 416: loaders, MMA issuer, scanners
constexpr int kSxParts = 8;                                    // partial |x|^2 sums kept per row (summed by the scanner)
constexpr uint32_t kAtomBytes = kTileRows * 128;    // 16 KB: one 64-column SWIZZLE_128B atom of a 128-row operand tile
constexpr float kBand = 2.5e-4f;                    // rigorous bound is ~9.3e-5 * |x||c| on the dot, x2 for the score

template <int kD>                  // sub-vector width: 128 (ProductQuantization, vq.json) or 64 (VectorQuantizerEMA)
struct PqSmem {
  static constexpr uint32_t kOpBytes = kTileRows * kD * 2;   // one bf16 operand tile (kD / 64 swizzle atoms)
  uint8_t a_ops[2][2][kOpBytes];   // query tile, double-buffered: [buffer][hi | lo] (each 1024-B aligned)
  uint8_t b_ops[2][kOpBytes];      // codebook chunk hi | lo
  float sc2[kK];                   // |c_j|^2 (fp32, deterministic order); +inf for the padding rows of a short chunk
  float sx2p[4][kSxParts][kTileRows];   // partial |x_r|^2, ring of 4 tiles (only used for the error band)
  float cmax_w[8];
  uint64_t a_full[2], a_empty[2], t_full[2], t_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ void split_store(uint8_t* hi_tile, uint8_t* lo_tile, uint32_t off, const float (&v)[8]) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) pt::split_bf16x2(v[2 * i], v[2 * i + 1], h[i], l[i]);
  *reinterpret_cast<uint4*>(hi_tile + off) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(lo_tile + off) = make_uint4(l[0], l[1], l[2], l[3]);
}

// 4 consecutive fp32 -> hi / lo bf16 pairs at element k0 of row r (K-major SW128 tile of 64-column atoms)
__device__ __forceinline__ void split_store4(uint8_t* hi_tile, uint8_t* lo_tile, int r, int k0, const float4& c) {
  uint2 hi, lo;
  pt::split_bf16x2(c.x, c.y, hi.x, lo.x);
  pt::split_bf16x2(c.z, c.w, hi.y, lo.y);
  const uint32_t off = (k0 >> 6) * kAtomBytes + pt::sw128_offset(r, (k0 & 63) >> 3) + (k0 & 7) * 2;
  *reinterpret_cast<uint2*>(hi_tile + off) = hi;
  *reinterpret_cast<uint2*>(lo_tile + off) = lo;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// sum over the kLanes consecutive lanes that share a row (kLanes = 32 or 16)
template <int kLanes>
__device__ __forceinline__ float row_sum(float v) {
#pragma unroll
  for (int o = kLanes / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// order-preserving map fp32 -> uint32 (for atomicMin on packed (distance, index) words)
__device__ __forceinline__ unsigned ordered_bits(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct PqArgs {
  const float* z;
  long long zs_b, zs_w, zs_c;   // element strides of z[b, c, w]
  int B, W, G, K;
  const float* codebooks;       // [G, K, D] fp32
  long long* idx_out;           // [B*W, G]
  float* gap_out;               // [B*W, G] or null
  float* score_dump;            // [B*W, G, K] or null (test hook)
  int* n_rescored;              // single counter or null (test hook)
  int rows_per_tile, n_tiles, ctas_per_group;
  int n_chunks;                 // ceil(K / 128): a CTA is bound to ONE (sub-codebook, chunk of 128 codes)
  int rows_fast;                // 1: channels-last rows, 16-byte aligned -> float4 row loader
  long long* trace;             // developer hook (PT_PQ_TRACE): clock64 stamps of CTA 0's roles, [3 roles][32]
};

// Codebooks with more than 128 codes (VectorQuantizerEMA, code_num = 2048): every chunk's CTAs find the chunk's fp32
// winner for each row and fold it into idx_out with atomicMin on the packed word (ordered fp32 distance << 32 | index)
// -- the global fp32 argmin is the minimum over the chunk argmins, equal distances resolve to the lower index like
// torch.argmin.  idx_out is preset to ~0 before the launch and unpacked by this kernel afterwards.
__global__ void pq_unpack_kernel(long long* __restrict__ idx, long long n) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) idx[i] = (long long)(static_cast<unsigned long long>(idx[i]) & 0xffffffffull);
}

template <int kD, bool kDump>
__global__ void __launch_bounds__(kThreads, 1) pq_search_kernel(PqArgs a) {
  using Smem = PqSmem<kD>;
  constexpr int kRowLanes = kD / 4;                  // lanes that share one row in the float4 loaders (32 or 16)
  constexpr int kRowsPerInstr = 32 / kRowLanes;      // rows one warp-wide float4 access covers (1 or 2)
  extern __shared__ uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int vg = blockIdx.x / a.ctas_per_group, part = blockIdx.x % a.ctas_per_group;
  const int g = vg / a.n_chunks, chunk = vg - g * a.n_chunks;
  const int code0 = chunk * kK;                      // first code of this CTA's chunk
  const int n_codes = min(kK, a.K - code0);          // real codes in the chunk (the rest is padding)
  const bool multi = a.n_chunks > 1;
  const long long total_rows = (long long)a.B * a.W;
  const float* cb = a.codebooks + ((size_t)g * a.K + code0) * kD;
  const int W = a.W;
  const int sub_row = lane / kRowLanes, sub_col = (lane % kRowLanes) * 4;
  int n_stamp = 0;
  auto stamp = [&](int role) {       // role 0 = loader warp 0, 1 = MMA issuer, 2 = scan warp 0
    if (a.trace && blockIdx.x == 0 && lane == 0 && n_stamp < 32) a.trace[role * 32 + n_stamp++] = clock64();
  };
  if (warp == 0) stamp(0);

  if (warp == kLoadWarps) {                       // the MMA warp also owns TMEM
    pt::tmem_alloc(&s.tmem_base, 256);
    pt::tmem_relinquish();
  }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&s.a_full[i], kLoadWarps); pt::mbar_init(&s.a_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&s.t_full[i], 1); pt::mbar_init(&s.t_empty[i], kScanWarps); }
    pt::fence_mbar_init();
  }
  // ---- prologue: the raw chunk (n_codes x kD fp32 <= 64 KB) goes global -> smem with cp.async, all copies of all threads in
  // flight together (register-staged loads serialise under this kernel's register budget); the second query-operand
  // buffer is idle until the loaders' second tile, which starts only after the __syncthreads below
  float* craw = reinterpret_cast<float*>(&s.a_ops[1][0][0]);
  for (int i = tid; i < n_codes * (kD / 4); i += kThreads)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(pt::smem_u32(craw + (size_t)i * 4)), "l"(cb + (size_t)i * 4) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (warp == 0) stamp(0);                                   // raw codebook chunk in shared memory
  if (warp < 8) {
    constexpr int kPerWarp = kK / 8 / kRowsPerInstr;
    float wmax = 0.f;
#pragma unroll 4
    for (int i = 0; i < kPerWarp; ++i) {
      const int j = (warp + i * 8) * kRowsPerInstr + sub_row;
      const float4 cv = (j < n_codes) ? *reinterpret_cast<const float4*>(craw + (size_t)j * kD + sub_col) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float sq = row_sum<kRowLanes>(fmaf(cv.w, cv.w, fmaf(cv.z, cv.z, fmaf(cv.y, cv.y, cv.x * cv.x))));
      wmax = fmaxf(wmax, sq);
      if (sub_col == 0) s.sc2[j] = (j < n_codes) ? sq : INFINITY;         // padding codes can never win
      split_store4(s.b_ops[0], s.b_ops[1], j, sub_col, cv);
    }
    wmax = fmaxf(wmax, __shfl_xor_sync(0xffffffffu, wmax, 16));           // kRowLanes = 16: two rows per instruction
    if (lane == 0) s.cmax_w[warp] = sqrtf(wmax);
  }
  pt::fence_async_smem();
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  if (warp == 0) stamp(0);                                   // codebook staged, set-up done
  if (warp == kLoadWarps) stamp(1);
  if (warp == kLoadWarps + 1) stamp(2);

  if (warp < kLoadWarps) {
    // ------------------------------------------------------------ loaders
    // Rows go global -> registers -> (hi, lo) bf16 -> K-major SWIZZLE_128B operand tile, no staging copy: with two operand
    // buffers the loaders run a whole tile ahead of the tensor core, and a warp keeps eight 512-byte row loads in flight.
    int it = 0;
    for (int t = part; t < a.n_tiles; t += a.ctas_per_group, ++it) {
      const int ab = it & 1;
      const uint32_t aph = (uint32_t)(it >> 1) & 1u;
      const long long row0 = (long long)t * a.rows_per_tile;
      const int nrows = (int)min((long long)a.rows_per_tile, total_rows - row0);
      const int b0 = (int)(row0 / W), w0 = (int)(row0 % W);
      pt::mbar_wait(&s.a_empty[ab], aph ^ 1);               // the MMAs that last read this operand buffer are done
      if (warp == 0) stamp(0);                               // operand buffer free
      uint8_t* hi = s.a_ops[ab][0];
      uint8_t* lo = s.a_ops[ab][1];
      float* sxp = &s.sx2p[it & 3][0][0];   // 4-deep: a slot is rewritten only after the scan of that tile has retired
      if (a.rows_fast) {
        constexpr int kBatch = 8;                            // row loads in flight per lane
        constexpr int kRowStep = kLoadWarps * kRowsPerInstr;
        // (warp-uniform trip count: with two rows per instruction the lanes of a warp hold different rows, and the
        // shuffles below need the whole warp)
        for (int rb0 = warp * kRowsPerInstr; rb0 < nrows; rb0 += kBatch * kRowStep) {
          const int rb = rb0 + sub_row;
          float4 x[kBatch];
#pragma unroll
          for (int i = 0; i < kBatch; ++i) {
            const int r = rb + i * kRowStep;
            x[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) {
              const int q = w0 + r, b = b0 + q / W, w = q % W;
              x[i] = __ldg(reinterpret_cast<const float4*>(a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD + sub_col));
            }
          }
#pragma unroll
          for (int i = 0; i < kBatch; ++i) {
            const int r = rb + i * kRowStep;
            float sq = fmaf(x[i].w, x[i].w, fmaf(x[i].z, x[i].z, fmaf(x[i].y, x[i].y, x[i].x * x[i].x)));
            sq += __shfl_xor_sync(0xffffffffu, sq, 1);
            sq += __shfl_xor_sync(0xffffffffu, sq, 2);
            if (r < nrows) {                                 // (rows beyond the tile keep stale operands: never read back)
              if ((lane & 3) == 0) sxp[((lane % kRowLanes) >> 2) * kTileRows + r] = sq;
              split_store4(hi, lo, r, sub_col, x[i]);
            }
          }
        }
      } else {
        // channel-first z: lanes = consecutive rows (w contiguous in memory); the two halves of the loader threads split k
        const int r = tid & (kTileRows - 1), kh = tid >> 7;
        if (r < nrows) {
          const int q = w0 + r, b = b0 + q / W, w = q % W;
          const float* prow = a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD * a.zs_c;
          float sq = 0.f;
#pragma unroll 2
          for (int c8 = kh * (kD / 16); c8 < (kh + 1) * (kD / 16); ++c8) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(prow + (size_t)(c8 * 8 + i) * a.zs_c);
#pragma unroll
            for (int i = 0; i < 8; ++i) sq = fmaf(v[i], v[i], sq);
            split_store(hi, lo, (c8 >> 3) * kAtomBytes + pt::sw128_offset(r, c8 & 7), v);
          }
          sxp[kh * kTileRows + r] = sq;
        }
      }
      pt::fence_async_smem();                                // generic-proxy smem writes -> tensor-core proxy
      __syncwarp();
      if (lane == 0) pt::mbar_arrive(&s.a_full[ab]);
      if (warp == 0) stamp(0);                               // converted
    }
  } else if (warp == kLoadWarps) {
    // ------------------------------------------------------------ MMA issuer: D = Ahi*Bhi^T + Ahi*Blo^T + Alo*Bhi^T
    if (lane == 0) {
      const uint32_t idesc = pt::umma_idesc_bf16(128, 128);
      const uint32_t b_hi = pt::smem_u32(s.b_ops[0]), b_lo = pt::smem_u32(s.b_ops[1]);
      int it = 0;
      for (int t = part; t < a.n_tiles; t += a.ctas_per_group, ++it) {
        const int buf = it & 1;                              // TMEM accumulator stage
        const uint32_t ph = (uint32_t)(it >> 1) & 1u;
        pt::mbar_wait(&s.a_full[buf], ph);                   // operand buffer and accumulator stage share the parity
        pt::mbar_wait(&s.t_empty[buf], ph ^ 1);
        pt::tc_fence_after();
        stamp(1);                                            // operands + accumulator ready
        const uint32_t a_hi = pt::smem_u32(s.a_ops[buf][0]), a_lo = pt::smem_u32(s.a_ops[buf][1]);
        const uint32_t pa[3] = {a_hi, a_hi, a_lo}, pb[3] = {b_hi, b_lo, b_hi};
        uint32_t acc = 0;
#pragma unroll
        for (int p = 0; p < 3; ++p) {
#pragma unroll
          for (int k = 0; k < kD / 16; ++k) {
            const uint32_t koff = (k >> 2) * kAtomBytes + (k & 3) * 32;
            pt::umma_bf16(tmem + buf * 128, pt::umma_desc_sw128(pa[p] + koff), pt::umma_desc_sw128(pb[p] + koff), idesc, acc);
            acc = 1;
          }
        }
        pt::umma_commit(&s.t_full[buf]);
        pt::umma_commit(&s.a_empty[buf]);
        stamp(1);                                            // issued
      }
    }
  } else {
    // ------------------------------------------------------------ scanners: thread = row, all 128 scores
    const int qd = warp & 3;                                 // TMEM lane quadrant (== warp % 4)
    const int r = qd * 32 + lane;
    float cmax = s.cmax_w[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) cmax = fmaxf(cmax, s.cmax_w[i]);
    int it = 0;
    for (int t = part; t < a.n_tiles; t += a.ctas_per_group, ++it) {
      const int buf = it & 1;
      const uint32_t ph = (uint32_t)(it >> 1) & 1u;
      const long long row0 = (long long)t * a.rows_per_tile;
      const int nrows = (int)min((long long)a.rows_per_tile, total_rows - row0);
      const int b0 = (int)(row0 / W), w0 = (int)(row0 % W);
      pt::mbar_wait(&s.t_full[buf], ph);
      pt::tc_fence_after();
      if (warp == kLoadWarps + 1) stamp(2);                  // scores ready
      const uint32_t t_row = tmem + buf * 128 + (static_cast<uint32_t>(qd * 32) << 16);
      // two independent (min, argmin, second-min) trackers (even / odd columns) for instruction-level parallelism
      float ma1 = INFINITY, ma2 = INFINITY, mb1 = INFINITY, mb2 = INFINITY;
      int ia = 0, ib = 1;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        pt::tmem_ld32(t_row + c * 32, v);
        pt::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const int j = c * 32 + i;
          const float sa = fmaf(-2.f, __uint_as_float(v[i]), s.sc2[j]);
          const float sb = fmaf(-2.f, __uint_as_float(v[i + 1]), s.sc2[j + 1]);
          if constexpr (kDump) {
            if (r < nrows) {
              float* dump = a.score_dump + ((row0 + r) * a.G + g) * a.K + code0;
              if (j < n_codes) dump[j] = sa;
              if (j + 1 < n_codes) dump[j + 1] = sb;
            }
          }
          ma2 = fminf(ma2, fmaxf(sa, ma1));          // second-min update, then min (strict < keeps the first index)
          if (sa < ma1) { ma1 = sa; ia = j; }
          mb2 = fminf(mb2, fmaxf(sb, mb1));
          if (sb < mb1) { mb1 = sb; ib = j + 1; }
        }
      }
      // merge the trackers; on equal scores the smaller index wins (first minimum, like torch.argmin)
      float best, second;
      int bi;
      if (mb1 < ma1 || (mb1 == ma1 && ib < ia)) { best = mb1; bi = ib; second = fminf(ma1, mb2); }
      else { best = ma1; bi = ia; second = fminf(mb1, ma2); }
      float x2 = 0.f;
      {
        const int nparts = a.rows_fast ? kRowLanes / 4 : 2;
        for (int pp = 0; pp < nparts; ++pp) x2 += s.sx2p[it & 3][pp][r];
      }
      const float band = 2.f * kBand * sqrtf(x2) * cmax + 4e-7f * fabsf(best);
      // a chunk of a larger codebook re-scores EVERY row: the chunk winners are merged by their fp32 distances
      const bool amb = (r < nrows) && (multi || second - best <= band);
      const unsigned amb_mask = __ballot_sync(0xffffffffu, amb);
      unsigned cand[4] = {0u, 0u, 0u, 0u};
      if (amb_mask) {                              // warp-uniform: second pass collects every code inside the band
        const float lim = best + band;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          pt::tmem_ld32(t_row + c * 32, v);
          pt::tmem_ld_wait();
          unsigned m = 0u;
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (fmaf(-2.f, __uint_as_float(v[i]), s.sc2[c * 32 + i]) <= lim) m |= 1u << i;
          cand[c] = m;
        }
      }
      pt::tc_fence_before();
      __syncwarp();
      if (lane == 0) pt::mbar_arrive(&s.t_empty[buf]);       // the accumulator may be overwritten from here on
      if (warp == kLoadWarps + 1) stamp(2);                  // scanned
      if (!amb && r < nrows) {
        a.idx_out[(row0 + r) * a.G + g] = bi;
        if (a.gap_out) a.gap_out[(row0 + r) * a.G + g] = second - best;
      }
      // fp32 re-score of the ambiguous rows of this warp, one row at a time, lanes split k;
      // reference formula (vqvae_modules.py:549-553)
      unsigned todo = amb_mask;
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int rr = qd * 32 + src;
        const long long row = row0 + rr;
        const int q = w0 + rr, b = b0 + q / W, w = q % W;
        const float* px = a.z + (long long)b * a.zs_b + (long long)w * a.zs_w + (size_t)g * kD * a.zs_c;
        float x[kD / 32];
#pragma unroll
        for (int i = 0; i < kD / 32; ++i) x[i] = __ldg(px + (size_t)(lane + 32 * i) * a.zs_c);
        float xs = 0.f;
#pragma unroll
        for (int i = 0; i < kD / 32; ++i) xs = fmaf(x[i], x[i], xs);
        const float sx2 = warp_sum(xs);
        float rbest = INFINITY, rsecond = INFINITY;
        int rbi = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          unsigned m = __shfl_sync(0xffffffffu, cand[c], src);
          while (m) {
            const int j = c * 32 + __ffs(m) - 1;
            m &= m - 1;
            const float* pc = cb + (size_t)j * kD;
            float dot = 0.f;
#pragma unroll
            for (int i = 0; i < kD / 32; ++i) dot = fmaf(x[i], __ldg(pc + lane + 32 * i), dot);
            dot = warp_sum(dot);
            const float d = (sx2 + s.sc2[j]) - 2.f * dot;
            if (d < rbest) { rsecond = rbest; rbest = d; rbi = j; }      // ascending j: first minimum wins ties
            else if (d < rsecond) { rsecond = d; }
          }
        }
        if (lane == 0) {
          if (multi) {
            const unsigned long long word = (static_cast<unsigned long long>(ordered_bits(rbest)) << 32) |
                                            static_cast<unsigned>(code0 + rbi);
            atomicMin(reinterpret_cast<unsigned long long*>(a.idx_out + row * a.G + g), word);
          } else {
            a.idx_out[row * a.G + g] = rbi;
            if (a.gap_out) a.gap_out[row * a.G + g] = rsecond - rbest;
          }
          if (a.n_rescored && !multi) atomicAdd(a.n_rescored, 1);
        }
      }
      if (warp == kLoadWarps + 1) stamp(2);                  // re-scored, written
    }
  }
  pt::tc_fence_before();
  __syncthreads();
  if (warp == kLoadWarps) pt::tmem_dealloc(tmem, 256);
}

// ---------------------------------------------------------------------------------------------
// gather: indices -> quantized vectors (ProductQuantization.quantize_all, vqvae_modules.py:557-564)
// out element (b, w, g*128 + d) at out + b*os_b + w*os_w + (g*128+d)*os_c ; fp32 or bf16
template <typename OutT, typename IdxT>
__global__ void pq_gather_kernel(const IdxT* __restrict__ idx, const float* __restrict__ cb, OutT* __restrict__ out,
                                 long long os_b, long long os_w, long long os_c, int B, int W, int G, int K, int D, int* bad) {
  const int d4 = D / 4;
  const long long n = (long long)B * W * G * d4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % d4);
    const long long e = i / d4;                // (row, g)
    const int g = (int)(e % G);
    const long long row = e / G;
    const long long b = row / W, w = row % W;
    long long j = (long long)idx[e];
    if (j < 0 || j >= K) { if (bad && q == 0) atomicAdd(bad, 1); j = 0; }
    const float4 v = __ldg(reinterpret_cast<const float4*>(cb + ((size_t)g * K + j) * D) + q);
    OutT* p = out + b * os_b + w * os_w + (long long)(g * D + q * 4) * os_c;
    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if constexpr (sizeof(OutT) == 4) p[d * os_c] = vv[d];
      else p[d * os_c] = __float2bfloat16_rn(vv[d]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// gather fused into the decoder's first (1x1) convolution.  aft_vq_conv is linear and its input is a concatenation of
// G codebook rows, so
//     out[r, n] = b[n] + sum_c W[n, c] e[r, c] = b[n] + sum_g T_g[idx[r, g]][n],   T_g = C_g . W[:, gD:(g+1)D]^T
// with the G tables (G x K x N fp32, 2 MB for vq.json: L2 resident) computed once per checkpoint.  No quantized
// tensor, no GEMM: G table rows are read per output row and one bf16 row (or a split-bf16 pair) is written.
// One thread = 8 consecutive output columns of one row (a 16-byte store); N / 8 threads cover a row.
template <typename IdxT>
__global__ void __launch_bounds__(256) pq_gather_conv_kernel(const IdxT* __restrict__ idx, const float* __restrict__ table,
                                                              const float* __restrict__ bias,
                                                              __nv_bfloat16* __restrict__ out_hi,
                                                              __nv_bfloat16* __restrict__ out_lo, long long oss,
                                                              long long ors, long long rows, int W, int G, int K, int N,
                                                              int* bad) {
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  const int per_row = N >> 3;
  const long long total = rows * per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / per_row;
    const int n = (int)(i - row * per_row) << 3;
    float acc[8];
    {
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + n)), b1 = __ldg(reinterpret_cast<const float4*>(bias + n) + 1);
      acc[0] = b0.x; acc[1] = b0.y; acc[2] = b0.z; acc[3] = b0.w; acc[4] = b1.x; acc[5] = b1.y; acc[6] = b1.z; acc[7] = b1.w;
    }
    for (int g = 0; g < G; ++g) {
      long long j = (long long)idx[row * G + g];
      if (j < 0 || j >= K) { if (bad && n == 0) atomicAdd(bad, 1); j = 0; }
      const float4* t = reinterpret_cast<const float4*>(table + ((size_t)g * K + j) * N + n);
      const float4 t0 = __ldg(t), t1 = __ldg(t + 1);
      acc[0] += t0.x; acc[1] += t0.y; acc[2] += t0.z; acc[3] += t0.w; acc[4] += t1.x; acc[5] += t1.y; acc[6] += t1.z; acc[7] += t1.w;
    }
    uint4 hi, lo;
    pt::split_bf16x2(acc[0], acc[1], hi.x, lo.x); pt::split_bf16x2(acc[2], acc[3], hi.y, lo.y);
    pt::split_bf16x2(acc[4], acc[5], hi.z, lo.z); pt::split_bf16x2(acc[6], acc[7], hi.w, lo.w);
    const long long b = row / W, w = row - b * W;
    const long long off = b * oss + w * ors + n;
    *reinterpret_cast<uint4*>(out_hi + off) = hi;
    if (out_lo) *reinterpret_cast<uint4*>(out_lo + off) = lo;
  }
}

}  // namespace

extern "C" int pt_pq_gather_conv1x1(const void* idx, int idx_is_i64, const float* table, const float* bias, void* out_bf16,
                                    void* out_lo, long long oss, long long ors, int B, int W, int G, int K, int N,
                                    int* bad_count, void* stream) {
  const long long rows = (long long)B * W;
  if (rows <= 0) return 0;
  if (!idx || !table || !bias || !out_bf16) return pt::fail("pt_pq_gather_conv1x1: null pointer");
  if (G < 1 || K < 1 || N < 8 || (N & 7)) return pt::fail("pt_pq_gather_conv1x1: need G, K >= 1 and N a multiple of 8 (got %d, %d, %d)", G, K, N);
  if (((reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(bias) | reinterpret_cast<uintptr_t>(out_bf16) |
        reinterpret_cast<uintptr_t>(out_lo)) & 15) || (oss & 7) || (ors & 7))
    return pt::fail("pt_pq_gather_conv1x1: table, bias, outputs and output strides must be 16-byte aligned");
  const long long total = rows * (N >> 3);
  const int blocks = (int)min((long long)pt::sm_count() * 8, (total + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (idx_is_i64)
    PT_CUDA(pt::launch_pdl(pq_gather_conv_kernel<long long>, dim3(blocks), dim3(256), 0, st, 1, (const long long*)idx, table, bias,
                           (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, oss, ors, rows, W, G, K, N, bad_count));
  else
    PT_CUDA(pt::launch_pdl(pq_gather_conv_kernel<int>, dim3(blocks), dim3(256), 0, st, 1, (const int*)idx, table, bias,
                           (__nv_bfloat16*)out_bf16, (__nv_bfloat16*)out_lo, oss, ors, rows, W, G, K, N, bad_count));
  PT_CUDA(cudaGetLastError());
  return 0;
}

namespace {
template <int kD>
int launch_search(const PqArgs& a, unsigned grid, bool dump, cudaStream_t st) {
  const size_t smem = sizeof(PqSmem<kD>) + 1024;
  if (dump) {
    PT_SMEM_OPT_IN((pq_search_kernel<kD, true>), smem);
    pq_search_kernel<kD, true><<<grid, kThreads, smem, st>>>(a);
  } else {
    PT_SMEM_OPT_IN((pq_search_kernel<kD, false>), smem);
    pq_search_kernel<kD, false><<<grid, kThreads, smem, st>>>(a);
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}
}  // namespace

extern "C" int pt_pq_search(const float* z, long long zs_b, long long zs_c, long long zs_w, int B, int W, int G,
                            int K, int D, const float* codebooks, long long* idx_out, float* gap_out,
                            float* score_dump, int* n_rescored, void* stream) {
  if (D != 64 && D != 128) return pt::fail("pt_pq_search: sub-vector width D must be 64 or 128 (got %d)", D);
  if (K < 16 || (K & 15) || K > 65536) return pt::fail("pt_pq_search: K must be a multiple of 16 in [16, 65536] (got %d)", K);
  if (G < 1 || G > 65535) return pt::fail("pt_pq_search: bad group count %d", G);
  if (B < 0 || W < 0) return pt::fail("pt_pq_search: negative shape");
  if ((long long)B * W == 0) return 0;
  if (!z || !codebooks || !idx_out) return pt::fail("pt_pq_search: null pointer");
  if (reinterpret_cast<uintptr_t>(codebooks) & 15) return pt::fail("pt_pq_search: codebooks must be 16-byte aligned");
  const int n_chunks = (K + kK - 1) / kK;
  if (n_chunks > 1 && gap_out)
    return pt::fail("pt_pq_search: gap_out is only available for K <= %d (got K=%d)", kK, K);
  const int sms = pt::sm_count();
  cudaStream_t st = (cudaStream_t)stream;
  PqArgs a;
  a.z = z; a.zs_b = zs_b; a.zs_w = zs_w; a.zs_c = zs_c; a.B = B; a.W = W; a.G = G; a.K = K;
  a.codebooks = codebooks; a.idx_out = idx_out; a.gap_out = gap_out; a.score_dump = score_dump;
  a.n_rescored = n_rescored; a.n_chunks = n_chunks;
  {
    const char* e = getenv("PT_PQ_TRACE");                 // developer hook: device address of 96 int64 (hex or decimal)
    a.trace = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
  }
  a.rows_fast = (zs_c == 1 && !(reinterpret_cast<uintptr_t>(z) & 15) && !(zs_b & 3) && !(zs_w & 3)) ? 1 : 0;
  const long long rows = (long long)B * W;
  const int vgroups = G * n_chunks;                  // (sub-codebook, chunk of 128 codes) pairs; a CTA serves one
  // persistent CTAs: `per_group` CTAs per (sub-codebook, chunk), each walking k tiles; tile height chosen so that the
  // tiles divide evenly (a tile costs ~ its rows, so 2 x 77 rows beats 128 + 25)
  int per_group = sms / vgroups;
  if (per_group < 1) per_group = 1;
  const long long full_tiles = (rows + kTileRows - 1) / kTileRows;
  if (full_tiles < per_group) per_group = (int)full_tiles;
  const long long k = (full_tiles + per_group - 1) / per_group;            // tiles per CTA
  long long R = (rows + k * per_group - 1) / (k * per_group);
  R = (R + 7) / 8 * 8;
  if (R > kTileRows) R = kTileRows;
  a.rows_per_tile = (int)R;
  a.n_tiles = (int)((rows + R - 1) / R);
  a.ctas_per_group = per_group;
  const unsigned grid = (unsigned)(per_group * vgroups);
  if (n_chunks > 1)      // packed (distance, index) words are folded into idx_out with atomicMin: start from all ones
    PT_CUDA(cudaMemsetAsync(idx_out, 0xff, sizeof(long long) * (size_t)rows * G, st));
  const int rc = (D == 128) ? launch_search<128>(a, grid, score_dump != nullptr, st)
                            : launch_search<64>(a, grid, score_dump != nullptr, st);
  if (rc) return rc;
  if (n_chunks > 1) {
    const long long n = rows * G;
    pq_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(idx_out, n);
    PT_CUDA(cudaGetLastError());
  }
  return 0;
}

extern "C" int pt_pq_gather(const void* idx, int idx_is_i64, const float* codebooks, void* out, int out_is_bf16,
                            long long os_b, long long os_c, long long os_w, int B, int W, int G, int K, int D,
                            int* bad_count, void* stream) {
  if (K < 1 || D < 4 || (D & 3)) return pt::fail("pt_pq_gather: need K >= 1 and D a positive multiple of 4 (got K=%d D=%d)", K, D);
  const long long n = (long long)B * W * G * (D / 4);
  if (n == 0) return 0;
  if (!idx || !codebooks || !out) return pt::fail("pt_pq_gather: null pointer");
  if (reinterpret_cast<uintptr_t>(codebooks) & 15) return pt::fail("pt_pq_gather: codebooks must be 16-byte aligned");
  const int threads = 256;
  const int blocks = (int)min((long long)148 * 8, (n + threads - 1) / threads);
  cudaStream_t st = (cudaStream_t)stream;
  if (idx_is_i64) {
    if (out_is_bf16) pq_gather_kernel<__nv_bfloat16, long long><<<blocks, threads, 0, st>>>((const long long*)idx, codebooks, (__nv_bfloat16*)out, os_b, os_w, os_c, B, W, G, K, D, bad_count);
    else pq_gather_kernel<float, long long><<<blocks, threads, 0, st>>>((const long long*)idx, codebooks, (float*)out, os_b, os_w, os_c, B, W, G, K, D, bad_count);
  } else {
    if (out_is_bf16) pq_gather_kernel<__nv_bfloat16, int><<<blocks, threads, 0, st>>>((const int*)idx, codebooks, (__nv_bfloat16*)out, os_b, os_w, os_c, B, W, G, K, D, bad_count);
    else pq_gather_kernel<float, int><<<blocks, threads, 0, st>>>((const int*)idx, codebooks, (float*)out, os_b, os_w, os_c, B, W, G, K, D, bad_count);
  }
  PT_CUDA(cudaGetLastError());
  return 0;
}
