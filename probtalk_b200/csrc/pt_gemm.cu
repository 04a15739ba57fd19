// probtalk_b200 -- the tensor-core contraction engine (sm_100a: TMA + tcgen05 + TMEM).
//
// One persistent, warp-specialised kernel computes
//     out[row(m), n] = epilogue( sum_{tap, c} A[m*1 + base_row + tap/s][(tap%s)*C + c] * W[n][tap*C + c] )
// which covers every dense op on ProbTalk's inference path with ONE code path:
//   * nn.Linear / 1x1 Conv1d               taps = 1
//   * Conv1d k3 s1 p1, k7 s1 p3            taps = k,  s = 1   (activations are channels-last with zero halo rows)
//   * Conv1d k4 s2 p1                      taps = 4,  s = 2   (A viewed as [rows/2, 2C]: tap j -> row j/2, column half j%2)
//   * ConvTranspose1d k4 s2 p1             two launches (even / odd output phase), taps = 2, output row stride 2
// (reference layers: nets/spg/vqvae_modules.py:107-245, nets/inpainting/gated_pixelcnn_1d.py:39-83,
//  nets/inpainting/PTransformer.py:284-372).  A is bf16 [rows, C] row-major (K-major), W is bf16 [N, taps*C]
// K-major; accumulation is fp32 in TMEM.
//
// Warp roles (384 threads, 1 CTA / SM, persistent over 128 x BN output tiles):
//   warp 0  TMA producer: per K-step one 128x64 A box + one BNx64 W box (SWIZZLE_128B) into a smem ring
//   warp 1  MMA issuer  : one thread, 4 x tcgen05.mma (K=16 each) per K-step, tcgen05.commit frees the slot
//   warp 2  TMEM allocator (2 accumulator stages x BN columns)
//   warps 4-11 epilogue : tcgen05.ld (lane = output row), bias / residual / activation, smem-staged 64-byte row stores
// so the epilogue of tile i overlaps the main loop of tile i+1.
#include <cuda.h>
#include <stdlib.h>
#include "pt_common.cuh"
#include "pt_host.h"
#include "pt_gemm.h"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle row
constexpr int kEpiWarp0 = 4;
constexpr int kRnMaxCls = 4;                              // row-norm classes whose tables are staged in shared memory
constexpr int kEpiWarps = 8;
constexpr int kThreads = (kEpiWarp0 + kEpiWarps) * 32;   // 384
// epilogue staging tile: 32 rows x 64 B, 16-byte pieces XOR-swizzled by the row pair so that both access patterns
// (thread = row writes its 4 pieces; 4 lanes x 8 rows read / write global 64-byte segments) are bank-conflict free
#define PT_TILE_OFF(r, piece) ((r) * 64 + ((((piece) ^ ((r) >> 1)) & 3) << 4))

template <int BN, bool k2 = false>
struct Cfg {
  // k2 (cta_group::2): each CTA of a pair stages its 128 rows of A and HALF of the BN weight rows
  static constexpr uint32_t kABytes = BM * BK * 2;          // 16 KB
  static constexpr uint32_t kBBytes = (k2 ? BN / 2 : BN) * BK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (kStageBytes == 48 * 1024) ? 4 : 6;
  static constexpr uint32_t kTmemCols = 2 * BN;             // double-buffered accumulator
  static constexpr size_t kSmem = 1024 + (size_t)kStages * kStageBytes + 256;   // + 24 KB static (epilogue staging)
};

#define GTRACE(role, tl, ev)                                                                                 \
  do {                                                                                                       \
    if (p.trace && blockIdx.x == 0 && (tl) < 16) p.trace[((role) * 16 + (tl)) * 8 + (ev)] = clock64();         \
  } while (0)

struct GemmKernelArgs {
  int M, N_tiles, K_iters;        // K_iters = taps * (C / 64)
  int c_iters;                    // C / 64
  int C;                          // channels per tap
  int s;                          // input stride (1 or 2)
  int base_row;
  int rows_per_sample, t_valid;
  int n_valid;
  long long oss, ors, ocs;        // output strides (elements): sample, row-in-sample, column
  long long rss, rrs;             // residual strides (elements): sample, row-in-sample (column stride 1)
  void* out;
  const float* bias;
  const void* residual;
  int out_f32, res_f32;
  int act;                        // pt_gemm.h: PT_ACT_*
  float slope;
  int m_tiles;
  // row-norm epilogue (cluster of 2 CTAs covers the 512-wide row): out = f( (v - mean) * rstd * A[c] + B[c] )
  int row_norm;                   // 0 none, 1 SiLU, 2 identity
  const float* norm_a;            // [n_classes, 512]
  const float* norm_b;            // [n_classes, 512]
  const int* labels;              // class per sample (row / label_rows) or null (class 0)
  int label_rows;
  int norm_classes;
  int a_col_tab[8];               // extra A column offset per 128-column output tile (grouped convolutions), zeros otherwise
  long long* trace;               // developer hook (PT_GEMM_TRACE): clock64 stamps of CTA 0, [3 roles][16 tiles][8]
  int res_mma;                    // row-norm: the bf16 residual tile is added by the tensor core (identity x residual), tmap_r
  int tma_store;                  // tmap_o describes the output as a dense bf16 [M, n_valid] matrix: 32 x 32 boxes are stored by TMA
  // split-bf16 ("bf16x3") arithmetic: the K loop runs three passes  A_hi W_hi + A_hi W_lo + A_lo W_hi  (tmap_r = the lo
  // plane of A, W rows are [hi | lo]); bf16 outputs / residuals are hi + lo plane pairs
  int split;
  int k_base;                     // K blocks of ONE pass = taps * (C / 64)
  void* out_lo;
  const void* residual_lo;
};

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  if (act == PT_ACT_RELU) return fmaxf(v, 0.f);
  if (act == PT_ACT_LEAKY) return v > 0.f ? v : v * slope;
  if (act == PT_ACT_GELU) {
    // erf GELU (nn.GELU()), erf from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7): two MUFU ops and ten FMA-pipe
    // instructions instead of erff's ~25; 0.5 v (1 + erf(v / sqrt 2)) = 0.5 v + 0.5 |v| - 0.5 |v| poly(t) exp(-z^2)
    const float z = fabsf(v) * 0.70710678118654752f;
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -1.4426950408889634f));
    float p = fmaf(t, 1.061405429f, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    p = p * t * e;
    const float h = 0.5f * fabsf(v);
    return fmaf(-h, p, fmaf(0.5f, v, h));
  }
  return v;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_peer(const void* local_smem, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(pt::smem_u32(local_smem)), "r"(rank));
  return r;
}
// packed fp32x2 add (sm_100): two bias adds per instruction
__device__ __forceinline__ void add2(float& a0, float& a1, float b0, float b1) {
  unsigned long long a, b, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(d));
}
// packed fp32x2 fused multiply-add: (a0, a1) = (x0, x1) * (y0, y1) + (z0, z1)
__device__ __forceinline__ void fma2(float& a0, float& a1, float x0, float x1, float y0, float y1, float z0, float z1) {
  unsigned long long x, y, z, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(x0), "f"(x1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(y) : "f"(y0), "f"(y1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(z0), "f"(z1));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(x), "l"(y), "l"(z));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(d));
}
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// erf GELU of two values in place (Abramowitz & Stegun 7.1.26, see apply_act): the polynomial runs as packed f32x2 FMAs
__device__ __forceinline__ void gelu2(float& v0, float& v1) {
  const float z0 = fabsf(v0) * 0.70710678118654752f, z1 = fabsf(v1) * 0.70710678118654752f;
  float t0, t1, e0, e1;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(fmaf(0.3275911f, z0, 1.f)));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(fmaf(0.3275911f, z1, 1.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(z0 * z0 * -1.4426950408889634f));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(z1 * z1 * -1.4426950408889634f));
  float p0, p1;
  fma2(p0, p1, t0, t1, 1.061405429f, 1.061405429f, -1.453152027f, -1.453152027f);
  fma2(p0, p1, p0, p1, t0, t1, 1.421413741f, 1.421413741f);
  fma2(p0, p1, p0, p1, t0, t1, -0.284496736f, -0.284496736f);
  fma2(p0, p1, p0, p1, t0, t1, 0.254829592f, 0.254829592f);
  p0 *= t0 * e0; p1 *= t1 * e1;
  const float h0 = 0.5f * fabsf(v0), h1 = 0.5f * fabsf(v1);
  float l0, l1;
  fma2(l0, l1, 0.5f, 0.5f, v0, v1, h0, h1);
  fma2(v0, v1, -h0, -h1, p0, p1, l0, l1);
}
// two fp32 -> packed bf16x2 (low half = first argument), optionally clamped at zero by the conversion itself
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi, bool relu) {
  uint32_t d;
  if (relu) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// Chan's parallel-variance merge of (n, mean, M2) partials
__device__ __forceinline__ void welford_merge(float& n, float& mean, float& m2, float nb, float mb, float m2b) {
  const float tot = n + nb, d = mb - mean;
  mean += d * (nb / tot);
  m2 += m2b + d * d * (n * nb / tot);
  n = tot;
}

// kRowNorm: launched as clusters of two CTAs; CTA `rank` owns columns [rank*256, rank*256+256) of the same 128
// rows, and the two exchange per-row (sum, sum of squares) partials through distributed shared memory so that the epilogue can
// apply a LayerNorm over the full 512-wide row (StylizationBlock_V3 / norm1-3 of the reference) in place.
// k2: pairs of CTAs (cta_group::2) compute a 256-row x BN tile together: the leader's MMA thread issues
// M = 256 instructions that read 128 A rows and BN/2 weight rows from EACH CTA's smem, so a CTA stages (and the tensor
// core reads) half the weight bytes per FLOP; stages shrink to 32 KB and the ring deepens from 4 to 6.
template <int BN, bool kRowNorm, int kCL, bool k2 = false>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
            const __grid_constant__ CUtensorMap tmap_o, const __grid_constant__ CUtensorMap tmap_r,
            const GemmKernelArgs p) {
  using C = Cfg<BN, k2>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)C::kStages * C::kStageBytes);
  uint64_t* full = bars;                         // [kStages]
  uint64_t* empty = bars + C::kStages;           // [kStages]
  uint64_t* tmem_full = bars + 2 * C::kStages;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  // epilogue staging lives in STATIC shared memory so the compiler emits LDS/STS (pointers carved from the
  // dynamic buffer lose their address space and become slow generic loads)
  __shared__ __align__(1024) uint8_t tile_s[kEpiWarps][32 * 64];   // PT_TILE_OFF == TMA SWIZZLE_64B, so a tile can be stored by TMA
  __shared__ __align__(16) float bias_s[kRowNorm ? 1 : kEpiWarps][BN / 2];   // this warp's 128 (or 64) bias columns of the current tile
  __shared__ long long rowoff_s[kRowNorm ? 1 : kEpiWarps][64];
  __shared__ __align__(16) float rn_bias_s[kRowNorm ? BN : 4];               // row-norm: the CTA's 256 bias columns
  __shared__ __align__(16) float rn_tab_s[2][kRowNorm ? kRnMaxCls : 1][kRowNorm ? BN : 4];   // [a|b][class][column]
  __shared__ float2 stats_s[kRowNorm ? 2 : 1][4][kRowNorm ? BM : 1];    // [tile parity][source warp-column][row] (mean, M2)
  __shared__ uint64_t stats_bar[2];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  static_assert(kCL == 1, "every CTA loads its own weight tiles (no clusters along M)");
  static_assert(!(k2 && (kRowNorm || kCL != 1)), "cta_group::2 is its own variant");
  constexpr int kClusterSize = (kRowNorm || k2) ? 2 : kCL;
  const uint32_t rank = (kClusterSize > 1) ? cluster_ctarank() : 0u;
  const int tile_first = (int)(blockIdx.x / kClusterSize);
  const int tile_stride = (int)(gridDim.x / kClusterSize);
  // kRowNorm: one "tile" = a 128-row block (the two CTAs split its 512 columns);
  // kCL > 1 : one "tile" = kCL consecutive 128-row blocks x one column tile (CTA `rank` owns block rank)
  constexpr int kRowBlocks = k2 ? 2 : kCL;             // 128-row blocks covered by one cluster step
  const int m_super = (p.m_tiles + kRowBlocks - 1) / kRowBlocks;
  const int n_tiles_total = kRowNorm ? p.m_tiles : m_super * p.N_tiles;
  auto tile_m0 = [&](int tile) { return kRowNorm ? tile * BM : ((tile / p.N_tiles) * kRowBlocks + (int)rank) * BM; };
  auto tile_n0 = [&](int tile) { return kRowNorm ? (int)rank * BN : (tile % p.N_tiles) * BN; };

  if (warp == 0 && lane == 0) {
    pt::tma_prefetch_desc(&tmap_a);
    pt::tma_prefetch_desc(&tmap_b);
    if (p.tma_store) pt::tma_prefetch_desc(&tmap_o);
    if (p.res_mma || p.split) pt::tma_prefetch_desc(&tmap_r);
  }
  if (warp == 1 && lane == 0) {
    const uint32_t full_count = 1u;
    for (int i = 0; i < C::kStages; ++i) { pt::mbar_init(&full[i], full_count); pt::mbar_init(&empty[i], kCL); }
    for (int i = 0; i < 2; ++i) { pt::mbar_init(&tmem_full[i], 1); pt::mbar_init(&tmem_empty[i], (k2 ? 2 : 1) * kEpiWarps); }
    if constexpr (kRowNorm) { pt::mbar_init(&stats_bar[0], kEpiWarps); pt::mbar_init(&stats_bar[1], kEpiWarps); }
    pt::fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (k2) { pt::tmem_alloc_2sm(tmem_slot, C::kTmemCols); pt::tmem_relinquish_2sm(); }
    else { pt::tmem_alloc(tmem_slot, C::kTmemCols); pt::tmem_relinquish(); }
  }
  pt::tc_fence_before();
  __syncthreads();
  pt::tc_fence_after();
  if constexpr (kClusterSize > 1) cluster_sync_all();  // the peers' barriers exist before anyone arrives on them
  const uint32_t tmem_base = *tmem_slot;
  // set-up done: the next kernel of the stream may start ITS set-up; everything below reads / writes dependent memory
  pt::pdl_launch_dependents();
  pt::pdl_wait();
  // Register re-balancing for the row-norm variant (setmaxnreg acts on a warpgroup of 4 warps): the producer / MMA /
  // idle warps keep 40 registers, the two epilogue warpgroups grow to 232 so that a thread can hold its row's 128 fp32
  // accumulators across the LayerNorm.  128 * 40 + 256 * 232 = 64512 = 384 * 168, what the kernel is launched with.
  if (warp < kEpiWarp0) {
   if constexpr (kRowNorm) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
   if (warp == 0 && kRowNorm && p.res_mma) {
    // ------------------------------------------------------------ TMA producer, residual-through-MMA flavour
    // After the K blocks of a tile come two more pipeline slots that make the tensor core add the bf16 residual:
    //   acc[m, n] += sum_k I[m, k] * R[k, n],  k = row of the 128-row tile (two blocks of 64).
    // The B slot gets the residual rows as four [64 rows x 64 columns] boxes, which is exactly the MN-major operand
    // layout (rows = K); the A slot gets the matching 128 x 64 block of the identity, written by THIS warp's 32 lanes before
    // the elected lane arms the barrier.  1.0 * bf16 accumulates exactly in fp32, and the epilogue no longer touches the
    // residual at all.  (A separate filler warp used to write the identity and kept in step with the ring by waiting on the
    // same `empty` barriers; a parity wait is only valid one phase ahead, and once the producer became fast a delayed
    // filler could miss a whole phase of a K-block slot and dead-lock the CTA -- seen once in ~10^5 launches.)
    // The WHOLE warp walks the loop and one elected lane issues (see the plain producer below).
    {
      const int sh = p.s - 1;                              // s is 1 or 2: tap % s = tap & sh, tap / s = tap >> sh
      int stage = 0; uint32_t phase = 0;
      for (int tile = tile_first; tile < n_tiles_total; tile += tile_stride) {
        const int m0 = tile_m0(tile), n0 = tile_n0(tile);
        int tap = 0, cc = 0;
        for (int kk = 0; kk < p.K_iters + 2; ++kk) {
          pt::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * C::kStageBytes;
          if (kk >= p.K_iters) {
            // identity block of this residual slot: row m = 64 kb + r holds its 1.0 at column r (SWIZZLE_128B, K-major)
            const int kb = kk - p.K_iters;
            for (int i = lane; i < (int)(C::kABytes / 16); i += 32) reinterpret_cast<uint4*>(sa)[i] = make_uint4(0, 0, 0, 0);
            __syncwarp();
            for (int r = lane; r < 64; r += 32) {
              const int m = 64 * kb + r;
              *reinterpret_cast<uint16_t*>(sa + m * 128 + ((((r >> 3) ^ (m & 7)) & 7) << 4) + (r & 7) * 2) = 0x3F80;
            }
            pt::fence_async_smem();                        // generic-proxy writes -> visible to the tensor core's reads
            __syncwarp();
          }
          if (pt::elect_one()) {
            if (kk < p.K_iters) {
              pt::mbar_expect_tx(&full[stage], C::kStageBytes);
              pt::tma_load_2d(sa, &tmap_a, &full[stage], cc * BK + (tap & sh) * p.C, m0 + p.base_row + (tap >> sh));
              pt::tma_load_2d(sa + C::kABytes, &tmap_b, &full[stage], kk * BK, n0);
            } else {
              const int kb = kk - p.K_iters;
              pt::mbar_expect_tx(&full[stage], C::kBBytes);
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                pt::tma_load_2d(sa + C::kABytes + j * 8192, &tmap_r, &full[stage], n0 + 64 * j, m0 + 64 * kb);
            }
          }
          __syncwarp();
          if (++cc == p.c_iters) { cc = 0; ++tap; }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
   } else if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    // The WHOLE warp walks the loop (every lane waits on the slot's barrier) and one ELECTED lane issues, with (tap, channel
    // block, pass) kept as counters.  The first version ran under `if (lane == 0)` and derived them by division: 156
    // dependent instructions per K step (two integer divisions, an election loop with six R2UR moves around every
    // UTMALDG), ~750 cycles -- longer than the 512 cycles the step's MMAs take.  PT_GEMM_TRACE showed the issuer waiting for
    // operands all tile long, with activations resident in L2 or not: the GEMM ran at the pace of this thread.
    {
      const int sh = p.s - 1;                              // s is 1 or 2: tap % s = tap & sh, tap / s = tap >> sh
      int stage = 0; uint32_t phase = 0;
      int tl = 0;
      for (int tile = tile_first; tile < n_tiles_total; tile += tile_stride, ++tl) {
        const int m0 = tile_m0(tile), n0 = tile_n0(tile);
        if (lane == 0) GTRACE(0, tl, 0);
        // split arithmetic: pass 0 = A_hi W_hi, pass 1 = A_hi W_lo (weight columns shifted by one K extent),
        // pass 2 = A_lo W_hi (A read through the lo plane's tensor map)
        int pass = 0, kq = 0, tap = 0, cc = 0;
        for (int kk = 0; kk < p.K_iters; ++kk) {
          if (lane == 0 && kk == p.K_iters - 1) GTRACE(0, tl, 1);
          pt::mbar_wait(&empty[stage], phase ^ 1);
          if (pt::elect_one()) {
            const int wcol = (pass == 1 ? kq + p.k_base : kq) * BK;
            const CUtensorMap* tma = (pass == 2) ? &tmap_r : &tmap_a;
            uint8_t* sa = smem + (size_t)stage * C::kStageBytes;
            const int acol = cc * BK + (tap & sh) * p.C + (BN == 128 ? p.a_col_tab[(n0 >> 7) & 7] : 0);
            const int arow = m0 + p.base_row + (tap >> sh);
            if constexpr (k2) {
              // both CTAs' boxes complete on the leader's barrier: the leader arms it for the pair's bytes
              if (rank == 0) pt::mbar_expect_tx(&full[stage], 2 * C::kStageBytes);
              pt::tma_load_2d_2sm(sa, tma, &full[stage], acol, arow);
              pt::tma_load_2d_2sm(sa + C::kABytes, &tmap_b, &full[stage], wcol, n0 + (int)rank * (BN / 2));
            } else {
              pt::mbar_expect_tx(&full[stage], C::kStageBytes);
              pt::tma_load_2d(sa, tma, &full[stage], acol, arow);
              pt::tma_load_2d(sa + C::kABytes, &tmap_b, &full[stage], wcol, n0);
            }
          }
          __syncwarp();
          ++kq;
          if (++cc == p.c_iters) { cc = 0; ++tap; }
          if (p.split && kq == p.k_base) { kq = 0; tap = 0; cc = 0; ++pass; }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if (lane == 0) GTRACE(0, tl, 2);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (k2: the leader CTA issues for the pair)
    // The WHOLE warp walks the loop (every lane waits on the same barriers) and one ELECTED lane issues: with
    // `if (lane == 0)` around the loop the compiler cannot keep descriptors in uniform registers and wraps every
    // tcgen05.mma in an election loop with five R2UR moves (~20 dependent instructions, ~70-100 cycles per MMA: as long
    // as a 128 x 128 x 16 MMA takes to execute).  Descriptors advance by adding to their low word.
    if (!k2 || rank == 0) {
      const uint32_t idesc = pt::umma_idesc_bf16(k2 ? 2 * BM : BM, BN);
      const uint64_t da0 = pt::umma_desc_sw128(pt::smem_u32(smem));
      const uint64_t db0 = pt::umma_desc_sw128(pt::smem_u32(smem) + C::kABytes);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      int tl = 0;
      for (int tile = tile_first; tile < n_tiles_total; tile += tile_stride, ++tl) {
        if (lane == 0) GTRACE(1, tl, 0);
        pt::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        if (lane == 0) GTRACE(1, tl, 1);
        pt::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kk = 0; kk < p.K_iters; ++kk) {
          pt::mbar_wait(&full[stage], phase);
          if (lane == 0 && kk == 0) GTRACE(1, tl, 2);
          if (lane == 0 && kk == p.K_iters - 1) GTRACE(1, tl, 3);
          pt::tc_fence_after();
          if (pt::elect_one()) {
            const uint64_t so = (uint64_t)((uint32_t)stage * (uint32_t)(C::kStageBytes >> 4));
            const uint64_t da = da0 + so, db = db0 + so;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              if constexpr (k2)
                pt::umma_bf16_2sm(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kk | k) ? 1u : 0u);
              else
                pt::umma_bf16(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kk | k) ? 1u : 0u);
            }
            if constexpr (k2) {
              pt::umma_commit_2sm(&empty[stage], 3);                  // both CTAs' slots are free
              if (kk == p.K_iters - 1) pt::umma_commit_2sm(&tmem_full[acc], 3);   // both CTAs' accumulators are ready
            } else {
              pt::umma_commit(&empty[stage]);                         // slot reusable once these MMAs have read it
              if (kk == p.K_iters - 1 && !(kRowNorm && p.res_mma)) pt::umma_commit(&tmem_full[acc]);
            }
          }
          __syncwarp();
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        if constexpr (kRowNorm) {
          if (p.res_mma) {
            // acc += I * R: A = identity block (K-major), B = residual rows (MN-major, one 64-column atom per MMA)
            constexpr uint32_t idesc_r = pt::umma_idesc_bf16(BM, BN) | (1u << 16);
            for (int kb = 0; kb < 2; ++kb) {
              pt::mbar_wait(&full[stage], phase);
              pt::tc_fence_after();
              if (pt::elect_one()) {
                const uint32_t sa = pt::smem_u32(smem + (size_t)stage * C::kStageBytes);
                const uint64_t da = pt::umma_desc_sw128(sa);
                // one N = 256 instruction per K step: the four 64-column atoms of B are 8 KB apart (LBO)
                const uint64_t db = pt::umma_desc_sw128_mn(sa + C::kABytes, 8192);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)
                  pt::umma_bf16(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * (2048 >> 4)), idesc_r, 1u);
                pt::umma_commit(&empty[stage]);
                if (kb == 1) pt::umma_commit(&tmem_full[acc]);
              }
              __syncwarp();
              if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            }
          }
        }
        if (lane == 0) GTRACE(1, tl, 4);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
   }
  } else {
    if constexpr (kRowNorm) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    // ------------------------------------------------------------ epilogue (8 warps: TMEM quadrant x column half)
    // TMEM hands every thread one output ROW.  Arithmetic (bias, residual, activation) is done in that layout;
    // global traffic goes through a per-warp padded smem tile (32 rows x 64 B) so that every LDG/STG of a warp
    // covers whole 64-byte row segments (4 lanes x 16 B per row) instead of 32 different cache lines.
    const int e = warp - kEpiWarp0;
    const int q = e & 3, half = e >> 2;                  // TMEM lane quadrant (== warp % 4) and column half
    uint8_t* tile = tile_s[e];                           // [32 rows][4 swizzled 16-byte pieces]
    long long* row_out = rowoff_s[kRowNorm ? 0 : e];                    // [32] output row offsets (or -1), [32] residual offsets
    long long* row_res = rowoff_s[kRowNorm ? 0 : e] + 32;
    int acc = 0; uint32_t acc_phase = 0;
    const bool row_major = (p.ocs == 1);
    const bool gate = (p.act == PT_ACT_GATE);
    // the common case (bf16 row-major output, no residual, identity / ReLU) gets a lean path: with K = 512 the
    // epilogue of a 128 x 256 tile has only ~4000 cycles before it starts to stall the tensor core
    const bool lean = row_major && !p.out_f32 && !p.residual && !kRowNorm && !p.split &&
                      (p.act == PT_ACT_NONE || p.act == PT_ACT_RELU || p.act == PT_ACT_GELU);
    const bool lean_relu = (p.act == PT_ACT_RELU), lean_gelu = (p.act == PT_ACT_GELU);
    const int pr = lane >> 2, pc = lane & 3;             // piece mapping: row pr + 8*j, 16-byte piece pc
    // release an accumulator stage to the MMA issuer (k2: the issuer lives in the leader CTA)
    auto release_acc = [&](int acc_stage) {
      if constexpr (k2) {
        if (rank == 0) pt::mbar_arrive(&tmem_empty[acc_stage]);
        // relaxed: the barrier only hands the accumulator (TMEM, ordered by tcgen05.wait::ld + fence::before_thread_sync)
        // back to the issuer.  A release at cluster scope would also wait for this thread's output stores to drain.
        else asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];"
                          ::"r"(map_to_peer(&tmem_empty[acc_stage], 0u)) : "memory");
      } else {
        pt::mbar_arrive(&tmem_empty[acc_stage]);
      }
    };
    // row-norm: this CTA's 256 columns never change, so its bias and per-class scale / shift rows are staged in
    // shared memory once (the host admits at most kRnMaxCls classes).  SiLU mode stores them pre-multiplied by 1/2.
    const bool silu = (p.row_norm == 1);
    uint8_t* tile_row[4];                                // this lane's own row of the staging tile, 4 x 16 bytes
    uint8_t* tile_in[4];                                 // the 16-byte pieces this lane moves to / from global memory
#pragma unroll
    for (int i = 0; i < 4; ++i) { tile_row[i] = tile + PT_TILE_OFF(lane, i); tile_in[i] = tile + PT_TILE_OFF(pr + 8 * i, pc); }
    if constexpr (kRowNorm) {
      const int col = threadIdx.x - kEpiWarp0 * 32, n0c = (int)rank * BN;      // 0..255
      const float sc = silu ? 0.5f : 1.f;
      rn_bias_s[col] = __ldg(p.bias + n0c + col);
      const int ncls = p.labels ? p.norm_classes : 1;
      for (int c = 0; c < ncls; ++c) {
        rn_tab_s[0][c][col] = sc * __ldg(p.norm_a + (size_t)c * 512 + n0c + col);
        rn_tab_s[1][c][col] = sc * __ldg(p.norm_b + (size_t)c * 512 + n0c + col);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
    }
    int it = 0;
    for (int tile_i = tile_first; tile_i < n_tiles_total; tile_i += tile_stride, ++it) {
      const int m0 = tile_m0(tile_i), n0 = tile_n0(tile_i);
      const int m = m0 + q * 32 + lane;
      const int b = m / p.rows_per_sample, t = m - b * p.rows_per_sample;
      const bool row_ok = (m < p.M) && (t < p.t_valid);
      const long long obase = (long long)b * p.oss + (long long)t * p.ors;
      if constexpr (kRowNorm) {
        // ================= fused row normalisation (LayerNorm / StylizationBlock) epilogue =================
        // The warp's 32 rows x 128 columns live in REGISTERS from the accumulator read to the store (setmaxnreg gives
        // the epilogue warps 232 registers): the accumulator is read from TMEM once (a second pass doubles the
        // tcgen05.ld / wait round trips of the chain) and goes back to the MMA issuer immediately.
        // Offsets are 32-bit element offsets (the host checks the ranges) to keep the integer work per tile small.
        constexpr int kChunks = BN / 2 / 32;
        constexpr unsigned kFull = 0xffffffffu;
        const int cb = half * (BN / 2);
        const int obase_m = row_ok ? (int)obase + n0 + cb : -1;
        const __nv_bfloat16* resp = p.res_mma ? nullptr : static_cast<const __nv_bfloat16*>(p.residual);
        // residual: all 16 loads of this lane (4 chunks x 4 row segments of 64 bytes) leave before the accumulator wait
        uint4 xr[kChunks][4];
        if (resp) {
          const int rbase = b * (int)p.rss + t * (int)p.rrs + n0 + cb;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int rs = __shfl_sync(kFull, rbase, pr + 8 * j);
            const bool ok = __shfl_sync(kFull, obase_m, pr + 8 * j) >= 0;
#pragma unroll
            for (int c = 0; c < kChunks; ++c)
              xr[c][j] = ok ? __ldg(reinterpret_cast<const uint4*>(resp + rs + 32 * c) + pc) : make_uint4(0, 0, 0, 0);
          }
          // and the residual rows of the NEXT row block (this warp's 32 rows x 256 bytes) are pulled into L2, so that
          // one tile later the loads above cost an L2 hit instead of a DRAM round trip
          const int mn = tile_m0(tile_i + tile_stride) + q * 32 + lane;
          if (mn < p.M) {
            const int bn_ = mn / p.rows_per_sample, tn_ = mn - bn_ * p.rows_per_sample;
            const char* rp = reinterpret_cast<const char*>(resp + bn_ * (int)p.rss + tn_ * (int)p.rrs + n0 + cb);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + 128));
          }
        }
        int cls = 0;
        if (p.labels) cls = min(max(__ldg(p.labels + min(m, p.M - 1) / p.label_rows), 0), p.norm_classes - 1);
        if (e == 0 && lane == 0) GTRACE(2, it, 0);
        pt::mbar_wait(&tmem_full[acc], acc_phase);
        if (e == 0 && lane == 0) GTRACE(2, it, 1);
        pt::tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        uint32_t vr[kChunks][32];
#pragma unroll
        for (int c = 0; c < kChunks; ++c) pt::tmem_ld32(t_row + cb + 32 * c, vr[c]);
        pt::tmem_ld_wait();
        pt::tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
        if (e == 0 && lane == 0) GTRACE(2, it, 2);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        float f[kChunks][32];
        // ---- v = acc + bias (+ residual); per-row sum and sum of squares over this warp's 128 columns
        // (fp32, packed f32x2 math; LayerNorm inputs here have |mean| <~ std, so E[v^2] - E[v]^2 is well conditioned)
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bb = *reinterpret_cast<const float4*>(rn_bias_s + cb + 32 * c + 4 * i);
            f[c][4 * i] = __uint_as_float(vr[c][4 * i]); f[c][4 * i + 1] = __uint_as_float(vr[c][4 * i + 1]);
            f[c][4 * i + 2] = __uint_as_float(vr[c][4 * i + 2]); f[c][4 * i + 3] = __uint_as_float(vr[c][4 * i + 3]);
            add2(f[c][4 * i], f[c][4 * i + 1], bb.x, bb.y);
            add2(f[c][4 * i + 2], f[c][4 * i + 3], bb.z, bb.w);
          }
          if (resp) {
            // 64-byte row segments -> swizzled smem tile -> this lane's own row
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(tile_in[j]) = xr[c][j];
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint4 x = *reinterpret_cast<const uint4*>(tile_row[i]);
              const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int j = 0; j < 4; ++j)            // bf16 -> fp32 is a 16-bit shift: low half << 16, high half masked
                add2(f[c][8 * i + 2 * j], f[c][8 * i + 2 * j + 1], __uint_as_float(w4[j] << 16),
                     __uint_as_float(w4[j] & 0xffff0000u));
            }
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            add2(s0, s1, f[c][2 * i], f[c][2 * i + 1]);
            fma2(q0, q1, f[c][2 * i], f[c][2 * i + 1], f[c][2 * i], f[c][2 * i + 1], q0, q1);
          }
        }
        const float wsum = s0 + s1, wsq = q0 + q1;
        // ---- publish (sum, sumsq) to both CTAs of the pair, then wait for all four partials of every row
        const int par = it & 1, src = (int)rank * 2 + half, row = q * 32 + lane;
        stats_s[par][src][row] = make_float2(wsum, wsq);
        // The peer's copy travels as an asynchronous remote store that completes 8 bytes on the PEER's barrier
        // (st.async ... mbarrier::complete_tx): no cluster-scope release or fence is needed, which matters because a
        // release at cluster scope stalls until this thread's earlier output stores have drained.  The local copy is
        // ordered by __syncwarp + a CTA-scope arrive; warp 0 of the epilogue also posts the 2048 bytes the peer's
        // eight warps will deliver.
        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                     ::"r"(map_to_peer(&stats_s[par][src][row], rank ^ 1u)), "f"(wsum), "f"(wsq),
                       "r"(map_to_peer(&stats_bar[par], rank ^ 1u)) : "memory");
        __syncwarp();
        if (lane == 0) {
          if (e == 0) pt::mbar_expect_tx(&stats_bar[par], kEpiWarps * 32 * 8);
          else pt::mbar_arrive(&stats_bar[par]);
        }
        // this row's class picks the staged scale / shift rows (SiLU mode: staged pre-multiplied by 1/2, see below)
        const float* na = rn_tab_s[0][cls] + cb;
        const float* nb = rn_tab_s[1][cls] + cb;
        int ro4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) ro4[j] = __shfl_sync(kFull, obase_m, pr + 8 * j);
        if (e == 0 && lane == 0) GTRACE(2, it, 3);
        pt::mbar_wait(&stats_bar[par], (uint32_t)(it >> 1) & 1u);   // partials land in THIS CTA's smem: CTA-scope acquire
        if (e == 0 && lane == 0) GTRACE(2, it, 4);
        float tsum = 0.f, tsq = 0.f;
#pragma unroll
        for (int sidx = 0; sidx < 4; ++sidx) {             // fixed order: all four owners of a row get identical statistics
          const float2 st = stats_s[par][sidx][row];
          tsum += st.x; tsq += st.y;
        }
        const float mean = tsum * (1.f / 512.f);
        const float rstd = rsqrtf(fmaxf(tsq * (1.f / 512.f) - mean * mean, 0.f) + 1e-5f);
        const float nmr = -mean * rstd;
        __nv_bfloat16* outp = static_cast<__nv_bfloat16*>(p.out);
        // ---- y = ((v - mean) * rstd) * A[c] + B[c]  (+ SiLU), packed math, transpose through smem, store
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 ga = *reinterpret_cast<const float4*>(na + 32 * c + 4 * i);
            const float4 gb = *reinterpret_cast<const float4*>(nb + 32 * c + 4 * i);
            float y0, y1, y2, y3;
            fma2(y0, y1, f[c][4 * i], f[c][4 * i + 1], rstd, rstd, nmr, nmr);
            fma2(y2, y3, f[c][4 * i + 2], f[c][4 * i + 3], rstd, rstd, nmr, nmr);
            fma2(y0, y1, y0, y1, ga.x, ga.y, gb.x, gb.y);
            fma2(y2, y3, y2, y3, ga.z, ga.w, gb.z, gb.w);
            if (silu) {
              // SiLU(2h) = h + h * tanh(h): the staged tables carry the factor 1/2, one MUFU op per element
              fma2(y0, y1, y0, y1, tanh_approx(y0), tanh_approx(y1), y0, y1);
              fma2(y2, y3, y2, y3, tanh_approx(y2), tanh_approx(y3), y2, y3);
            }
            pk[2 * i] = cvt_bf16x2(y0, y1, false);
            pk[2 * i + 1] = cvt_bf16x2(y2, y3, false);
          }
          if (p.tma_store) {
            if (e == 0 && lane == 0 && c == 1) GTRACE(2, it, 6);
            if (lane == 0) pt::bulk_wait_read_all();     // the previous box has been read out of the tile
            __syncwarp();
            if (e == 0 && lane == 0 && c == 1) GTRACE(2, it, 7);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(tile_row[i]) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            pt::fence_async_smem();
            __syncwarp();
            if (lane == 0) { pt::tma_store_2d(&tmap_o, tile, n0 + cb + 32 * c, m0 + q * 32); pt::bulk_commit(); }
            continue;
          }
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(tile_row[i]) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (ro4[j] >= 0)
              *(reinterpret_cast<uint4*>(outp + ro4[j] + 32 * c) + pc) = *reinterpret_cast<const uint4*>(tile_in[j]);
          }
        }
        if (resp && p.tma_store) {                       // the next tile's residual transposition reuses the tile
          if (lane == 0) pt::bulk_wait_read_all();
          __syncwarp();
        }
        if (e == 0 && lane == 0) GTRACE(2, it, 5);
        continue;
      }
      __syncwarp();
      row_out[lane] = row_ok ? obase : -1;
      row_res[lane] = (long long)b * p.rss + (long long)t * p.rrs;
      // this warp's bias columns of the tile go to smem now, so their latency hides behind the accumulator wait
      float* bias_w = bias_s[kRowNorm ? 0 : e];
      if (p.bias) {
        if (lane < BN / 8)
          *reinterpret_cast<float4*>(bias_w + lane * 4) = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + half * (BN / 2)) + lane);
      }
      __syncwarp();
      long long ro4[4], rs4[4];                            // offsets of the rows this lane moves to / from global memory
#pragma unroll
      for (int j = 0; j < 4; ++j) { ro4[j] = row_out[pr + 8 * j]; rs4[j] = row_res[pr + 8 * j]; }
      // f[0..31] += residual[row, n .. n+31] (bf16, row-major): coalesced 64-byte row segments -> smem tile -> own row
      // the residual of a 32-column chunk: res_fetch issues this lane's four 16-byte loads (coalesced 64-byte row
      // segments), res_apply transposes them through the smem tile and adds them to the lane's own row
      auto res_fetch = [&](uint4 (&xr)[4], int n) {
  #pragma unroll
        for (int j = 0; j < 4; ++j) {
          xr[j] = make_uint4(0, 0, 0, 0);
          if (ro4[j] >= 0 && n + pc * 8 < p.n_valid)
            xr[j] = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + rs4[j] + n) + pc);
        }
      };
      auto res_apply = [&](float (&f)[32], const uint4 (&xr)[4]) {
        __syncwarp();
  #pragma unroll
        for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(tile + PT_TILE_OFF(pr + 8 * j, pc)) = xr[j];
        __syncwarp();
  #pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint4 x = *reinterpret_cast<const uint4*>(tile + PT_TILE_OFF(lane, i));
          const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
  #pragma unroll
          for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 hh = *reinterpret_cast<const __nv_bfloat162*>(&w4[j]);
            f[8 * i + 2 * j] += __bfloat162float(hh.x);
            f[8 * i + 2 * j + 1] += __bfloat162float(hh.y);
          }
        }
      };
      pt::mbar_wait(&tmem_full[acc], acc_phase);
      pt::tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c0 = half * (BN / 2); c0 < (half + 1) * (BN / 2); c0 += 32) {
        const int n = n0 + c0;
        if (n >= p.n_valid) break;                       // warp-uniform
        uint32_t v[32];
        pt::tmem_ld32(t_row + c0, v);
        pt::tmem_ld_wait();
        if (lean && n + 32 <= p.n_valid) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float a0 = __uint_as_float(v[4 * i]), a1 = __uint_as_float(v[4 * i + 1]);
            float a2 = __uint_as_float(v[4 * i + 2]), a3 = __uint_as_float(v[4 * i + 3]);
            if (p.bias) {
              const float4 bb = *reinterpret_cast<const float4*>(bias_w + (c0 - half * (BN / 2)) + 4 * i);
              add2(a0, a1, bb.x, bb.y);
              add2(a2, a3, bb.z, bb.w);
            }
            if (lean_gelu) { gelu2(a0, a1); gelu2(a2, a3); }     // wav2vec2's convolutions and FF1 (erf GELU)
            pk[2 * i] = cvt_bf16x2(a0, a1, lean_relu);
            pk[2 * i + 1] = cvt_bf16x2(a2, a3, lean_relu);
          }
          if (p.tma_store) {
            // the tile leaves through the TMA: no LDS / STG, no per-row predicates (the tensor map clips at M)
            if (lane == 0) pt::bulk_wait_read_all();     // the previous box has been read out of the tile
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(tile_row[i]) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            pt::fence_async_smem();
            __syncwarp();
            if (lane == 0) { pt::tma_store_2d(&tmap_o, tile, n, m0 + q * 32); pt::bulk_commit(); }
            continue;
          }
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(tile_row[i]) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (ro4[j] >= 0)
              *(reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + ro4[j] + n) + pc) =
                  *reinterpret_cast<const uint4*>(tile_in[j]);
          }
          continue;
        }
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bb = *reinterpret_cast<const float4*>(bias_w + (c0 - half * (BN / 2)) + 4 * i);
            f[4 * i] += bb.x; f[4 * i + 1] += bb.y; f[4 * i + 2] += bb.z; f[4 * i + 3] += bb.w;
          }
        }
        if (!row_major) {
          // channel-major output (final projections): lanes = consecutive rows -> already coalesced per column
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (n + i < p.n_valid) {
                const float x = apply_act(f[i], p.act, p.slope);
                const long long off = obase + (long long)(n + i) * p.ocs;
                if (p.out_f32) static_cast<float*>(p.out)[off] = x;
                else static_cast<__nv_bfloat16*>(p.out)[off] = __float2bfloat16_rn(x);
              }
            }
          }
          continue;
        }
        // ---- residual: coalesced 64-byte row segments -> smem tile -> own row
        if (p.residual) {
          if (p.res_f32) {
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {             // 16 fp32 columns (64 B) at a time
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int rr = pr + 8 * j;
                float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ro4[j] >= 0 && n + h2 * 16 + pc * 4 < p.n_valid)
                  x = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) + rs4[j] + n + h2 * 16) + pc);
                *reinterpret_cast<float4*>(tile + PT_TILE_OFF(rr, pc)) = x;
              }
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 x = *reinterpret_cast<const float4*>(tile + PT_TILE_OFF(lane, i));
                f[h2 * 16 + 4 * i] += x.x; f[h2 * 16 + 4 * i + 1] += x.y; f[h2 * 16 + 4 * i + 2] += x.z; f[h2 * 16 + 4 * i + 3] += x.w;
              }
            }
          } else {
            // bf16 residual; in split arithmetic it is a (hi, lo) plane pair and both planes are added
            const int planes = p.split ? 2 : 1;
            for (int pl = 0; pl < planes; ++pl) {
              const __nv_bfloat16* rp = static_cast<const __nv_bfloat16*>(pl ? p.residual_lo : p.residual);
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int rr = pr + 8 * j;
                uint4 x = make_uint4(0, 0, 0, 0);
                if (ro4[j] >= 0 && n + pc * 8 < p.n_valid)
                  x = __ldg(reinterpret_cast<const uint4*>(rp + rs4[j] + n) + pc);
                *reinterpret_cast<uint4*>(tile + PT_TILE_OFF(rr, pc)) = x;
              }
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 x = *reinterpret_cast<const uint4*>(tile + PT_TILE_OFF(lane, i));
                const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const __nv_bfloat162 hh = *reinterpret_cast<const __nv_bfloat162*>(&w4[j]);
                  f[8 * i + 2 * j] += __bfloat162float(hh.x);
                  f[8 * i + 2 * j + 1] += __bfloat162float(hh.y);
                }
              }
            }
          }
        }
        // ---- activation + store through the tile
        if (gate) {
          // interleaved columns (a0,b0,a1,b1,...) -> tanh(a) * sigmoid(b): 16 bf16 outputs (32 B) per row
          float g[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) g[i] = tanhf(f[2 * i]) * (1.f / (1.f + __expf(-f[2 * i + 1])));
          const int planes = p.split ? 2 : 1;
          for (int pl = 0; pl < planes; ++pl) {
            __nv_bfloat16* outp = static_cast<__nv_bfloat16*>(pl ? p.out_lo : p.out);
            uint32_t g8[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) g8[i] = pt::pack_bf16x2(g[2 * i], g[2 * i + 1]);
            __syncwarp();
            *reinterpret_cast<uint4*>(tile + PT_TILE_OFF(lane, 0)) = make_uint4(g8[0], g8[1], g8[2], g8[3]);
            *reinterpret_cast<uint4*>(tile + PT_TILE_OFF(lane, 1)) = make_uint4(g8[4], g8[5], g8[6], g8[7]);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const int rr = (lane >> 1) + 16 * j, pp = lane & 1;
              const long long ro = row_out[rr];
              if (ro >= 0)
                *reinterpret_cast<uint4*>(outp + ro + (n >> 1) + pp * 8) =
                    *reinterpret_cast<const uint4*>(tile + PT_TILE_OFF(rr, pp));
            }
            if (planes == 2 && pl == 0) {                   // what the hi plane could not represent goes to the lo plane
#pragma unroll
              for (int i = 0; i < 16; ++i) g[i] -= __bfloat162float(__float2bfloat16_rn(g[i]));
            }
          }
        } else if (p.out_f32) {
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<float4*>(tile + PT_TILE_OFF(lane, i)) =
                  make_float4(apply_act(f[h2 * 16 + 4 * i], p.act, p.slope), apply_act(f[h2 * 16 + 4 * i + 1], p.act, p.slope),
                              apply_act(f[h2 * 16 + 4 * i + 2], p.act, p.slope), apply_act(f[h2 * 16 + 4 * i + 3], p.act, p.slope));
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int rr = pr + 8 * j;
              const long long ro = ro4[j];
              if (ro >= 0 && n + h2 * 16 + pc * 4 < p.n_valid)
                *(reinterpret_cast<float4*>(static_cast<float*>(p.out) + ro + n + h2 * 16) + pc) =
                    *reinterpret_cast<const float4*>(tile + PT_TILE_OFF(rr, pc));
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = apply_act(f[i], p.act, p.slope);
          const int planes = p.split ? 2 : 1;
          for (int pl = 0; pl < planes; ++pl) {
            __nv_bfloat16* outp = static_cast<__nv_bfloat16*>(pl ? p.out_lo : p.out);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(tile + PT_TILE_OFF(lane, i)) = make_uint4(
                  pt::pack_bf16x2(f[8 * i], f[8 * i + 1]), pt::pack_bf16x2(f[8 * i + 2], f[8 * i + 3]),
                  pt::pack_bf16x2(f[8 * i + 4], f[8 * i + 5]), pt::pack_bf16x2(f[8 * i + 6], f[8 * i + 7]));
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int rr = pr + 8 * j;
              const long long ro = ro4[j];
              if (ro >= 0 && n + pc * 8 < p.n_valid)
                *(reinterpret_cast<uint4*>(outp + ro + n) + pc) = *reinterpret_cast<const uint4*>(tile + PT_TILE_OFF(rr, pc));
            }
            if (planes == 2 && pl == 0) {                   // lo plane = what the hi plane could not represent
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] -= __bfloat162float(__float2bfloat16_rn(f[i]));
            }
          }
        }
      }
      pt::tc_fence_before();
      __syncwarp();
      if (lane == 0) release_acc(acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (p.tma_store && lane == 0) pt::bulk_wait_all();   // this warp's TMA stores are complete before the CTA may exit
  }
  pt::tc_fence_before();
  __syncthreads();
  if constexpr (kClusterSize > 1) cluster_sync_all();  // no CTA exits while a peer may still write into its smem
  if (warp == 2) {
    if constexpr (k2) pt::tmem_dealloc_2sm(tmem_base, C::kTmemCols);
    else pt::tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 row-major [rows, cols] matrix with row pitch `pitch` elements; box = 64 columns x box_rows rows
int make_tmap(CUtensorMap* tm, const void* base, long long rows, long long cols, long long pitch, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return pt::fail("cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return pt::fail("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld pitch=%lld base=%p", (int)r, rows, cols,
                    pitch, base);
  return 0;
}

// bf16 [rows, cols] matrix with `pitch` elements per row, boxes of box_cols x box_rows:
//   * output map for the epilogue's TMA stores: 32 x 32 boxes in the SWIZZLE_64B layout that PT_TILE_OFF produces
//   * residual map for the residual-through-MMA path: 64 x 64 boxes, SWIZZLE_128B
int make_tmap_box(CUtensorMap* tm, const void* base, long long rows, long long cols, long long pitch, int box_cols,
                  int box_rows, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return pt::fail("cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return pt::fail("cuTensorMapEncodeTiled failed (%d) for a %dx%d-box map rows=%lld cols=%lld pitch=%lld", (int)r,
                    box_cols, box_rows, rows, cols, pitch);
  return 0;
}
int make_tmap_out(CUtensorMap* tm, void* base, long long rows, long long cols, long long pitch) {
  return make_tmap_box(tm, base, rows, cols, pitch, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B);
}

template <int BN, int CL>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& tr, const GemmKernelArgs& ka, int sms, cudaStream_t st) {
  PT_SMEM_OPT_IN((gemm_kernel<BN, false, CL>), Cfg<BN>::kSmem);
  const int m_super = (ka.m_tiles + CL - 1) / CL;
  const int tiles = m_super * ka.N_tiles;
  int clusters = sms / CL;
  if (tiles < clusters) clusters = tiles;
  PT_CUDA(pt::launch_pdl(gemm_kernel<BN, false, CL>, dim3(CL * clusters), dim3(kThreads), Cfg<BN>::kSmem, st, CL, ta, tb, to, tr, ka));
  return 0;
}

// cta_group::2 variant: pairs of CTAs, one 256-row x 256-column tile per pair step
int launch_2cta(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& tr, const GemmKernelArgs& ka, int sms, cudaStream_t st) {
  using Cf = Cfg<256, true>;
  PT_SMEM_OPT_IN((gemm_kernel<256, false, 1, true>), Cf::kSmem);
  const int tiles = ((ka.m_tiles + 1) / 2) * ka.N_tiles;
  int pairs = sms / 2;
  if (tiles < pairs) pairs = tiles;
  PT_CUDA(pt::launch_pdl(gemm_kernel<256, false, 1, true>, dim3(2 * pairs), dim3(kThreads), Cf::kSmem, st, 2, ta, tb, to, tr, ka));
  return 0;
}

// row-norm variant: clusters of two CTAs, one 128-row tile per cluster step
int launch_rownorm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& tr, const GemmKernelArgs& ka, int sms, cudaStream_t st) {
  PT_SMEM_OPT_IN((gemm_kernel<256, true, 1>), Cfg<256>::kSmem);
  int clusters = sms / 2;
  if (ka.m_tiles < clusters) clusters = ka.m_tiles;
  PT_CUDA(pt::launch_pdl(gemm_kernel<256, true, 1>, dim3(2 * clusters), dim3(kThreads), Cfg<256>::kSmem, st, 2, ta, tb, to, tr, ka));
  return 0;
}

}  // namespace

extern "C" int pt_gemm_bf16(const pt_gemm_desc* d, void* stream) {
  if (!d) return pt::fail("pt_gemm_bf16: null descriptor");
  if (d->M <= 0 || d->N <= 0) return 0;
  if (!d->a || !d->w || !d->out) return pt::fail("pt_gemm_bf16: null operand");
  if (d->C <= 0 || d->C % BK) return pt::fail("pt_gemm_bf16: C (%d) must be a positive multiple of 64", d->C);
  if (d->taps < 1 || (d->s != 1 && d->s != 2)) return pt::fail("pt_gemm_bf16: bad taps/stride %d/%d", d->taps, d->s);
  if (d->N % 128) return pt::fail("pt_gemm_bf16: N (%d) must be padded to a multiple of 128", d->N);
  if (d->n_valid <= 0 || d->n_valid > d->N) return pt::fail("pt_gemm_bf16: bad n_valid %d", d->n_valid);
  if (d->rows_per_sample <= 0) return pt::fail("pt_gemm_bf16: rows_per_sample must be positive");
  if (d->act == PT_ACT_GATE && (d->out_f32 || d->ocs != 1 || d->residual || (d->n_valid & 31)))
    return pt::fail("pt_gemm_bf16: gated epilogue needs bf16 row-major output, no residual, n_valid %% 32 == 0");
  if ((reinterpret_cast<uintptr_t>(d->a) | reinterpret_cast<uintptr_t>(d->w)) & 15)
    return pt::fail("pt_gemm_bf16: operands must be 16-byte aligned");
  if (d->bias && (reinterpret_cast<uintptr_t>(d->bias) & 15)) return pt::fail("pt_gemm_bf16: bias must be 16-byte aligned");
  if (d->ocs == 1) {
    // row-major outputs / residuals move in 16-byte pieces
    const int vec = d->out_f32 ? 4 : 8;
    const int nv = d->act == PT_ACT_GATE ? d->n_valid / 2 : d->n_valid;
    if ((nv % vec) || (d->ors % vec) || (d->oss % vec) || (reinterpret_cast<uintptr_t>(d->out) & 15))
      return pt::fail("pt_gemm_bf16: row-major output needs n_valid, strides and base aligned to 16 bytes");
    if (d->residual) {
      const int rvec = d->res_f32 ? 4 : 8;
      if ((d->n_valid % rvec) || (d->rrs % rvec) || (d->rss % rvec) || (reinterpret_cast<uintptr_t>(d->residual) & 15))
        return pt::fail("pt_gemm_bf16: residual needs n_valid, strides and base aligned to 16 bytes");
    }
  } else if (d->residual) {
    return pt::fail("pt_gemm_bf16: a residual needs a row-major output");
  }
  const int sms = pt::sm_count();

  CUtensorMap ta, tb;
  // A viewed as [a_rows / s, s * C] (tap j of a stride-2 conv = row j/2, column half j%2)
  if (d->a_rows % d->s) return pt::fail("pt_gemm_bf16: a_rows (%lld) not divisible by stride", d->a_rows);
  if (d->a_col_table) {
    if (d->row_norm || d->split || d->s != 1 || d->N / 128 > 8)
      return pt::fail("pt_gemm_bf16: a_col_table needs s = 1, N <= 1024, no row_norm, no split arithmetic");
    for (int j = 0; j < d->N / 128; ++j)
      if (d->a_col_table[j] < 0 || (d->a_col_table[j] & 7) || d->a_col_table[j] + d->C > d->a_pitch)
        return pt::fail("pt_gemm_bf16: a_col_table[%d] = %d must be a multiple of 8 with offset + C <= a_pitch", j, d->a_col_table[j]);
  }
  // (with a column table the map spans the whole row pitch: every tile's slab lies inside it)
  if (make_tmap(&ta, d->a, d->a_rows / d->s, d->a_col_table ? d->a_pitch : (long long)d->s * d->C, (long long)d->s * d->a_pitch, BM))
    return 1;
  if (d->split) {
    if (d->row_norm) return pt::fail("pt_gemm_bf16: split arithmetic has no row_norm epilogue");
    if (!d->a_lo || (reinterpret_cast<uintptr_t>(d->a_lo) & 15)) return pt::fail("pt_gemm_bf16: split needs a 16-byte aligned a_lo plane");
    if (!d->out_f32 && (!d->out_lo || d->ocs != 1 || (reinterpret_cast<uintptr_t>(d->out_lo) & 15)))
      return pt::fail("pt_gemm_bf16: split with a bf16 output needs a row-major, 16-byte aligned out_lo plane");
    if (d->residual && !d->res_f32 && (!d->residual_lo || (reinterpret_cast<uintptr_t>(d->residual_lo) & 15)))
      return pt::fail("pt_gemm_bf16: split with a bf16 residual needs a 16-byte aligned residual_lo plane");
  }
  if (d->row_norm) {
    if (d->row_norm < 1 || d->row_norm > 2) return pt::fail("pt_gemm_bf16: bad row_norm %d", d->row_norm);
    if (d->N != 512 || d->n_valid != 512 || d->ocs != 1 || d->out_f32 || (d->residual && d->res_f32) || d->act != PT_ACT_NONE)
      return pt::fail("pt_gemm_bf16: row_norm needs N = 512, bf16 row-major output, bf16 residual (or none), no activation");
    if (!d->bias || !d->norm_a || !d->norm_b) return pt::fail("pt_gemm_bf16: row_norm needs bias, norm_a, norm_b");
    if (d->labels && d->label_rows <= 0) return pt::fail("pt_gemm_bf16: label_rows must be positive");
    if (d->labels && (d->norm_classes < 1 || d->norm_classes > kRnMaxCls))
      return pt::fail("pt_gemm_bf16: row_norm with labels needs 1 <= norm_classes <= %d (got %d)", kRnMaxCls, d->norm_classes);
    {
      // the row-norm epilogue addresses output and residual with 32-bit element offsets
      const long long samples = (d->M + d->rows_per_sample - 1) / d->rows_per_sample;
      const long long omax = samples * llabs(d->oss) + (long long)d->rows_per_sample * llabs(d->ors) + 512;
      const long long rmax = d->residual ? samples * llabs(d->rss) + (long long)d->rows_per_sample * llabs(d->rrs) + 512 : 0;
      if (omax >= (1LL << 31) || rmax >= (1LL << 31) || d->oss < 0 || d->ors < 0 || d->rss < 0 || d->rrs < 0)
        return pt::fail("pt_gemm_bf16: row_norm needs non-negative strides and offsets below 2^31 elements");
    }
    if ((reinterpret_cast<uintptr_t>(d->norm_a) | reinterpret_cast<uintptr_t>(d->norm_b)) & 15)
      return pt::fail("pt_gemm_bf16: norm tables must be 16-byte aligned");
  }
  int BN = d->row_norm ? 256
           : (d->N % 256 == 0 && ((long long)((d->M + BM - 1) / BM) * (d->N / 256)) >= sms) ? 256 : 128;
  {
    const char* e = getenv("PT_GEMM_BN");                 // developer override: force 128- / 256-column tiles
    if (e && !d->row_norm && atoi(e) == 128) BN = 128;
    if (e && !d->row_norm && atoi(e) == 256 && d->N % 256 == 0) BN = 256;
    if (d->a_col_table) BN = 128;                          // the table is indexed by 128-column tile
  }
  constexpr int CL = 1;
  // cta_group::2 pairs for the big plain GEMMs (PT_GEMM_2CTA=0 disables, for A/B measurements)
  bool two_cta = false;
  {
    static const bool off = getenv("PT_GEMM_2CTA") && atoi(getenv("PT_GEMM_2CTA")) == 0;
    two_cta = !off && !d->row_norm && BN == 256 && CL == 1 &&
              ((long long)((d->M + BM - 1) / BM) * (d->N / 256)) >= 2LL * sms;
  }
  // split arithmetic: weight rows are [hi (taps*C) | lo (taps*C)]
  const long long wcols = (long long)(d->split ? 2 : 1) * d->taps * d->C;
  if (make_tmap(&tb, d->w, d->N, wcols, wcols, two_cta ? 128 : BN)) return 1;

  GemmKernelArgs ka;
  ka.M = d->M; ka.N_tiles = d->N / BN; ka.c_iters = d->C / BK; ka.k_base = d->taps * ka.c_iters;
  ka.split = d->split ? 1 : 0; ka.K_iters = (ka.split ? 3 : 1) * ka.k_base;
  ka.out_lo = d->out_lo; ka.residual_lo = d->residual_lo;
  ka.C = d->C; ka.s = d->s; ka.base_row = d->base_row;
  ka.rows_per_sample = d->rows_per_sample; ka.t_valid = d->t_valid; ka.n_valid = d->n_valid;
  ka.oss = d->oss; ka.ors = d->ors; ka.ocs = d->ocs; ka.rss = d->rss; ka.rrs = d->rrs;
  ka.out = d->out; ka.bias = d->bias; ka.residual = d->residual; ka.out_f32 = d->out_f32; ka.res_f32 = d->res_f32;
  ka.act = d->act; ka.slope = d->slope; ka.m_tiles = (d->M + BM - 1) / BM;
  {
    const char* e = getenv("PT_GEMM_TRACE");              // developer hook: device address of 384 int64 (hex or decimal)
    ka.trace = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
  }
  for (int j = 0; j < 8; ++j) ka.a_col_tab[j] = (d->a_col_table && j < d->N / 128) ? d->a_col_table[j] : 0;
  ka.row_norm = d->row_norm; ka.norm_a = d->norm_a; ka.norm_b = d->norm_b; ka.labels = d->labels; ka.label_rows = d->label_rows; ka.norm_classes = d->norm_classes;
  cudaStream_t st = (cudaStream_t)stream;
  // Dense bf16 row-major outputs (every row of every sample valid, one pitch) are stored by TMA from the epilogue's
  // swizzled staging tiles; anything else (ragged rows, fp32, channel-major) keeps the LDS/STG path.
  CUtensorMap to = ta;
  ka.tma_store = 0;
  {
    static const bool off = getenv("PT_GEMM_TMA_STORE") && atoi(getenv("PT_GEMM_TMA_STORE")) == 0;
    const bool dense = !d->out_f32 && d->ocs == 1 && d->t_valid == d->rows_per_sample &&
                       (d->M <= d->rows_per_sample || d->oss == (long long)d->rows_per_sample * d->ors) &&
                       d->ors >= d->n_valid && (d->ors % 8) == 0 &&
                       (d->n_valid % 8) == 0 && (reinterpret_cast<uintptr_t>(d->out) & 15) == 0;
    if (!off && dense && !d->split) {
      if (make_tmap_out(&to, d->out, d->M, d->n_valid, d->ors)) return 1;
      ka.tma_store = 1;
    }
  }
  // Row-norm with a dense bf16 residual: the residual tile rides through the pipeline and is added by the tensor core
  CUtensorMap tr = ta;
  ka.res_mma = 0;
  if (d->row_norm && d->residual && !d->res_f32) {
    static const bool off = getenv("PT_GEMM_RES_MMA") && atoi(getenv("PT_GEMM_RES_MMA")) == 0;
    const bool dense = d->t_valid == d->rows_per_sample &&
                       (d->M <= d->rows_per_sample || d->rss == (long long)d->rows_per_sample * d->rrs) &&
                       d->rrs >= 512 && (d->rrs % 8) == 0 && (reinterpret_cast<uintptr_t>(d->residual) & 15) == 0;
    if (!off && dense) {
      if (make_tmap_box(&tr, d->residual, d->M, 512, d->rrs, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return 1;
      ka.res_mma = 1;
    }
  }
  if (d->split && make_tmap(&tr, d->a_lo, d->a_rows / d->s, (long long)d->s * d->C, (long long)d->s * d->a_pitch, BM)) return 1;
  if (d->row_norm) return launch_rownorm(ta, tb, to, tr, ka, sms, st);
  if (two_cta) return launch_2cta(ta, tb, to, tr, ka, sms, st);
  if (BN == 128) return launch<128, 1>(ta, tb, to, tr, ka, sms, st);
  return launch<256, 1>(ta, tb, to, tr, ka, sms, st);
}
