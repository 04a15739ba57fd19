/* probtalk_b200 -- C ABI of the B200-native ProbTalk token-level inference path.
 *
 * The reference (feifeifeiliu/probtalk) is pure Python/PyTorch and has NO plugin / FFI layer
 * (SURVEY.md section 8b); its de-facto operator boundary is the set of nn.Module.forward calls on
 * the hot path.  Each entry point below names the reference function it replaces (file:line,
 * relative to the reference root).  The Python package `probtalk_b200` binds these with ctypes and
 * mirrors the reference's `nets/` module tree on top.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory unless stated otherwise;
 *   - `stream` is a cudaStream_t (0 = default stream); calls enqueue work and return, they never
 *     synchronise, allocate or free device memory;
 *   - return value 0 = success, non-zero = failure with a message in pt_last_error() (thread local);
 *   - "rows" are time steps / tokens, activations are channels-last bf16 unless stated otherwise.
 */
#ifndef PROBTALK_B200_H_
#define PROBTALK_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

/* library version (major * 10000 + minor * 100 + patch) and last error text */
int pt_version(void);
const char* pt_last_error(void);

/* ---------------------------------------------------------------------------------------------
 * Product-quantizer nearest-code search.
 * Replaces ProductQuantization.get_code_indices and the per-group loop of
 * ProductQuantization.forward (eval): nets/spg/vqvae_modules.py:546-555 and :470-491
 * (same math as VectorQuantizerEMA.get_code_indices :350-358 for G = 1).
 *
 *   z          fp32, logical shape [B, G*D, W]; element (b, c, w) at z[b*zs_b + c*zs_c + w*zs_w]
 *              (the reference's channel-first tensor is zs_b = G*D*W, zs_c = W, zs_w = 1; a
 *              channels-last tensor is zs_c = 1)
 *   codebooks  fp32 [G, K, D] (the stacked `embeddings_{g}` buffers), 16-byte aligned
 *   idx_out    int64 [B, W, G]   -- what the reference returns in QuantizationOutput.loss (eval)
 *   gap_out    fp32  [B, W, G] or NULL: distance gap between the best and the second-best code
 *   score_dump fp32  [B, W, G, K] or NULL (test hook: the tensor-core scores |c|^2 - 2 x.c)
 *   n_rescored int32 [1] or NULL (test hook: incremented once per row decided by fp32 re-scoring)
 * Only K = D = 128 (config/vq.json) is built. */
int pt_pq_search(const float* z, long long zs_b, long long zs_c, long long zs_w, int B, int W, int G, int K, int D,
                 const float* codebooks, long long* idx_out, float* gap_out, float* score_dump, int* n_rescored,
                 void* stream);

/* Codebook gather.  Replaces ProductQuantization.quantize_all / .quantize:
 * nets/spg/vqvae_modules.py:557-569.
 *   idx   [B, W, G] int64 (idx_is_i64 != 0) or int32
 *   out   element (b, w, g*D + d) at out[b*os_b + (g*D+d)*os_c + w*os_w], fp32 or bf16
 *   bad_count int32 [1] or NULL: incremented for every index outside [0, K) (such rows read code 0) */
int pt_pq_gather(const void* idx, int idx_is_i64, const float* codebooks, void* out, int out_is_bf16,
                 long long os_b, long long os_c, long long os_w, int B, int W, int G, int K, int D, int* bad_count,
                 void* stream);

/* ---------------------------------------------------------------------------------------------
 * Tensor-core contraction engine (TMA + tcgen05/TMEM).  One descriptor covers the dense ops of
 *   ConvNormRelu / Res_CNR_Stack (BN folded)    nets/spg/vqvae_modules.py:107-245
 *   GatedConv                                    nets/inpainting/gated_pixelcnn_1d.py:39-83
 *   nn.Linear / 1x1 Conv1d of the transformers   nets/inpainting/PTransformer.py:284-372, :657-694
 *                                                nets/inpainting/RTransformer.py:209-230
 *   out[row(m), n] = act( sum_{tap<taps, c<C} A[m + base_row + tap/s][(tap%s)*C + c] * W[n][tap*C + c]
 *                         + bias[n] + residual[row(m), n] )
 * A is bf16 with `a_rows` rows of pitch `a_pitch` elements (viewed as [a_rows/s, s*C] when s = 2),
 * W is bf16 [N, taps*C] row-major, accumulation fp32.
 * row(m): b = m / rows_per_sample, t = m % rows_per_sample; rows with t >= t_valid are not stored;
 * element (m, n) is stored at out[b*oss + t*ors + n*ocs] (residual: [b*rss + t*rrs + n]). */
enum { PT_ACT_NONE = 0, PT_ACT_RELU = 1, PT_ACT_LEAKY = 2, PT_ACT_GATE = 3 };

typedef struct pt_gemm_desc {
  const void* a;          /* bf16 activations */
  long long a_rows;       /* rows in the A buffer (TMA zero-fills beyond) */
  long long a_pitch;      /* elements between consecutive A rows (>= C, multiple of 8) */
  const void* w;          /* bf16 [N, taps*C] */
  int M, N, C, taps, s;   /* N multiple of 128 (zero-padded weights), C multiple of 64 */
  int base_row;           /* may be negative (TMA zero-fills) */
  int rows_per_sample, t_valid, n_valid;
  void* out;
  long long oss, ors, ocs;
  int out_f32;            /* 0 = bf16, 1 = fp32 */
  const float* bias;      /* fp32 [N] or NULL */
  const void* residual;   /* or NULL */
  long long rss, rrs;
  int res_f32;
  int act;                /* PT_ACT_*; PT_ACT_GATE: columns interleaved (a0,b0,a1,b1..) -> tanh(a)*sigmoid(b) */
  float slope;            /* LeakyReLU slope */
} pt_gemm_desc;

int pt_gemm_bf16(const pt_gemm_desc* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PROBTALK_B200_H_ */
