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
 * Shapes: D = 64 or 128, K a multiple of 16, G >= 1 -- config/vq.json's ProductQuantization (G = 4, K = D = 128) and
 * the single-codebook VectorQuantizerEMA of nets/smplx_body_vq.py:48-53 (G = 1, D = 64, K = code_num = 2048).
 * Codebooks with more than 128 codes are searched in chunks of 128 whose fp32 winners are merged through idx_out
 * itself (packed atomicMin, unpacked before the call's work on the stream ends); gap_out must be NULL for K > 128. */
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

/* Training-side update of the product quantizer: the EMA codebook update and dead-code reset that
 * ProductQuantization.forward performs in training mode, nets/spg/vqvae_modules.py:495-534 (share_code = False) with
 * ExponentialMovingAverage.forward :271-276, plus the commitment term F.mse_loss(x, quantized) of :537.
 *   z, zs_*      the encoder output of this batch, addressed like pt_pq_search;  idx [B, W, G] int64 its codes
 *   codebooks [G, K, D], ema_dw [G, K, D] (`ema_dw.{g}.hidden`), ema_cluster_size [G, K] (`ema_cluster_size.{g}.hidden`):
 *                fp32 state, updated in place;  step = the reference's EMA counter AFTER its increment (1, 2, ...)
 *   mse_out      fp64 [1] or NULL: mean |x - codebook_before[idx]|^2 (multiply by commitment_cost for the loss)
 *   dead_out     int32 [G] or NULL: codes re-seeded from the batch this step (normalised EMA cluster size < 1)
 *   workspace    device memory, 8-byte aligned, pt_pq_ema_workspace_bytes(G, K, D) bytes (contents clobbered)
 * Needs B*W >= K (the reference re-seeds from the first K rows of the batch; with fewer rows it adds torch.randn noise,
 * which is refused here).  Deterministic: fixed summation order, no floating-point atomics. */
long long pt_pq_ema_workspace_bytes(int G, int K, int D);
int pt_pq_ema_update(const float* z, long long zs_b, long long zs_c, long long zs_w, int B, int W, int G, int K, int D,
                     const long long* idx, float* codebooks, float* ema_dw, float* ema_cluster_size, double decay,
                     double epsilon, int step, double* mse_out, int* dead_out, void* workspace, long long workspace_bytes,
                     void* stream);

/* Codebook gather fused into the decoder's first convolution.  Replaces quantize_all (nets/spg/vqvae_modules.py:557-564)
 * followed by DecoderSC.aft_vq_conv, the 1x1 Conv1d(G*D -> N) at nets/inpainting/vqvae_1d_sc.py:259-261:
 *   out[b, w, n] = bias[n] + sum_g table[g][idx[b, w, g]][n],   table[g] = C_g . W[:, g*D:(g+1)*D]^T  (fp32 [G, K, N],
 *   prepared once per checkpoint by the caller), fp32 sums in group order.
 *   idx [B, W, G] int64 / int32;  out bf16 at out[b*oss + w*ors + n] (N multiple of 8, 16-byte aligned rows);
 *   out_lo: NULL, or the lo plane of a split-bf16 pair (see pt_gemm_desc.split);
 *   bad_count int32 [1] or NULL: incremented for every index outside [0, K) (such entries read code 0). */
int pt_pq_gather_conv1x1(const void* idx, int idx_is_i64, const float* table, const float* bias, void* out_bf16,
                         void* out_lo, long long oss, long long ors, int B, int W, int G, int K, int N, int* bad_count,
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
enum { PT_ACT_NONE = 0, PT_ACT_RELU = 1, PT_ACT_LEAKY = 2, PT_ACT_GATE = 3, PT_ACT_GELU = 4 /* erf GELU (wav2vec2) */ };

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
  /* fused row normalisation over a 512-wide output row (N must be 512, bf16 row-major output):
   *   out = f( (v - mean(v)) * rstd(v) * norm_a[c] + norm_b[c] ),  v = A.W^T + bias (+ bf16 residual, which may
   *   alias out),  eps 1e-5,
   *   c = labels[row / label_rows] (labels NULL: c = 0);  row_norm 0 = off, 1 = SiLU, 2 = identity.
   * row_norm 1 is StylizationBlock_V3's  SiLU(LN(h)(1+scale)+shift)  (PTransformer.py:134-139) with
   * norm_a = gamma*(1+scale_c), norm_b = beta*(1+scale_c)+shift_c;  row_norm 2 is a plain nn.LayerNorm. */
  int row_norm;
  const float* norm_a;    /* fp32 [n_classes, 512] */
  const float* norm_b;    /* fp32 [n_classes, 512] */
  const int* labels;      /* int32 per sample or NULL */
  int label_rows;
  int norm_classes;       /* rows of norm_a / norm_b that labels may select: 1..4 when labels is not NULL (the tables are
                           * staged on chip; labels are clamped to the range) */
  /* split-bf16 ("bf16x3") arithmetic, the tolerance-tight mode (fp32 reference: every nn.Conv1d / nn.Linear above).
   * An fp32 value v travels as two bf16 planes  hi = bf16(v), lo = bf16(v - hi)  (16 mantissa bits).  split != 0:
   *   A    = the plane pair (a, a_lo), same rows / pitch;
   *   W    = [N, 2*taps*C]: each row is its hi plane followed by its lo plane;
   *   acc  = A_hi W_hi + A_hi W_lo + A_lo W_hi  (three tensor-core passes, fp32 accumulation);
   *   bf16 outputs are written as the pair (out, out_lo), bf16 residuals are read as (residual, residual_lo);
   *   fp32 outputs / residuals are unchanged; row_norm is not available. */
  int split;
  const void* a_lo;
  void* out_lo;
  const void* residual_lo;
  /* grouped convolutions whose groups do not align with the 128-column output tiles (wav2vec2's positional convolution,
   * transformers modeling_wav2vec2 Wav2Vec2PositionalConvEmbedding: Conv1d(768, 768, 128, groups = 16)): output tile j
   * (columns [128 j, 128 j + 128)) reads its C input channels starting at column a_col_table[j] of A instead of 0, so
   * that ONE launch covers every tile (W holds the tiles' zero-padded [128, taps*C] matrices one after the other).
   * NULL, or a HOST array of N / 128 <= 8 ints, multiples of 8 with a_col_table[j] + C <= a_pitch; forces 128-column
   * tiles; not with row_norm, split or s = 2. */
  const int* a_col_table;
} pt_gemm_desc;

int pt_gemm_bf16(const pt_gemm_desc* desc, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Multi-head attention core (softmax(q k^T * scale) v), bf16 in/out, fp32 softmax and accumulation.
 * Replaces the scaled-dot-product part of nn.MultiheadAttention as called at
 * nets/inpainting/PTransformer.py:340 (self) and :350 (cross); in/out projections are pt_gemm_bf16.
 *   q   rows b*Nq + i, head h at columns [h*head_dim, (h+1)*head_dim), row stride q_stride elements
 *   k,v rows b*Nk + j, same column rule, strides k_stride / v_stride
 *   out rows b*Nq + i, row stride out_stride
 * head_dim must be 64.  Nk <= 256: tcgen05 kernel; up to 1024 keys / queries: K, V, Q of one (sample, head) resident in
 * shared memory (mma.sync); beyond that (wav2vec2's ~3000 frames per minute of audio) the keys are streamed. */
int pt_attention(const void* q, long long q_stride, const void* k, long long k_stride, const void* v,
                 long long v_stride, void* out, long long out_stride, int B, int heads, int head_dim, int Nq, int Nk,
                 float scale, void* stream);

/* The same attention core for the split-bf16 (tolerance-tight) mode: q, k, v fp32 (same row / head addressing, strides in
 * elements), fp32 CUDA-core arithmetic, softmax normalised before P.V like the reference; the output is bf16, or a
 * split-bf16 pair when out_lo is not NULL.  head_dim 64, 1 <= Nk <= 256. */
int pt_attention_f32(const float* q, long long q_stride, const float* k, long long k_stride, const float* v,
                     long long v_stride, void* out_bf16, void* out_lo, long long out_stride, int B, int heads,
                     int head_dim, int Nq, int Nk, float scale, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Layout kernels between the reference's channel-first fp32 tensors [B, C, T] and the engine's
 * channels-last bf16 rows [B, Tp, Cp] (H zero halo rows in front of every sample).
 * pt_pack_rows writes channels [c_off, c_off + C) of rows H .. H+T-1:
 *   mode 0: src                     (audio features, predictor.py:428 / refiner.py:544)
 *   mode 1: src * mask              (input_poses = gt_poses * mask, predictor.py:425)
 *   mode 2: src * mask + interp(pred, T) * (1 - mask)     (refiner.py:524-529, linear F.interpolate)
 * dst_lo: NULL, or the lo plane of a split-bf16 pair (same layout as dst_bf16; see pt_gemm_desc.split).
 * pt_pack_mask writes the mask channel (onehot = 0; predictornet.py:85) or the two one-hot channels
 * (onehot = 1; RTransformer.py:218). */
int pt_pack_rows(const float* src, const float* mask, const float* pred, int pred_T, void* dst_bf16, void* dst_lo, int B,
                 int C, int T, int Tp, int H, int Cp, int c_off, int mode, void* stream);
int pt_pack_mask(const float* mask, void* dst_bf16, int B, int T, int Tp, int H, int Cp, int c_off, int onehot,
                 void* stream);

/* Feature statistics for the Frechet gesture distance: mean (np.mean(axis=0)) and unbiased covariance
 * (np.cov(rowvar=False)) of N feature vectors, fp64, deterministic (fixed reduction order, no atomics).
 * Replaces the statistics half of EmbeddingSpaceEvaluator.get_scores / frechet_distance (evaluation/FGD.py:172-176);
 * the 128 x 128 matrix square root of calculate_frechet_distance (:190-257) stays on the host.
 *   feats [N, D] fp32 row-major, 2 <= N, 1 <= D <= 128;  mean_out [D], cov_out [D, D] fp64 (device memory).
 * The per-chunk partial sums need scratch memory: the caller provides `workspace` (device memory, 8-byte aligned,
 * at least pt_feature_stats_workspace_bytes(D) bytes; contents need not be initialised and are clobbered). */
long long pt_feature_stats_workspace_bytes(int D);
int pt_feature_stats(const float* feats, int N, int D, double* mean_out, double* cov_out, void* workspace,
                     long long workspace_bytes, void* stream);

/* pt_feature_stats without a workspace argument and without any allocation: the partial sums go into a fixed
 * 2 113 536-byte array inside the library (one per device).  NO COUNTERPART IN THE REFERENCE -- it exists for callers
 * that may not allocate device memory (see csrc/pt_stats.cu).  D > 128 is refused.  The array is shared: calls that can
 * overlap in time (other streams / threads) must be serialised by the caller; make one plain call per device before
 * capturing a stream, because the first call on a device looks the array's address up through the CUDA runtime.
 * The array used is the CURRENT device's: the device that owns `feats`, the outputs and `stream` must be current. */
int pt_feature_stats_static(const float* feats, int N, int D, double* mean_out, double* cov_out, void* stream);

/* Motion metrics of the evaluator, fp64 results, deterministic, no scratch beyond the outputs.
 * pt_lvd: landmark velocity difference, LVD -> Batch_LVD with symmetrical = False, weight = False
 *   (evaluation/metrics.py:79-94, :26-77; called at scripts/test_holisticbody.py:120,130):
 *   gt [T, J, D], pr [B, T, J, D] fp32 -> per_sample [B] = sum_t sum_j | |d pr|_2 - |d gt|_2 | / (T - 1), mean_out [1].
 * pt_diversity: mean pairwise L1 distance between samples (evaluation/metrics.py:96-110):
 *   kps [B, L] fp32 -> per_pair [B (B - 1) / 2] (pairs i < j, i outer) = mean_l |x_i - x_j|, mean_out [1]. */
int pt_lvd(const float* gt, const float* pr, int B, int T, int J, int D, double* per_sample, double* mean_out, void* stream);
int pt_diversity(const float* kps, int B, long long L, double* per_pair, double* mean_out, void* stream);

/* Pose post-processing: normalised network output -> SMPL-X axis-angle parameters in one pass.
 * Replaces the sequence of scripts/demo.py:183-205:
 *   denormalize (nets/utils.py:27-43; mean / std already reordered to the 'fbhe' channel order, NULL = skip)
 *   -> per joint rotation_6d_to_matrix -> matrix_to_axis_angle (data_utils/rotation_conversion.py:512-534, 433-447,
 *      97-119, 481-509) -> cat(expression) -> part2full (data_utils/lower_body.py:66-83, lower27 = its `lp` row).
 *   poses [B, 6*n_joints + n_expr, T] fp32 channel-first  ->  out [B, T, 3*n_joints + n_expr + 27] fp32. */
int pt_pose_postprocess(const float* poses, const float* mean, const float* stdv, const float* lower27, float* out, int B,
                        int T, int n_joints, int n_expr, void* stream);

/* out = known * mask^mask_pow + interp(pred, T) * (1 - mask), all fp32 channel-first [B, C, T] (pred [B, C, pred_T]).
 * mask_pow = 2: predictor.py:445-446 ((gt*mask)*mask + ...);  mask_pow = 1: refiner.py:553. */
int pt_interp_blend(const float* known, const float* mask, const float* pred, int pred_T, float* out, int B, int C,
                    int T, int mask_pow, void* stream);

/* out = LayerNorm512(in + residual), eps 1e-5: the `x = norm_i(x + sublayer(x))` of the post-norm layer
 * (PTransformer.py:329-331).  in fp32 or bf16, residual bf16 or NULL (may alias out_bf16), fp32 statistics,
 * fp32 and/or bf16 out.  residual_lo / out_lo: NULL, or the lo planes of split-bf16 pairs (pt_gemm_desc.split). */
int pt_layernorm512(const void* in, int in_f32, const void* residual_bf16, const void* residual_lo, const float* gamma,
                    const float* beta, float* out_f32, void* out_bf16, void* out_lo, int M, void* stream);

/* StylizationBlock_V3 'sb' branch before its output Linear (PTransformer.py:134-139):
 *   out = SiLU( LN(in) * (1 + scale) + shift ),  [scale | shift] = mod[labels[row / rows_per_sample]]  (mod: [n_classes, 1024] fp32,
 *   the emb_layers_2 Linear applied to SiLU(class_embed) -- precomputed per class by the host). */
int pt_ln_mod_silu512(const void* in, int in_f32, const float* gamma, const float* beta, const float* mod,
                      const int* labels, int rows_per_sample, void* out_bf16, void* out_lo, int M, void* stream);

/* Token embedding + positional encodings (PTransformer.py:659-679): row r -> float32(emb[tokens[r]] + pe64[r % N]),
 * the fp64 add + fp32 round of the reference.  bad_count (or NULL) counts tokens outside [0, n_emb). */
int pt_embed_tokens512(const int* tokens, const float* emb, int n_emb, const double* pe64, int N, float* out_f32,
                       void* out_bf16, void* out_lo, int M, int* bad_count, void* stream);

/* One MaskGIT refinement step: PTransformer.sample_good loop body (PTransformer.py:603-638) and
 * mask_by_random_topk (:696-704).  logits fp32 [B, N, n_codes]; cur_ids/next_ids int32 [B, N] (may alias);
 * u uniform(0,1) and gumbel Gumbel(0,1) noise [B, N] supplied by the caller (the reference draws them from
 * torch RNG streams that are not reproducible across devices); base_mask_len = floor(N * cos(pi/2 * (step+1)/T)). */
int pt_maskgit_step(const float* logits, const int* cur_ids, const float* u, const float* gumbel, int* next_ids,
                    int B, int N, int n_codes, int mask_id, int base_mask_len, float temperature, void* stream);

/* ---- wav2vec2 audio features: scripts/demo.py:105-111 -------------------------------------------
 * `aud_m(**aud_p(audio)).last_hidden_state` (transformers' Wav2Vec2Model, wav2vec2-base geometry; the reference's own
 * copy of the model is nets/spg/wav2vec.py:64-151) followed by data_utils/utils.py:14-16 linear_interpolation.
 *
 * pt_w2v_frontend: the processor's zero-mean / unit-variance normalisation (eps 1e-7, population variance) + feature
 * encoder layer 0 (Conv1d(1, 512, 10, stride 5, no bias) -> GroupNorm(512, 512, eps 1e-5) -> GELU) in one pass.
 *   wave [B, L] fp32 (all utterances of a call have the same length; the reference runs batch 1),  conv0_w [512, 10],
 *   out bf16 rows [B, out_rows_per_sample, 512], rows t < T0 = (L - 10) / 5 + 1 written.  workspace: see _bytes(B). */
long long pt_w2v_frontend_workspace_bytes(int B);
int pt_w2v_frontend(const float* wave, int B, long long L, const float* conv0_w, const float* gn_gamma,
                    const float* gn_beta, void* out_bf16, long long out_rows_per_sample, void* workspace,
                    long long workspace_bytes, void* stream);

/* F.interpolate(x.transpose(1, 2), size = Tout, mode = 'linear') of token-major rows: in [B, in_rows_per_sample, C]
 * (bf16, or fp32 when in_f32), rows t < Tin used -> out [B, C, Tout] fp32 (data_utils/utils.py:14-16). */
int pt_w2v_interp(const void* in, int in_f32, long long in_rows_per_sample, int B, int Tin, int C, float* out,
                  int Tout, void* stream);

/* out = LayerNorm_C(in + residual) over rows of C bf16 values (C a multiple of 128, at most 1024), fp32 statistics:
 * wav2vec2's feature-projection LayerNorm(512) and the post-norm LayerNorm(768)s of its encoder
 * (transformers modeling_wav2vec2 Wav2Vec2EncoderLayer / nets/spg/wav2vec.py).  residual NULL or bf16 [M, C]; out may
 * alias in or residual. */
int pt_layernorm_rows(const void* in_bf16, const void* residual_bf16, const float* gamma, const float* beta, float eps,
                      void* out_bf16, int M, int C, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PROBTALK_B200_H_ */
