/*
 * acai_b200.h — C ABI of libacai_b200.so, the sm_100a kernels behind the
 * `acai_omr_b200.models` drop-in for jsnchon/acai-omr's ViTOMR hot path.
 *
 * The reference (pure Python/PyTorch, no FFI of its own) does all arithmetic on
 * this path through torch ops; each entry point below names the reference call
 * site(s) (file:line in jsnchon/acai-omr @ 5819ae3) whose arithmetic it
 * replaces.  The binding a maintainer would add on the reference side is the
 * ctypes stub in INTEGRATION.md (it is what acai_omr_b200/_lib.py does).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its comment says "host";
 *   - row-major, innermost dimension contiguous; `ld*` / `*_stride` are in ELEMENTS;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and
 *     nothing synchronises; the library allocates nothing (workspaces are passed in);
 *   - return value: ACAI_OK (0) or a negative acai_status; acai_last_error() gives the
 *     message for the calling thread.  There is no CPU fallback anywhere.
 */
#ifndef ACAI_B200_H_
#define ACAI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  ACAI_OK = 0,
  ACAI_ERR_INVALID = -1,     /* bad argument (shape, alignment, null pointer)          */
  ACAI_ERR_UNSUPPORTED = -2, /* shape/dtype combination this build has no kernel for   */
  ACAI_ERR_CUDA = -3         /* a CUDA runtime/driver call failed; see acai_last_error */
} acai_status;

typedef enum { ACAI_F32 = 0, ACAI_BF16 = 1 } acai_dtype;
typedef enum { ACAI_ACT_NONE = 0, ACAI_ACT_GELU = 1 /* exact erf */ } acai_act;

int acai_version(void);
const char* acai_last_error(void);
/* 1 if a CUDA device of compute capability 10.x is present and the kernels can launch. */
int acai_device_ok(void);

/* ---- input pipeline: DynamicResize --------------------------------------------------------------------------------
 * Replaces v2.ToImage + v2.ToDtype(float32, scale=True) + DynamicResize (omr_teacher_force_train.py:278-282,
 * pre_train.py:162-166, utils/utils.py:334-370) for a ragged batch of single-channel images: torchvision
 * F.resize(BICUBIC, antialias=True) to (dst_h, dst_w) - torch's separable antialiased bicubic (a = -0.5, support
 * 2 * max(in / out, 1), normalised weights; width first, then height) - the optional centre crop as an output window
 * [crop_top, crop_top + out_h) x [crop_left, crop_left + out_w) of the resized image, and the final clamp to [0, 1].
 * `src` is uint8 (scaled by 1/255 on load) or fp32; `tmp` holds src_h * out_w floats per image; `dst` is what
 * acai_image_desc.pixels points at.  The host computes the target sizes (pure shape logic, DynamicResize.forward). */
enum { ACAI_RESIZE_F32 = 0, ACAI_RESIZE_U8 = 1 };
typedef struct {
  const void* src;      /* (H, W) rows, uint8 or fp32                                   */
  int32_t src_dtype;    /* ACAI_RESIZE_F32 / ACAI_RESIZE_U8                             */
  int32_t src_h, src_w;
  int64_t src_stride;   /* elements between source rows                                */
  float* tmp;           /* [src_h, out_w] scratch of this image                        */
  float* dst;           /* [out_h, out_w] fp32, row pitch dst_stride                   */
  int32_t dst_h, dst_w; /* resize target (before the crop)                             */
  int32_t crop_top, crop_left, out_h, out_w;
  int64_t dst_stride;
} acai_resize_desc;
/* max_taps: an upper bound of the filter window (taps) over the batch and both axes, e.g. 2 * ceil(2 * max(in / out, 1)) + 2:
 * the weights of a strip of outputs are then tabulated once per CTA in shared memory; 0 = per-thread weights (any window). */
int acai_dynamic_resize(const acai_resize_desc* descs /* device array [num_images] */, int32_t num_images,
                        int32_t max_src_h, int32_t max_out_h, int32_t max_out_w, int32_t max_taps, void* stream);

/* ---- K1: fused patchify + embed + 2-D position -----------------------------------------
 * Replaces Encoder.batchify / OMREncoder.batchify / MAEEncoder.batchify (models.py:36-66,
 * 304-332, 128-173): nn.Unfold im2col, nn.Linear(C*P*P -> E), + pos_embedding[:h_p,:w_p]
 * (or its bilinear resize, models.py:291-302, when h_p/w_p exceed the grid and
 * `interpolate` != 0), over a ragged list of images, written PACKED (no padding rows). */
typedef struct {
  const float* pixels; /* (C, H, W) contiguous fp32                                  */
  int32_t height, width;
  int32_t hp, wp;      /* patches per column / row = H / P, W / P                     */
} acai_image_desc;

/* `workspace` (device, >= acai_patch_embed_workspace_bytes, or NULL): with a workspace, bf16 output, interpolate == 0
 * and C*P*P % 64 == 0 the projection runs on the tensor cores: an im2col pass writes the patches as bf16 rows plus each
 * token's position-grid row, then the tcgen05 GEMM adds bias and the gathered position embedding in its epilogue.
 * Otherwise (fp32 output = the token-exact mode, interpolated grids, no workspace) one fp32 SIMT kernel does it all. */
size_t acai_patch_embed_workspace_bytes(int32_t total_tokens, int32_t patch_size, int32_t channels, int32_t embed_dim);
int acai_patch_embed_fwd(const acai_image_desc* images, /* device array [num_images]             */
                         int32_t num_images,
                         const int32_t* cu_seqlens,      /* [num_images+1] packed token offsets   */
                         const int32_t* patch_index,     /* [total_tokens] token -> patch id in its image,
                                                            or NULL for identity (MAE keeps 25 %) */
                         int32_t total_tokens, int32_t patch_size, int32_t channels,
                         const float* weight,            /* [E, C*P*P] fp32 (nn.Linear layout)    */
                         const float* bias,              /* [E] or NULL                            */
                         const float* pos_embedding,     /* [pe_h, pe_w, E] fp32                   */
                         int32_t pe_h, int32_t pe_w, int32_t interpolate,
                         void* out, int32_t out_dtype, int32_t embed_dim,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ---- K2: dense projections ----------------------------------------------------------------
 * out[M,N] = act(a[M,K] . w[N,K]^T + bias) + residual.   Replaces every nn.Linear / F.linear on
 * the path: QKV in_proj, out_proj, linear1/linear2, transition head (models.py:655-660), cross
 * K/V projection (kv_caching.py:241-245), q projections (kv_caching.py:193,215), unembed
 * (models.py:482,527).  bf16 x bf16 problems with K % 64 == 0 and 16-byte aligned rows run on
 * tcgen05 tensor cores (TMA-fed, TMEM accumulators); everything else runs the fp32-accumulating
 * SIMT kernel.  `residual` may alias nothing; `out` may alias `residual`. */
int acai_linear_fwd(const void* a, int32_t a_dtype, int64_t lda,
                    const void* w, int32_t w_dtype, int64_t ldw,
                    const float* bias,
                    const void* residual, int32_t res_dtype, int64_t ldr,
                    void* out, int32_t out_dtype, int64_t ldo,
                    int32_t M, int32_t N, int32_t K, int32_t act, void* stream);

/* Small-M projections of the decode step, fused with what follows them in the layer (kv_caching.py:193-223:
 * x = norm1(x + out_proj(attn)), q = W_q x + b_q, x = norm2(x + out_proj(cross)), x = norm3(x + linear2(gelu(linear1 x)))):
 *   out[M,N] = LayerNorm( act(a[M,K] . w[N,K]^T + bias) + residual )      (LayerNorm only when ln_gamma != NULL)
 * for M <= 256 bf16 rows.  A skinny GEMM offers too few output tiles to fill 148 SMs, so K is SPLIT across CTAs (every SM
 * streams a slice of the weight matrix), fp32 partial tiles go to `workspace`, and one reduce kernel - one warp per row,
 * partials still in L2 - sums them and applies bias, activation, residual and the LayerNorm (two-pass variance in fp32).
 * Returns ACAI_ERR_UNSUPPORTED, nothing written, when the problem does not qualify (operands not 16-byte aligned /
 * K % 64, M > 256, no K slices to gain, N % 4, or a LayerNorm row that is not a multiple of 128 <= 1024 columns): the
 * caller then runs acai_linear_fwd + acai_layernorm_fwd.  `workspace` holds acai_linear_splitk_workspace_bytes(M, N, K).
 * `out2`: under torch.autocast the reference keeps the residual stream in fp32 (LayerNorm and the residual adds run in
 * fp32, vitomr_inference.py:82-84) and only the Linear operands are bf16: out = fp32 row (the next residual), out2 = the
 * same row in bf16 (the next GEMM's operand), written by the same reduce kernel. */
size_t acai_linear_splitk_workspace_bytes(int32_t M, int32_t N, int32_t K);
int acai_linear_splitk(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias,
                       const void* residual, int32_t res_dtype, int64_t ldr,
                       const float* ln_gamma, const float* ln_beta, float ln_eps,
                       void* out, int32_t out_dtype, int64_t ldo,
                       void* out2, int32_t out2_dtype, int64_t ldo2,   /* optional second copy of the result (NULL: none) */
                       int32_t M, int32_t N, int32_t K, int32_t act,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ---- K2 + K5 fused: out = LayerNorm(a . w^T + bias + residual) * gamma + beta ------------------------------------------
 * The tail of every post-norm sub-layer (x = norm1(x + out_proj(attn)), x = norm2(x + linear2(...)): nn.TransformerEncoderLayer
 * / DecoderLayer with norm_first = False, models.py:31,423; spelled out at kv_caching.py:208,220,222) as ONE tcgen05 kernel:
 * a thread-block cluster of N / 256 CTAs owns a 128-row tile, each CTA computes 256 columns into TMEM, adds bias and residual
 * in place, and the row moments are merged across the cluster through distributed shared memory (two-pass variance per
 * thread, Chan's formula across threads and CTAs), so the pre-norm sum never goes to HBM.  Writes the bf16 copy (`out_bf16`,
 * the next GEMM's operand) and optionally the fp32 copy (`out_f32`, the residual stream under autocast semantics; NULL: none).
 * bf16 operands, N in {512, 768, 1024}, K % 64 == 0, 16-byte aligned rows; otherwise ACAI_ERR_UNSUPPORTED and nothing is
 * launched (the caller runs acai_linear_fwd + acai_layernorm_fwd).  `out_*` must not alias `residual`. */
int acai_linear_layernorm(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias,
                          const void* residual, int32_t res_dtype, int64_t ldr,
                          const float* gamma, const float* beta, float eps,
                          void* out_bf16, int64_t ldo16, void* out_f32, int64_t ldo32,
                          int32_t M, int32_t N, int32_t K, void* stream);
/* clusters of `cluster_size` (= N / 256: 2, 3, 4) CTAs of that kernel the device can hold at once (its grid size) */
int acai_linear_layernorm_resident_clusters(int32_t cluster_size);

/* ---- K5: LayerNorm(x + residual) ---------------------------------------------------------
 * Post-norm sublayer tail `norm(x + sublayer(x))` (kv_caching.py:208,220,222; the
 * nn.TransformerEncoderLayer/DecoderLayer internals) and the final stack norm (eps 1e-6).
 * Optionally writes a second copy `out2` (e.g. bf16 operand for the next GEMM). */
int acai_layernorm_fwd(const void* x, int32_t x_dtype, const void* residual, int32_t res_dtype,
                       const float* gamma, const float* beta, float eps,
                       void* out, int32_t out_dtype, void* out2, int32_t out2_dtype,
                       int32_t rows, int32_t cols, void* stream);

/* ---- K3: packed variable-length attention ----------------------------------------------------
 * softmax(q k^T * scale [causal]) v per (sequence, head) over packed rows delimited by
 * cu_seqlens, no padding rows.  Replaces the nn.MultiheadAttention / SDPA inside
 * nn.TransformerEncoder (models.py:30-34,78,186-195,356-366) and nn.TransformerDecoder
 * (models.py:422-426,480): key-padding masks become sequence boundaries.  q/k/v rows are
 * [heads*head_dim] slices of wider rows (e.g. the fused QKV buffer), hence the strides; total_q/total_k
 * are the packed row counts (= cu_seqlens[batch]).  bf16 with head_dim 64 runs the tcgen05 kernel. */
int acai_attn_varlen_fwd(const void* q, const void* k, const void* v, int32_t dtype,
                         int64_t q_stride, int64_t k_stride, int64_t v_stride,
                         const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                         int32_t batch, int32_t max_seqlen_q, int32_t total_q, int32_t total_k,
                         int32_t heads, int32_t head_dim,
                         int32_t causal, float scale,
                         void* out, int32_t out_dtype, int64_t o_stride,
                         float* lse, /* optional [heads, total_q]: log2-domain log-sum-exp per query row (saved for
                                        acai_attn_varlen_bwd); NULL in inference */
                         float dropout_p, uint64_t dropout_seed, /* attention-probability dropout, 0 = off (see below) */
                         void* stream);

/* Backward of the above (autograd of nn.MultiheadAttention / SDPA in the teacher-forced and MAE training steps,
 * omr_teacher_force_train.py:120, pre_train.py:60): flash-style recomputation from `lse`; dq/dk/dv are written (not
 * accumulated) as [heads*head_dim] slices of rows with the given strides, e.g. straight into a fused d(QKV) buffer.
 * `delta_ws`: [heads * total_q] floats of scratch. */
int acai_attn_varlen_bwd(const void* q, const void* k, const void* v, int32_t dtype,
                         int64_t q_stride, int64_t k_stride, int64_t v_stride,
                         const void* out, int64_t o_stride, const void* dout, int64_t do_stride, const float* lse,
                         const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                         int32_t batch, int32_t max_seqlen_q, int32_t max_seqlen_k, int32_t total_q, int32_t total_k,
                         int32_t heads, int32_t head_dim, int32_t causal, float scale,
                         void* dq, void* dk, void* dv, int32_t grad_dtype,
                         int64_t dq_stride, int64_t dk_stride, int64_t dv_stride, float* delta_ws,
                         float dropout_p, uint64_t dropout_seed, void* stream);

/* ---- K4: decode step ---------------------------------------------------------------------------
 * One query row per (sequence, head) against a KV cache.  Replaces
 * CachedMultiheadAttention.cached_forward's SDPA (kv_caching.py:123-136) for both the
 * self-attention cache and the encoder-memory cache.
 *
 * acai_decode_attn: strided (B,H,T,d) caches exactly as the reference KVCache lays them out
 * (kv_caching.py:35-41).  `kv_len` keys are attended; `key_padding_mask` ([B, mask_stride] bytes,
 * nonzero = ignore; kv_caching.py:126-128) may be NULL. */
int acai_decode_attn(const void* q, int32_t q_dtype, int64_t q_batch_stride,
                     const void* k, const void* v, int32_t kv_dtype,
                     int64_t kv_batch_stride, int64_t kv_head_stride, int64_t kv_tok_stride,
                     int32_t batch, int32_t heads, int32_t head_dim, int32_t kv_len,
                     const uint8_t* key_padding_mask, int64_t mask_stride, float scale,
                     void* out, int32_t out_dtype, int64_t o_batch_stride,
                     void* workspace, size_t workspace_bytes, void* stream);

/* KVCache.update (kv_caching.py:57-109): writes `n_new` new positions k_new/v_new (logical
 * (B,H,n_new,d), d contiguous, strides given) at tokens pos..pos+n_new-1 of the (B,H,T,d) caches. */
int acai_kv_append(const void* k_new, const void* v_new, int32_t src_dtype, int64_t src_batch_stride,
                   int64_t src_head_stride, int64_t src_tok_stride, int32_t n_new,
                   void* k_cache, void* v_cache, int32_t cache_dtype,
                   int64_t kv_batch_stride, int64_t kv_head_stride, int64_t kv_tok_stride,
                   int32_t batch, int32_t heads, int32_t head_dim, int32_t pos, void* stream);

/* acai_decode_attn_paged: the same attention over a TOKEN-MAJOR cache whose rows are [K(H*64) | V(H*64)] - the
 * layout the fused QKV / cross-KV projections emit - addressed through a page table (64 tokens per page; the
 * self-attention pool) or as seq_start[b] + t (the packed encoder-memory cache, no padding rows).  If `new_kv` is
 * given, this step's [K|V] row of every sequence is appended at position len (KVCache.update, kv_caching.py:57-109)
 * and attended to.  len = kv_lens[b], or *step_ptr + len_offset (device-side step counter, so that one captured
 * CUDA graph serves every step).  head_dim must be 64; q/new_kv/cache share `dtype`.  Split-KV over 64..256-token chunks;
 * `workspace` (acai_decode_workspace_bytes) holds one int32 arrival counter per (sequence, head) followed by the chunk
 * partials.  bf16 caches: the last warp of a (sequence, head) to publish its partial folds them inside the same kernel;
 * the counters must be ZERO before the first call with a given workspace (cudaMemset once) and every call leaves them
 * zero.  fp32 caches: a separate combine launch. */
size_t acai_decode_workspace_bytes(int32_t batch, int32_t heads, int32_t max_kv_len);
int acai_decode_attn_paged(const void* q, int64_t q_stride, const void* new_kv, int64_t new_kv_stride,
                           void* kv_rows, int32_t dtype, const int32_t* page_table, int32_t table_stride,
                           const int32_t* seq_start, const int32_t* kv_lens, const int32_t* step_ptr,
                           int32_t len_offset, int32_t max_kv_len, int32_t batch, int32_t heads, int32_t head_dim,
                           float scale, void* out, int32_t out_dtype, int64_t o_stride,
                           void* workspace, size_t workspace_bytes,
                           int64_t kv_rows_total, /* rows allocated behind kv_rows (bounds the TMA descriptor of the
                                                     bf16 streaming kernel); 0 = unknown -> register kernel */
                           void* stream);

/* GRPO rollout step: GRPOViTOMR.cached_forward_rollout_policy models.py:1003-1029 (torch.topk, softmax(top_k_logits /
 * temperature), torch.multinomial, log_softmax(top_k_logits).gather) for every row of `logits` [rows, vocab], device-step
 * variant like acai_greedy_update: writes seqs[row, *step_ptr], log_probs[row, *step_ptr], sets finished[row] on <eos>.
 * Randomness: `uniforms` (fp32 [rows, uniforms_stride], entry [row, *step_ptr] in [0, 1)) when given, else a counter-based
 * hash of (*seed_ptr, row, *step_ptr) - the seed lives on the device so that a captured CUDA graph can be replayed with a
 * new seed per rollout.  The token is drawn by inverse CDF over the kept tokens in vocabulary order. */
int acai_topk_sample_update(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab, int32_t top_k,
                            float temperature, const float* uniforms, int64_t uniforms_stride, const uint64_t* seed_ptr,
                            const int32_t* step_ptr, int64_t* seqs, int64_t seq_stride, float* log_probs, uint8_t* finished,
                            int32_t eos_idx, void* stream);

/* Greedy-loop bookkeeping driven by the device-side step counter t = *step_ptr (models.py:575-583, 603-610):
 *   acai_decode_embed : out[b] = table[seqs[b, t-1]] + pos[t]   (time_step = t for the token at t-1, models.py:576)
 *   acai_greedy_update: seqs[b, t] = argmax(logits[b]); log_probs[b, t] = log_softmax(logits[b])[argmax];
 *                       finished[b] |= (argmax == eos)
 *   acai_advance_step : *step_ptr += 1 */
int acai_decode_embed(const int64_t* seqs, int64_t seq_stride, const int32_t* step_ptr, int32_t batch,
                      const float* table, const float* pos, void* out, int32_t out_dtype, void* out2, int32_t out2_dtype,
                      int32_t dim, void* stream);                  /* out2: optional second copy (NULL: none) */
int acai_greedy_update(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab,
                       const int32_t* step_ptr, int64_t* seqs, int64_t seq_stride, float* log_probs,
                       uint8_t* finished, int32_t eos_idx, void* stream);
int acai_advance_step(int32_t* step_ptr, void* stream);

/* Token embedding + learned 1-D position (models.py:460-466, 492-495, 522-524):
 * out[r,:] = table[ids[r],:] + pos[pos_ids ? pos_ids[r] : pos_offset + (r % period), :]. */
int acai_embed_tokens(const int64_t* ids, int32_t rows, const float* table, const float* pos,
                      const int32_t* pos_ids, int32_t pos_offset, int32_t period,
                      void* out, int32_t out_dtype, int32_t dim, void* stream);

/* Greedy epilogue (models.py:579-581): argmax over the vocab (first max wins, as torch.argmax),
 * and log_softmax(logits)[argmax], per row. */
int acai_argmax_logprob(const void* logits, int32_t dtype, int64_t ld, int32_t rows, int32_t vocab,
                        int64_t* next_token, float* chosen_logprob, void* stream);

/* Packed rows -> padded (B, L_max, dim) with `fill` (per-feature vector, or NULL = 0) on the
 * padding rows: the `to_padded_tensor` half of the API (models.py:55-61) and the eval fast
 * path's LayerNorm(0) = norm.bias padding rows. */
int acai_unpack_padded(const void* packed, int32_t in_dtype, const int32_t* cu_seqlens, int32_t batch,
                       int32_t max_len, int32_t dim, const float* fill,
                       void* out, int32_t out_dtype, void* stream);
/* Padded (B, L_max, dim) -> packed rows (drops padding rows). */
int acai_pack_padded(const void* padded, int32_t in_dtype, const int32_t* cu_seqlens, int32_t batch,
                     int32_t max_len, int32_t dim, void* out, int32_t out_dtype, void* stream);

/* ---- MAE glue ----------------------------------------------------------------------------------------------------
 * acai_patchify: nn.Unfold(P, P) patches of a ragged image list as packed rows [total, C*P*P] (the reconstruction
 *   targets, models.py:265-267).
 * acai_mae_restore: MAE.prepare_for_decoder (models.py:219-241) on packed rows: row i of sequence b is
 *   latent[cu_keep[b] + r] if r = ids_restore[cu_full[b] + i] < keep_b else mask_token, plus the decoder position
 *   embedding of patch i (grid row i / patches_per_row[b], column i % patches_per_row[b]).
 * acai_mae_loss_fwd: MAELoss (models.py:271-288): row_loss[r] = mean_f (pred - (target - mean_r) / sqrt(var_r + 1e-6))^2
 *   with the unbiased per-patch variance, for rows with loss_mask != 0; sums = {sum of those, their count} reduced in a
 *   fixed order (the loss is sums[0] / sums[1]). */
int acai_patchify(const acai_image_desc* images, int32_t num_images, const int32_t* cu_seqlens, int32_t total_tokens,
                  int32_t patch_size, int32_t channels, void* out, int32_t out_dtype, void* stream);
int acai_mae_restore(const void* latent, int32_t latent_dtype, const int32_t* cu_keep, const int32_t* cu_full,
                     const int64_t* ids_restore, const int32_t* patches_per_row, int32_t batch, int32_t total_rows,
                     const float* mask_token, const float* dec_pos_embedding, int32_t pe_w,
                     void* out, int32_t out_dtype, int32_t dim, void* stream);
int acai_mae_loss_fwd(const void* pred, int32_t pred_dtype, const float* target, const uint8_t* loss_mask,
                      int32_t rows, int32_t feats, float* row_loss, float* sums, void* stream);

/* ---- training step: backward building blocks --------------------------------------------------------------------
 * The reference trains through torch autograd (loss.backward(), omr_teacher_force_train.py:120, pre_train.py:60); these
 * are the kernels behind the autograd Functions of acai_omr_b200/autograd.py.  Linear backward = two acai_linear_fwd
 * calls on transposed operands (dX = dY . W,  dW = dY^T . X) + acai_colsum for the bias.
 *   acai_transpose : out[c, r] = in[r, c], zero-padded up to ld_out columns (K % 64 for the tensor-core GEMM)
 *   acai_colsum    : out[c] (+)= sum_r x[r, c] in a fixed order (bias gradients)
 *   acai_gelu_fwd / acai_gelu_bwd : exact-erf GELU on a saved pre-activation (training keeps z)
 *   acai_layernorm_bwd : dv for v = x + residual, dgamma / dbeta (+)= column reductions (fixed order)
 *   acai_ce_loss   : OMRCELoss (models.py:784-796 = nn.CrossEntropyLoss(ignore_index, label_smoothing)): with
 *                    q = (1 - label_smoothing) onehot + label_smoothing / vocab, row_loss = -sum_c q_c log softmax_c,
 *                    valid, sums = {sum, count} and/or dlogits = *scale * (softmax - q) for non-ignored rows
 *   acai_scatter_add_rows / acai_gather_rows : embedding, position-table and MAE-restore gradients
 *   acai_mae_loss_bwd : dpred = mask * *scale * 2 (pred - normalised target) / feats */
int acai_transpose(const void* in, int32_t dtype, int64_t ld_in, int32_t rows, int32_t cols, void* out, int64_t ld_out,
                   void* stream);
size_t acai_colsum_workspace_bytes(int32_t rows, int32_t cols);
int acai_colsum(const void* x, int32_t dtype, int64_t ld, int32_t rows, int32_t cols, float* out, int32_t accumulate,
                void* workspace, size_t workspace_bytes, void* stream);
/* acai_colsum_static: acai_colsum without a workspace argument.  NO COUNTERPART IN THE REFERENCE - added
 * for callers that cannot allocate scratch memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  The slab partials live in a file-scope `static __device__` array of
 * acai_colsum_static_max_workspace_bytes() = 4 MiB (every GPU that loads the library holds it until the process exits);
 * the entry point calls no cudaMalloc / cudaFree / allocating library routine.  Limit: acai_colsum_workspace_bytes(rows,
 * cols) = ceil(rows / 128) * cols * 4 must not exceed 4 MiB (e.g. 65 536 rows x 2 048 columns); larger problems return
 * ACAI_ERR_INVALID with the sizes in acai_last_error().  Same kernels, same fixed summation order, bit-identical results
 * to acai_colsum.  Every call shares the one array per device (it keeps the previous call's partials; nothing needs to
 * start at zero): calls that overlap on the GPU - different streams, or concurrent graph branches - RACE on it and must
 * be serialised by the caller.  The device current at the call must be the one `x` / `out` live on. */
size_t acai_colsum_static_max_workspace_bytes(void);
int acai_colsum_static(const void* x, int32_t dtype, int64_t ld, int32_t rows, int32_t cols, float* out, int32_t accumulate,
                       void* stream);
int acai_gelu_fwd(const void* z, int32_t z_dtype, void* out, int32_t out_dtype, int64_t n, float dropout_p,
                  uint64_t dropout_seed, void* stream);     /* out = dropout(gelu(z)): the feed-forward block's nn.Dropout fused */
int acai_gelu_bwd(const void* dy, int32_t dy_dtype, const void* z, int32_t z_dtype, void* dz, int32_t dz_dtype, int64_t n,
                  float dropout_p, uint64_t dropout_seed, void* stream);
size_t acai_layernorm_bwd_workspace_bytes(int32_t rows, int32_t cols);
int acai_layernorm_bwd(const void* dy, int32_t dy_dtype, const void* x, int32_t x_dtype, const void* residual,
                       int32_t res_dtype, const float* gamma, float eps, void* dv, int32_t dv_dtype,
                       float* dgamma, float* dbeta, int32_t accumulate, int32_t rows, int32_t cols,
                       void* workspace, size_t workspace_bytes, void* stream);
int acai_ce_loss(const void* logits, int32_t dtype, int64_t ld, const int64_t* targets, int32_t rows, int32_t vocab,
                 int32_t ignore_index, float* row_loss, uint8_t* valid, float* sums,
                 void* dlogits, int32_t dlogits_dtype, int64_t ldd, const float* scale, float label_smoothing, void* stream);
int acai_scatter_add_rows(const void* src, int32_t dtype, int64_t ld, const int64_t* index, int32_t rows, int32_t dim,
                          float* dst, float scale, void* stream);
int acai_gather_rows(const void* src, int32_t dtype, const int64_t* index, int32_t rows, int32_t dim, void* out,
                     int32_t out_dtype, void* stream);
int acai_mae_loss_bwd(const void* pred, int32_t pred_dtype, const float* target, const uint8_t* loss_mask, int32_t rows,
                      int32_t feats, const float* scale, void* dpred, int32_t dpred_dtype, void* stream);

/* Attention-probability dropout (nn.MultiheadAttention(dropout=p) in training mode): acai_attn_varlen_fwd / _bwd take
 * (dropout_p, dropout_seed); P is masked and scaled by 1/(1-p) before P V, the softmax normaliser is unaffected.  The
 * mask is a pure function of (seed, head, global query row, key position) that forward and backward regenerate;
 * acai_attn_dropout_mask writes it out (uint8 keep[heads, total_q, max_k]) for tests. */
int acai_attn_dropout_mask(float dropout_p, uint64_t dropout_seed, int32_t heads, int32_t total_q, int32_t max_k,
                           uint8_t* keep, void* stream);

/* Dropout on hidden states (nn.Dropout inside nn.TransformerEncoder/DecoderLayer and the transition head, training mode):
 *   out[i] = residual[i] + (keep(seed, i) ? x[i] / (1 - p) : 0)      (residual may be NULL)
 * keep() is a counter-based hash of (seed, element index) - torch's Philox stream cannot be reproduced, the keep rate
 * and the 1/(1-p) scaling are nn.Dropout's.  The backward is the same call on dY with the same seed; no mask is stored.
 * acai_dropout_mask writes the keep decisions (uint8) for tests. */
int acai_dropout(const void* x, int32_t dtype, const void* residual, int32_t res_dtype, int64_t n, float p, uint64_t seed,
                 void* out, int32_t out_dtype, void* stream);
int acai_dropout_mask(int64_t n, float p, uint64_t seed, uint8_t* keep, void* stream);

/* Scheduled sampling (ScheduledSamplingViTOMR.sample_and_mix_seqs, models.py:801-817).
 *   acai_gumbel_softmax_fwd : F.gumbel_softmax(logits, tau, hard) with caller-supplied Gumbel noise [rows, vocab] fp32;
 *                             writes y_soft (fp32, kept for the backward) and y (one-hot of the first maximum when hard)
 *   acai_gumbel_softmax_bwd : dlogits = (1/tau) y_soft (dy - <y_soft, dy>)   (straight-through for hard)
 *   acai_mix_rows           : out[r] = (sel && sel[r] >= 0 ? a[sel[r]] : b[r]) + (pos ? pos[r % period] : 0) - the
 *                             torch.where / cat / right-shift of :809-816 and the position add of OMRDecoder.forward :469-470 */
int acai_gumbel_softmax_fwd(const void* logits, int32_t dtype, int64_t ld, const float* noise, int32_t rows, int32_t vocab,
                            float tau, int32_t hard, float* y_soft, void* y, int32_t y_dtype, void* stream);
int acai_gumbel_softmax_bwd(const void* dy, int32_t dy_dtype, const float* y_soft, int32_t rows, int32_t vocab, float tau,
                            void* dlogits, int32_t dlogits_dtype, void* stream);
int acai_mix_rows(const void* a, int32_t a_dtype, const int64_t* sel, const void* b, int32_t b_dtype, const float* pos,
                  int32_t period, int32_t rows, int32_t dim, void* out, int32_t out_dtype, void* stream);

/* out = residual + dropout(a @ w^T + bias): output projection + nn.Dropout + residual add of a post-norm sub-layer
 * (reference: self.dropout1(self.self_attn(...)) + x and self.dropout2(self.linear2(...)) + x inside
 * nn.TransformerEncoderLayer / nn.TransformerDecoderLayer, which the reference's models instantiate) in one GEMM.  bf16
 * operands and output, dense [M, N] output (ldo == N): the mask is acai_dropout's on the same tensor (counter = row * N +
 * column), so acai_dropout(dy, dropout_p, dropout_seed) is its backward.  residual may be NULL.  Returns
 * ACAI_ERR_UNSUPPORTED, nothing written, when the problem does not qualify (then: acai_linear_fwd + acai_dropout). */
int acai_linear_dropout(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* residual,
                        int32_t res_dtype, int64_t ldr, void* out, int64_t ldo, int32_t M, int32_t N, int32_t K, float dropout_p,
                        uint64_t dropout_seed, void* stream);
/* Feed-forward block, training (models.py:30-34 / nn.TransformerEncoderLayer's linear1 -> gelu -> dropout): one tcgen05
 * GEMM whose epilogue stores the pre-activation z = a @ w^T + bias (bf16 [M, N], kept for the backward) AND
 * h = dropout(gelu(z)) (bf16 [M, N]); the dropout mask is that of acai_gelu_fwd on a dense [M, N] tensor with the same
 * (dropout_p, dropout_seed).  Returns ACAI_ERR_UNSUPPORTED (nothing written) when the operands do not qualify for the
 * TMA-store epilogue; the caller then runs acai_linear_fwd + acai_gelu_fwd. */
int acai_linear_gelu_train(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* z, int64_t ldz,
                           void* h, int64_t ldh, int32_t M, int32_t N, int32_t K, float dropout_p, uint64_t dropout_seed,
                           void* stream);

/* Backward of the same block through the activation: dz = (dy @ wt^T) * gelu'(z) * dropout_mask (bf16 [M, N]), with
 * dy [M, K] bf16, wt [N, K] bf16 (the transposed down-projection weight) and z the tensor acai_linear_gelu_train kept.
 * Same fallback contract (acai_linear_fwd + acai_gelu_bwd). */
int acai_linear_gelu_bwd(const void* dy, int64_t ldd, const void* wt, int64_t ldw, const void* z, int64_t ldz, void* dz,
                         int64_t ldo, int32_t M, int32_t N, int32_t K, float dropout_p, uint64_t dropout_seed, void* stream);

/* Input gradient of a linear layer from the weight AS THE FORWARD HOLDS IT:  dx[M, N] = dz[M, K] . w[K, N]  with w the
 * layer's [out_features = K, in_features = N] bf16 operand copy (autograd of F.linear inside every nn.Transformer*Layer,
 * omr_teacher_force_train.py:120) - the contraction index is w's ROW, which tcgen05 reads as an "MN-major" B operand, so no
 * transposed weight copy is made per step (acai_transpose + acai_linear_fwd on w^T was 139 launches of the teacher-forced
 * step).  Epilogue: + residual (bf16 [M, N], optional: the residual-stream gradient of the block), or, with z (bf16 [M, N],
 * the pre-activation acai_linear_gelu_train kept), x gelu'(z) x the dropout mask of (dropout_p, dropout_seed) - the
 * acai_linear_gelu_bwd epilogue; z and residual are exclusive.  bf16 in / out.  ACAI_ERR_UNSUPPORTED (nothing launched)
 * unless K % 64 == 0, N >= 128, N % 8 == 0 and every operand has 16-byte aligned rows; the caller then transposes. */
int acai_linear_dgrad(const void* dz, int64_t ldz, const void* w, int64_t ldw, const void* residual, int64_t ldr, const void* z,
                      int64_t ldzz, void* dx, int64_t ldo, int32_t M, int32_t N, int32_t K, float dropout_p, uint64_t dropout_seed,
                      void* stream);

/* Weight gradient of a linear layer, accumulated in place:  dw[n_out, k_in] += dz[tokens, n_out]^T . x[tokens, k_in]
 * (bf16 operands exactly as the forward/backward left them - no transposed copies; fp32 dw, typically the parameter's
 * slice of the flat gradient buffer).  tcgen05 GEMM with both operands MN-major, split over token ranges, partial
 * tiles combined by TMA reduce-add.  Returns ACAI_ERR_UNSUPPORTED (nothing written) for non-bf16 operands or rows that
 * are not 16-byte aligned; the caller then uses acai_transpose + acai_linear_fwd. */
int acai_wgrad_accumulate(const void* dz, int64_t ldz, const void* x, int64_t ldx, int32_t dtype, float* dw, int64_t ldw,
                          int32_t n_out, int32_t k_in, int32_t tokens, void* stream);

/* The same with the BIAS gradient from the same pass:  dbias[n_out] += column sums of dz (fp32, accumulated with atomic
 * adds: like the split-K reduce-adds of dw, not in a fixed order).  Autograd of nn.Linear's bias inside every
 * nn.Transformer*Layer of the reference's training step (omr_teacher_force_train.py:120, pre_train.py:57): the dz boxes
 * the weight-gradient pipeline stages in shared memory are summed there by four extra warps, so dz is not read a second
 * time by acai_colsum.  Same ACAI_ERR_UNSUPPORTED contract (nothing written; the caller runs acai_wgrad_accumulate's
 * fallback and acai_colsum). */
int acai_wgrad_bias_accumulate(const void* dz, int64_t ldz, const void* x, int64_t ldx, int32_t dtype, float* dw, int64_t ldw,
                               float* dbias, int32_t n_out, int32_t k_in, int32_t tokens, void* stream);

/* One AdamW update (torch.optim.AdamW semantics; omr_teacher_force_train.py:207, mae_pre_train.py) of a flat fp32
 * parameter segment, `grad_scale` applied to the gradient first (1/world_size after a summed all-reduce).
 * `param_bf16` (may be NULL) receives the bf16 copy of the updated parameters that the tensor-core GEMMs read. */
int acai_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, void* param_bf16, int64_t n,
                    float lr, float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale,
                    void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ACAI_B200_H_ */
