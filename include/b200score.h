/* b200score.h -- C ABI of libb200score.so: B200 (sm_100a) relative-probability scorer + bootstrap.
 *
 * The reference (jonathanhchoi/llm-interpretation-replication) is pure Python and has no FFI of
 * its own; its de-facto boundary for this path is two Python callables.  Each entry
 * point below names the reference code it replaces.  All pointers are borrowed: the caller (torch
 * in our host layer) owns every allocation; the library allocates nothing on the device unless
 * stated, never frees caller memory, and keeps only the weight pointers registered at
 * b2s_model_create (they must outlive the handle).  Every function returns 0 on success or a
 * negative B2S_E* code and records a thread-local message for b2s_last_error(); nothing throws
 * across the boundary.  All device work is enqueued on the caller's stream (a cudaStream_t passed
 * as void*); functions return after enqueueing unless they say they synchronise.
 */
#ifndef B200SCORE_H
#define B200SCORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2S_ABI_VERSION 4   /* 4: BF16_SPLIT mode, B2S_ERANGE, compacted look-ahead (gen_ids -1 at positions a prompt did not need) */

#if defined(__GNUC__)
#define B2S_API __attribute__((visibility("default")))
#else
#define B2S_API
#endif

enum { B2S_OK = 0, B2S_EINVAL = -1, B2S_ECUDA = -2, B2S_ENOMEM = -3, B2S_EUNSUPPORTED = -4,
       B2S_ERANGE = -5 /* results were produced, but some rows are not finite (see b2s_score_host) */ };

enum { B2S_ARCH_FALCON = 0, B2S_ARCH_GPT_NEOX = 1, B2S_ARCH_LLAMA = 2, B2S_ARCH_OPT = 3, B2S_ARCH_BLOOM = 4 };
/* position encoding: rotary; learned table added to the token embedding (HF opt, rows pos + pos_offset); ALiBi
 * (HF bloom: slope[head] * key position added to the scaled scores) */
enum { B2S_POS_ROPE = 0, B2S_POS_LEARNED = 1, B2S_POS_ALIBI = 2 };
enum { B2S_ACT_GELU = 0, B2S_ACT_RELU = 1, B2S_ACT_GELU_TANH = 2 };   /* MLP activation when mlp_type == B2S_MLP_GELU */
enum { B2S_NORM_LAYERNORM = 0, B2S_NORM_RMS = 1 };   /* RMSNorm: x * rsqrt(mean(x^2) + eps) * g, *_B slots null (HF llama LlamaRMSNorm) */
enum { B2S_MLP_GELU = 0, B2S_MLP_SWIGLU = 1 /* SwiGLU: down(silu(gate(x)) * up(x)); W_UP slot = [2*d_ff, d_model], gate rows first */,
       B2S_MLP_SWIGLU_FUSED = 2 /* same MLP, W_UP rows interleaved (row 2i = gate_proj row i, row 2i+1 = up_proj row i): the
                                   tensor-core modes form silu(gate)*up in the projection's epilogue, no elementwise pass */ };
enum { B2S_QKV_FALCON_MQA = 0 /* [q heads..., k heads..., v heads...]: Falcon MQA, cat(q,k,v)_proj of the Llama family */,
       B2S_QKV_NEOX_INTERLEAVED = 1 /* per head [q, k, v] */ };
/* compute modes (tensor-core modes: fp32 accumulation in TMEM, fp32 residual stream, LayerNorm and softmax):
 *   BF16_FAST   bf16 weight matrices; activations (every GEMM A operand, Q/K/V/P) rounded to bf16.  Fastest; at 32
 *               layers its |delta relative_prob| sits at the bf16 noise floor (max ~3e-3: a share of the prompts misses
 *               the 2e-3 gate);
 *   FP16        fp16 weight matrices (bf16 checkpoints convert exactly for |w| >= 2^-17) and activations rounded
 *               to fp16: 3 more mantissa bits at the same MMA rate (tcgen05 kind::f16 needs A and B alike, a
 *               mixed fp16 x bf16 MMA is an illegal instruction); holds the 2e-3 gate at 32 layers.  fp16 overflows
 *               beyond 65504: such prompts come back flagged (B2S_ERANGE, NaN rows), never as numbers;
 *   BF16_SPLIT  bf16 weight matrices untouched; every GEMM A operand is kept as two bf16 halves hi = bf16(x),
 *               lo = bf16(x - hi) (16 mantissa bits, bf16 range) and both halves meet the same weights in the same fp32
 *               accumulator: holds the gate with bf16's range in every linear layer at twice the tensor work.  Attention has
 *               no split form: q/k/v, the KV cache and P are fp16 there (an overflow surfaces as B2S_ERANGE like in the
 *               FP16 mode; q/k/v are LayerNorm-fed projections, not where decoder outliers live).  Needs d_model, d_ff
 *               and n_head*head_dim to be multiples of 64;
 *   FP32_CHECK  fp32 weights and activations on CUDA cores (parity gate 1e-5);
 *   INT8_EMU    the reference's own GPU configuration, bitsandbytes LLM.int8 (run_base_vs_instruct_100q.py:84,433-437
 *               USE_8BIT / BitsAndBytesConfig(load_in_8bit)), emulated on the FP16 kernels: the host hands over every
 *               layer matrix already on its row-wise int8 grid (fp16(rint(W * 127 / s_row) * s_row / 127); embeddings and
 *               lm_head untouched, as bitsandbytes leaves them), and the library puts the input rows of every linear layer
 *               on theirs first -- per prompt, columns holding an |x| >= 6.0 stay fp16 (the mixed-precision decomposition),
 *               every other element becomes rint(x * 127 / s) * s / 127 with s the row's largest |x| below the threshold.
 *               fp16 x fp16 -> fp32 then multiplies the same numbers as int8 x int8 -> int32 followed by the dequantisation,
 *               up to the fp16 rounding of the dequantised operands (2^-11, 16 times finer than the int8 step).  Decoder-only
 *               handles only.  bitsandbytes is not in this image: the mode is pinned against a restatement of its published
 *               MatMul8bitLt (oracle/int8_ref.py), not against the library itself;
 *   FP16_FFN_SPLIT (encoder-decoder handles only) FP16 everywhere except the one place where the T5 family is known to leave
 *               fp16's range, the activation that feeds the feed-forward's second projection (HF keeps that layer, `wo`, in fp32
 *               for fp16 inference: t5 `_keep_in_fp32_modules`): the first projection's epilogue forms act(wi_0 x) * wi_1 x on the
 *               fp32 accumulators and stores it as bf16 hi + lo halves, and the W_DOWN matrices are handed over in bf16 and
 *               meet both halves (as in BF16_SPLIT).  The residual stream, which receives that projection's output, is fp32
 *               in every mode.  Costs the doubled K of one GEMM in four instead of all of them;
 *   BF16_SPLIT_ATTN32 (decoder-only handles) BF16_SPLIT with q/k/v, the KV cache and all of attention in fp32 on the CUDA-core
 *               kernels: nothing between two fp32 points is rounded below 16 mantissa bits.  For models whose scores are more
 *               sensitive to operand rounding than the BASELINE configs -- the Llama family at 32 layers leaves the 2e-3 gate on
 *               a fifth of the prompts in every 16-bit-attention mode, in the torch emulation of the roundings as on the GPU
 *                -- at the price of the slow attention path. */
enum { B2S_MODE_BF16_FAST = 0, B2S_MODE_FP16 = 1, B2S_MODE_BF16_SPLIT = 2, B2S_MODE_FP32_CHECK = 3, B2S_MODE_INT8_EMU = 4,
       B2S_MODE_FP16_FFN_SPLIT = 5, B2S_MODE_BF16_SPLIT_ATTN32 = 6 };

/* weight_ptrs slots: 10 global + 16 per layer (see llm_interpretation_replication_b200/weights.py); every non-null
 * pointer must be 16-byte aligned (B2S_EINVAL otherwise) */
enum { B2S_W_WTE = 0, B2S_W_LNF_G = 1, B2S_W_LNF_B = 2, B2S_W_LM_HEAD = 3, B2S_W_ROPE_COS = 4,
       B2S_W_ROPE_SIN = 5, B2S_W_WPE = 6 /* [max_pos + pos_offset, d] */, B2S_W_EMB_LN_G = 7, B2S_W_EMB_LN_B = 8,
       B2S_W_ALIBI = 9 /* fp32 [n_head] */, B2S_W_NGLOBAL = 10, B2S_W_PER_LAYER = 16 };
enum { B2S_WL_LN1_G = 0, B2S_WL_LN1_B, B2S_WL_LN2_G, B2S_WL_LN2_B, B2S_WL_W_QKV, B2S_WL_B_QKV,
       B2S_WL_W_DENSE, B2S_WL_B_DENSE, B2S_WL_W_UP, B2S_WL_B_UP, B2S_WL_W_DOWN, B2S_WL_B_DOWN };

typedef struct {
  int32_t abi_version;      /* B2S_ABI_VERSION */
  int32_t arch;             /* B2S_ARCH_* */
  int32_t n_layer, d_model, n_head, n_kv_head, head_dim, d_ff, vocab, rotary_dim;
  float   norm_eps;
  int32_t parallel_residual;   /* HF gpt_neox use_parallel_residual / falcon parallel_attn */
  int32_t n_ln_per_layer;      /* 1 (falcon-7b) or 2 */
  int32_t has_bias;            /* 0 none; 1 every linear layer (B_QKV, B_DENSE, B_UP, B_DOWN); 2 the q|k|v projection only (Qwen) */
  int32_t qkv_layout;          /* B2S_QKV_* */
  int32_t max_pos;             /* rows of the rope cos/sin tables */
  int32_t compute_mode;        /* B2S_MODE_*: BF16_FAST and BF16_SPLIT expect bf16 matrices, FP16 fp16, FP32_CHECK fp32 */
  int32_t norm_type;           /* B2S_NORM_*  (ABI 2) */
  int32_t mlp_type;            /* B2S_MLP_*   (ABI 2) */
  int32_t pos_type;            /* B2S_POS_*   (ABI 3) */
  int32_t pos_offset;          /* learned positions: table row of position 0 (OPT: 2) */
  int32_t act_type;            /* B2S_ACT_*   (ABI 3) */
  int32_t emb_norm;            /* LayerNorm on the token embeddings with B2S_W_EMB_LN_* (BLOOM) */
} b2s_model_desc;

typedef struct b2s_model b2s_model;

B2S_API const char* b2s_last_error(void);
B2S_API int         b2s_abi_version(void);

/* ---- scorer ------------------------------------------------------------------------------
 * Replaces, for a whole batch of prompts at once, the per-prompt body of the reference's
 * get_yes_no_logprobs (analysis/run_base_vs_instruct_100q.py:329-363; compare_instruct_models.py:
 * 228-272; compare_base_vs_instruct.py:242-286): HF model.generate(...) prefill + greedy look-ahead
 * (HF generation/utils.py:2736-2805), softmax, top-k membership test of the two candidate ids over
 * the first max_lookahead positions, fallback to position 0. */
B2S_API int    b2s_model_create(const b2s_model_desc* desc, const void* const* weight_ptrs, int32_t n_ptrs,
                        int32_t device, b2s_model** out);
B2S_API void   b2s_model_destroy(b2s_model* m);
B2S_API size_t b2s_workspace_bytes(const b2s_model* m, int64_t max_tokens, int32_t max_seqs, int32_t max_lookahead);

/* out rows: [p_a, p_b, rel, lse, logit_a, logit_b, pos_found, found]; gen_ids[N,max_lookahead]
 * holds the greedy tokens (-1 after EOS, and at positions a prompt did not need).  topk is 5 (100q script) or 2
 * (compare_* scripts).  tokens/cu_seqlens/cand_ids/out/gen_ids are DEVICE pointers.  Position 0 (the prompt pass) is
 * only enqueued.  Every later look-ahead position costs one 4-byte read-back and a stream synchronisation: the number
 * of prompts that still need a position sizes that step's launches -- a one-token step runs only the unresolved,
 * still-alive prompts (in completion mode: every prompt that has not produced EOS), gathered into consecutive rows --
 * and the call returns early once none is left.  A prompt's row does not depend on which other prompts share its
 * batch or its step.
 * Range check: a prompt whose logits are not finite at a position it is scored on (in the FP16 mode an activation
 * beyond 65504 overflows to infinity) gets p_a = p_b = rel = lse = NaN and found = -1, and stops taking steps; it is
 * never reported as a number.  b2s_score leaves it at that; the _host calls return B2S_ERANGE after copying all
 * rows back. */
B2S_API int b2s_score(b2s_model* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs,
              int32_t total_tokens, int32_t max_seqlen, const int32_t* cand_ids /*[N,2]*/,
              int32_t topk, int32_t max_lookahead, int32_t eos_id,
              float* out /*[N,8]*/, int32_t* gen_ids /*[N,max_lookahead] or NULL*/,
              void* workspace, size_t workspace_bytes, void* stream);

/* Same call with HOST buffers (pinned or pageable): copies inputs host->device into the workspace,
 * runs b2s_score, copies the result rows back and synchronises the stream.  This is the call the
 * reference-side binding makes and the one bench.py times as "e2e". */
B2S_API int b2s_score_host(b2s_model* m, const int32_t* tokens_h, const int32_t* cu_seqlens_h, int32_t n_seqs,
                   const int32_t* cand_ids_h, int32_t topk, int32_t max_lookahead, int32_t eos_id,
                   float* out_h, int32_t* gen_ids_h /*or NULL*/,
                   void* workspace, size_t workspace_bytes, void* stream);

/* Score + complete ("next" row N1 of SURVEY.md 8f): the reference always generates max_new_tokens=50 and scans only
 * the first 10 positions (run_base_vs_instruct_100q.py:338-346,348-363,365-368; compare_instruct_models.py:274,290).
 * Same as b2s_score / b2s_score_host, except that greedy decoding continues to gen_tokens (>= scan_positions, <= 64)
 * for every prompt that has not produced eos_id, whether or not it is already resolved; the top-k membership scan
 * still covers positions < scan_positions only.  gen_ids is [N, gen_tokens] (-1 after EOS); the workspace must be
 * sized with b2s_workspace_bytes(..., max_lookahead = gen_tokens).  b2s_score(max_lookahead) ==
 * b2s_score_complete(scan_positions = gen_tokens = max_lookahead). */
B2S_API int b2s_score_complete(b2s_model* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs,
              int32_t total_tokens, int32_t max_seqlen, const int32_t* cand_ids /*[N,2]*/,
              int32_t topk, int32_t scan_positions, int32_t gen_tokens, int32_t eos_id,
              float* out /*[N,8]*/, int32_t* gen_ids /*[N,gen_tokens] or NULL*/,
              void* workspace, size_t workspace_bytes, void* stream);
B2S_API int b2s_score_complete_host(b2s_model* m, const int32_t* tokens_h, const int32_t* cu_seqlens_h, int32_t n_seqs,
                   const int32_t* cand_ids_h, int32_t topk, int32_t scan_positions, int32_t gen_tokens, int32_t eos_id,
                   float* out_h, int32_t* gen_ids_h /*or NULL*/,
                   void* workspace, size_t workspace_bytes, void* stream);

/* ---- encoder-decoder scorer (T5 family) -----------------------------------------------------------
 * Replaces the `is_encoder_decoder` branch of the same reference functions (compare_instruct_models.py:178-226,
 * run_base_vs_instruct_100q.py:290-325, compare_base_vs_instruct.py:192-237: tk-instruct-3b-def, T0_3B, t5-v1_1-base,
 * flan-t5-base): AutoModelForSeq2SeqLM.generate(input_ids, max_new_tokens, output_scores) = one encoder pass over the
 * prompt (HF t5 T5Stack, bidirectional self-attention with the bucketed relative position bias of the first layer
 * shared by all layers, no 1/sqrt(d) scaling, T5LayerNorm = RMSNorm, no biases), then greedy decoder steps from
 * decoder_start_id (self-attention over the generated prefix with the decoder's relative bias, cross-attention over
 * the encoder output without bias, ReLU or gated tanh-GELU feed-forward), the LM head on every step, and the same
 * top-k look-ahead scan and position-0 fallback as the decoder-only branch.
 *
 * weight_ptrs: B2S_T5_NGLOBAL global slots, then n_enc_layer x B2S_T5_ENC_PER_LAYER, then n_dec_layer x
 * B2S_T5_DEC_PER_LAYER.  Matrices [out, in] row-major in the mode's weight type, vectors and bias tables fp32.
 *   ENC_BIAS / DEC_BIAS: fp32 [n_head, 2*max_pos - 1]; entry [h][(key position - query position) + max_pos - 1] =
 *   relative_attention_bias.weight[bucket(key - query)][h] with HF's T5Attention._relative_position_bucket
 *   (bidirectional for the encoder, not for the decoder).  The host expands the bucket table once per model, the way
 *   it builds the rotary tables of the decoder-only families (llm_interpretation_replication_b200/t5.py bias_table).
 *   W_QKV = rows of q, then k, then v ([3*n_head*head_dim, d_model]); W_CKV = cross-attention k rows then v rows.
 *   W_UP: ff_type 0 = wi [d_ff, d_model] with ReLU; 1 = gated, wi_0 (tanh-GELU gate) rows then wi_1 rows
 *   [2*d_ff, d_model]; 2 = gated with the rows interleaved (2i = wi_0 row i, 2i+1 = wi_1 row i; tensor-core modes:
 *   the gate is formed in the projection's epilogue).
 * A model with tied embeddings (original t5-*: logits = lm_head(hidden * d_model^-0.5)) is given with DEC_LNF_G
 * pre-multiplied by d_model^-0.5; the v1.1 / Flan / LM-adapted checkpoints of the reference's lists are untied. */
enum { B2S_T5_WTE = 0, B2S_T5_ENC_LNF_G = 1, B2S_T5_DEC_LNF_G = 2, B2S_T5_LM_HEAD = 3, B2S_T5_ENC_BIAS = 4, B2S_T5_DEC_BIAS = 5,
       B2S_T5_NGLOBAL = 8, B2S_T5_ENC_PER_LAYER = 8, B2S_T5_DEC_PER_LAYER = 12 };
enum { B2S_T5E_LN1_G = 0, B2S_T5E_W_QKV, B2S_T5E_W_O, B2S_T5E_LN2_G, B2S_T5E_W_UP, B2S_T5E_W_DOWN };
enum { B2S_T5D_LN1_G = 0, B2S_T5D_W_QKV, B2S_T5D_W_O, B2S_T5D_LN2_G, B2S_T5D_W_CQ, B2S_T5D_W_CKV, B2S_T5D_W_CO, B2S_T5D_LN3_G,
       B2S_T5D_W_UP, B2S_T5D_W_DOWN };
enum { B2S_T5_FF_RELU = 0, B2S_T5_FF_GATED_GELU = 1, B2S_T5_FF_GATED_GELU_FUSED = 2 };

typedef struct {
  int32_t abi_version;       /* B2S_ABI_VERSION */
  int32_t n_enc_layer, n_dec_layer, d_model, n_head, head_dim /* d_kv: 64 */, d_ff, vocab;
  float   norm_eps;
  int32_t ff_type;           /* B2S_T5_FF_* */
  int32_t max_pos;           /* longest prompt / generated prefix the bias tables cover */
  int32_t compute_mode;      /* B2S_MODE_* (same meaning as for the decoder-only handle) */
  int32_t decoder_start_id;  /* config.decoder_start_token_id (T5: the pad id, 0) */
} b2s_encdec_desc;

typedef struct b2s_encdec b2s_encdec;

B2S_API int    b2s_encdec_create(const b2s_encdec_desc* desc, const void* const* weight_ptrs, int32_t n_ptrs, int32_t device,
                         b2s_encdec** out);
B2S_API void   b2s_encdec_destroy(b2s_encdec* m);
/* The cross-attention keys / values of every decoder layer stay resident for the whole call:
 * max_tokens * n_dec_layer * 2 * n_head * head_dim elements dominate this figure (size the batch with it). */
B2S_API size_t b2s_encdec_workspace_bytes(const b2s_encdec* m, int64_t max_tokens, int32_t max_seqs, int32_t max_gen_tokens);
/* Same arguments, rows, compaction, range check and return codes as b2s_score_complete / b2s_score_complete_host.
 * tokens are the ENCODER input ids (the T5 tokenizer's trailing </s> included); gen_ids the decoder's greedy tokens
 * (the start id is not part of them). */
B2S_API int b2s_encdec_score(b2s_encdec* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs, int32_t total_tokens,
                     int32_t max_seqlen, const int32_t* cand_ids /*[N,2]*/, int32_t topk, int32_t scan_positions,
                     int32_t gen_tokens, int32_t eos_id, float* out /*[N,8]*/, int32_t* gen_ids /*[N,gen_tokens] or NULL*/,
                     void* workspace, size_t workspace_bytes, void* stream);
B2S_API int b2s_encdec_score_host(b2s_encdec* m, const int32_t* tokens_h, const int32_t* cu_seqlens_h, int32_t n_seqs,
                          const int32_t* cand_ids_h, int32_t topk, int32_t scan_positions, int32_t gen_tokens, int32_t eos_id,
                          float* out_h, int32_t* gen_ids_h /*or NULL*/, void* workspace, size_t workspace_bytes, void* stream);

/* Building block of B2S_MODE_INT8_EMU: puts the fp16 rows x[rows, K] (device, in place) on bitsandbytes' LLM.int8 grid the way
 * MatMul8bitLt.forward treats the input of one Linear8bitLt call per prompt (prompt i = rows [q_start[i], q_start[i] + q_len[i]),
 * packed in row order): columns where any row of the prompt has |x| >= threshold stay fp16, every other element becomes
 * rint(x * 127 / s) * s / 127 with s = the row's largest |x| below the threshold.  mask_scratch: n_seqs * K bytes on the device.
 * threshold <= 0: no outlier columns. */
B2S_API int b2s_int8_fakequant_f16(void* x, int32_t rows, int32_t K, const int32_t* q_start, const int32_t* q_len, int32_t n_seqs,
                           void* mask_scratch, float threshold, void* stream);

/* Number of kernels this library launched since the last call with reset != 0 (bench "gpu_launches"). */
B2S_API int64_t b2s_launch_count(int32_t reset);

/* Measurement hook (no counterpart in the reference): when enabled, every tensor-core GEMM launch is
 * bracketed by CUDA events on its stream; _read() waits for them and returns the summed device time, the
 * number of launches and their algorithmic FLOPs (2*M*N*K) since the previous read. */
B2S_API int b2s_profile_gemm(int32_t enable);
B2S_API int b2s_profile_gemm_read(double* total_ms, int64_t* launches, double* flops);

/* Tuning / A-B knobs (no counterpart in the reference; process-wide; none of them changes a result bit except
 * "gemm_serpentine", which changes the fp32 summation order).  Defaults in brackets.
 *   "gemm_cta_group"     [0] 0 = per-launch policy; 1 = one CTA per 128x256 tile; 2 = CTA pairs (tcgen05
 *                        cta_group::2) on 256x256 tiles; 3 = CTA pairs on 512x256 tiles when M >= 4096; 4 = always
 *   "gemm_policy"        [1] policy used by gemm_cta_group 0: 1 = 512x256 pair tiles for K >= 8192 and M >= 4096, 256x256
 *                        pairs otherwise; 0 = the earlier policy (single-CTA tiles for long-K and GELU launches)
 *   "gemm_wave_sync"     [4] the producer threads of all CTAs meet at a global counter every N tiles so that CTAs sharing
 *                        an A or W panel walk K together and hit in L2 (0 = never; bounded wait, cannot dead-lock)
 *   "gemm_group_m"       [0] rasterisation group in 128-row blocks (0 = heuristic); "gemm_small_w_group" [8] group used
 *                        when W itself is L2-sized
 *   "gemm_band_n"        [-1] CTA-pair 256x256 kernel: tiles walked in column bands (all row blocks inside a band) so that a band of
 *                        W stays L2-resident; -1 = only for launches whose W is L2-sized (QKV, dense), k > 0 = bands of k tiles, 0 = off
 *   "gemm_prefetch"      [0] producers issue TMA L2 prefetches this many k-blocks ahead (measured: -11 % end to end; kept for A/B)
 *   "gemm_serpentine"    [0] K walked backwards by odd column tiles (1) or odd waves (2)
 *   "gemm_l2_hint"       [0] TMA L2 eviction hints 0..7; "gemm_tma_store" [1] epilogue through TMA bulk stores;
 *                        "gemm_pair_stride" [1] experiment: permutation of the CTA-pair -> tile-sequence assignment
 *   "attn_step"          [1] one-token attention of head_dim 80 / 128, ALiBi models and the fp32 mode on the streaming kernel
 *                        (0 = the varlen kernel that also runs their prompt pass)
 *   "int8_threshold_milli" [6000] INT8_EMU: bitsandbytes' llm_int8_threshold in thousandths (0 = no outlier decomposition)
 *   "lookahead_compact"  [1] one-token steps run only the prompts that still need a position (0 = every prompt of the
 *                        batch runs every step, the round-1 behaviour; rows are bit-identical either way)
 *   "attn_ts"            [1] prompt-pass / one-token attention with P kept in tensor memory (0 = first-generation kernels)
 *   "attn_hold_s"        [0] first-generation attention only: keep S in registers between the two softmax passes */
B2S_API int b2s_set_option(const char* name, int32_t value);

/* ---- building blocks (exported so the parity tests can exercise each kernel through the ABI) ----
 * C[M,N] = A[M,K] . W[N,K]^T (+bias[N]) with the epilogues the decoder layers need; replaces the
 * FalconLinear / nn.Linear calls (HF falcon:60-65, gpt_neox:41-42,200-201).
 *  epilogue: 0 store bf16, 1 GELU(erf) then store bf16, 2 out_f32 += acc (residual), 4 store fp32,
 *  6 SwiGLU: W rows alternate (gate_i, up_i), out is bf16 [M, N/2] = silu(acc[:, 2i]) * acc[:, 2i+1] (no bias) */
B2S_API int b2s_gemm_bf16(const void* A /*bf16 [M,K]*/, const void* W /*bf16 [N,K]*/, const float* bias,
                  void* out, int32_t M, int32_t N, int32_t K, int32_t epilogue, void* stream);
/* same as b2s_gemm_bf16 with fp16 operands (and fp16 16-bit outputs) */
B2S_API int b2s_gemm_f16(const void* A /*fp16 [M,K]*/, const void* W /*fp16 [N,K]*/, const float* bias,
                  void* out, int32_t M, int32_t N, int32_t K, int32_t epilogue, void* stream);
B2S_API int b2s_gemm_f32(const float* A, const float* W, const float* bias, float* out,
                 int32_t M, int32_t N, int32_t K, int32_t epilogue, void* stream);

/* Varlen causal attention over a KV cache (replaces F.scaled_dot_product_attention(..., is_causal=True),
 * HF falcon:349-362 / gpt_neox:162-185).  qkv bf16 [total_rows, qkv_dim] holds the (already rotated) queries of
 * head h at columns q_base + h*q_hstride; kcache/vcache bf16 [cache_slots, n_kv_head*head_dim]; sequence i has
 * q_len[i] query rows starting at q_start[i] and kv_len[i] keys starting at slot kv_start[i]; query j sees keys
 * [0, kv_len - q_len + j].  out bf16 [total_rows, n_head*head_dim].  use_tensor_cores != 0 selects the tcgen05
 * kernels (<= 256 keys; cache slots past a sequence must hold finite values; prompt pass: head_dim 64, 80 or 128;
 * max_qlen == 1: head_dim 64 with contiguous 64-wide query heads), 0 the CUDA-core varlen kernel.  (Inside b2s_score the
 * one-token steps of the other shapes run a third, streaming kernel: "attn_step" in b2s_set_option.) */
B2S_API int b2s_attention_bf16(const void* qkv, const void* kcache, const void* vcache, void* out,
                       const int32_t* q_start, const int32_t* q_len, const int32_t* kv_start, const int32_t* kv_len,
                       int32_t n_seqs, int32_t total_rows, int32_t cache_slots, int32_t max_qlen, int32_t max_kv,
                       int32_t n_head, int32_t n_kv_head, int32_t head_dim, int32_t qkv_dim, int32_t q_base,
                       int32_t q_hstride, int32_t use_tensor_cores, void* stream);

/* Last-position LM head fused with online log-sum-exp, running top-5, argmax and candidate gather
 * (replaces lm_head + .float() + softmax + topk + 2x .item(): HF falcon:905-906, generation/utils.py:
 * 2762; run_base_vs_instruct_100q.py:351-356).  hidden bf16 [M,d], W bf16 [V,d].
 * stats rows: [lse, top1..top5, argmax(as float), logit_a, logit_b, 0...] (16 floats per row).
 * scratch: b2s_lm_head_scratch_bytes(M, V) bytes of device memory (per-tile partials). */
B2S_API size_t b2s_lm_head_scratch_bytes(int32_t M, int32_t V);
B2S_API int b2s_lm_head_bf16(const void* hidden, const void* W, const int32_t* cand_ids /*[M,2]*/,
                     int32_t M, int32_t V, int32_t d, float* stats /*[M,16]*/,
                     void* scratch, size_t scratch_bytes, void* stream);

/* No counterpart in the reference.  Same computation and same stats layout as b2s_lm_head_bf16, for callers
 * that cannot allocate device memory: plug-ins living inside an application whose memory pool owns the whole
 * GPU, latency-sensitive loops that must avoid the device-wide synchronisation of cudaMalloc/cudaFree, and
 * code that is handed only input/output pointers by a foreign runtime.  The per-tile partial records go into
 * a fixed array compiled into the library (b2s_lm_head_static_capacity() bytes = 4 MiB + 4 KiB, reserved on
 * every GPU that loads the library until the process exits); the call performs no cudaMalloc/cudaFree and calls
 * no allocating library routine.  Limits and rules:
 *   - b2s_lm_head_scratch_bytes(M, V) must not exceed the capacity (e.g. M <= 512 at V = 65024); larger
 *     problems are rejected with B2S_EINVAL and a message, nothing is launched;
 *   - the first call on each device looks the array's device address up (the device must be current); make
 *     that first call before starting any stream capture on that device;
 *   - every call on a device shares the one array and finds the previous call's contents in it (the call
 *     clears the part that must start at zero itself): two calls whose GPU work can overlap -- different
 *     streams, or different host threads -- race on it and must be serialised by the caller. */
B2S_API size_t b2s_lm_head_static_capacity(void);
B2S_API int b2s_lm_head_bf16_static(const void* hidden, const void* W, const int32_t* cand_ids /*[M,2]*/,
                            int32_t M, int32_t V, int32_t d, float* stats /*[M,16]*/, void* stream);

/* ---- bootstrap ---------------------------------------------------------------------------
 * NumPy legacy RNG (MT19937 + masked rejection), bit-exact: np.random.seed(s) followed by repeated
 * np.random.choice(n, size=..., replace=True), as used at run_base_vs_instruct_100q.py:616,657;
 * analyze_base_vs_instruct_mae_100q.py:296,312,357,382; analyze_results_base_versus_instruct.py:304,
 * 312,349.  state = 624 words + position (625 uint32, device).
 * b2s_mt19937_choice takes its temporary (the raw word stream of the call, about 4 * count * (mask + 1) / n bytes,
 * at most ~0.6 GB per 2^26-draw segment) from a memory pool the library creates per device on first use
 * (cudaMallocFromPoolAsync / cudaFreeAsync on the caller's stream: stream-ordered, no device-wide synchronisation; the
 * pool keeps up to 1 GiB between calls and never touches the application's default pool). */
B2S_API int b2s_mt19937_seed(uint32_t seed, uint32_t* state /*dev [625]*/, void* stream);
B2S_API int b2s_mt19937_choice(uint32_t* state /*dev [625], in/out*/, int32_t n, int64_t count,
                       int32_t* idx_out /*dev [count]*/, void* stream);

/* Resample statistics: vals [F,n] fp64 (rows < split are "base", rows >= split "instruct"; all F rows
 * share the resampled column indices, as analyze_base_vs_instruct_mae_100q.py:380-388 does).
 * idx [B,n] from b2s_mt19937_choice, or NULL for the counter-based Philox4x32-10 stream keyed by
 * (philox_seed, b, j) (not NumPy-compatible; for unseeded reference loops).
 * boot[B,3] = {mean over rows<split, mean over rows>=split (NaN if none), mean |v| over all rows}. */
B2S_API int b2s_bootstrap_resample(const double* vals, int32_t F, int32_t split, int32_t n, int64_t B,
                           const int32_t* idx, uint64_t philox_seed, double* boot /*dev [B,3]*/, void* stream);

/* The index table of the Philox mode, for consumers that need the indices themselves (b2s_corr_bootstrap): idx_out [B, n],
 * idx[b, j] = the index b2s_bootstrap_resample(idx = NULL, philox_seed = seed) uses for draw j of resample b.  Stands in for
 * the reference loops that resample from an UNSEEDED generator (survey_analysis/analyze_llm_agreement_simple_bootstrap.py:
 * 176-178; survey_analysis_consolidated.py:228,323,386,499,635), where no particular stream can be reproduced. */
B2S_API int b2s_philox_indices(uint64_t seed, int32_t n, int64_t B, int32_t* idx_out /*dev [B,n]*/, void* stream);

/* Question-resampled agreement of one model with the human averages: the body of the loop of survey_analysis/
 * analyze_llm_agreement_simple_bootstrap.py:169-180 with calculate_bootstrap_metrics (:89-148).  human / model [n] fp64 (model
 * NaN = no answer for that question), idx [B, n] sampled question numbers (b2s_philox_indices: the script is unseeded).  A
 * resample uses every question that was drawn at least once AND has a model value (the script's `prompt in sampled_questions`).
 * out [B, 4] = MAE, MSE, MAPE in percent (finite terms only), Pearson r (NaN without spread); a row is all NaN when fewer than
 * min_matched (the script: 10) questions matched. */
B2S_API int b2s_agreement_bootstrap(const double* human, const double* model, int32_t n, const int32_t* idx, int64_t B,
                            int32_t min_matched, double* out /*dev [B,4]*/, void* stream);

/* Percentile CI + tail fractions of one statistic of boot (np.percentile linear interpolation):
 * stat 0 = boot[:,0], 1 = boot[:,1]-boot[:,0], 2 = boot[:,2].
 * summary[8] = {pct(q_lo), pct(q_hi), frac(x<=0), frac(x>=0), mean(x), frac(x>=observed), min, max}.
 * scratch: B rounded up to a power of two doubles (device). */
B2S_API size_t b2s_bootstrap_scratch_bytes(int64_t B);
B2S_API int b2s_bootstrap_summary(const double* boot, int64_t B, int32_t stat, double q_lo, double q_hi,
                          double observed, double* summary /*dev [8]*/, double* stat_out /*dev [B] or NULL*/,
                          void* scratch, size_t scratch_bytes, void* stream);

/* Prompt-resampled inter-model correlation bootstrap: the body of the loop of
 * analysis/model_comparison_graph.py:261-291 -- per resample b, DataFrame.corr(method) of the rows idx[b, :] of the
 * [P prompts, M models] table X (fp64, NaN = missing; pairwise-complete observations, min_periods 1 as pandas) for
 * the listed model pairs, then np.mean / np.median / np.std of the non-NaN pair values.
 * method 0 = pearson, 1 = spearman (average-tie ranks of the rows both models answered; P <= 2048).
 * idx [B, P] int32 row numbers (b2s_mt19937_choice reproduces np.random.choice(P, P) of a seeded stream),
 * pairs [n_pairs, 2] int32 column numbers.  corr [B, n_pairs] receives every pair correlation (NaN where pandas
 * yields NaN); stats (optional, n_pairs <= 2048) [B, 4] = mean, median, std (ddof 0), number of non-NaN pairs. */
B2S_API int b2s_corr_bootstrap(const double* X, int32_t P, int32_t M, const int32_t* idx, int32_t B,
                       const int32_t* pairs, int32_t n_pairs, int32_t method, double* corr, double* stats /*or NULL*/,
                       void* stream);

/* Same, with the respondent filter of the reference's cross-prompt bootstraps (survey_analysis_consolidated.py:520-523, 748-751):
 * with min_valid > 0 (pearson only) a column takes part in a resample only if at least min_valid of its resampled rows are not
 * NaN; the pairs of a dropped column are NaN for that resample.  n_pairs may exceed 2048 here: stats then holds mean, NaN
 * (no median), std and count. */
B2S_API int b2s_corr_bootstrap_ex(const double* X, int32_t P, int32_t M, const int32_t* idx, int32_t B,
                          const int32_t* pairs, int32_t n_pairs, int32_t method, int32_t min_valid, double* corr,
                          double* stats /*or NULL*/, void* stream);

/* Simulated-individual correlation bootstrap: the body of the 10,000-iteration loop of
 * survey_analysis/bootstrap_confidence_intervals.py:54-99,126-148.  Seed s = seed0 + i runs
 * calculate_individual_correlations(model_responses, n_samples, seed=s) on a NumPy-legacy MT19937 of its own
 * (randint(1,6), normal(mean,std) with the cached polar-method Gaussian, scipy pearsonr) and yields the mean
 * correlation of base and of instruct models.  avail/rel [n_models,5,10] (groups x questions 1..11 without 8),
 * mean/std [5,10] on the 0-1 scale, model_type 0 = base / 1 = instruct, models in CSV order.
 * corrs_out (optional) receives the per-sample correlations of the first seed; final_state (optional) the MT19937
 * state after the last seed, so that the script's later np.random.choice calls continue the same stream. */
B2S_API int b2s_ci_simulate(const uint8_t* avail, const double* rel, const double* mean, const double* stdv,
                    const int32_t* model_type, int32_t n_models, int32_t n_samples, uint32_t seed0, int32_t n_seeds,
                    double* means_out /*dev [n_seeds,2]*/, double* corrs_out /*dev [n_models,n_samples] or NULL*/,
                    uint32_t* final_state /*dev [625] or NULL*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SCORE_H */
