// Argument blocks and launchers for the non-GEMM decoder kernels (b2s_decoder.cu).
#pragma once
#include "b2s_common.cuh"

namespace b2s {

constexpr int ATT_THREADS = 256;
constexpr int ATT_HG = 8;   // query heads per CTA (sharing one KV head) for multi-row sequences

struct BatchMeta {           // all device pointers, sized for max_seqs / max_tokens
  int32_t* q_start;   // [N] first row of the sequence in the current activation buffers
  int32_t* q_len;     // [N]
  int32_t* kv_start;  // [N] first cache slot of the sequence
  int32_t* kv_len;    // [N] keys visible to the last query row
  int32_t* last_row;  // [N] row whose hidden state feeds the LM head
  int32_t* pos;       // [T] rotary position of each row
  int32_t* slot;      // [T] cache slot each row's K/V is appended to
};

struct LnArgs {
  float* x;                 // [rows_in, d] residual stream (written when tokens != null)
  const int32_t* tokens;    // gather rows of wte instead of reading x (layer-0 fusion), or null
  const void* wte;          // embedding table (bf16 or f32)
  const int32_t* rows;      // optional row gather: output row i <- input row rows[i]
  const float *g1, *b1, *g2, *b2;
  void *y1, *y2;            // outputs (AT); y2 optional (second LayerNorm of a parallel-residual layer)
  int d;
  float eps;
  int rms;                  // RMSNorm: no mean subtraction, no bias (b1/b2 null)
  const void* wpe;          // learned position table added to the gathered embedding (tokens != null), or null
  const int32_t* pos;       // position of every row (with wpe)
  int pos_offset;
  int split;                // BF16_SPLIT: output rows are [hi(d) | lo(d)] bf16, lo = bf16(y - bf16(y))
};

struct RopeArgs {
  void* qkv;                // [T, qkv_dim] (AT), q rotated in place
  void *kcache, *vcache;    // this layer's cache: [slots, nkv*hd] (AT)
  const int32_t *pos, *slot;
  const float *cos, *sin;   // [max_pos, rd/2]
  int qkv_dim, nh, nkv, hd, rd;
  int q_base, q_hstride, k_base, v_base, kv_hstride;
};

struct AttnArgs {
  const void* qkv;          // rotated queries live here
  const void *kcache, *vcache;
  void* out;                // [T, nh*hd] (AT)
  BatchMeta meta;
  int qkv_dim, nh, nkv, hd, rep /* nh / nkv */;
  int q_base, q_hstride;
  int max_kv;
  int f16;             // 16-bit buffers hold fp16 instead of bf16 (tensor-core kernel)
  const float* alibi;  // per-head ALiBi slopes (CUDA-core kernel only), or null
  int heads_per_cta;   // set by the launcher: ATT_HG for prefill, all heads of the KV head for single-token steps
  float scale;
  int split;           // BF16_SPLIT: output rows are [hi(nh*hd) | lo(nh*hd)] bf16
  // encoder-decoder (T5) forms, CUDA-core kernel only:
  const float* relbias;  // [nh, 2*rel_R + 1] additive bias by (key position - query position), or null
  int rel_R;
  int noncausal;         // every query row sees all kv_len keys (encoder self-attention, cross-attention)
  int kv_stride;         // elements between consecutive keys in kcache / vcache (0 = nkv*hd): K/V read in place from a
                         // fused projection output instead of a cache
};

struct ScanState {
  int32_t* alive;      // [N]
  int32_t* resolved;   // [N]
  int32_t* next_tok;   // [N] greedy token = input of the next decode step
  int32_t* gen_ids;    // [N, lookahead] or null
  int32_t* n_active;   // [2]: prompts still needing another position; prompts whose logits were not finite (this call)
  float* out;          // [N, 8]
  // Compaction of the look-ahead: row r of a one-token step is prompt act_in[r] (null: row == prompt).  The scan writes
  // the prompts that need another position, in ascending order, with their next input token and candidate pair.
  const int32_t* act_in;   // [rows of this step] or null
  int32_t* act_out;        // [N]
  int32_t* tok_out;        // [N]
  int32_t* cand_out;       // [N, 2]
  const int32_t* cand;     // [N, 2] candidate ids per prompt
};

// step 0: prefill bookkeeping for n_rows = all prompts.  step >= 1: one row per prompt of `act` (null: every prompt).
int launch_prep(const int32_t* cu, int n_rows, int lookahead, int step, const int32_t* act, const BatchMeta& m, cudaStream_t st);
template <typename AT, typename WT> int launch_ln(const LnArgs& a, int n_rows, cudaStream_t st);
template <typename AT> int launch_rope_cache(const RopeArgs& a, int n_rows, cudaStream_t st);
template <typename AT> int launch_attn(const AttnArgs& a, int n_seqs, cudaStream_t st);
// One-token steps (every sequence has q_len 1) for the shapes the tensor-core one-token kernel does not take (head_dim 80 /
// 128, ALiBi, the fp32 mode): HBM-bound streaming kernel, one warp per (sequence, head).
template <typename AT> int launch_attn_step(const AttnArgs& a, int n_seqs, cudaStream_t st);
// act[r, j] = silu(gu[r, j]) * gu[r, ff + j]   (HF llama LlamaMLP: act_fn(gate_proj(x)) * up_proj(x))
// gate_kind B2S_ACT_GELU_TANH: tanh-GELU gate instead of SiLU (HF t5 T5DenseGatedActDense: act(wi_0(x)) * wi_1(x))
template <typename AT> int launch_swiglu(const void* gu, void* act, int64_t rows, int ff, cudaStream_t st, int gate_kind = 0);
// x = relu(x) (kind 1) or tanh-GELU(x) (kind 2) in place (HF opt ReLU MLP; HF bloom bloom_gelu_forward)
template <typename AT> int launch_act(void* x, int64_t n, int kind, cudaStream_t st);
// BF16_SPLIT: the MLP pre-activations arrive in fp32 (pre [rows, ff], or [rows, 2*ff] = gate | up for kind 0) and the
// activation leaves as [rows, hi(ff) | lo(ff)] bf16.  kind 0 = SwiGLU, 1 = ReLU, 2 = tanh-GELU, 3 = tanh-GELU gate * up.
int launch_act_split(const float* pre, void* act, int64_t rows, int ff, int kind, cudaStream_t st);
// tcgen05 prefill attention (b2s_attn_tc.cu): head_dim 64, <= 256 keys per sequence
bool attn_tc_supported(int hd, int max_kv);
bool attn_tc_decode_supported(int hd, int max_kv, int qkv_dim, int q_base, int q_hstride);
int launch_attn_tc(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st);
// tcgen05 multi-head bidirectional attention with a relative-bias table (T5 encoder; b2s_attn_tc2.cu): q|k|v read in place
// from a.qkv = [rows, 3*nh*64], <= 256 keys per prompt
bool attn_enc_supported(const AttnArgs& a);
int launch_attn_enc(const AttnArgs& a, int n_seqs, int total_rows, int max_qlen, cudaStream_t st);
// LLM.int8 emulation (B2S_MODE_INT8_EMU): the fp16 rows x[rows, K] that feed a linear layer are replaced in place by what
// bitsandbytes' MatMul8bitLt makes of them.  A column is an outlier column of a prompt when any of the prompt's rows holds
// |x| >= threshold there (prompt = rows [q_start, q_start + q_len) of `meta`; one row per prompt in one-token steps): those
// columns stay fp16.  Every other element becomes rint(x * 127 / s) * s / 127 with s = the row's largest |x| below the threshold.
// `mask` is scratch of n_seqs * K bytes.
int launch_int8_fakequant(void* x, int rows, int K, int n_seqs, const BatchMeta& meta, uint8_t* mask, float threshold, cudaStream_t st);
int launch_lmhead_partials_from_logits(const float* logits, int M, int V, const int32_t* cand, float* partials,
                                       float* cand_logits, cudaStream_t st);
int launch_lmhead_merge(const float* partials, const float* cand_logits, int M, int tiles_n, float* stats,
                        cudaStream_t st);
int launch_scan_update(const float* stats, int n, int pos, int topk, int eos_id, int scan_positions, int gen_tokens,
                       const ScanState& s, cudaStream_t st);

}  // namespace b2s
