// Model handle, workspace carving, the decoder forward and the scorer entry points.
//
// b2s_score() is the batched replacement of the reference's per-prompt get_yes_no_logprobs body
// (analysis/run_base_vs_instruct_100q.py:329-363): prefill over varlen-packed prompts, LM head on the last
// row of each prompt, then up to max_lookahead-1 batched single-token greedy steps over a KV cache while
// the look-ahead scan (:348-363) is unresolved.
#include "b2s_decoder.cuh"
#include <stdarg.h>
#include <algorithm>
#include <type_traits>
#include <vector>

namespace b2s {

thread_local std::string g_last_error;
std::atomic<int64_t> g_launch_count{0};
int g_opt_compact = 1;   // b2s_set_option "lookahead_compact": one-token steps run only the prompts that still need a position
int g_opt_attn_step = 1;   // b2s_set_option "attn_step": streaming one-token attention (0 = the varlen kernel, for A/B and parity checks)
float g_opt_int8_threshold = 6.0f;   // b2s_set_option "int8_threshold_milli": bitsandbytes' llm_int8_threshold (default 6.0)

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
}

}  // namespace b2s

using namespace b2s;

struct b2s_model {
  b2s_model_desc d;
  std::vector<const void*> w;
  int device;
  int qkv_dim;
  int q_base, q_hstride, k_base, v_base, kv_hstride;
  bool bf16;      // tensor-core mode (bf16 weights); false = fp32 check mode
  int act_f16;    // tensor-core mode with fp16 operands (weights handed over as fp16)
  int split;      // BF16_SPLIT: every GEMM A operand is stored as [hi | lo] bf16 halves (row width doubled)
  int attn32;     // BF16_SPLIT_ATTN32: as BF16_SPLIT, with q/k/v, the KV cache and the whole of attention in fp32 (CUDA-core kernels)
  int int8;       // INT8_EMU: FP16 kernels, every linear layer's input put on bitsandbytes' int8 grid first (launch_int8_fakequant)
  size_t at_size;
  const void* gw(int slot) const { return w[slot]; }
  const void* lw(int layer, int slot) const { return w[B2S_W_NGLOBAL + layer * B2S_W_PER_LAYER + slot]; }
  const float* lwf(int layer, int slot) const { return reinterpret_cast<const float*>(lw(layer, slot)); }
};

namespace {

struct Workspace {
  BatchMeta meta;
  int32_t *tokens, *cu, *cand, *gen;
  int32_t *alist[2], *tok_step, *cand_step;   // look-ahead compaction: active prompt lists (ping-pong), their tokens and candidates
  float* out;
  ScanState scan;
  float* resid;
  void *ln1, *ln2, *qkv, *attn, *act, *gu, *hfin;
  void *kcache, *vcache;
  size_t cache_layer_bytes;
  size_t cache_slots;
  float *partials, *cand_logits, *stats, *logits;
  uint8_t* qmask;   // INT8_EMU: outlier-column flags, [N, widest linear input]
  size_t total;
};

// Carve `base` (may be null: size query) into the buffers above.
Workspace carve(const b2s_model* m, int64_t T, int32_t N, int32_t LA, uint8_t* base) {
  Workspace w{};
  size_t off = 0;
  auto take = [&](size_t bytes) -> uint8_t* {
    uint8_t* p = base ? base + off : nullptr;
    off += align_up(bytes, 256);
    return p;
  };
  const b2s_model_desc& d = m->d;
  const size_t at = m->at_size;
  const int64_t rows = T > N ? T : N;   // decode steps run N rows through the same buffers
  w.meta.q_start = reinterpret_cast<int32_t*>(take(4 * N));
  w.meta.q_len = reinterpret_cast<int32_t*>(take(4 * N));
  w.meta.kv_start = reinterpret_cast<int32_t*>(take(4 * N));
  w.meta.kv_len = reinterpret_cast<int32_t*>(take(4 * N));
  w.meta.last_row = reinterpret_cast<int32_t*>(take(4 * N));
  w.meta.pos = reinterpret_cast<int32_t*>(take(4 * rows));
  w.meta.slot = reinterpret_cast<int32_t*>(take(4 * rows));
  w.tokens = reinterpret_cast<int32_t*>(take(4 * rows));
  w.cu = reinterpret_cast<int32_t*>(take(4 * (N + 1)));
  w.cand = reinterpret_cast<int32_t*>(take(8 * N));
  w.gen = reinterpret_cast<int32_t*>(take(4 * static_cast<size_t>(N) * LA));
  w.out = reinterpret_cast<float*>(take(32 * N));
  w.scan.alive = reinterpret_cast<int32_t*>(take(4 * N));
  w.scan.resolved = reinterpret_cast<int32_t*>(take(4 * N));
  w.scan.next_tok = reinterpret_cast<int32_t*>(take(4 * N));
  w.scan.n_active = reinterpret_cast<int32_t*>(take(256));
  w.alist[0] = reinterpret_cast<int32_t*>(take(4 * N));
  w.alist[1] = reinterpret_cast<int32_t*>(take(4 * N));
  w.tok_step = reinterpret_cast<int32_t*>(take(4 * N));
  w.cand_step = reinterpret_cast<int32_t*>(take(8 * N));
  w.scan.gen_ids = nullptr;
  w.scan.out = nullptr;
  w.resid = reinterpret_cast<float*>(take(4 * rows * d.d_model));
  const size_t am = m->split ? 2 : 1;   // [hi | lo] rows in BF16_SPLIT
  w.ln1 = take(at * am * rows * d.d_model);
  w.ln2 = d.n_ln_per_layer == 2 ? take(at * am * rows * d.d_model) : nullptr;
  const size_t qt = m->attn32 ? 4 : at;     // q/k/v and the KV cache: 16-bit, or fp32 in BF16_SPLIT_ATTN32
  w.qkv = take(qt * rows * m->qkv_dim);
  w.attn = take(at * am * rows * d.n_head * d.head_dim);
  w.act = take(at * am * rows * d.d_ff);
  // MLP pre-activations that a separate kernel turns into the activation: gate | up before SwiGLU; in BF16_SPLIT also
  // the ReLU / tanh-GELU pre-activations, and then in fp32 so that nothing is rounded before the [hi | lo] split
  const bool pre_buf = d.mlp_type == B2S_MLP_SWIGLU;
  w.gu = pre_buf ? take((m->split ? 4 : at) * rows * (d.mlp_type == B2S_MLP_SWIGLU ? 2 : 1) * d.d_ff) : nullptr;
  w.hfin = take(at * am * static_cast<size_t>(N) * d.d_model);
  const size_t slots = static_cast<size_t>(T) + static_cast<size_t>(N) * LA;
  w.cache_slots = slots;
  w.cache_layer_bytes = align_up(qt * slots * d.n_kv_head * d.head_dim, 256);
  w.kcache = take(w.cache_layer_bytes * d.n_layer);
  w.vcache = take(w.cache_layer_bytes * d.n_layer);
  const size_t tiles_n = static_cast<size_t>(ceil_div(d.vocab, LMH_TILE_N));
  w.partials = reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * tiles_n * LMH_PART));
  w.cand_logits = reinterpret_cast<float*>(take(8 * N));
  w.stats = reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * LMH_STATS));
  w.logits = m->bf16 ? nullptr : reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * d.vocab));
  const size_t widest = static_cast<size_t>(std::max(std::max(d.d_model, d.d_ff), d.n_head * d.head_dim));
  w.qmask = m->int8 ? take(static_cast<size_t>(N) * widest) : nullptr;
  w.total = off;
  return w;
}

// C = A . W^T with the mode's GEMM.  `out` is AT except for EPI_RESID_F32 (always the fp32 residual).
int linear(const b2s_model* m, const void* A, const void* W, const float* bias, void* out, int M, int N, int K, int epi,
           cudaStream_t st, int out_split = 0) {
  if (m->bf16) return gemm_tc_launch(A, W, bias, out, M, N, K, epi, nullptr, st, m->act_f16, nullptr, m->split, out_split);
  return gemm_f32_launch(reinterpret_cast<const float*>(A), reinterpret_cast<const float*>(W), bias,
                         reinterpret_cast<float*>(out), M, N, K, epi, st);
}

// AT: type of the GEMM A operands (LayerNorm / activation outputs); WT: weight matrices; QT: q/k/v, the KV cache and P
// inside attention.  QT == AT except in BF16_SPLIT, where the linear layers see bf16 [hi | lo] halves and attention, which
// has no split form, works on fp16 (3 more mantissa bits than bf16; an overflow there surfaces as B2S_ERANGE).
template <typename AT, typename WT, typename QT = AT>
int forward_t(const b2s_model* m, const Workspace& w, const int32_t* tokens, int rows, int n_seqs, int max_kv,
              int max_qlen, cudaStream_t st) {
  const int q_f16 = std::is_same<QT, __half>::value ? 1 : 0;
  const b2s_model_desc& d = m->d;
  const int dm = d.d_model;
  const bool two_ln = d.n_ln_per_layer == 2;
  for (int l = 0; l < d.n_layer; ++l) {
    LnArgs ln{};
    ln.x = w.resid; ln.d = dm; ln.eps = d.norm_eps; ln.rms = d.norm_type == B2S_NORM_RMS; ln.split = m->split;
    ln.tokens = (l == 0) ? tokens : nullptr;
    ln.wte = m->gw(B2S_W_WTE);
    if (d.pos_type == B2S_POS_LEARNED) { ln.wpe = m->gw(B2S_W_WPE); ln.pos = w.meta.pos; ln.pos_offset = d.pos_offset; }
    if (l == 0 && d.emb_norm) {
      // BLOOM: the residual stream starts at LayerNorm(embedding); gather + normalise into the fp32 residual in place
      // (the kernel writes the raw row before its barriers and the normalised row after them), then layer 0 reads it
      LnArgs le = ln;
      le.g1 = reinterpret_cast<const float*>(m->gw(B2S_W_EMB_LN_G)); le.b1 = reinterpret_cast<const float*>(m->gw(B2S_W_EMB_LN_B));
      le.y1 = w.resid; le.rms = 0; le.split = 0;
      B2S_TRY((launch_ln<float, WT>(le, rows, st)));
      ln.tokens = nullptr;
    }
    ln.g1 = m->lwf(l, B2S_WL_LN1_G); ln.b1 = m->lwf(l, B2S_WL_LN1_B);
    ln.y1 = w.ln1;
    if (two_ln && d.parallel_residual) {   // both LayerNorms read the same residual: one pass, two affines
      ln.g2 = m->lwf(l, B2S_WL_LN2_G); ln.b2 = m->lwf(l, B2S_WL_LN2_B); ln.y2 = w.ln2;
    }
    B2S_TRY((launch_ln<AT, WT>(ln, rows, st)));
    // INT8_EMU: what bitsandbytes' Linear8bitLt does to the input of every linear layer (one prompt = one of its calls)
    auto q8 = [&](void* x, int K) -> int {
      return m->int8 ? launch_int8_fakequant(x, rows, K, n_seqs, w.meta, w.qmask, g_opt_int8_threshold, st) : B2S_OK;
    };
    B2S_TRY(q8(w.ln1, dm));
    if (ln.y2) B2S_TRY(q8(w.ln2, dm));

    uint8_t* kc = reinterpret_cast<uint8_t*>(w.kcache) + w.cache_layer_bytes * l;
    uint8_t* vc = reinterpret_cast<uint8_t*>(w.vcache) + w.cache_layer_bytes * l;
    // Falcon layout with 64-wide fully rotated heads: RoPE and the cache append ride in the QKV GEMM epilogue
    constexpr bool q32 = std::is_same<QT, float>::value && !std::is_same<AT, float>::value;   // BF16_SPLIT_ATTN32
    const bool fused_rope = m->bf16 && !q32 && d.qkv_layout == B2S_QKV_FALCON_MQA && d.head_dim == 64 && d.rotary_dim == 64;
    if (fused_rope) {
      RopeEpiArgs re{};
      re.pos = w.meta.pos; re.slot = w.meta.slot;
      re.cos = reinterpret_cast<const float*>(m->gw(B2S_W_ROPE_COS));
      re.sin = reinterpret_cast<const float*>(m->gw(B2S_W_ROPE_SIN));
      re.kcache = kc; re.vcache = vc; re.nh = d.n_head; re.nkv = d.n_kv_head;
      B2S_TRY(gemm_tc_launch(w.ln1, m->lw(l, B2S_WL_W_QKV), m->lwf(l, B2S_WL_B_QKV), w.qkv, rows, m->qkv_dim, dm,
                             EPI_QKV_ROPE, nullptr, st, m->act_f16, &re, m->split, 0, q_f16));
    } else {
      if (m->bf16)
        B2S_TRY(gemm_tc_launch(w.ln1, m->lw(l, B2S_WL_W_QKV), m->lwf(l, B2S_WL_B_QKV), w.qkv, rows, m->qkv_dim, dm,
                               q32 ? EPI_STORE_F32 : EPI_STORE_BF16, nullptr, st, m->act_f16, nullptr, m->split, 0, q_f16));
      else
        B2S_TRY(linear(m, w.ln1, m->lw(l, B2S_WL_W_QKV), m->lwf(l, B2S_WL_B_QKV), w.qkv, rows, m->qkv_dim, dm, EPI_STORE_BF16, st));
      RopeArgs ro{};
      ro.qkv = w.qkv; ro.kcache = kc; ro.vcache = vc; ro.pos = w.meta.pos; ro.slot = w.meta.slot;
      ro.cos = reinterpret_cast<const float*>(m->gw(B2S_W_ROPE_COS));
      ro.sin = reinterpret_cast<const float*>(m->gw(B2S_W_ROPE_SIN));
      ro.qkv_dim = m->qkv_dim; ro.nh = d.n_head; ro.nkv = d.n_kv_head; ro.hd = d.head_dim; ro.rd = d.rotary_dim;
      ro.q_base = m->q_base; ro.q_hstride = m->q_hstride; ro.k_base = m->k_base; ro.v_base = m->v_base;
      ro.kv_hstride = m->kv_hstride;
      B2S_TRY(launch_rope_cache<QT>(ro, rows, st));
    }

    AttnArgs at{};
    at.qkv = w.qkv; at.kcache = kc; at.vcache = vc; at.out = w.attn; at.meta = w.meta;
    at.qkv_dim = m->qkv_dim; at.nh = d.n_head; at.nkv = d.n_kv_head; at.hd = d.head_dim; at.rep = d.n_head / d.n_kv_head;
    at.q_base = m->q_base; at.q_hstride = m->q_hstride; at.max_kv = max_kv;
    at.heads_per_cta = (max_qlen == 1) ? at.rep : ATT_HG;
    at.f16 = q_f16;
    at.split = m->split;
    at.alibi = d.pos_type == B2S_POS_ALIBI ? reinterpret_cast<const float*>(m->gw(B2S_W_ALIBI)) : nullptr;
    at.scale = 1.0f / sqrtf(static_cast<float>(d.head_dim));
    // tensor-core kernel when the shape allows it (head_dim 64, <= 256 keys), CUDA-core kernel otherwise
    // (ALiBi: prompt pass only -- the one-token tensor-core kernel has no bias term)
    // (the one-token tensor-core kernel makes a tile out of the query heads that share a KV head: worth it for multi-query
    // models -- 71 rows for Falcon -- not for 1..4 heads per KV head, which stream better)
    const bool tc_ok = m->bf16 && !q32 && (at.alibi == nullptr || max_qlen > 1) && (max_qlen > 1 ? attn_tc_supported(d.head_dim, max_kv)
                                                : (attn_tc_decode_supported(d.head_dim, max_kv, m->qkv_dim, m->q_base, m->q_hstride) &&
                                                   (at.rep >= 8 || !g_opt_attn_step)));
    if (tc_ok) {
      B2S_TRY(launch_attn_tc(at, n_seqs, rows, static_cast<int>(w.cache_slots), max_qlen, st));
    } else if (max_qlen == 1 && g_opt_attn_step) {    // one-token steps of head_dim 80 / 128, ALiBi models, the fp32 mode
      B2S_TRY(launch_attn_step<QT>(at, n_seqs, st));
    } else {
      B2S_TRY(launch_attn<QT>(at, n_seqs, st));
    }

    B2S_TRY(q8(w.attn, d.n_head * d.head_dim));
    B2S_TRY(linear(m, w.attn, m->lw(l, B2S_WL_W_DENSE), m->lwf(l, B2S_WL_B_DENSE), w.resid, rows, dm,
                   d.n_head * d.head_dim, EPI_RESID_F32, st));
    const void* mlp_in = w.ln1;
    if (two_ln) {
      mlp_in = w.ln2;
      if (!d.parallel_residual) {   // sequential residual: second LayerNorm reads x + attn
        LnArgs l2{};
        l2.x = w.resid; l2.d = dm; l2.eps = d.norm_eps; l2.rms = d.norm_type == B2S_NORM_RMS; l2.split = m->split;
        l2.g1 = m->lwf(l, B2S_WL_LN2_G); l2.b1 = m->lwf(l, B2S_WL_LN2_B); l2.y1 = w.ln2;
        B2S_TRY((launch_ln<AT, WT>(l2, rows, st)));
        B2S_TRY(q8(w.ln2, dm));
      }
    }
    if (d.mlp_type == B2S_MLP_SWIGLU_FUSED) {   // gate / up rows interleaved: silu(gate) * up formed in the projection's epilogue
      B2S_TRY(gemm_tc_launch(mlp_in, m->lw(l, B2S_WL_W_UP), nullptr, w.act, rows, 2 * d.d_ff, dm, EPI_SWIGLU, nullptr, st, m->act_f16,
                             nullptr, m->split, m->split));
    } else if (m->split && d.mlp_type == B2S_MLP_SWIGLU) {
      // BF16_SPLIT: fp32 pre-activations, then one kernel forms the activation and its [hi | lo] halves
      const bool glu = d.mlp_type == B2S_MLP_SWIGLU;
      B2S_TRY(linear(m, mlp_in, m->lw(l, B2S_WL_W_UP), glu ? nullptr : m->lwf(l, B2S_WL_B_UP), w.gu, rows, (glu ? 2 : 1) * d.d_ff, dm,
                     EPI_STORE_F32, st));
      B2S_TRY(launch_act_split(reinterpret_cast<const float*>(w.gu), w.act, rows, d.d_ff, glu ? 0 : d.act_type, st));
    } else if (d.mlp_type == B2S_MLP_SWIGLU) {   // fused gate|up projection, then act = silu(gate) * up
      B2S_TRY(linear(m, mlp_in, m->lw(l, B2S_WL_W_UP), nullptr, w.gu, rows, 2 * d.d_ff, dm, EPI_STORE_BF16, st));
      B2S_TRY(launch_swiglu<AT>(w.gu, w.act, rows, d.d_ff, st));
    } else if (d.act_type != B2S_ACT_GELU && !m->bf16) {   // fp32 check mode: plain store, then ReLU / tanh-GELU in place
      B2S_TRY(linear(m, mlp_in, m->lw(l, B2S_WL_W_UP), m->lwf(l, B2S_WL_B_UP), w.act, rows, d.d_ff, dm, EPI_STORE_BF16, st));
      B2S_TRY(launch_act<AT>(w.act, static_cast<int64_t>(rows) * d.d_ff, d.act_type, st));
    } else if (m->bf16) {                                  // GELU / ReLU / tanh-GELU fused into the up-projection's epilogue
      B2S_TRY(gemm_tc_launch(mlp_in, m->lw(l, B2S_WL_W_UP), m->lwf(l, B2S_WL_B_UP), w.act, rows, d.d_ff, dm, EPI_GELU_BF16, nullptr, st,
                             m->act_f16, nullptr, m->split, m->split, -1, d.act_type));
    } else {
      B2S_TRY(linear(m, mlp_in, m->lw(l, B2S_WL_W_UP), m->lwf(l, B2S_WL_B_UP), w.act, rows, d.d_ff, dm, EPI_GELU_BF16, st, m->split));
    }
    B2S_TRY(q8(w.act, d.d_ff));
    B2S_TRY(linear(m, w.act, m->lw(l, B2S_WL_W_DOWN), m->lwf(l, B2S_WL_B_DOWN), w.resid, rows, dm, d.d_ff,
                   EPI_RESID_F32, st));
  }
  // final LayerNorm on the last row of every prompt only (HF generate sets logits_to_keep=1)
  LnArgs lf{};
  lf.x = w.resid; lf.d = dm; lf.eps = d.norm_eps; lf.rows = w.meta.last_row; lf.rms = d.norm_type == B2S_NORM_RMS; lf.split = m->split;
  lf.g1 = reinterpret_cast<const float*>(m->gw(B2S_W_LNF_G)); lf.b1 = reinterpret_cast<const float*>(m->gw(B2S_W_LNF_B));
  lf.y1 = w.hfin;
  B2S_TRY((launch_ln<AT, WT>(lf, n_seqs, st)));
  return B2S_OK;
}

int forward(const b2s_model* m, const Workspace& w, const int32_t* tokens, int rows, int n_seqs, int max_kv,
            int max_qlen, cudaStream_t st) {
  if (m->bf16 && m->split && m->attn32) return forward_t<__nv_bfloat16, __nv_bfloat16, float>(m, w, tokens, rows, n_seqs, max_kv, max_qlen, st);
  if (m->bf16 && m->split) return forward_t<__nv_bfloat16, __nv_bfloat16, __half>(m, w, tokens, rows, n_seqs, max_kv, max_qlen, st);
  if (m->bf16 && m->act_f16) return forward_t<__half, __half>(m, w, tokens, rows, n_seqs, max_kv, max_qlen, st);
  if (m->bf16) return forward_t<__nv_bfloat16, __nv_bfloat16>(m, w, tokens, rows, n_seqs, max_kv, max_qlen, st);
  return forward_t<float, float>(m, w, tokens, rows, n_seqs, max_kv, max_qlen, st);
}

int lm_head(const b2s_model* m, const Workspace& w, const int32_t* cand, int n_seqs, cudaStream_t st) {
  const b2s_model_desc& d = m->d;
  const int tiles_n = static_cast<int>(ceil_div(d.vocab, LMH_TILE_N));
  B2S_CHECK_CUDA(cudaMemsetAsync(w.cand_logits, 0, 8 * static_cast<size_t>(n_seqs), st));
  if (m->bf16) {
    LmHeadArgs a{cand, w.partials, w.cand_logits};
    B2S_TRY(gemm_tc_launch(w.hfin, m->gw(B2S_W_LM_HEAD), nullptr, nullptr, n_seqs, d.vocab, d.d_model, EPI_LMHEAD, &a, st,
                           m->act_f16, nullptr, m->split, 0));
  } else {
    B2S_TRY(gemm_f32_launch(reinterpret_cast<const float*>(w.hfin), reinterpret_cast<const float*>(m->gw(B2S_W_LM_HEAD)),
                            nullptr, w.logits, n_seqs, d.vocab, d.d_model, EPI_STORE_F32, st));
    B2S_TRY(launch_lmhead_partials_from_logits(w.logits, n_seqs, d.vocab, cand, w.partials, w.cand_logits, st));
  }
  return launch_lmhead_merge(w.partials, w.cand_logits, n_seqs, tiles_n, w.stats, st);
}

}  // namespace

// =================================================================================================
// extern "C"
// =================================================================================================
extern "C" {

const char* b2s_last_error(void) { return g_last_error.c_str(); }
int b2s_abi_version(void) { return B2S_ABI_VERSION; }

int64_t b2s_launch_count(int32_t reset) {
  return reset ? g_launch_count.exchange(0) : g_launch_count.load();
}

int b2s_model_create(const b2s_model_desc* desc, const void* const* weight_ptrs, int32_t n_ptrs, int32_t device,
                     b2s_model** out) {
  B2S_REQUIRE(desc && weight_ptrs && out, "model_create: null argument");
  B2S_REQUIRE(desc->abi_version == B2S_ABI_VERSION, "model_create: ABI version %d, library is %d", desc->abi_version,
              B2S_ABI_VERSION);
  const b2s_model_desc& d = *desc;
  B2S_REQUIRE(d.n_layer > 0 && d.d_model > 0 && d.n_head > 0 && d.n_kv_head > 0 && d.vocab > 0 && d.d_ff > 0,
              "model_create: non-positive dimension");
  B2S_REQUIRE(d.n_head % d.n_kv_head == 0, "model_create: n_head %d not a multiple of n_kv_head %d", d.n_head, d.n_kv_head);
  B2S_REQUIRE(d.head_dim == 64 || d.head_dim == 80 || d.head_dim == 128, "model_create: head_dim %d unsupported (64, 80, 128)",
              d.head_dim);
  B2S_REQUIRE(d.pos_type == B2S_POS_ROPE || d.pos_type == B2S_POS_LEARNED || d.pos_type == B2S_POS_ALIBI,
              "model_create: unknown pos_type %d", d.pos_type);
  B2S_REQUIRE(d.act_type == B2S_ACT_GELU || d.act_type == B2S_ACT_RELU || d.act_type == B2S_ACT_GELU_TANH,
              "model_create: unknown act_type %d", d.act_type);
  B2S_REQUIRE(d.pos_type == B2S_POS_ROPE ? (d.rotary_dim > 0 && d.rotary_dim <= d.head_dim && d.rotary_dim % 2 == 0) : d.rotary_dim == 0,
              "model_create: bad rotary_dim %d",
              d.rotary_dim);
  B2S_REQUIRE(d.n_ln_per_layer == 1 || d.n_ln_per_layer == 2, "model_create: n_ln_per_layer must be 1 or 2");
  B2S_REQUIRE(d.norm_type == B2S_NORM_LAYERNORM || d.norm_type == B2S_NORM_RMS, "model_create: unknown norm_type %d", d.norm_type);
  B2S_REQUIRE(d.mlp_type == B2S_MLP_GELU || d.mlp_type == B2S_MLP_SWIGLU || d.mlp_type == B2S_MLP_SWIGLU_FUSED,
              "model_create: unknown mlp_type %d", d.mlp_type);
  B2S_REQUIRE(d.has_bias >= 0 && d.has_bias <= 2, "model_create: has_bias %d outside [0,2]", d.has_bias);
  B2S_REQUIRE(d.mlp_type == B2S_MLP_GELU || d.has_bias != 1, "model_create: SwiGLU with MLP biases is not supported (q|k|v bias only: has_bias = 2)");
  B2S_REQUIRE(d.mlp_type != B2S_MLP_SWIGLU_FUSED || d.compute_mode != B2S_MODE_FP32_CHECK,
              "model_create: interleaved gate/up rows (B2S_MLP_SWIGLU_FUSED) are for the tensor-core modes");
  B2S_REQUIRE(d.n_ln_per_layer == 2 || d.parallel_residual, "model_create: a single LayerNorm implies parallel residual");
  B2S_REQUIRE(d.compute_mode == B2S_MODE_BF16_FAST || d.compute_mode == B2S_MODE_FP16 || d.compute_mode == B2S_MODE_BF16_SPLIT ||
                  d.compute_mode == B2S_MODE_FP32_CHECK || d.compute_mode == B2S_MODE_INT8_EMU || d.compute_mode == B2S_MODE_BF16_SPLIT_ATTN32,
              "model_create: unknown compute mode %d", d.compute_mode);
  B2S_REQUIRE((d.compute_mode != B2S_MODE_BF16_SPLIT && d.compute_mode != B2S_MODE_BF16_SPLIT_ATTN32) || (d.d_model % 64 == 0 && d.d_ff % 64 == 0 && (d.n_head * d.head_dim) % 64 == 0),
              "model_create: BF16_SPLIT needs d_model, d_ff and n_head*head_dim to be multiples of 64");
  B2S_REQUIRE(d.d_model % 8 == 0 && d.d_ff % 8 == 0, "model_create: d_model and d_ff must be multiples of 8");
  const int need = B2S_W_NGLOBAL + B2S_W_PER_LAYER * d.n_layer;
  B2S_REQUIRE(n_ptrs >= need, "model_create: %d weight pointers given, %d needed", n_ptrs, need);
  for (int i = 0; i < need; ++i)   // TMA descriptors and the vectorised LayerNorm need 16-byte aligned tensors (null = absent)
    B2S_REQUIRE((reinterpret_cast<uintptr_t>(weight_ptrs[i]) & 15u) == 0, "model_create: weight pointer %d is not 16-byte aligned", i);
  int ndev = 0;
  B2S_CHECK_CUDA(cudaGetDeviceCount(&ndev));
  B2S_REQUIRE(device >= 0 && device < ndev, "model_create: device %d out of range (%d visible)", device, ndev);
  cudaDeviceProp prop;
  B2S_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("model_create: device %d is sm_%d%d; this library only contains sm_100a code", device, prop.major, prop.minor);
    return B2S_EUNSUPPORTED;
  }
  b2s_model* m = new b2s_model();
  m->d = d;
  m->w.assign(weight_ptrs, weight_ptrs + need);
  m->device = device;
  m->bf16 = d.compute_mode != B2S_MODE_FP32_CHECK;
  m->int8 = d.compute_mode == B2S_MODE_INT8_EMU ? 1 : 0;
  m->act_f16 = (d.compute_mode == B2S_MODE_FP16 || m->int8) ? 1 : 0;
  m->attn32 = d.compute_mode == B2S_MODE_BF16_SPLIT_ATTN32 ? 1 : 0;
  m->split = (d.compute_mode == B2S_MODE_BF16_SPLIT || m->attn32) ? 1 : 0;
  m->at_size = m->bf16 ? 2 : 4;
  m->qkv_dim = (d.n_head + 2 * d.n_kv_head) * d.head_dim;
  if (d.qkv_layout == B2S_QKV_FALCON_MQA) {
    m->q_base = 0; m->q_hstride = d.head_dim;
    m->k_base = d.n_head * d.head_dim; m->v_base = (d.n_head + d.n_kv_head) * d.head_dim; m->kv_hstride = d.head_dim;
  } else if (d.qkv_layout == B2S_QKV_NEOX_INTERLEAVED) {
    if (d.n_head != d.n_kv_head) { delete m; set_error("model_create: NeoX qkv layout needs n_kv_head == n_head"); return B2S_EINVAL; }
    m->q_base = 0; m->q_hstride = 3 * d.head_dim; m->k_base = d.head_dim; m->v_base = 2 * d.head_dim;
    m->kv_hstride = 3 * d.head_dim;
  } else {
    delete m;
    set_error("model_create: unknown qkv layout %d", d.qkv_layout);
    return B2S_EINVAL;
  }
  const bool rms = d.norm_type == B2S_NORM_RMS;   // RMSNorm has gains only
  const int required[] = {B2S_W_WTE, B2S_W_LNF_G, B2S_W_LNF_B, B2S_W_LM_HEAD, B2S_W_ROPE_COS, B2S_W_ROPE_SIN};
  for (int s : required) {
    const bool rope_slot = s == B2S_W_ROPE_COS || s == B2S_W_ROPE_SIN;
    if (m->w[s] == nullptr && !(rms && s == B2S_W_LNF_B) && !(rope_slot && d.pos_type != B2S_POS_ROPE)) { delete m; set_error("model_create: global weight slot %d is null", s); return B2S_EINVAL; }
  }
  if ((d.pos_type == B2S_POS_LEARNED && m->w[B2S_W_WPE] == nullptr) || (d.pos_type == B2S_POS_ALIBI && m->w[B2S_W_ALIBI] == nullptr) ||
      (d.emb_norm && (m->w[B2S_W_EMB_LN_G] == nullptr || m->w[B2S_W_EMB_LN_B] == nullptr))) {
    delete m; set_error("model_create: position table / ALiBi slopes / embedding LayerNorm slot is null"); return B2S_EINVAL;
  }
  for (int l = 0; l < d.n_layer; ++l) {
    const int mats[] = {B2S_WL_LN1_G, B2S_WL_LN1_B, B2S_WL_W_QKV, B2S_WL_W_DENSE, B2S_WL_W_UP, B2S_WL_W_DOWN};
    for (int s : mats) {
      if (m->lw(l, s) == nullptr && !(rms && s == B2S_WL_LN1_B)) { delete m; set_error("model_create: layer %d slot %d is null", l, s); return B2S_EINVAL; }
    }
    if (d.n_ln_per_layer == 2 && (m->lw(l, B2S_WL_LN2_G) == nullptr || (!rms && m->lw(l, B2S_WL_LN2_B) == nullptr))) {
      delete m; set_error("model_create: layer %d second LayerNorm is null", l); return B2S_EINVAL;
    }
  }
  *out = m;
  return B2S_OK;
}

void b2s_model_destroy(b2s_model* m) { delete m; }

size_t b2s_workspace_bytes(const b2s_model* m, int64_t max_tokens, int32_t max_seqs, int32_t max_lookahead) {
  if (!m || max_tokens <= 0 || max_seqs <= 0 || max_lookahead <= 0) return 0;
  return carve(m, max_tokens, max_seqs, max_lookahead, nullptr).total + 256;
}

int b2s_score_complete(b2s_model* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs, int32_t total_tokens,
                       int32_t max_seqlen, const int32_t* cand_ids, int32_t topk, int32_t scan_positions, int32_t gen_tokens,
                       int32_t eos_id, float* out, int32_t* gen_ids, void* workspace, size_t workspace_bytes, void* stream) {
  B2S_REQUIRE(m && tokens && cu_seqlens && cand_ids && out && workspace, "score: null argument");
  B2S_REQUIRE(n_seqs > 0 && total_tokens >= n_seqs && max_seqlen > 0, "score: empty batch (n_seqs=%d total_tokens=%d)", n_seqs,
              total_tokens);
  B2S_REQUIRE(topk >= 1 && topk <= 5, "score: topk=%d outside [1,5]", topk);
  B2S_REQUIRE(scan_positions >= 1 && scan_positions <= 64, "score: max_lookahead=%d outside [1,64]", scan_positions);
  B2S_REQUIRE(gen_tokens >= scan_positions && gen_tokens <= 64, "score: gen_tokens=%d outside [max_lookahead=%d, 64]", gen_tokens,
              scan_positions);
  B2S_REQUIRE(max_seqlen + gen_tokens <= m->d.max_pos, "score: prompt of %d tokens + %d generated exceeds max_pos %d",
              max_seqlen, gen_tokens, m->d.max_pos);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* base = reinterpret_cast<uint8_t*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  Workspace w = carve(m, total_tokens, n_seqs, gen_tokens, base);
  B2S_REQUIRE(w.total + (base - reinterpret_cast<uint8_t*>(workspace)) <= workspace_bytes,
              "score: workspace of %zu bytes is too small, %zu needed", workspace_bytes, w.total + 256);
  B2S_CHECK_CUDA(cudaSetDevice(m->device));
  w.scan.out = out;
  w.scan.gen_ids = gen_ids;
  if (gen_ids) B2S_CHECK_CUDA(cudaMemsetAsync(gen_ids, 0xFF, 4 * static_cast<size_t>(n_seqs) * gen_tokens, st));

  // The tensor-core attention kernel loads whole 256-slot K/V boxes; slots it masks out must hold finite
  // numbers (0 * NaN would poison P.V), so the cache starts from zeros on every call.
  B2S_CHECK_CUDA(cudaMemsetAsync(w.kcache, 0, w.cache_layer_bytes * m->d.n_layer, st));
  B2S_CHECK_CUDA(cudaMemsetAsync(w.vcache, 0, w.cache_layer_bytes * m->d.n_layer, st));
  B2S_TRY(launch_prep(cu_seqlens, n_seqs, gen_tokens, 0, nullptr, w.meta, st));
  B2S_TRY(forward(m, w, tokens, total_tokens, n_seqs, max_seqlen, max_seqlen, st));
  B2S_TRY(lm_head(m, w, cand_ids, n_seqs, st));
  B2S_CHECK_CUDA(cudaMemsetAsync(w.scan.n_active, 0, 8, st));
  const bool compact = g_opt_compact != 0;
  w.scan.cand = cand_ids; w.scan.tok_out = w.tok_step; w.scan.cand_out = w.cand_step;
  w.scan.act_in = nullptr; w.scan.act_out = w.alist[0];
  B2S_TRY(launch_scan_update(w.stats, n_seqs, 0, topk, eos_id, scan_positions, gen_tokens, w.scan, st));
  for (int step = 1; step < gen_tokens; ++step) {
    // The number of prompts that need another position sizes every launch of the step (grids, tensor maps), so it
    // is read back: 4 bytes and one stream synchronisation per position, after which the step's ~200 launches are
    // enqueued while the GPU is already working on the first of them.
    int32_t active = 0;
    B2S_CHECK_CUDA(cudaMemcpyAsync(&active, w.scan.n_active, 4, cudaMemcpyDeviceToHost, st));
    B2S_CHECK_CUDA(cudaStreamSynchronize(st));
    if (active == 0) break;
    const int rows = compact ? active : n_seqs;
    const int32_t* act = compact ? w.alist[(step - 1) & 1] : nullptr;
    B2S_TRY(launch_prep(cu_seqlens, rows, gen_tokens, step, act, w.meta, st));
    B2S_TRY(forward(m, w, compact ? w.tok_step : w.scan.next_tok, rows, rows, max_seqlen + step, 1, st));
    B2S_TRY(lm_head(m, w, compact ? w.cand_step : cand_ids, rows, st));
    w.scan.act_in = act; w.scan.act_out = w.alist[step & 1];
    B2S_TRY(launch_scan_update(w.stats, rows, step, topk, eos_id, scan_positions, gen_tokens, w.scan, st));
  }
  return B2S_OK;
}

int b2s_score(b2s_model* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs, int32_t total_tokens,
              int32_t max_seqlen, const int32_t* cand_ids, int32_t topk, int32_t max_lookahead, int32_t eos_id, float* out,
              int32_t* gen_ids, void* workspace, size_t workspace_bytes, void* stream) {
  return b2s_score_complete(m, tokens, cu_seqlens, n_seqs, total_tokens, max_seqlen, cand_ids, topk, max_lookahead, max_lookahead,
                            eos_id, out, gen_ids, workspace, workspace_bytes, stream);
}

int b2s_score_complete_host(b2s_model* m, const int32_t* tokens_h, const int32_t* cu_h, int32_t n_seqs, const int32_t* cand_h,
                            int32_t topk, int32_t scan_positions, int32_t gen_tokens, int32_t eos_id, float* out_h,
                            int32_t* gen_h, void* workspace, size_t workspace_bytes, void* stream) {
  B2S_REQUIRE(m && tokens_h && cu_h && cand_h && out_h && workspace, "score_host: null argument");
  B2S_REQUIRE(n_seqs > 0, "score_host: empty batch");
  B2S_REQUIRE(scan_positions >= 1 && scan_positions <= 64, "score_host: max_lookahead=%d outside [1,64]", scan_positions);
  B2S_REQUIRE(gen_tokens >= scan_positions && gen_tokens <= 64, "score_host: gen_tokens=%d outside [max_lookahead=%d, 64]",
              gen_tokens, scan_positions);
  int64_t total = cu_h[n_seqs];
  int max_len = 0;
  B2S_REQUIRE(cu_h[0] == 0, "score_host: cu_seqlens[0] must be 0");
  for (int i = 0; i < n_seqs; ++i) {
    const int L = cu_h[i + 1] - cu_h[i];
    B2S_REQUIRE(L > 0, "score_host: prompt %d is empty", i);
    if (L > max_len) max_len = L;
  }
  for (int64_t t = 0; t < total; ++t)
    B2S_REQUIRE(tokens_h[t] >= 0 && tokens_h[t] < m->d.vocab, "score_host: token %d at %lld outside the vocabulary", tokens_h[t],
                static_cast<long long>(t));
  for (int i = 0; i < 2 * n_seqs; ++i)
    B2S_REQUIRE(cand_h[i] >= 0 && cand_h[i] < m->d.vocab, "score_host: candidate id %d outside the vocabulary", cand_h[i]);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* base = reinterpret_cast<uint8_t*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  Workspace w = carve(m, total, n_seqs, gen_tokens, base);
  B2S_REQUIRE(w.total + 256 <= workspace_bytes, "score_host: workspace of %zu bytes is too small, %zu needed", workspace_bytes,
              w.total + 256);
  B2S_CHECK_CUDA(cudaSetDevice(m->device));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.tokens, tokens_h, 4 * total, cudaMemcpyHostToDevice, st));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.cu, cu_h, 4 * (n_seqs + 1), cudaMemcpyHostToDevice, st));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.cand, cand_h, 8 * n_seqs, cudaMemcpyHostToDevice, st));
  // During the prefill the token buffer is only read by layer 0; decode steps read scan.next_tok.
  B2S_TRY(b2s_score_complete(m, w.tokens, w.cu, n_seqs, static_cast<int32_t>(total), max_len, w.cand, topk, scan_positions,
                             gen_tokens, eos_id, w.out, gen_h ? w.gen : nullptr, workspace, workspace_bytes, stream));
  B2S_CHECK_CUDA(cudaMemcpyAsync(out_h, w.out, 32 * static_cast<size_t>(n_seqs), cudaMemcpyDeviceToHost, st));
  if (gen_h)
    B2S_CHECK_CUDA(cudaMemcpyAsync(gen_h, w.gen, 4 * static_cast<size_t>(n_seqs) * gen_tokens, cudaMemcpyDeviceToHost, st));
  int32_t not_finite = 0;
  B2S_CHECK_CUDA(cudaMemcpyAsync(&not_finite, w.scan.n_active + 1, 4, cudaMemcpyDeviceToHost, st));
  B2S_CHECK_CUDA(cudaStreamSynchronize(st));
  if (not_finite > 0) {   // the rows are all there (NaN, found = -1 for the affected prompts); the caller decides what to do with them
    set_error("score_host: %d of %d prompts produced non-finite logits%s", not_finite, n_seqs,
              m->act_f16 ? " (an fp16 activation overflowed; rerun this model in a bf16 mode)" : "");
    return B2S_ERANGE;
  }
  return B2S_OK;
}

int b2s_score_host(b2s_model* m, const int32_t* tokens_h, const int32_t* cu_h, int32_t n_seqs, const int32_t* cand_h,
                   int32_t topk, int32_t max_lookahead, int32_t eos_id, float* out_h, int32_t* gen_h, void* workspace,
                   size_t workspace_bytes, void* stream) {
  return b2s_score_complete_host(m, tokens_h, cu_h, n_seqs, cand_h, topk, max_lookahead, max_lookahead, eos_id, out_h, gen_h,
                                 workspace, workspace_bytes, stream);
}

// ---- building blocks ------------------------------------------------------------------------------
int b2s_gemm_bf16(const void* A, const void* W, const float* bias, void* out, int32_t M, int32_t N, int32_t K, int32_t epilogue,
                  void* stream) {
  B2S_REQUIRE(A && W && out, "gemm_bf16: null argument");
  B2S_REQUIRE(epilogue != EPI_LMHEAD, "gemm_bf16: use b2s_lm_head_bf16 for the LM-head epilogue");
  return gemm_tc_launch(A, W, bias, out, M, N, K, epilogue, nullptr, reinterpret_cast<cudaStream_t>(stream));
}

int b2s_gemm_f16(const void* A, const void* W, const float* bias, void* out, int32_t M, int32_t N, int32_t K, int32_t epilogue,
                  void* stream) {
  B2S_REQUIRE(A && W && out, "gemm_f16: null argument");
  B2S_REQUIRE(epilogue != EPI_LMHEAD, "gemm_f16: LM-head epilogue not available here");
  return gemm_tc_launch(A, W, bias, out, M, N, K, epilogue, nullptr, reinterpret_cast<cudaStream_t>(stream), 1);
}

int b2s_gemm_f32(const float* A, const float* W, const float* bias, float* out, int32_t M, int32_t N, int32_t K, int32_t epilogue,
                 void* stream) {
  B2S_REQUIRE(A && W && out, "gemm_f32: null argument");
  return gemm_f32_launch(A, W, bias, out, M, N, K, epilogue, reinterpret_cast<cudaStream_t>(stream));
}

int b2s_attention_bf16(const void* qkv, const void* kcache, const void* vcache, void* out, const int32_t* q_start,
                       const int32_t* q_len, const int32_t* kv_start, const int32_t* kv_len, int32_t n_seqs, int32_t total_rows,
                       int32_t cache_slots, int32_t max_qlen, int32_t max_kv, int32_t n_head, int32_t n_kv_head, int32_t head_dim,
                       int32_t qkv_dim, int32_t q_base, int32_t q_hstride, int32_t use_tensor_cores, void* stream) {
  B2S_REQUIRE(qkv && kcache && vcache && out && q_start && q_len && kv_start && kv_len, "attention: null argument");
  B2S_REQUIRE(n_seqs > 0 && n_head > 0 && n_kv_head > 0 && n_head % n_kv_head == 0, "attention: bad head counts");
  AttnArgs a{};
  a.qkv = qkv; a.kcache = kcache; a.vcache = vcache; a.out = out;
  a.meta.q_start = const_cast<int32_t*>(q_start); a.meta.q_len = const_cast<int32_t*>(q_len);
  a.meta.kv_start = const_cast<int32_t*>(kv_start); a.meta.kv_len = const_cast<int32_t*>(kv_len);
  a.qkv_dim = qkv_dim; a.nh = n_head; a.nkv = n_kv_head; a.hd = head_dim; a.rep = n_head / n_kv_head;
  a.q_base = q_base; a.q_hstride = q_hstride; a.max_kv = max_kv;
  a.scale = 1.0f / sqrtf(static_cast<float>(head_dim));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (use_tensor_cores) return launch_attn_tc(a, n_seqs, total_rows, cache_slots, max_qlen, st);
  return launch_attn<__nv_bfloat16>(a, n_seqs, st);
}

int b2s_int8_fakequant_f16(void* x, int32_t rows, int32_t K, const int32_t* q_start, const int32_t* q_len, int32_t n_seqs,
                           void* mask_scratch, float threshold, void* stream) {
  B2S_REQUIRE(x && q_start && q_len && mask_scratch, "int8_fakequant: null argument");
  BatchMeta meta{};
  meta.q_start = const_cast<int32_t*>(q_start); meta.q_len = const_cast<int32_t*>(q_len);
  return launch_int8_fakequant(x, rows, K, n_seqs, meta, reinterpret_cast<uint8_t*>(mask_scratch), threshold,
                               reinterpret_cast<cudaStream_t>(stream));
}

size_t b2s_lm_head_scratch_bytes(int32_t M, int32_t V) {
  if (M <= 0 || V <= 0) return 0;
  const size_t tiles_n = static_cast<size_t>(ceil_div(V, LMH_TILE_N));
  return align_up(4 * static_cast<size_t>(M) * tiles_n * LMH_PART, 256) + align_up(8 * static_cast<size_t>(M), 256);
}

int b2s_lm_head_bf16(const void* hidden, const void* W, const int32_t* cand_ids, int32_t M, int32_t V, int32_t d, float* stats,
                     void* scratch, size_t scratch_bytes, void* stream) {
  B2S_REQUIRE(hidden && W && cand_ids && stats && scratch, "lm_head: null argument");
  B2S_REQUIRE(M > 0 && V > 0 && d > 0, "lm_head: empty problem");
  B2S_REQUIRE(scratch_bytes >= b2s_lm_head_scratch_bytes(M, V), "lm_head: scratch of %zu bytes is too small, %zu needed",
              scratch_bytes, b2s_lm_head_scratch_bytes(M, V));
  B2S_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 255) == 0, "lm_head: scratch must be 256-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int tiles_n = static_cast<int>(ceil_div(V, LMH_TILE_N));
  float* partials = reinterpret_cast<float*>(scratch);
  float* cand_logits = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) +
                                                align_up(4 * static_cast<size_t>(M) * tiles_n * LMH_PART, 256));
  B2S_CHECK_CUDA(cudaMemsetAsync(cand_logits, 0, 8 * static_cast<size_t>(M), st));
  LmHeadArgs a{cand_ids, partials, cand_logits};
  B2S_TRY(gemm_tc_launch(hidden, W, nullptr, nullptr, M, V, d, EPI_LMHEAD, &a, st));
  return launch_lmhead_merge(partials, cand_logits, M, tiles_n, stats, st);
}

}  // extern "C"
