// Encoder-decoder (T5 family) scorer: handle, workspace carving, the encoder pass, the greedy decoder steps and the
// entry points b2s_encdec_*.
//
// Batched replacement of the `is_encoder_decoder` branch of the reference's get_yes_no_logprobs
// (analysis/compare_instruct_models.py:178-226, run_base_vs_instruct_100q.py:290-325, compare_base_vs_instruct.py:192-237):
// AutoModelForSeq2SeqLM.generate(..., output_scores=True) + the top-k look-ahead scan.  One encoder pass over the
// varlen-packed prompts, the cross-attention keys / values of every decoder layer projected once from the encoder
// output, then one-token decoder steps (all prompts at step 0, afterwards only the prompts the scan has not resolved).
// The layers are HF:models/t5/modeling_t5.py T5LayerNorm (:46-69), T5DenseActDense / T5DenseGatedActDense (:84-132),
// T5Attention (:153-345: no 1/sqrt(d) scaling, position_bias added to the scores, fp32 softmax), T5Block (:411-498),
// T5Stack (:617-790) and T5ForConditionalGeneration.forward (:1040-1137).  Every linear layer is the tensor-core GEMM
// of b2s_gemm_tc.cu (fp32 check mode: b2s_gemm_f32.cu) with the residual add, the ReLU or the gated tanh-GELU in its
// epilogue.  Attention: the encoder runs attn_enc_kernel (b2s_attn_tc2.cu: tcgen05, P in tensor memory, one item per head
// with its own Q/K/V tiles); the decoder's self-attention over the few generated positions runs the CUDA-core varlen
// kernel of b2s_decoder.cu (bias-table form), its cross-attention the streaming kernel below; the fp32 check mode and
// prompts beyond 256 tokens run the CUDA-core kernel everywhere (non-causal, strided-K/V forms).
#include "b2s_decoder.cuh"
#include <type_traits>
#include <vector>

using namespace b2s;

struct b2s_encdec {
  b2s_encdec_desc d;
  std::vector<const void*> w;
  int device;
  int inner;      // n_head * head_dim
  bool bf16;      // tensor-core mode
  int act_f16, split;
  int ffn_split;  // FP16_FFN_SPLIT: fp16 everywhere, the feed-forward activation as bf16 hi | lo halves against bf16 W_DOWN
  size_t at_size;
  const void* gw(int slot) const { return w[slot]; }
  const float* gwf(int slot) const { return reinterpret_cast<const float*>(w[slot]); }
  const void* ew(int l, int slot) const { return w[B2S_T5_NGLOBAL + l * B2S_T5_ENC_PER_LAYER + slot]; }
  const void* dw(int l, int slot) const {
    return w[B2S_T5_NGLOBAL + d.n_enc_layer * B2S_T5_ENC_PER_LAYER + l * B2S_T5_DEC_PER_LAYER + slot];
  }
};

namespace {

struct EdWorkspace {
  BatchMeta meta;                  // encoder pass: prompt bookkeeping; decoder steps: self-attention over the generated prefix
  int32_t *ckv_start, *ckv_len;    // decoder steps: first encoder row / length of the prompt each row belongs to
  int32_t *tokens, *cu, *cand, *gen, *start_tok;
  int32_t *alist[2], *tok_step, *cand_step;
  float* out;
  ScanState scan;
  float* resid;
  void *ln, *qkv, *attn, *act, *gu, *cq, *hfin, *enc_out;
  void *ckv, *kcache, *vcache;
  size_t ckv_layer_bytes, cache_layer_bytes;
  float *partials, *cand_logits, *stats, *logits;
  size_t total;
};

EdWorkspace carve(const b2s_encdec* m, int64_t T, int32_t N, int32_t G, uint8_t* base) {
  EdWorkspace w{};
  size_t off = 0;
  auto take = [&](size_t bytes) -> uint8_t* {
    uint8_t* p = base ? base + off : nullptr;
    off += align_up(bytes, 256);
    return p;
  };
  auto take_i32 = [&](size_t n) { return reinterpret_cast<int32_t*>(take(4 * n)); };
  const b2s_encdec_desc& d = m->d;
  const size_t at = m->at_size, am = m->split ? 2 : 1;
  const size_t rows = static_cast<size_t>(T > N ? T : N);
  const size_t inner = static_cast<size_t>(m->inner);
  w.meta.q_start = take_i32(N); w.meta.q_len = take_i32(N); w.meta.kv_start = take_i32(N); w.meta.kv_len = take_i32(N);
  w.meta.last_row = take_i32(N); w.meta.pos = take_i32(rows); w.meta.slot = take_i32(rows);
  w.ckv_start = take_i32(N); w.ckv_len = take_i32(N);
  w.tokens = take_i32(rows); w.cu = take_i32(N + 1); w.cand = take_i32(2 * static_cast<size_t>(N));
  w.gen = take_i32(static_cast<size_t>(N) * G); w.start_tok = take_i32(N);
  w.out = reinterpret_cast<float*>(take(32 * static_cast<size_t>(N)));
  w.scan.alive = take_i32(N); w.scan.resolved = take_i32(N); w.scan.next_tok = take_i32(N);
  w.scan.n_active = reinterpret_cast<int32_t*>(take(256));
  w.alist[0] = take_i32(N); w.alist[1] = take_i32(N); w.tok_step = take_i32(N); w.cand_step = take_i32(2 * static_cast<size_t>(N));
  w.resid = reinterpret_cast<float*>(take(4 * rows * d.d_model));
  w.ln = take(at * am * rows * d.d_model);
  w.qkv = take(at * rows * 3 * inner);
  w.attn = take(at * am * rows * inner);
  w.act = take(at * (m->ffn_split ? 2 : am) * rows * d.d_ff);
  w.gu = d.ff_type == B2S_T5_FF_GATED_GELU ? take((m->split ? 4 : at) * rows * 2 * d.d_ff) : nullptr;
  w.cq = take(at * static_cast<size_t>(N) * inner);
  w.hfin = take(at * am * static_cast<size_t>(N) * d.d_model);
  w.enc_out = take(at * am * static_cast<size_t>(T) * d.d_model);
  w.ckv_layer_bytes = align_up(at * static_cast<size_t>(T) * 2 * inner, 256);
  w.ckv = take(w.ckv_layer_bytes * d.n_dec_layer);
  w.cache_layer_bytes = align_up(at * static_cast<size_t>(N) * G * inner, 256);
  w.kcache = take(w.cache_layer_bytes * d.n_dec_layer);
  w.vcache = take(w.cache_layer_bytes * d.n_dec_layer);
  const size_t tiles_n = static_cast<size_t>(ceil_div(d.vocab, LMH_TILE_N));
  w.partials = reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * tiles_n * LMH_PART));
  w.cand_logits = reinterpret_cast<float*>(take(8 * static_cast<size_t>(N)));
  w.stats = reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * LMH_STATS));
  w.logits = m->bf16 ? nullptr : reinterpret_cast<float*>(take(4 * static_cast<size_t>(N) * d.vocab));
  w.total = off;
  return w;
}

__global__ void fill_i32_kernel(int32_t* p, int n, int32_t v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// Decoder step `step` (>= 0): row r carries the token at decoder position `step` of prompt act[r] (act null: prompt r).
// Self-attention sees the step + 1 cached positions of that prompt, cross-attention its L encoder rows.
__global__ void encdec_prep_kernel(const int32_t* __restrict__ cu, int n_rows, int gen_tokens, int step,
                                   const int32_t* __restrict__ act, BatchMeta m, int32_t* __restrict__ ckv_start,
                                   int32_t* __restrict__ ckv_len) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const int i = act ? act[r] : r;
  m.q_start[r] = r; m.q_len[r] = 1; m.kv_start[r] = i * gen_tokens; m.kv_len[r] = step + 1; m.last_row[r] = r;
  m.pos[r] = step; m.slot[r] = i * gen_tokens + step;
  ckv_start[r] = cu[i]; ckv_len[r] = cu[i + 1] - cu[i];
}

// -------------------------------------------------------------------------------------------------
// Attention of a one-token decoder step, both kinds (HF t5:153-345): one warp per (decoder row, head), four heads of a
// row per CTA, so that a CTA reads 512 contiguous bytes of every key row.
//   cross (SELF = false)  keys / values = the prompt's rows of this layer's projected encoder output; no mask, no bias, no
//                         scaling.  HBM-bound: every step re-reads 2 * L * inner elements per row and layer.
//   self  (SELF = true)   the warp first appends this position's k and v (read from the fused q|k|v row) to the cache,
//                         then attends over the step + 1 cached positions with the decoder's relative bias
//                         (bias[head][key - query], query = last position).
// 8 lanes share a key (16 bytes = 8 dims each of its K row and of its V row), the four 8-lane groups of the warp walk the
// keys in four interleaved streams with an online softmax each (running maximum, sum, 8 output dims per lane; no shared
// memory, no barrier) and are merged with shuffles at the end.  (A three-pass version -- scores parked in shared memory,
// statistics, then P.V -- ran at the same speed: the kernel is bound by CTA turnover and HBM latency, not by instructions.)
// The generic varlen kernel spent 470 us (cross) and 35 us (self, after a separate 12 us append kernel) per layer and step
// here: one CTA per (row, head) staging K and V into shared memory as fp32 and then using one of its eight warps.
// -------------------------------------------------------------------------------------------------
struct StepAttnArgs {
  const void* q;          // [rows, q_pitch] QT: query of head h at column h*64 (self: the fused q|k|v row)
  const void *k, *v;      // key / value rows, kv_pitch elements apart; head h at column h*64
  void* out;              // [rows, inner] QT-sized 16-bit / fp32, or [rows, 2*inner] bf16 hi | lo (split)
  const int32_t *kv_start, *kv_len;
  int q_pitch, kv_pitch;
  int nh, inner, max_kv, split;
  // self-attention only
  const int32_t* slot;    // cache slot this row's k / v go to
  void *k_cache, *v_cache;   // == k, v (written)
  const float* relbias;   // [nh, 2*R + 1]
  int R;
};

template <typename QT, bool SELF>
__global__ void __launch_bounds__(128)
step_attn_kernel(StepAttnArgs a) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x * 4 + warp, r = blockIdx.y;
  if (h >= a.nh) return;                                // warp-uniform; no block-wide barrier below
  const int L = a.kv_len[r];
  const size_t pitch = static_cast<size_t>(a.kv_pitch);
  const QT* qrow = reinterpret_cast<const QT*>(a.q) + static_cast<size_t>(r) * a.q_pitch;
  if constexpr (SELF) {     // append this position's k, v (columns inner + h*64 and 2*inner + h*64 of the fused row)
    const size_t dst = static_cast<size_t>(a.slot[r]) * pitch + h * 64 + 2 * lane;
    using V2 = typename std::conditional<std::is_same<QT, float>::value, float2, uint32_t>::type;
    *reinterpret_cast<V2*>(reinterpret_cast<QT*>(a.k_cache) + dst) = *reinterpret_cast<const V2*>(qrow + a.inner + h * 64 + 2 * lane);
    *reinterpret_cast<V2*>(reinterpret_cast<QT*>(a.v_cache) + dst) = *reinterpret_cast<const V2*>(qrow + 2 * a.inner + h * 64 + 2 * lane);
    __syncwarp();           // the warp reads these rows back below
  }
  const QT* kbase = reinterpret_cast<const QT*>(a.k) + static_cast<size_t>(a.kv_start[r]) * pitch + h * 64;
  const QT* vbase = reinterpret_cast<const QT*>(a.v) + static_cast<size_t>(a.kv_start[r]) * pitch + h * 64;
  const float* tb = SELF ? a.relbias + static_cast<size_t>(h) * (2 * a.R + 1) + (a.R - (L - 1)) : nullptr;
  const int sub = lane & 7, grp = lane >> 3;
  float q[8];
  load8(qrow + h * 64 + 8 * sub, q);
  // One pass, online softmax, four independent streams: the 8 lanes of group g own keys g, g + 4, g + 8, ... -- 16-byte
  // slices of the K row and of the V row of two keys in flight per lane (K and V of 8 keys per warp instruction pair),
  // running maximum / sum / 8 output dims per lane; the four groups are merged with shuffles at the end.
  float m = -INFINITY, l = 0.f;
  float o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = 0.f;
  // K / V slices reach the lanes through a per-warp cp.async ring in shared memory (as in attn_step_kernel, b2s_decoder.cu):
  // three iterations of two keys per lane group in flight while one is consumed; a lane copies exactly what it consumes.
  constexpr int U = 2, STG = 3, NG = 4;
  constexpr int VB = 8 * sizeof(QT);
  extern __shared__ __align__(16) uint8_t ring_raw[];
  uint8_t* ring = ring_raw + static_cast<size_t>(warp) * (STG * U * 2 * 32 * VB);
  const uint32_t ring_s = static_cast<uint32_t>(__cvta_generic_to_shared(ring)) + static_cast<uint32_t>(lane) * VB;
  const int n_iter = (L + U * NG - 1) / (U * NG);
  auto issue = [&](int it) {
    if (it < n_iter) {
      const uint32_t dst = ring_s + static_cast<uint32_t>(it % STG) * (U * 2 * 32 * VB);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = it * (U * NG) + NG * u + grp;
        if (j < L) {
          const uint8_t* kp = reinterpret_cast<const uint8_t*>(kbase + static_cast<size_t>(j) * pitch + 8 * sub);
          const uint8_t* vp = reinterpret_cast<const uint8_t*>(vbase + static_cast<size_t>(j) * pitch + 8 * sub);
#pragma unroll
          for (int c = 0; c < VB / 16; ++c) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (2 * u) * 32 * VB + 16 * c), "l"(kp + 16 * c) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (2 * u + 1) * 32 * VB + 16 * c), "l"(vp + 16 * c) : "memory");
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int s0 = 0; s0 < STG; ++s0) issue(s0);
  for (int it = 0; it < n_iter; ++it) {
    asm volatile("cp.async.wait_group %0;" ::"n"(STG - 1) : "memory");
    const QT* slot = reinterpret_cast<const QT*>(ring + static_cast<size_t>(it % STG) * (U * 2 * 32 * VB) + static_cast<size_t>(lane) * VB);
    const int j0 = it * (U * NG);
    float k[2][8], v[2][8], d[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = j0 + 4 * u + grp;
      if (j < L) {
        load8(slot + (2 * u) * 32 * 8, k[u]);
        load8(slot + (2 * u + 1) * 32 * 8, v[u]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) { k[u][i] = 0.f; v[u][i] = 0.f; }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t = fmaf(q[i], k[u][i], t);
      t += __shfl_xor_sync(0xffffffffu, t, 1);
      t += __shfl_xor_sync(0xffffffffu, t, 2);
      t += __shfl_xor_sync(0xffffffffu, t, 4);
      const int j = j0 + 4 * u + grp;
      if constexpr (SELF) { if (j < L) t += tb[j]; }
      d[u] = j < L ? t : -INFINITY;
    }
    const float mn = fmaxf(m, fmaxf(d[0], d[1]));
    if (mn > -INFINITY) {                               // (a group with no key yet keeps its empty state)
      const float corr = __expf(m - mn), p0 = __expf(d[0] - mn), p1 = __expf(d[1] - mn);
      l = fmaf(l, corr, p0 + p1);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = fmaf(o[i], corr, fmaf(p0, v[0][i], p1 * v[1][i]));
      m = mn;
    }
    issue(it + STG);       // refill the slot just consumed
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  // merge the four groups
  float mall = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
  mall = fmaxf(mall, __shfl_xor_sync(0xffffffffu, mall, 16));
  const float w = m > -INFINITY ? __expf(m - mall) : 0.f;
  l *= w;
  l += __shfl_xor_sync(0xffffffffu, l, 8);
  l += __shfl_xor_sync(0xffffffffu, l, 16);
  const float inv = 1.f / l;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    o[i] *= w;
    o[i] += __shfl_xor_sync(0xffffffffu, o[i], 8);
    o[i] += __shfl_xor_sync(0xffffffffu, o[i], 16);
    o[i] *= inv;
  }
  if (grp == 0) {
    const size_t col = static_cast<size_t>(h) * 64 + 8 * sub;
    if (a.split) {
      __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(a.out) + static_cast<size_t>(r) * (2 * static_cast<size_t>(a.inner)) + col;
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 hv = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
        const __nv_bfloat162 lv = __floats2bfloat162_rn(o[2 * i] - __low2float(hv), o[2 * i + 1] - __high2float(hv));
        hi[i] = *reinterpret_cast<const uint32_t*>(&hv);
        lo[i] = *reinterpret_cast<const uint32_t*>(&lv);
      }
      *reinterpret_cast<uint4*>(ob) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(ob + a.inner) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    } else {
      QT* op = reinterpret_cast<QT*>(a.out) + static_cast<size_t>(r) * a.inner + col;
      if constexpr (std::is_same<QT, float>::value) {
        *reinterpret_cast<float4*>(op) = make_float4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<float4*>(op + 4) = make_float4(o[4], o[5], o[6], o[7]);
      } else if constexpr (std::is_same<QT, __half>::value) {
        *reinterpret_cast<uint4*>(op) = make_uint4(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3]), pack_f16x2(o[4], o[5]), pack_f16x2(o[6], o[7]));
      } else {
        *reinterpret_cast<uint4*>(op) = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      }
    }
  }
}

template <typename QT, bool SELF>
int launch_step_attn(const StepAttnArgs& a, int rows, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>(ceil_div(a.nh, 4)), static_cast<unsigned>(rows));
  const size_t smem = 4 * 3 * 2 * 2 * 32 * 8 * sizeof(QT);      // 4 warps x STG x U x (K, V) x 32 lanes x one 8-element slice
  step_attn_kernel<QT, SELF><<<grid, 128, smem, st>>>(a);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// AT: GEMM A operands (norm / activation outputs); WT: weight matrices; QT: q/k/v, the caches and the cross K/V
// (QT == AT except in BF16_SPLIT, where attention works on fp16: see forward_t in b2s_model.cu).
template <typename AT, typename WT, typename QT>
struct Run {
  const b2s_encdec* m;
  const EdWorkspace& w;
  cudaStream_t st;
  static constexpr int q_f16 = std::is_same<QT, __half>::value ? 1 : 0;

  int lin(const void* A, const void* W, void* out, int M, int N, int K, int epi, int out_split = 0, int out_f16 = -1,
          int act_kind = 0) const {
    if (m->bf16)
      return gemm_tc_launch(A, W, nullptr, out, M, N, K, epi, nullptr, st, m->act_f16, nullptr, m->split, out_split, out_f16, act_kind);
    return gemm_f32_launch(reinterpret_cast<const float*>(A), reinterpret_cast<const float*>(W), nullptr,
                           reinterpret_cast<float*>(out), M, N, K, epi, st);
  }

  // T5LayerNorm of the residual rows (optionally gathering the token embeddings into the residual first)
  int norm(const void* gain, void* y, int rows, const int32_t* tokens = nullptr, const int32_t* gather = nullptr) const {
    LnArgs a{};
    a.x = w.resid; a.d = m->d.d_model; a.eps = m->d.norm_eps; a.rms = 1; a.split = m->split;
    a.tokens = tokens; a.wte = m->gw(B2S_T5_WTE); a.rows = gather;
    a.g1 = reinterpret_cast<const float*>(gain); a.y1 = y;
    return launch_ln<AT, WT>(a, rows, st);
  }

  // x += wo(act(wi(ln)))   (T5LayerFF)
  int ffn(const void* w_up, const void* w_down, int rows) const {
    const b2s_encdec_desc& d = m->d;
    const int dm = d.d_model, ff = d.d_ff;
    if (m->ffn_split) {
      // fp16 operands in, act(...) formed on the fp32 accumulators and stored as bf16 [hi | lo]; then the bf16 second
      // projection over both halves (fp32 accumulate, reduce-add into the fp32 residual)
      if (d.ff_type == B2S_T5_FF_GATED_GELU_FUSED)
        B2S_TRY(gemm_tc_launch(w.ln, w_up, nullptr, w.act, rows, 2 * ff, dm, EPI_SWIGLU, nullptr, st, 1, nullptr, 0, 1, 0, B2S_ACT_GELU_TANH));
      else
        B2S_TRY(gemm_tc_launch(w.ln, w_up, nullptr, w.act, rows, ff, dm, EPI_GELU_BF16, nullptr, st, 1, nullptr, 0, 1, 0, B2S_ACT_RELU));
      return gemm_tc_launch(w.act, w_down, nullptr, w.resid, rows, dm, ff, EPI_RESID_F32, nullptr, st, 0, nullptr, 1, 0);
    }
    if (d.ff_type == B2S_T5_FF_GATED_GELU_FUSED) {
      B2S_TRY(lin(w.ln, w_up, w.act, rows, 2 * ff, dm, EPI_SWIGLU, m->split, -1, B2S_ACT_GELU_TANH));
    } else if (d.ff_type == B2S_T5_FF_GATED_GELU) {
      if (m->split) {
        B2S_TRY(lin(w.ln, w_up, w.gu, rows, 2 * ff, dm, EPI_STORE_F32));
        B2S_TRY(launch_act_split(reinterpret_cast<const float*>(w.gu), w.act, rows, ff, 3, st));
      } else {
        B2S_TRY(lin(w.ln, w_up, w.gu, rows, 2 * ff, dm, EPI_STORE_BF16));
        B2S_TRY(launch_swiglu<AT>(w.gu, w.act, rows, ff, st, B2S_ACT_GELU_TANH));
      }
    } else if (m->bf16) {
      B2S_TRY(lin(w.ln, w_up, w.act, rows, ff, dm, EPI_GELU_BF16, m->split, -1, B2S_ACT_RELU));
    } else {
      B2S_TRY(lin(w.ln, w_up, w.act, rows, ff, dm, EPI_STORE_BF16));
      B2S_TRY(launch_act<AT>(w.act, static_cast<int64_t>(rows) * ff, B2S_ACT_RELU, st));
    }
    return lin(w.act, w_down, w.resid, rows, dm, ff, EPI_RESID_F32);
  }

  AttnArgs attn_args(int max_kv) const {
    AttnArgs a{};
    const b2s_encdec_desc& d = m->d;
    a.out = w.attn; a.meta = w.meta;
    a.nh = d.n_head; a.nkv = d.n_head; a.hd = d.head_dim; a.rep = 1; a.heads_per_cta = 1;
    a.q_base = 0; a.q_hstride = d.head_dim; a.max_kv = max_kv;
    a.f16 = q_f16; a.split = m->split;
    a.scale = 1.0f;                      // T5 folds 1/sqrt(d_kv) into the initialisation: scores = q.k + bias
    a.rel_R = d.max_pos - 1;
    return a;
  }

  // Encoder over the packed prompts, then the cross-attention K/V of every decoder layer.
  int encoder(const int32_t* tokens, const int32_t* cu, int T, int n_seqs, int max_len) const {
    const b2s_encdec_desc& d = m->d;
    const int dm = d.d_model, inner = m->inner;
    B2S_TRY(launch_prep(cu, n_seqs, 0, 0, nullptr, w.meta, st));   // look-ahead 0: "cache slot" = row of the packed batch
    for (int l = 0; l < d.n_enc_layer; ++l) {
      B2S_TRY(norm(m->ew(l, B2S_T5E_LN1_G), w.ln, T, l == 0 ? tokens : nullptr));
      B2S_TRY(lin(w.ln, m->ew(l, B2S_T5E_W_QKV), w.qkv, T, 3 * inner, dm, EPI_STORE_BF16, 0, q_f16));
      AttnArgs a = attn_args(max_len);
      a.qkv = w.qkv; a.qkv_dim = 3 * inner;
      a.kcache = reinterpret_cast<const QT*>(w.qkv) + inner;        // keys and values read in place from the projection
      a.vcache = reinterpret_cast<const QT*>(w.qkv) + 2 * inner;
      a.kv_stride = 3 * inner;
      a.noncausal = 1;
      a.relbias = m->gwf(B2S_T5_ENC_BIAS);
      if (m->bf16 && attn_enc_supported(a)) {      // tensor-core kernel for <= 256-token prompts, CUDA-core kernel beyond
        B2S_TRY(launch_attn_enc(a, n_seqs, T, max_len, st));
      } else {
        B2S_TRY(launch_attn<QT>(a, n_seqs, st));
      }
      B2S_TRY(lin(w.attn, m->ew(l, B2S_T5E_W_O), w.resid, T, dm, inner, EPI_RESID_F32));
      B2S_TRY(norm(m->ew(l, B2S_T5E_LN2_G), w.ln, T));
      B2S_TRY(ffn(m->ew(l, B2S_T5E_W_UP), m->ew(l, B2S_T5E_W_DOWN), T));
    }
    B2S_TRY(norm(m->gw(B2S_T5_ENC_LNF_G), w.enc_out, T));
    for (int l = 0; l < d.n_dec_layer; ++l) {
      uint8_t* ckv = reinterpret_cast<uint8_t*>(w.ckv) + w.ckv_layer_bytes * l;
      B2S_TRY(lin(w.enc_out, m->dw(l, B2S_T5D_W_CKV), ckv, T, 2 * inner, dm, EPI_STORE_BF16, 0, q_f16));
    }
    return B2S_OK;
  }

  // One decoder position for `rows` prompts: hidden states of the new token in w.hfin.
  int decoder_step(const int32_t* tokens, int rows, int step, int max_len) const {
    const b2s_encdec_desc& d = m->d;
    const int dm = d.d_model, inner = m->inner;
    for (int l = 0; l < d.n_dec_layer; ++l) {
      uint8_t* kc = reinterpret_cast<uint8_t*>(w.kcache) + w.cache_layer_bytes * l;
      uint8_t* vc = reinterpret_cast<uint8_t*>(w.vcache) + w.cache_layer_bytes * l;
      const QT* ckv = reinterpret_cast<const QT*>(reinterpret_cast<const uint8_t*>(w.ckv) + w.ckv_layer_bytes * l);
      // self-attention over the generated prefix
      B2S_TRY(norm(m->dw(l, B2S_T5D_LN1_G), w.ln, rows, l == 0 ? tokens : nullptr));
      B2S_TRY(lin(w.ln, m->dw(l, B2S_T5D_W_QKV), w.qkv, rows, 3 * inner, dm, EPI_STORE_BF16, 0, q_f16));
      StepAttnArgs sa{};                   // appends this position's k / v to the cache, then attends over the prefix
      sa.q = w.qkv; sa.q_pitch = 3 * inner; sa.k = kc; sa.v = vc; sa.kv_pitch = inner; sa.out = w.attn;
      sa.kv_start = w.meta.kv_start; sa.kv_len = w.meta.kv_len; sa.nh = d.n_head; sa.inner = inner; sa.max_kv = step + 1;
      sa.split = m->split; sa.slot = w.meta.slot; sa.k_cache = kc; sa.v_cache = vc;
      sa.relbias = m->gwf(B2S_T5_DEC_BIAS); sa.R = d.max_pos - 1;
      B2S_TRY((launch_step_attn<QT, true>(sa, rows, st)));
      B2S_TRY(lin(w.attn, m->dw(l, B2S_T5D_W_O), w.resid, rows, dm, inner, EPI_RESID_F32));
      // cross-attention over the prompt's encoder output (no position bias: HF t5:310-317)
      B2S_TRY(norm(m->dw(l, B2S_T5D_LN2_G), w.ln, rows));
      B2S_TRY(lin(w.ln, m->dw(l, B2S_T5D_W_CQ), w.cq, rows, inner, dm, EPI_STORE_BF16, 0, q_f16));
      StepAttnArgs c{};
      c.q = w.cq; c.q_pitch = inner; c.k = ckv; c.v = ckv + inner; c.kv_pitch = 2 * inner; c.out = w.attn;
      c.kv_start = w.ckv_start; c.kv_len = w.ckv_len; c.nh = d.n_head; c.inner = inner; c.max_kv = max_len; c.split = m->split;
      B2S_TRY((launch_step_attn<QT, false>(c, rows, st)));
      B2S_TRY(lin(w.attn, m->dw(l, B2S_T5D_W_CO), w.resid, rows, dm, inner, EPI_RESID_F32));
      B2S_TRY(norm(m->dw(l, B2S_T5D_LN3_G), w.ln, rows));
      B2S_TRY(ffn(m->dw(l, B2S_T5D_W_UP), m->dw(l, B2S_T5D_W_DOWN), rows));
    }
    return norm(m->gw(B2S_T5_DEC_LNF_G), w.hfin, rows, nullptr, w.meta.last_row);
  }
};

int lm_head(const b2s_encdec* m, const EdWorkspace& w, const int32_t* cand, int n_rows, cudaStream_t st) {
  const b2s_encdec_desc& d = m->d;
  const int tiles_n = static_cast<int>(ceil_div(d.vocab, LMH_TILE_N));
  B2S_CHECK_CUDA(cudaMemsetAsync(w.cand_logits, 0, 8 * static_cast<size_t>(n_rows), st));
  if (m->bf16) {
    LmHeadArgs a{cand, w.partials, w.cand_logits};
    B2S_TRY(gemm_tc_launch(w.hfin, m->gw(B2S_T5_LM_HEAD), nullptr, nullptr, n_rows, d.vocab, d.d_model, EPI_LMHEAD, &a, st,
                           m->act_f16, nullptr, m->split, 0));
  } else {
    B2S_TRY(gemm_f32_launch(reinterpret_cast<const float*>(w.hfin), reinterpret_cast<const float*>(m->gw(B2S_T5_LM_HEAD)), nullptr,
                            w.logits, n_rows, d.vocab, d.d_model, EPI_STORE_F32, st));
    B2S_TRY(launch_lmhead_partials_from_logits(w.logits, n_rows, d.vocab, cand, w.partials, w.cand_logits, st));
  }
  return launch_lmhead_merge(w.partials, w.cand_logits, n_rows, tiles_n, w.stats, st);
}

template <typename AT, typename WT, typename QT>
int score_t(const b2s_encdec* m, EdWorkspace& w, const int32_t* tokens, const int32_t* cu, int n_seqs, int total_tokens,
            int max_seqlen, const int32_t* cand_ids, int topk, int scan_positions, int gen_tokens, int eos_id, cudaStream_t st) {
  Run<AT, WT, QT> run{m, w, st};
  B2S_TRY(run.encoder(tokens, cu, total_tokens, n_seqs, max_seqlen));
  fill_i32_kernel<<<static_cast<unsigned>(ceil_div(n_seqs, 256)), 256, 0, st>>>(w.start_tok, n_seqs, m->d.decoder_start_id);
  B2S_CHECK_LAUNCH();
  B2S_CHECK_CUDA(cudaMemsetAsync(w.scan.n_active, 0, 8, st));
  const bool compact = g_opt_compact != 0;
  w.scan.cand = cand_ids; w.scan.tok_out = w.tok_step; w.scan.cand_out = w.cand_step;
  for (int step = 0; step < gen_tokens; ++step) {
    int rows = n_seqs;
    const int32_t* act = nullptr;
    if (step > 0) {   // the number of prompts that need another position sizes the step's launches (see b2s_score)
      int32_t active = 0;
      B2S_CHECK_CUDA(cudaMemcpyAsync(&active, w.scan.n_active, 4, cudaMemcpyDeviceToHost, st));
      B2S_CHECK_CUDA(cudaStreamSynchronize(st));
      if (active == 0) break;
      if (compact) { rows = active; act = w.alist[(step - 1) & 1]; }
    }
    const int32_t* tok_in = step == 0 ? w.start_tok : (compact ? w.tok_step : w.scan.next_tok);
    encdec_prep_kernel<<<static_cast<unsigned>(ceil_div(rows, 128)), 128, 0, st>>>(cu, rows, gen_tokens, step, act, w.meta,
                                                                                   w.ckv_start, w.ckv_len);
    B2S_CHECK_LAUNCH();
    B2S_TRY(run.decoder_step(tok_in, rows, step, max_seqlen));
    B2S_TRY(lm_head(m, w, (step > 0 && compact) ? w.cand_step : cand_ids, rows, st));
    w.scan.act_in = act; w.scan.act_out = w.alist[step & 1];
    B2S_TRY(launch_scan_update(w.stats, rows, step, topk, eos_id, scan_positions, gen_tokens, w.scan, st));
  }
  return B2S_OK;
}

}  // namespace

extern "C" {

int b2s_encdec_create(const b2s_encdec_desc* desc, const void* const* weight_ptrs, int32_t n_ptrs, int32_t device,
                      b2s_encdec** out) {
  B2S_REQUIRE(desc && weight_ptrs && out, "encdec_create: null argument");
  B2S_REQUIRE(desc->abi_version == B2S_ABI_VERSION, "encdec_create: ABI version %d, library is %d", desc->abi_version,
              B2S_ABI_VERSION);
  const b2s_encdec_desc& d = *desc;
  B2S_REQUIRE(d.n_enc_layer > 0 && d.n_dec_layer > 0 && d.d_model > 0 && d.n_head > 0 && d.vocab > 0 && d.d_ff > 0 && d.max_pos > 0,
              "encdec_create: non-positive dimension");
  B2S_REQUIRE(d.head_dim == 64, "encdec_create: head_dim (d_kv) %d unsupported (64)", d.head_dim);
  B2S_REQUIRE(d.ff_type == B2S_T5_FF_RELU || d.ff_type == B2S_T5_FF_GATED_GELU || d.ff_type == B2S_T5_FF_GATED_GELU_FUSED,
              "encdec_create: unknown ff_type %d", d.ff_type);
  B2S_REQUIRE(d.ff_type != B2S_T5_FF_GATED_GELU_FUSED || d.compute_mode != B2S_MODE_FP32_CHECK,
              "encdec_create: interleaved wi_0 / wi_1 rows (B2S_T5_FF_GATED_GELU_FUSED) are for the tensor-core modes");
  B2S_REQUIRE(d.compute_mode == B2S_MODE_BF16_FAST || d.compute_mode == B2S_MODE_FP16 || d.compute_mode == B2S_MODE_BF16_SPLIT ||
                  d.compute_mode == B2S_MODE_FP32_CHECK || d.compute_mode == B2S_MODE_FP16_FFN_SPLIT,
              "encdec_create: unknown compute mode %d", d.compute_mode);
  B2S_REQUIRE((d.compute_mode != B2S_MODE_BF16_SPLIT && d.compute_mode != B2S_MODE_FP16_FFN_SPLIT) || (d.d_model % 64 == 0 && d.d_ff % 64 == 0),
              "encdec_create: the split modes need d_model and d_ff to be multiples of 64");
  B2S_REQUIRE(d.compute_mode != B2S_MODE_FP16_FFN_SPLIT || d.ff_type != B2S_T5_FF_GATED_GELU,
              "encdec_create: FP16_FFN_SPLIT forms the gate in the projection's epilogue: hand W_UP over interleaved (B2S_T5_FF_GATED_GELU_FUSED)");
  B2S_REQUIRE(d.d_model % 8 == 0 && d.d_ff % 8 == 0, "encdec_create: d_model and d_ff must be multiples of 8");
  B2S_REQUIRE(d.decoder_start_id >= 0 && d.decoder_start_id < d.vocab, "encdec_create: decoder_start_id %d outside the vocabulary",
              d.decoder_start_id);
  const int need = B2S_T5_NGLOBAL + B2S_T5_ENC_PER_LAYER * d.n_enc_layer + B2S_T5_DEC_PER_LAYER * d.n_dec_layer;
  B2S_REQUIRE(n_ptrs >= need, "encdec_create: %d weight pointers given, %d needed", n_ptrs, need);
  for (int i = 0; i < need; ++i)
    B2S_REQUIRE((reinterpret_cast<uintptr_t>(weight_ptrs[i]) & 15u) == 0, "encdec_create: weight pointer %d is not 16-byte aligned", i);
  int ndev = 0;
  B2S_CHECK_CUDA(cudaGetDeviceCount(&ndev));
  B2S_REQUIRE(device >= 0 && device < ndev, "encdec_create: device %d out of range (%d visible)", device, ndev);
  cudaDeviceProp prop;
  B2S_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("encdec_create: device %d is sm_%d%d; this library only contains sm_100a code", device, prop.major, prop.minor);
    return B2S_EUNSUPPORTED;
  }
  b2s_encdec* m = new b2s_encdec();
  m->d = d;
  m->w.assign(weight_ptrs, weight_ptrs + need);
  m->device = device;
  m->inner = d.n_head * d.head_dim;
  m->bf16 = d.compute_mode != B2S_MODE_FP32_CHECK;
  m->ffn_split = d.compute_mode == B2S_MODE_FP16_FFN_SPLIT ? 1 : 0;
  m->act_f16 = (d.compute_mode == B2S_MODE_FP16 || m->ffn_split) ? 1 : 0;
  m->split = d.compute_mode == B2S_MODE_BF16_SPLIT ? 1 : 0;
  m->at_size = m->bf16 ? 2 : 4;
  auto fail = [&](const char* what, int l, int s) {
    set_error("encdec_create: %s layer %d slot %d is null", what, l, s);
    delete m;
    return B2S_EINVAL;
  };
  for (int s = 0; s <= B2S_T5_DEC_BIAS; ++s)
    if (m->w[s] == nullptr) return fail("global", -1, s);
  for (int l = 0; l < d.n_enc_layer; ++l)
    for (int s = 0; s <= B2S_T5E_W_DOWN; ++s)
      if (m->ew(l, s) == nullptr) return fail("encoder", l, s);
  for (int l = 0; l < d.n_dec_layer; ++l)
    for (int s = 0; s <= B2S_T5D_W_DOWN; ++s)
      if (m->dw(l, s) == nullptr) return fail("decoder", l, s);
  *out = m;
  return B2S_OK;
}

void b2s_encdec_destroy(b2s_encdec* m) { delete m; }

size_t b2s_encdec_workspace_bytes(const b2s_encdec* m, int64_t max_tokens, int32_t max_seqs, int32_t max_gen_tokens) {
  if (!m || max_tokens <= 0 || max_seqs <= 0 || max_gen_tokens <= 0) return 0;
  return carve(m, max_tokens, max_seqs, max_gen_tokens, nullptr).total + 256;
}

int b2s_encdec_score(b2s_encdec* m, const int32_t* tokens, const int32_t* cu_seqlens, int32_t n_seqs, int32_t total_tokens,
                     int32_t max_seqlen, const int32_t* cand_ids, int32_t topk, int32_t scan_positions, int32_t gen_tokens,
                     int32_t eos_id, float* out, int32_t* gen_ids, void* workspace, size_t workspace_bytes, void* stream) {
  B2S_REQUIRE(m && tokens && cu_seqlens && cand_ids && out && workspace, "encdec_score: null argument");
  B2S_REQUIRE(n_seqs > 0 && total_tokens >= n_seqs && max_seqlen > 0, "encdec_score: empty batch (n_seqs=%d total_tokens=%d)", n_seqs,
              total_tokens);
  B2S_REQUIRE(topk >= 1 && topk <= 5, "encdec_score: topk=%d outside [1,5]", topk);
  B2S_REQUIRE(scan_positions >= 1 && scan_positions <= 64, "encdec_score: max_lookahead=%d outside [1,64]", scan_positions);
  B2S_REQUIRE(gen_tokens >= scan_positions && gen_tokens <= 64, "encdec_score: gen_tokens=%d outside [max_lookahead=%d, 64]", gen_tokens,
              scan_positions);
  B2S_REQUIRE(max_seqlen <= m->d.max_pos && gen_tokens <= m->d.max_pos, "encdec_score: prompt of %d tokens / %d generated exceeds max_pos %d",
              max_seqlen, gen_tokens, m->d.max_pos);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* base = reinterpret_cast<uint8_t*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  EdWorkspace w = carve(m, total_tokens, n_seqs, gen_tokens, base);
  B2S_REQUIRE(w.total + (base - reinterpret_cast<uint8_t*>(workspace)) <= workspace_bytes,
              "encdec_score: workspace of %zu bytes is too small, %zu needed", workspace_bytes, w.total + 256);
  B2S_CHECK_CUDA(cudaSetDevice(m->device));
  w.scan.out = out;
  w.scan.gen_ids = gen_ids;
  if (gen_ids) B2S_CHECK_CUDA(cudaMemsetAsync(gen_ids, 0xFF, 4 * static_cast<size_t>(n_seqs) * gen_tokens, st));
  if (m->bf16 && m->split)
    return score_t<__nv_bfloat16, __nv_bfloat16, __half>(m, w, tokens, cu_seqlens, n_seqs, total_tokens, max_seqlen, cand_ids, topk,
                                                         scan_positions, gen_tokens, eos_id, st);
  if (m->bf16 && m->act_f16)
    return score_t<__half, __half, __half>(m, w, tokens, cu_seqlens, n_seqs, total_tokens, max_seqlen, cand_ids, topk, scan_positions,
                                           gen_tokens, eos_id, st);
  if (m->bf16)
    return score_t<__nv_bfloat16, __nv_bfloat16, __nv_bfloat16>(m, w, tokens, cu_seqlens, n_seqs, total_tokens, max_seqlen, cand_ids,
                                                                topk, scan_positions, gen_tokens, eos_id, st);
  return score_t<float, float, float>(m, w, tokens, cu_seqlens, n_seqs, total_tokens, max_seqlen, cand_ids, topk, scan_positions,
                                      gen_tokens, eos_id, st);
}

int b2s_encdec_score_host(b2s_encdec* m, const int32_t* tokens_h, const int32_t* cu_h, int32_t n_seqs, const int32_t* cand_h,
                          int32_t topk, int32_t scan_positions, int32_t gen_tokens, int32_t eos_id, float* out_h, int32_t* gen_h,
                          void* workspace, size_t workspace_bytes, void* stream) {
  B2S_REQUIRE(m && tokens_h && cu_h && cand_h && out_h && workspace, "encdec_score_host: null argument");
  B2S_REQUIRE(n_seqs > 0, "encdec_score_host: empty batch");
  B2S_REQUIRE(scan_positions >= 1 && scan_positions <= 64, "encdec_score_host: max_lookahead=%d outside [1,64]", scan_positions);
  B2S_REQUIRE(gen_tokens >= scan_positions && gen_tokens <= 64, "encdec_score_host: gen_tokens=%d outside [max_lookahead=%d, 64]",
              gen_tokens, scan_positions);
  const int64_t total = cu_h[n_seqs];
  int max_len = 0;
  B2S_REQUIRE(cu_h[0] == 0, "encdec_score_host: cu_seqlens[0] must be 0");
  for (int i = 0; i < n_seqs; ++i) {
    const int L = cu_h[i + 1] - cu_h[i];
    B2S_REQUIRE(L > 0, "encdec_score_host: prompt %d is empty", i);
    if (L > max_len) max_len = L;
  }
  for (int64_t t = 0; t < total; ++t)
    B2S_REQUIRE(tokens_h[t] >= 0 && tokens_h[t] < m->d.vocab, "encdec_score_host: token %d at %lld outside the vocabulary", tokens_h[t],
                static_cast<long long>(t));
  for (int i = 0; i < 2 * n_seqs; ++i)
    B2S_REQUIRE(cand_h[i] >= 0 && cand_h[i] < m->d.vocab, "encdec_score_host: candidate id %d outside the vocabulary", cand_h[i]);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* base = reinterpret_cast<uint8_t*>(align_up(reinterpret_cast<uintptr_t>(workspace), 256));
  EdWorkspace w = carve(m, total, n_seqs, gen_tokens, base);
  B2S_REQUIRE(w.total + 256 <= workspace_bytes, "encdec_score_host: workspace of %zu bytes is too small, %zu needed", workspace_bytes,
              w.total + 256);
  B2S_CHECK_CUDA(cudaSetDevice(m->device));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.tokens, tokens_h, 4 * total, cudaMemcpyHostToDevice, st));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.cu, cu_h, 4 * (n_seqs + 1), cudaMemcpyHostToDevice, st));
  B2S_CHECK_CUDA(cudaMemcpyAsync(w.cand, cand_h, 8 * n_seqs, cudaMemcpyHostToDevice, st));
  B2S_TRY(b2s_encdec_score(m, w.tokens, w.cu, n_seqs, static_cast<int32_t>(total), max_len, w.cand, topk, scan_positions, gen_tokens,
                           eos_id, w.out, gen_h ? w.gen : nullptr, workspace, workspace_bytes, stream));
  B2S_CHECK_CUDA(cudaMemcpyAsync(out_h, w.out, 32 * static_cast<size_t>(n_seqs), cudaMemcpyDeviceToHost, st));
  if (gen_h)
    B2S_CHECK_CUDA(cudaMemcpyAsync(gen_h, w.gen, 4 * static_cast<size_t>(n_seqs) * gen_tokens, cudaMemcpyDeviceToHost, st));
  int32_t not_finite = 0;
  B2S_CHECK_CUDA(cudaMemcpyAsync(&not_finite, w.scan.n_active + 1, 4, cudaMemcpyDeviceToHost, st));
  B2S_CHECK_CUDA(cudaStreamSynchronize(st));
  if (not_finite > 0) {
    set_error("encdec_score_host: %d of %d prompts produced non-finite logits%s", not_finite, n_seqs,
              m->act_f16 ? " (an fp16 activation overflowed; rerun this model in a bf16 mode)" : "");
    return B2S_ERANGE;
  }
  return B2S_OK;
}

}  // extern "C"
