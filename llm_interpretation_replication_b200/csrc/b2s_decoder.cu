// Non-GEMM kernels of the decoder forward: batch bookkeeping, embedding gather + LayerNorm,
// RoPE + KV-cache append, varlen causal attention, LM-head reduction/merge and the look-ahead
// scan state machine.  All HBM-bound; templated on the activation type
// (bf16 in BF16_FAST mode, float in FP32_CHECK mode).
#include "b2s_decoder.cuh"
#include <type_traits>

namespace b2s {

// =================================================================================================
// batch bookkeeping: positions, cache slots, last rows  (replaces HF position_ids / cache_position,
// HF falcon:760-775)
// =================================================================================================
__global__ void prep_prefill_kernel(const int32_t* __restrict__ cu, int n_seqs, int lookahead, BatchMeta m) {
  const int i = blockIdx.x;
  if (i >= n_seqs) return;
  const int s = cu[i], e = cu[i + 1];
  const int kv0 = s + i * lookahead;
  if (threadIdx.x == 0) {
    m.q_start[i] = s; m.q_len[i] = e - s; m.kv_start[i] = kv0; m.kv_len[i] = e - s; m.last_row[i] = e - 1;
  }
  for (int t = s + threadIdx.x; t < e; t += blockDim.x) { m.pos[t] = t - s; m.slot[t] = kv0 + (t - s); }
}

// decode step `step` (>= 1): row r holds the token generated at position step-1 of prompt act[r] (act null: prompt r)
__global__ void prep_decode_kernel(const int32_t* __restrict__ cu, int n_rows, int lookahead, int step,
                                   const int32_t* __restrict__ act, BatchMeta m) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const int i = act ? act[r] : r;
  const int L = cu[i + 1] - cu[i];
  const int kv0 = cu[i] + i * lookahead;
  const int p = L + step - 1;
  m.q_start[r] = r; m.q_len[r] = 1; m.kv_start[r] = kv0; m.kv_len[r] = p + 1; m.last_row[r] = r;
  m.pos[r] = p; m.slot[r] = kv0 + p;
}

int launch_prep(const int32_t* cu, int n_rows, int lookahead, int step, const int32_t* act, const BatchMeta& m, cudaStream_t st) {
  if (step == 0) prep_prefill_kernel<<<n_rows, 128, 0, st>>>(cu, n_rows, lookahead, m);
  else prep_decode_kernel<<<static_cast<unsigned>(ceil_div(n_rows, 128)), 128, 0, st>>>(cu, n_rows, lookahead, step, act, m);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// =================================================================================================
// LayerNorm (+ embedding gather, + second affine output, + row gather)
//   HF falcon:578,598,686,744,819; gpt_neox:251-252,321,345,376.  torch.nn.functional.layer_norm:
//   mean, biased variance, (x-mean)*rsqrt(var+eps)*g+b, all in fp32.
// One CTA per row; the row lives in shared memory between the two passes.
// =================================================================================================
__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, float a, float b, float c, float d) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, d));
}
__device__ __forceinline__ void store4(__half* p, float a, float b, float c, float d) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_f16x2(a, b), pack_f16x2(c, d));
}
// store four values; with lo_off != 0 (BF16_SPLIT, bf16 only) also the rounding remainders lo_off elements further on
__device__ __forceinline__ float bf16_rem(float x) { return x - __bfloat162float(__float2bfloat16_rn(x)); }
template <typename T>
__device__ __forceinline__ void store4s(T* p, float a, float b, float c, float d, int lo_off) {
  store4(p, a, b, c, d);
  if constexpr (std::is_same<T, __nv_bfloat16>::value) {
    if (lo_off) store4(p + lo_off, bf16_rem(a), bf16_rem(b), bf16_rem(c), bf16_rem(d));
  }
}
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < nw) ? red[l] : 0.f;
  t = warp_sum(t);
  return t;
}

template <typename AT, typename WT>
__global__ void __launch_bounds__(256)
ln_kernel(LnArgs a) {
  extern __shared__ float srow[];
  __shared__ float red[8];
  const int orow = blockIdx.x;
  const int irow = a.rows ? a.rows[orow] : orow;
  const int d = a.d;
  float sum = 0.f;
  if (a.tokens != nullptr) {
    const WT* e = reinterpret_cast<const WT*>(a.wte) + static_cast<size_t>(a.tokens[irow]) * d;
    float* xr = a.x + static_cast<size_t>(irow) * d;
    const WT* pe = a.wpe ? reinterpret_cast<const WT*>(a.wpe) + static_cast<size_t>(a.pos[irow] + a.pos_offset) * d : nullptr;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
      const float v = pe ? to_f32(e[i]) + to_f32(pe[i]) : to_f32(e[i]);
      srow[i] = v; xr[i] = v; sum += v;
    }
  } else {
    const float4* xr = reinterpret_cast<const float4*>(a.x + static_cast<size_t>(irow) * d);
    for (int i = threadIdx.x; i < d / 4; i += blockDim.x) {
      const float4 v = xr[i];
      reinterpret_cast<float4*>(srow)[i] = v;
      sum += (v.x + v.y) + (v.z + v.w);
    }
  }
  const float row_sum = block_sum(sum, red);                  // (also the barrier that publishes srow)
  const float mean = a.rms ? 0.f : row_sum / d;               // RMSNorm: second moment about zero
  // passes 2 and 3 walk the row four elements at a time (d % 4 == 0): 16-byte smem / gain / bias loads, 8-byte
  // (16-bit outputs) or 16-byte (fp32) stores
  const float4* s4 = reinterpret_cast<const float4*>(srow);
  const int d4 = d >> 2;
  float sq = 0.f;
  for (int i = threadIdx.x; i < d4; i += blockDim.x) {
    const float4 v = s4[i];
    const float c0 = v.x - mean, c1 = v.y - mean, c2 = v.z - mean, c3 = v.w - mean;
    sq += (c0 * c0 + c1 * c1) + (c2 * c2 + c3 * c3);
  }
  const float rstd = rsqrtf(block_sum(sq, red) / d + a.eps);
  const size_t ypitch = static_cast<size_t>(d) * (a.split ? 2 : 1);
  AT* y1 = reinterpret_cast<AT*>(a.y1) + static_cast<size_t>(orow) * ypitch;
  AT* y2 = a.y2 ? reinterpret_cast<AT*>(a.y2) + static_cast<size_t>(orow) * ypitch : nullptr;
  const float4* g1 = reinterpret_cast<const float4*>(a.g1);
  const float4* b1 = reinterpret_cast<const float4*>(a.b1);
  const float4* g2 = reinterpret_cast<const float4*>(a.g2);
  const float4* b2 = reinterpret_cast<const float4*>(a.b2);
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = threadIdx.x; i < d4; i += blockDim.x) {
    const float4 v = s4[i];
    const float n0 = (v.x - mean) * rstd, n1 = (v.y - mean) * rstd, n2 = (v.z - mean) * rstd, n3 = (v.w - mean) * rstd;
    const float4 g = g1[i], b = b1 ? b1[i] : zero;
    store4s(y1 + 4 * i, n0 * g.x + b.x, n1 * g.y + b.y, n2 * g.z + b.z, n3 * g.w + b.w, a.split ? d : 0);
    if (y2) {
      const float4 h = g2[i], c = b2 ? b2[i] : zero;
      store4s(y2 + 4 * i, n0 * h.x + c.x, n1 * h.y + c.y, n2 * h.z + c.z, n3 * h.w + c.w, a.split ? d : 0);
    }
  }
}

template <typename AT, typename WT>
int launch_ln(const LnArgs& a, int n_rows, cudaStream_t st) {
  B2S_REQUIRE(a.d % 4 == 0, "layernorm: d=%d must be a multiple of 4", a.d);
  const size_t smem = static_cast<size_t>(a.d) * sizeof(float);
  if (smem > 48 * 1024) {
    B2S_CHECK_CUDA(cudaFuncSetAttribute(ln_kernel<AT, WT>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  }
  ln_kernel<AT, WT><<<n_rows, 256, smem, st>>>(a);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
template int launch_ln<float, float>(const LnArgs&, int, cudaStream_t);
template int launch_ln<__nv_bfloat16, __nv_bfloat16>(const LnArgs&, int, cudaStream_t);
template int launch_ln<__half, __half>(const LnArgs&, int, cudaStream_t);
template int launch_ln<float, __nv_bfloat16>(const LnArgs&, int, cudaStream_t);   // embedding LayerNorm into the fp32 residual
template int launch_ln<float, __half>(const LnArgs&, int, cudaStream_t);

// =================================================================================================
// SwiGLU gate (Llama family): act = silu(gate) * up on the [rows, 2*ff] output of the fused gate/up projection
//   HF:models/llama/modeling_llama.py LlamaMLP.forward: down_proj(act_fn(gate_proj(x)) * up_proj(x)), act_fn = SiLU.
// HBM-bound: rows * ff * 3 elements; four elements per thread, fp32 math.
// =================================================================================================
template <typename AT>
__global__ void __launch_bounds__(256)
swiglu_kernel(const AT* __restrict__ gu, AT* __restrict__ act, int64_t rows, int ff, int gate_kind) {
  const int ff4 = ff >> 2;
  const int64_t n = rows * ff4;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / ff4;
    const int c = static_cast<int>(i - r * ff4) * 4;
    const AT* g = gu + r * (2 * static_cast<int64_t>(ff)) + c;
    const AT* u = g + ff;
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float x = to_f32(g[k]);
      const float gate = gate_kind == B2S_ACT_GELU_TANH ? x * 0.5f * (1.0f + tanhf(0.79788456f * x * (1.0f + 0.044715f * x * x)))
                                                        : x / (1.f + expf(-x));
      o[k] = gate * to_f32(u[k]);
    }
    store4(act + r * ff + c, o[0], o[1], o[2], o[3]);
  }
}

template <typename AT>
int launch_swiglu(const void* gu, void* act, int64_t rows, int ff, cudaStream_t st, int gate_kind) {
  B2S_REQUIRE(ff % 4 == 0, "swiglu: d_ff=%d must be a multiple of 4", ff);
  const int64_t n = rows * (ff >> 2);
  const int64_t blocks = ceil_div(n, 256);
  const unsigned grid = static_cast<unsigned>(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);
  swiglu_kernel<AT><<<grid, 256, 0, st>>>(reinterpret_cast<const AT*>(gu), reinterpret_cast<AT*>(act), rows, ff, gate_kind);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
template <typename AT>
__global__ void __launch_bounds__(256)
act_kernel(AT* __restrict__ x, int64_t n4, int kind) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    AT* p = x + 4 * i;
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float v = to_f32(p[k]);
      o[k] = kind == 1 ? fmaxf(v, 0.f) : v * 0.5f * (1.0f + tanhf(0.79788456f * v * (1.0f + 0.044715f * v * v)));
    }
    store4(p, o[0], o[1], o[2], o[3]);
  }
}

template <typename AT>
int launch_act(void* x, int64_t n, int kind, cudaStream_t st) {
  B2S_REQUIRE(n % 4 == 0 && (kind == 1 || kind == 2), "activation: n=%lld kind=%d", static_cast<long long>(n), kind);
  const int64_t blocks = ceil_div(n / 4, 256);
  const unsigned grid = static_cast<unsigned>(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);
  act_kernel<AT><<<grid, 256, 0, st>>>(reinterpret_cast<AT*>(x), n / 4, kind);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
// BF16_SPLIT: fp32 pre-activations -> [hi | lo] bf16 activation (see b2s_decoder.cuh)
__global__ void __launch_bounds__(256)
act_split_kernel(const float* __restrict__ pre, __nv_bfloat16* __restrict__ act, int64_t rows, int ff, int kind) {
  const int ff4 = ff >> 2;
  const int64_t n = rows * ff4;
  const bool glu = kind == 0 || kind == 3;
  const int64_t in_pitch = glu ? 2 * static_cast<int64_t>(ff) : ff;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / ff4;
    const int c = static_cast<int>(i - r * ff4) * 4;
    const float4 g = *reinterpret_cast<const float4*>(pre + r * in_pitch + c);
    const float x[4] = {g.x, g.y, g.z, g.w};
    float o[4];
    if (glu) {
      const float4 u4 = *reinterpret_cast<const float4*>(pre + r * in_pitch + ff + c);
      const float u[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
      for (int k = 0; k < 4; ++k)
        o[k] = (kind == 0 ? x[k] / (1.f + expf(-x[k])) : x[k] * 0.5f * (1.0f + tanhf(0.79788456f * x[k] * (1.0f + 0.044715f * x[k] * x[k])))) * u[k];
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        o[k] = kind == 1 ? fmaxf(x[k], 0.f) : x[k] * 0.5f * (1.0f + tanhf(0.79788456f * x[k] * (1.0f + 0.044715f * x[k] * x[k])));
    }
    store4s(act + r * (2 * static_cast<int64_t>(ff)) + c, o[0], o[1], o[2], o[3], ff);
  }
}

int launch_act_split(const float* pre, void* act, int64_t rows, int ff, int kind, cudaStream_t st) {
  B2S_REQUIRE(ff % 4 == 0 && kind >= 0 && kind <= 3, "split activation: d_ff=%d kind=%d", ff, kind);
  const int64_t blocks = ceil_div(rows * (ff >> 2), 256);
  const unsigned grid = static_cast<unsigned>(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);
  act_split_kernel<<<grid, 256, 0, st>>>(pre, reinterpret_cast<__nv_bfloat16*>(act), rows, ff, kind);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

template int launch_act<float>(void*, int64_t, int, cudaStream_t);
template int launch_act<__nv_bfloat16>(void*, int64_t, int, cudaStream_t);
template int launch_act<__half>(void*, int64_t, int, cudaStream_t);

template int launch_swiglu<float>(const void*, void*, int64_t, int, cudaStream_t, int);
template int launch_swiglu<__nv_bfloat16>(const void*, void*, int64_t, int, cudaStream_t, int);
template int launch_swiglu<__half>(const void*, void*, int64_t, int, cudaStream_t, int);

// =================================================================================================
// LLM.int8 emulation: per-prompt outlier columns, per-row absmax int8 round trip (see b2s_decoder.cuh).
//   bitsandbytes MatMul8bitLt.forward: double_quant(A, threshold) -> row statistics without the outlier elements, int8 rows with
//   the outlier columns zeroed, the outlier columns multiplied in fp16 against the dequantised weight columns, the int32
//   product dequantised with SCA[m] * SCB[n] / 127^2.  Here the integer grid is applied to the fp16 operand itself and the
//   tensor core multiplies fp16 x fp16 into fp32: the same numbers up to the fp16 rounding of the dequantised operands
//   (2^-11 relative, 16 times finer than the int8 step).
// HBM-bound elementwise work: x is read twice and written once.
// =================================================================================================
__global__ void __launch_bounds__(256)
int8_colmask_kernel(const __half* __restrict__ x, int K, BatchMeta meta, uint8_t* __restrict__ mask, float thr) {
  const int seq = blockIdx.y;
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) * 2;        // two columns per thread (half2 loads)
  if (k >= K) return;
  const int r0 = meta.q_start[seq], n = meta.q_len[seq];
  bool o0 = false, o1 = false;
  for (int r = 0; r < n; ++r) {
    const float2 v = __half22float2(*reinterpret_cast<const __half2*>(x + static_cast<size_t>(r0 + r) * K + k));
    o0 |= fabsf(v.x) >= thr;
    o1 |= fabsf(v.y) >= thr;
  }
  mask[static_cast<size_t>(seq) * K + k] = o0;
  mask[static_cast<size_t>(seq) * K + k + 1] = o1;
}

// row -> prompt: prompts are packed back to back in row order, so a binary search over q_start finds the owner
__global__ void __launch_bounds__(256)
int8_fakequant_kernel(__half* __restrict__ x, int K, int n_seqs, BatchMeta meta, const uint8_t* __restrict__ mask, float thr) {
  __shared__ float red[8];
  const int row = blockIdx.x;
  int lo = 0, hi = n_seqs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (meta.q_start[mid] <= row) lo = mid; else hi = mid - 1;
  }
  const uint8_t* mk = mask + static_cast<size_t>(lo) * K;
  __half* xr = x + static_cast<size_t>(row) * K;
  float amax = 0.f;
  for (int k = threadIdx.x * 2; k < K; k += blockDim.x * 2) {
    const float2 v = __half22float2(*reinterpret_cast<const __half2*>(xr + k));
    const float a0 = fabsf(v.x), a1 = fabsf(v.y);
    if (a0 < thr) amax = fmaxf(amax, a0);
    if (a1 < thr) amax = fmaxf(amax, a1);
  }
  amax = warp_max(amax);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = amax;
  __syncthreads();
  amax = warp_max(l < (blockDim.x >> 5) ? red[l] : 0.f);
  if (amax == 0.f) return;                                          // an all-zero (or all-outlier) row quantises to itself
  const float q = 127.0f / amax, dq = amax / 127.0f;
  for (int k = threadIdx.x * 2; k < K; k += blockDim.x * 2) {
    float2 v = __half22float2(*reinterpret_cast<const __half2*>(xr + k));
    if (!mk[k]) v.x = rintf(v.x * q) * dq;
    if (!mk[k + 1]) v.y = rintf(v.y * q) * dq;
    *reinterpret_cast<__half2*>(xr + k) = __floats2half2_rn(v.x, v.y);
  }
}

int launch_int8_fakequant(void* x, int rows, int K, int n_seqs, const BatchMeta& meta, uint8_t* mask, float threshold, cudaStream_t st) {
  B2S_REQUIRE(K % 2 == 0 && rows > 0 && n_seqs > 0, "int8 emulation: K=%d rows=%d prompts=%d", K, rows, n_seqs);
  if (!(threshold > 0.f)) threshold = INFINITY;      // bitsandbytes: threshold 0 = no mixed-precision decomposition
  dim3 g1(static_cast<unsigned>(ceil_div(K / 2, 256)), static_cast<unsigned>(n_seqs));
  int8_colmask_kernel<<<g1, 256, 0, st>>>(reinterpret_cast<const __half*>(x), K, meta, mask, threshold);
  B2S_CHECK_LAUNCH();
  int8_fakequant_kernel<<<rows, 256, 0, st>>>(reinterpret_cast<__half*>(x), K, n_seqs, meta, mask, threshold);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// =================================================================================================
// RoPE (rotate_half form on the first rotary_dim dims) + KV-cache append.
//   HF falcon:69-99,154-165 / gpt_neox:126-159: q' = q*cos + rotate_half(q)*sin with cos/sin of shape
//   [pos, rd] = cat(freqs, freqs); here cos/sin tables hold the first rd/2 columns.
// One CTA per token row; q rotated in place, k rotated into the cache, v copied into the cache.
// =================================================================================================
template <typename AT>
__global__ void __launch_bounds__(256)
rope_cache_kernel(RopeArgs a) {
  const int t = blockIdx.x;
  const int pos = a.pos[t];
  const size_t slot = static_cast<size_t>(a.slot[t]);
  AT* row = reinterpret_cast<AT*>(a.qkv) + static_cast<size_t>(t) * a.qkv_dim;
  const int hd = a.hd, half = a.rd / 2;
  const float* cs = a.cos + static_cast<size_t>(pos) * half;
  const float* sn = a.sin + static_cast<size_t>(pos) * half;
  // q heads: pairs (i, i+half)
  const int nq = a.nh * half;
  for (int w = threadIdx.x; w < nq; w += blockDim.x) {
    const int h = w / half, i = w - h * half;
    AT* q = row + a.q_base + h * a.q_hstride;
    const float x1 = to_f32(q[i]), x2 = to_f32(q[i + half]);
    const float c = cs[i], s = sn[i];
    q[i] = from_f32<AT>(x1 * c - x2 * s);
    q[i + half] = from_f32<AT>(x2 * c + x1 * s);
  }
  AT* kc = reinterpret_cast<AT*>(a.kcache) + slot * (a.nkv * hd);
  AT* vc = reinterpret_cast<AT*>(a.vcache) + slot * (a.nkv * hd);
  const int nk = a.nkv * hd;
  for (int w = threadIdx.x; w < nk; w += blockDim.x) {
    const int g = w / hd, i = w - g * hd;
    const AT* k = row + a.k_base + g * a.kv_hstride;
    const AT* v = row + a.v_base + g * a.kv_hstride;
    float kv;
    if (i < half) kv = to_f32(k[i]) * cs[i] - to_f32(k[i + half]) * sn[i];
    else if (i < a.rd) kv = to_f32(k[i]) * cs[i - half] + to_f32(k[i - half]) * sn[i - half];
    else kv = to_f32(k[i]);
    kc[w] = from_f32<AT>(kv);
    vc[w] = v[i];
  }
}

// The same, eight elements (16 bytes of a 16-bit type) at a time: a work item is one 8-wide slice of a rotated pair
// (x[i..i+8), x[i+half..i+half+8)) of a query or key head, of a key head's unrotated tail, or of a value head.  The scalar
// kernel above (2-byte accesses, a division per element) ran at a quarter of the HBM rate: 6 % of a Llama-2-7B pass
// (profiles/r2_llama_launches.summary.txt).  Needs rotary_dim / 2, head_dim and every head offset to be multiples of 8.
template <typename AT>
__global__ void __launch_bounds__(256)
rope_cache_vec_kernel(RopeArgs a) {
  const int t = blockIdx.x;
  const int pos = a.pos[t];
  const size_t slot = static_cast<size_t>(a.slot[t]);
  AT* row = reinterpret_cast<AT*>(a.qkv) + static_cast<size_t>(t) * a.qkv_dim;
  const int hd = a.hd, half = a.rd / 2;
  const int hv = half / 8, tv = (hd - a.rd) / 8, dv = hd / 8;      // 8-wide slices per head: rotated pairs, tail, whole head
  const float* cs = a.cos + static_cast<size_t>(pos) * half;
  const float* sn = a.sin + static_cast<size_t>(pos) * half;
  AT* kc = reinterpret_cast<AT*>(a.kcache) + slot * (a.nkv * hd);
  AT* vc = reinterpret_cast<AT*>(a.vcache) + slot * (a.nkv * hd);
  const int nq = a.nh * hv, nk = a.nkv * hv, nkt = a.nkv * tv, nv = a.nkv * dv;
  for (int w = threadIdx.x; w < nq + nk + nkt + nv; w += blockDim.x) {
    if (w < nq + nk) {                       // a rotated pair of a query head (in place) or of a key head (into the cache)
      const bool is_q = w < nq;
      const int u = is_q ? w : w - nq;
      const int h = u / hv, i = (u - h * hv) * 8;
      const AT* src = row + (is_q ? a.q_base + h * a.q_hstride : a.k_base + h * a.kv_hstride);
      AT* dst = is_q ? row + a.q_base + h * a.q_hstride : kc + h * hd;
      float x1[8], x2[8], c[8], s[8], o1[8], o2[8];
      load8(src + i, x1); load8(src + i + half, x2);
      load8(cs + i, c); load8(sn + i, s);
#pragma unroll
      for (int e = 0; e < 8; ++e) { o1[e] = x1[e] * c[e] - x2[e] * s[e]; o2[e] = x2[e] * c[e] + x1[e] * s[e]; }
      store8(dst + i, o1); store8(dst + i + half, o2);
    } else if (w < nq + nk + nkt) {          // unrotated tail of a key head (partial rotary)
      const int u = w - nq - nk;
      const int h = u / tv, i = a.rd + (u - h * tv) * 8;
      float x[8];
      load8(row + a.k_base + h * a.kv_hstride + i, x);
      store8(kc + h * hd + i, x);
    } else {                                 // value head
      const int u = w - nq - nk - nkt;
      const int h = u / dv, i = (u - h * dv) * 8;
      float x[8];
      load8(row + a.v_base + h * a.kv_hstride + i, x);
      store8(vc + h * hd + i, x);
    }
  }
}

template <typename AT>
int launch_rope_cache(const RopeArgs& a, int n_rows, cudaStream_t st) {
  const bool vec = a.rd > 0 && (a.rd / 2) % 8 == 0 && a.hd % 8 == 0 && a.rd % 8 == 0 && a.qkv_dim % 8 == 0 && a.q_base % 8 == 0 &&
                   a.q_hstride % 8 == 0 && a.k_base % 8 == 0 && a.v_base % 8 == 0 && a.kv_hstride % 8 == 0;
  if (vec) {
    rope_cache_vec_kernel<AT><<<n_rows, 256, 0, st>>>(a);
    B2S_CHECK_LAUNCH();
    return B2S_OK;
  }
  rope_cache_kernel<AT><<<n_rows, 256, 0, st>>>(a);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
template int launch_rope_cache<float>(const RopeArgs&, int, cudaStream_t);
template int launch_rope_cache<__nv_bfloat16>(const RopeArgs&, int, cudaStream_t);
template int launch_rope_cache<__half>(const RopeArgs&, int, cudaStream_t);

// =================================================================================================
// Varlen causal attention over the KV cache (prefill: q_len = L; decode: q_len = 1).
//   HF falcon:349-362 (SDPA, is_causal, scale 1/sqrt(hd)); gpt_neox:162-185,224-238.  fp32 softmax.
// One CTA per (sequence, group of <= ATT_HG query heads sharing one KV head).  K and V of the whole
// sequence for that KV head are staged in shared memory once (rows padded to dodge bank conflicts);
// each warp then owns (query, head) items: lanes split keys for Q.K^T, exchange the softmax statistics
// with shuffles, and split head dims for P.V.
// =================================================================================================
// KV_GLOBAL: the prompt's K/V do not fit in shared memory as fp32 (e.g. head_dim 128 with more than 215 keys: a
// 190-token prompt plus a 50-token completion); keys and values are then read from global memory (L2) instead.
template <typename AT, int HD, bool KV_GLOBAL>
__global__ void __launch_bounds__(ATT_THREADS)
attn_kernel(AttnArgs a) {
  constexpr int HDP = HD + 4;              // padded row (floats)
  constexpr int DPL = (HD + 31) / 32;      // output dims per lane
  extern __shared__ float smem[];
  const int seq = blockIdx.x;
  const int HG = a.heads_per_cta;
  const int groups_per_kv = (a.rep + HG - 1) / HG;
  const int g = blockIdx.y / groups_per_kv;           // kv head
  const int hg = blockIdx.y - g * groups_per_kv;      // head group inside this kv head
  const int h0 = g * a.rep + hg * HG;
  const int nh_here = min(HG, a.rep - hg * HG);
  const int q0 = a.meta.q_start[seq], ql = a.meta.q_len[seq];
  const int kv0 = a.meta.kv_start[seq], kl = a.meta.kv_len[seq];
  const size_t kv_rows = KV_GLOBAL ? 0 : static_cast<size_t>(a.max_kv);
  float* sK = smem;
  float* sV = sK + kv_rows * HDP;
  float* sQ = sV + kv_rows * HDP;                           // [warps][HD]
  float* sP = sQ + (ATT_THREADS / 32) * HD;                 // [warps][32]
  const size_t kv_stride = a.kv_stride ? static_cast<size_t>(a.kv_stride) : static_cast<size_t>(a.nkv) * HD;
  const AT* kc = reinterpret_cast<const AT*>(a.kcache) + static_cast<size_t>(kv0) * kv_stride + g * HD;
  const AT* vc = reinterpret_cast<const AT*>(a.vcache) + static_cast<size_t>(kv0) * kv_stride + g * HD;
  if (!KV_GLOBAL) {
    for (int idx = threadIdx.x; idx < kl * HD; idx += blockDim.x) {
      const int j = idx / HD, dd = idx - j * HD;
      sK[j * HDP + dd] = to_f32(kc[j * kv_stride + dd]);
      sV[j * HDP + dd] = to_f32(vc[j * kv_stride + dd]);
    }
    __syncthreads();
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float* q = sQ + warp * HD;
  float* pbuf = sP + warp * 32;
  const int n_items = ql * nh_here;
  const AT* qkv = reinterpret_cast<const AT*>(a.qkv);
  AT* out = reinterpret_cast<AT*>(a.out);
  for (int item = warp; item < n_items; item += nw) {
    const int qi = item / nh_here, h = h0 + (item - qi * nh_here);
    const size_t row = static_cast<size_t>(q0 + qi);
    const AT* qp = qkv + row * a.qkv_dim + a.q_base + h * a.q_hstride;
    for (int dd = lane; dd < HD; dd += 32) q[dd] = to_f32(qp[dd]) * a.scale;
    __syncwarp();
    const int qpos = kl - ql + qi;                       // position of this query among the keys
    const int nkeys = a.noncausal ? kl : qpos + 1;       // causal: query qi sees keys [0, qpos]
    const float* rb = a.relbias ? a.relbias + static_cast<size_t>(h) * (2 * a.rel_R + 1) + (a.rel_R - qpos) : nullptr;
    float m = -INFINITY, l = 0.f;
    float acc[DPL];
#pragma unroll
    for (int r = 0; r < DPL; ++r) acc[r] = 0.f;
    for (int j0 = 0; j0 < nkeys; j0 += 32) {
      const int j = j0 + lane;
      float s = -INFINITY;
      if (j < nkeys) {
        const float4* qr = reinterpret_cast<const float4*>(q);
        float d0 = 0.f, d1 = 0.f;
        if (KV_GLOBAL) {
          const AT* kg = kc + static_cast<size_t>(j) * kv_stride;
#pragma unroll
          for (int c = 0; c < HD / 4; ++c) {      // same summation order as the shared-memory path
            const float4 qq = qr[c];
            d0 = fmaf(to_f32(kg[4 * c]), qq.x, d0); d1 = fmaf(to_f32(kg[4 * c + 1]), qq.y, d1);
            d0 = fmaf(to_f32(kg[4 * c + 2]), qq.z, d0); d1 = fmaf(to_f32(kg[4 * c + 3]), qq.w, d1);
          }
        } else {
          const float4* kr = reinterpret_cast<const float4*>(sK + j * HDP);
#pragma unroll
          for (int c = 0; c < HD / 4; ++c) {
            const float4 kk = kr[c], qq = qr[c];
            d0 = fmaf(kk.x, qq.x, d0); d1 = fmaf(kk.y, qq.y, d1);
            d0 = fmaf(kk.z, qq.z, d0); d1 = fmaf(kk.w, qq.w, d1);
          }
        }
        s = d0 + d1;
        if (a.alibi) s += a.alibi[h] * static_cast<float>(j);     // HF bloom: alibi = slope[head] * key position
        if (rb) s += rb[j];                                        // HF t5: position_bias[head][query, key]
      }
      const float mn = fmaxf(m, warp_max(s));
      const float p = (j < nkeys) ? __expf(s - mn) : 0.f;
      const float corr = __expf(m - mn);
      l = l * corr + warp_sum(p);
      m = mn;
      pbuf[lane] = p;
      __syncwarp();
      const int cnt = min(32, nkeys - j0);
#pragma unroll
      for (int r = 0; r < DPL; ++r) {
        const int dd = lane + 32 * r;
        float o = acc[r] * corr;
        if (dd < HD) {
          if (KV_GLOBAL) {
            for (int jj = 0; jj < cnt; ++jj) o = fmaf(pbuf[jj], to_f32(vc[static_cast<size_t>(j0 + jj) * kv_stride + dd]), o);
          } else {
            for (int jj = 0; jj < cnt; ++jj) o = fmaf(pbuf[jj], sV[(j0 + jj) * HDP + dd], o);
          }
        }
        acc[r] = o;
      }
      __syncwarp();
    }
    const float inv = 1.f / l;
    const size_t owidth = static_cast<size_t>(a.nh) * HD;
    AT* op = out + row * owidth + h * HD;
    // BF16_SPLIT: the dense projection reads the row as [hi | lo] bf16 halves, whatever format Q/K/V are stored in
    __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(a.out) + row * (2 * owidth) + h * HD;
#pragma unroll
    for (int r = 0; r < DPL; ++r) {
      const int dd = lane + 32 * r;
      if (dd < HD) {
        const float o = acc[r] * inv;
        if (a.split) {
          const __nv_bfloat16 hi = __float2bfloat16_rn(o);
          ob[dd] = hi;
          ob[owidth + dd] = __float2bfloat16_rn(o - __bfloat162float(hi));
        } else {
          op[dd] = from_f32<AT>(o);
        }
      }
    }
    __syncwarp();
  }
}

template <typename AT>
int launch_attn(const AttnArgs& a_in, int n_seqs, cudaStream_t st) {
  AttnArgs a = a_in;
  // single-token steps: one CTA per (sequence, KV head) stages K/V once for every query head that shares it
  if (a.heads_per_cta <= 0) a.heads_per_cta = ATT_HG;
  if (a.heads_per_cta > a.rep) a.heads_per_cta = a.rep;
  const int hd = a.hd;
  size_t smem = (2 * static_cast<size_t>(a.max_kv) * (hd + 4) + (ATT_THREADS / 32) * (hd + 32)) * sizeof(float);
  const bool kv_global = smem > 227 * 1024;     // K/V too long to stage: read them through L2
  if (kv_global) smem = static_cast<size_t>(ATT_THREADS / 32) * (hd + 32) * sizeof(float);
  const int groups_per_kv = (a.rep + a.heads_per_cta - 1) / a.heads_per_cta;
  dim3 grid(n_seqs, a.nkv * groups_per_kv);
#define B2S_ATT_CASE(HDV)                                                                                         \
  case HDV:                                                                                                      \
    if (kv_global) {                                                                                             \
      attn_kernel<AT, HDV, true><<<grid, ATT_THREADS, smem, st>>>(a);                                            \
    } else {                                                                                                     \
      B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_kernel<AT, HDV, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          static_cast<int>(smem)));                                               \
      attn_kernel<AT, HDV, false><<<grid, ATT_THREADS, smem, st>>>(a);                                           \
    }                                                                                                            \
    break;
  switch (hd) {
    B2S_ATT_CASE(64)
    B2S_ATT_CASE(80)
    B2S_ATT_CASE(128)
    default: set_error("attention: unsupported head_dim %d (64, 80, 128)", hd); return B2S_EUNSUPPORTED;
  }
#undef B2S_ATT_CASE
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
template int launch_attn<float>(const AttnArgs&, int, cudaStream_t);
template int launch_attn<__nv_bfloat16>(const AttnArgs&, int, cudaStream_t);
template int launch_attn<__half>(const AttnArgs&, int, cudaStream_t);

// =================================================================================================
// One-token attention, streaming form (q_len == 1 for every sequence): HF falcon:349-362 / gpt_neox:224-238 / llama / bloom with
// one query row.  HBM-bound: each step re-reads the K and V rows of every sequence.  One warp per (sequence, head), four heads
// per CTA (with grouped-query heads they share a KV head and its rows come from L1 after the first warp).  LPK lanes share a
// key (16-byte slices of its K row and of its V row: 8 lanes for head_dim 64, 16 for 80 / 128), the 32 / LPK lane groups walk
// the keys in interleaved streams with an online softmax each and are merged with shuffles at the end; the slices reach the
// lanes through a cp.async ring in shared memory (three iterations in flight per warp), no barrier.  The varlen kernel above, which stages the whole K and V of a sequence in shared memory as fp32 before one warp per
// head uses them, took 10x longer on these steps (profiles/r2_t5_launches_before.summary.txt: 472 us vs 47 us per layer).
// =================================================================================================
template <typename AT, int HD>
__global__ void __launch_bounds__(128)
attn_step_kernel(AttnArgs a) {
  constexpr int LPK = HD == 64 ? 8 : 16;       // lanes per key
  constexpr int NG = 32 / LPK;                 // key streams per warp
  constexpr int NSUB = HD / 8;                 // lanes of a group that hold data (10 of 16 for head_dim 80)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x * 4 + warp, seq = blockIdx.y;
  if (h >= a.nh) return;
  const int sub = lane % LPK, grp = lane / LPK;
  const bool live = sub < NSUB;
  const int L = a.meta.kv_len[seq];
  const size_t row = static_cast<size_t>(a.meta.q_start[seq]);
  const size_t pitch = a.kv_stride ? static_cast<size_t>(a.kv_stride) : static_cast<size_t>(a.nkv) * HD;
  const size_t kv_off = static_cast<size_t>(a.meta.kv_start[seq]) * pitch + static_cast<size_t>(h / a.rep) * HD + 8 * (live ? sub : 0);
  const AT* kbase = reinterpret_cast<const AT*>(a.kcache) + kv_off;
  const AT* vbase = reinterpret_cast<const AT*>(a.vcache) + kv_off;
  float q[8];
  load8(reinterpret_cast<const AT*>(a.qkv) + row * a.qkv_dim + a.q_base + h * a.q_hstride + 8 * (live ? sub : 0), q);
#pragma unroll
  for (int i = 0; i < 8; ++i) q[i] = live ? q[i] * a.scale : 0.f;
  const float slope = a.alibi ? a.alibi[h] : 0.f;
  float m = -INFINITY, l = 0.f;
  float o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] = 0.f;
  // K / V slices travel through a per-warp ring in shared memory (cp.async, 16 bytes per lane and key for K and for V): STG
  // iterations of U keys per lane group are in flight while one is consumed -- 6 KB per warp instead of the 2 KB that fit in
  // registers.  A lane copies exactly the slices it consumes itself, so the only synchronisation is its own wait_group.
  constexpr int U = 2, STG = 3;
  constexpr int VB = 8 * sizeof(AT);                       // bytes of one 8-element slice
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
          const AT* kp = kbase + static_cast<size_t>(j) * pitch;
          const AT* vp = vbase + static_cast<size_t>(j) * pitch;
#pragma unroll
          for (int c = 0; c < VB / 16; ++c) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (2 * u) * 32 * VB + 16 * c),
                         "l"(reinterpret_cast<const uint8_t*>(kp) + 16 * c) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (2 * u + 1) * 32 * VB + 16 * c),
                         "l"(reinterpret_cast<const uint8_t*>(vp) + 16 * c) : "memory");
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");     // (an empty group keeps the count uniform)
  };
#pragma unroll
  for (int s0 = 0; s0 < STG; ++s0) issue(s0);
  for (int it = 0; it < n_iter; ++it) {
    asm volatile("cp.async.wait_group %0;" ::"n"(STG - 1) : "memory");
    const AT* slot = reinterpret_cast<const AT*>(ring + static_cast<size_t>(it % STG) * (U * 2 * 32 * VB) + static_cast<size_t>(lane) * VB);
    const int j0 = it * (U * NG);
    float k[U][8], v[U][8], d[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = j0 + NG * u + grp;
      if (j < L) {
        load8(slot + (2 * u) * 32 * 8, k[u]);
        load8(slot + (2 * u + 1) * 32 * 8, v[u]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) { k[u][i] = 0.f; v[u][i] = 0.f; }
      }
    }
    float dmax = -INFINITY;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t = fmaf(q[i], k[u][i], t);      // (q is zero on the idle lanes of a head_dim-80 group)
#pragma unroll
      for (int s = 1; s < LPK; s <<= 1) t += __shfl_xor_sync(0xffffffffu, t, s);
      const int j = j0 + NG * u + grp;
      t += slope * static_cast<float>(j);                           // HF bloom: alibi = slope[head] * key position
      d[u] = j < L ? t : -INFINITY;
      dmax = fmaxf(dmax, d[u]);
    }
    const float mn = fmaxf(m, dmax);
    if (mn > -INFINITY) {
      const float corr = __expf(m - mn);
      float psum = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] *= corr;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float pu = __expf(d[u] - mn);
        psum += pu;
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaf(pu, v[u][i], o[i]);
      }
      l = fmaf(l, corr, psum);
      m = mn;
    }
    issue(it + STG);       // refill the slot just consumed (its values are in registers: the loads above have returned)
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  float mall = m;
#pragma unroll
  for (int s = LPK; s < 32; s <<= 1) mall = fmaxf(mall, __shfl_xor_sync(0xffffffffu, mall, s));
  const float w = m > -INFINITY ? __expf(m - mall) : 0.f;
  l *= w;
#pragma unroll
  for (int s = LPK; s < 32; s <<= 1) l += __shfl_xor_sync(0xffffffffu, l, s);
  const float inv = 1.f / l;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    o[i] *= w;
#pragma unroll
    for (int s = LPK; s < 32; s <<= 1) o[i] += __shfl_xor_sync(0xffffffffu, o[i], s);
    o[i] *= inv;
  }
  if (grp == 0 && live) {
    const size_t owidth = static_cast<size_t>(a.nh) * HD;
    const size_t col = static_cast<size_t>(h) * HD + 8 * sub;
    if (a.split) {
      __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(a.out) + row * (2 * owidth) + col;
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 hv = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
        const __nv_bfloat162 lv = __floats2bfloat162_rn(o[2 * i] - __low2float(hv), o[2 * i + 1] - __high2float(hv));
        hi[i] = *reinterpret_cast<const uint32_t*>(&hv);
        lo[i] = *reinterpret_cast<const uint32_t*>(&lv);
      }
      *reinterpret_cast<uint4*>(ob) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(ob + owidth) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    } else {
      AT* op = reinterpret_cast<AT*>(a.out) + row * owidth + col;
      if constexpr (std::is_same<AT, float>::value) {
        *reinterpret_cast<float4*>(op) = make_float4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<float4*>(op + 4) = make_float4(o[4], o[5], o[6], o[7]);
      } else if constexpr (std::is_same<AT, __half>::value) {
        *reinterpret_cast<uint4*>(op) = make_uint4(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3]), pack_f16x2(o[4], o[5]), pack_f16x2(o[6], o[7]));
      } else {
        *reinterpret_cast<uint4*>(op) = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      }
    }
  }
}

template <typename AT>
int launch_attn_step(const AttnArgs& a, int n_seqs, cudaStream_t st) {
  B2S_REQUIRE(a.relbias == nullptr && !a.noncausal, "attn_step: the relative-bias / non-causal forms belong to the varlen kernel");
  B2S_REQUIRE(a.q_base % 8 == 0 && a.q_hstride % 8 == 0 && a.qkv_dim % 8 == 0, "attn_step: query heads must start at 16-byte boundaries");
  dim3 grid(static_cast<unsigned>(ceil_div(a.nh, 4)), static_cast<unsigned>(n_seqs));
  const size_t smem = 4 * 3 * 2 * 2 * 32 * 8 * sizeof(AT);      // 4 warps x STG x U x (K, V) x 32 lanes x one 8-element slice
  switch (a.hd) {
    case 64: attn_step_kernel<AT, 64><<<grid, 128, smem, st>>>(a); break;
    case 80: attn_step_kernel<AT, 80><<<grid, 128, smem, st>>>(a); break;
    case 128: attn_step_kernel<AT, 128><<<grid, 128, smem, st>>>(a); break;
    default: set_error("attn_step: unsupported head_dim %d (64, 80, 128)", a.hd); return B2S_EUNSUPPORTED;
  }
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
template int launch_attn_step<float>(const AttnArgs&, int, cudaStream_t);
template int launch_attn_step<__nv_bfloat16>(const AttnArgs&, int, cudaStream_t);
template int launch_attn_step<__half>(const AttnArgs&, int, cudaStream_t);

// =================================================================================================
// LM head: reduce full fp32 logits to per-tile partials (FP32_CHECK mode; the bf16 path produces the
// same partial records inside the GEMM epilogue), merge partials per row, then advance the look-ahead
// scan state.
// =================================================================================================
// grid (tiles_n, M); one warp per (row, 256-column tile)
__global__ void __launch_bounds__(32)
lmhead_partials_from_logits_kernel(const float* __restrict__ logits, int M, int V, int tiles_n,
                                   const int32_t* __restrict__ cand, float* __restrict__ partials,
                                   float* __restrict__ cand_logits) {
  const int nb = blockIdx.x, row = blockIdx.y, lane = threadIdx.x;
  const float* lr = logits + static_cast<size_t>(row) * V;
  float m = -INFINITY, best = -INFINITY;
  int arg = 0;
  float top[5] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
  const int c0 = nb * LMH_TILE_N;
  const int ca = cand[2 * row], cb = cand[2 * row + 1];
  for (int c = c0 + lane; c < min(c0 + LMH_TILE_N, V); c += 32) {
    const float x = lr[c];
    if (x > best) { best = x; arg = c; }
    m = fmaxf(m, x);
    top5_insert(top, x);
    if (c == ca) cand_logits[2 * row] = x;
    if (c == cb) cand_logits[2 * row + 1] = x;
  }
  const float wm = warp_max(m);
  float s = 0.f;
  for (int c = c0 + lane; c < min(c0 + LMH_TILE_N, V); c += 32) s += expf(lr[c] - wm);
  s = warp_sum(s);
  // merge top-5 and argmax across lanes through shared memory
  __shared__ float st[32 * 5];
  __shared__ float sb[32];
  __shared__ int sa[32];
#pragma unroll
  for (int j = 0; j < 5; ++j) st[lane * 5 + j] = top[j];
  sb[lane] = best; sa[lane] = arg;
  __syncwarp();
  if (lane == 0) {
    float t5[5] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < 160; ++i) top5_insert(t5, st[i]);
    float bb = -INFINITY; int ba = 0;
    for (int i = 0; i < 32; ++i) if (sb[i] > bb || (sb[i] == bb && sa[i] < ba)) { bb = sb[i]; ba = sa[i]; }
    float* o = partials + (static_cast<size_t>(row) * tiles_n + nb) * LMH_PART;
    o[0] = wm; o[1] = s;
    for (int j = 0; j < 5; ++j) o[2 + j] = t5[j];
    o[7] = __int_as_float(ba);
  }
}

int launch_lmhead_partials_from_logits(const float* logits, int M, int V, const int32_t* cand, float* partials,
                                       float* cand_logits, cudaStream_t st) {
  const int tiles_n = static_cast<int>(ceil_div(V, LMH_TILE_N));
  dim3 grid(tiles_n, M);
  lmhead_partials_from_logits_kernel<<<grid, 32, 0, st>>>(logits, M, V, tiles_n, cand, partials, cand_logits);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// One warp per row: merge the per-tile partial records into the public 16-float stats row
//   [lse, top1..top5, argmax, logit_a, logit_b, 0..]
__global__ void __launch_bounds__(128)
lmhead_merge_kernel(const float* __restrict__ partials, const float* __restrict__ cand_logits, int M, int tiles_n,
                    float* __restrict__ stats) {
  __shared__ float st[4][32 * 5];
  __shared__ float sb[4][32];
  __shared__ int sa[4][32];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + w;
  if (row >= M) return;
  float m = -INFINITY, best = -INFINITY;
  int arg = 0x7fffffff;
  float top[5] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
  const float* pr = partials + static_cast<size_t>(row) * tiles_n * LMH_PART;
  for (int t = lane; t < tiles_n; t += 32) m = fmaxf(m, pr[t * LMH_PART]);
  m = warp_max(m);
  float s = 0.f;
  for (int t = lane; t < tiles_n; t += 32) {
    const float* p = pr + t * LMH_PART;
    s += p[1] * expf(p[0] - m);
    const int a = __float_as_int(p[7]);
    if (p[2] > best || (p[2] == best && a < arg)) { best = p[2]; arg = a; }   // p[2] is the tile max
#pragma unroll
    for (int j = 0; j < 5; ++j) top5_insert(top, p[2 + j]);
  }
  s = warp_sum(s);
#pragma unroll
  for (int j = 0; j < 5; ++j) st[w][lane * 5 + j] = top[j];
  sb[w][lane] = best; sa[w][lane] = arg;
  __syncwarp();
  if (lane == 0) {
    float t5[5] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < 160; ++i) top5_insert(t5, st[w][i]);
    float bb = -INFINITY; int ba = 0x7fffffff;
    for (int i = 0; i < 32; ++i) if (sb[w][i] > bb || (sb[w][i] == bb && sa[w][i] < ba)) { bb = sb[w][i]; ba = sa[w][i]; }
    float* o = stats + static_cast<size_t>(row) * LMH_STATS;
    o[0] = m + logf(s);
    for (int j = 0; j < 5; ++j) o[1 + j] = t5[j];
    o[6] = static_cast<float>(ba);
    o[7] = cand_logits[2 * row];
    o[8] = cand_logits[2 * row + 1];
    for (int j = 9; j < LMH_STATS; ++j) o[j] = 0.f;
    reinterpret_cast<int*>(o)[15] = ba;   // exact integer copy of the argmax
  }
}

int launch_lmhead_merge(const float* partials, const float* cand_logits, int M, int tiles_n, float* stats,
                        cudaStream_t st) {
  lmhead_merge_kernel<<<static_cast<unsigned>(ceil_div(M, 4)), 128, 0, st>>>(partials, cand_logits, M, tiles_n, stats);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// Look-ahead scan (reference analysis/run_base_vs_instruct_100q.py:348-363; compare_instruct_models.py:
// 249-272) as a per-prompt state machine advanced once per scored position:
//   - a prompt is "alive" at position s if no EOS was generated before s (HF generate stops there, so
//     outputs.scores has no entry for later positions);
//   - first alive position where either candidate logit is >= the k-th largest logit wins;
//   - otherwise the position-0 probabilities stand.
// The same pass compacts the look-ahead: the prompts that need another position (unresolved and alive; in completion
// mode every prompt that has not produced EOS) are written in ascending order to act_out with their next token and
// candidate pair, so the next one-token step runs exactly those rows.  One CTA walks the rows in chunks of its size
// (a batch holds a few thousand prompts; the work per row is a dozen loads) and keeps the running count, which makes
// the order stable without a second kernel.
// A prompt whose logits are not finite at a position it is scored on (an fp16 activation overflowed somewhere in the
// decoder) gets an all-NaN row with found = -1, is counted in n_active[1] and takes no further steps.
constexpr int SCAN_THREADS = 1024;
__global__ void __launch_bounds__(SCAN_THREADS)
scan_update_kernel(const float* __restrict__ stats, int n_rows, int pos, int topk, int eos_id,
                   int scan_positions, int gen_tokens, ScanState s) {
  __shared__ int warp_tot[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int base = 0, bad_total = 0;
  for (int c0 = 0; c0 < n_rows; c0 += SCAN_THREADS) {
    const int r = c0 + threadIdx.x;
    int need = 0, i = 0, arg = 0, bad = 0;
    if (r < n_rows) {
      i = s.act_in ? s.act_in[r] : r;
      const float* q = stats + static_cast<size_t>(r) * LMH_STATS;
      arg = reinterpret_cast<const int*>(q)[15];
      int alive = (pos == 0) ? 1 : s.alive[i];
      int resolved = (pos == 0) ? 0 : s.resolved[i];
      if (alive && !resolved && pos < scan_positions) {
        const float lse = q[0], la = q[7], lb = q[8], kth = q[topk];   // q[1..5] = top1..top5
        float* o = s.out + static_cast<size_t>(i) * 8;
        if (!(isfinite(lse) && isfinite(la) && isfinite(lb))) {
          const float nan = __int_as_float(0x7fc00000);
          o[0] = nan; o[1] = nan; o[2] = nan; o[3] = nan; o[4] = nan; o[5] = nan; o[6] = static_cast<float>(pos); o[7] = -1.f;
          resolved = 1; alive = 0; bad = 1;
        } else {
          const float pa = expf(la - lse), pb = expf(lb - lse);
          const bool hit = (la >= kth) || (lb >= kth);
          if (pos == 0 || hit) {
            o[0] = pa; o[1] = pb; o[2] = pa / (pa + pb); o[3] = lse; o[4] = la; o[5] = lb;
            o[6] = static_cast<float>(pos); o[7] = hit ? 1.f : 0.f;
          }
          if (hit) resolved = 1;
        }
      }
      if (s.gen_ids) s.gen_ids[static_cast<size_t>(i) * gen_tokens + pos] = alive ? arg : -1;
      s.next_tok[i] = arg;
      if (alive && arg == eos_id) alive = 0;
      s.alive[i] = alive; s.resolved[i] = resolved;
      // another step is needed while a prompt is unresolved (scoring only), or, in completion mode (gen_tokens >
      // scan_positions: the reference's max_new_tokens=50), while it has not produced EOS
      need = (alive && (gen_tokens > scan_positions ? pos + 1 < gen_tokens : !resolved)) ? 1 : 0;
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, need != 0);
    const unsigned bad_ballot = __ballot_sync(0xffffffffu, bad != 0);
    __syncthreads();                       // warp_tot of the previous chunk has been read by everyone
    if (lane == 0) warp_tot[w] = __popc(ballot) | (__popc(bad_ballot) << 16);
    __syncthreads();
    int before = 0, total = 0;
    for (int k = 0; k < SCAN_THREADS / 32; ++k) {
      const int t = warp_tot[k];
      if (k < w) before += t & 0xffff;
      total += t & 0xffff;
      bad_total += t >> 16;
    }
    if (need) {
      const int o = base + before + __popc(ballot & ((1u << lane) - 1u));
      s.act_out[o] = i; s.tok_out[o] = arg;
      s.cand_out[2 * o] = s.cand[2 * i]; s.cand_out[2 * o + 1] = s.cand[2 * i + 1];
    }
    base += total;
  }
  if (threadIdx.x == 0) { s.n_active[0] = base; if (bad_total) atomicAdd(s.n_active + 1, bad_total); }
}

int launch_scan_update(const float* stats, int n_rows, int pos, int topk, int eos_id, int scan_positions, int gen_tokens,
                       const ScanState& s, cudaStream_t st) {
  scan_update_kernel<<<1, SCAN_THREADS, 0, st>>>(stats, n_rows, pos, topk, eos_id, scan_positions, gen_tokens, s);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // namespace b2s
