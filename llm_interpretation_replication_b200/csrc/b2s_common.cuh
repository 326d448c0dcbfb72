// Shared host/device helpers for libb200score (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include <string>

#include "../../include/b200score.h"

namespace b2s {

// ---- error plumbing --------------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern thread_local std::string g_last_error;
extern std::atomic<int64_t> g_launch_count;   // kernels launched by this library, any thread (b2s_launch_count)
extern int g_opt_attn_step;                   // b2s_set_option "attn_step" (b2s_model.cu)
extern float g_opt_int8_threshold;            // b2s_set_option "int8_threshold_milli" (b2s_model.cu)
extern int g_opt_compact;                     // b2s_set_option "lookahead_compact" (b2s_model.cu)

#define B2S_CHECK_CUDA(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      b2s::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B2S_ECUDA;                                                                       \
    }                                                                                         \
  } while (0)

#define B2S_CHECK_LAUNCH()                                                                    \
  do {                                                                                        \
    b2s::g_launch_count.fetch_add(1, std::memory_order_relaxed);                                                                   \
    cudaError_t _e = cudaGetLastError();                                                      \
    if (_e != cudaSuccess) {                                                                  \
      b2s::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return B2S_ECUDA;                                                                       \
    }                                                                                         \
  } while (0)

#define B2S_REQUIRE(cond, ...)                                                                \
  do {                                                                                        \
    if (!(cond)) { b2s::set_error(__VA_ARGS__); return B2S_EINVAL; }                          \
  } while (0)

#define B2S_TRY(expr)                                                                         \
  do { int _r = (expr); if (_r != B2S_OK) return _r; } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// ---- epilogue codes (b200score.h) ---------------------------------------------------------------
enum { EPI_STORE_BF16 = 0, EPI_GELU_BF16 = 1, EPI_RESID_F32 = 2, EPI_LMHEAD = 3, EPI_STORE_F32 = 4, EPI_QKV_ROPE = 5,
       EPI_SWIGLU = 6 /* W rows interleaved (gate_i, up_i): out[M, N/2] = silu(acc[2i]) * acc[2i+1], 16-bit */ };

// LM-head partial record per (row, vocab tile): 8 floats
//   [0] max  [1] sum exp(x-max)  [2..6] top-5 (desc)  [7] argmax column (int bits)
constexpr int LMH_PART = 8;
constexpr int LMH_TILE_N = 256;
constexpr int LMH_STATS = 16;   // floats per row in the public stats layout

// ---- internal launchers (defined across the .cu files) -----------------------------------------
struct LmHeadArgs {
  const int32_t* cand_ids;  // [M,2]
  float* partials;          // [M, tiles_n, LMH_PART]
  float* cand_logits;       // [M,2]
};

// QKV projection fused with RoPE and the KV-cache append (EPI_QKV_ROPE): head_dim 64 = rotary_dim, query heads
// first then key heads then value heads (Falcon layout).  Query heads are rotated and stored to `out`; key heads are
// rotated and stored to the cache slot of their row; value heads are copied to the cache.
struct RopeEpiArgs {
  const int32_t *pos, *slot;     // [M]
  const float *cos, *sin;        // [max_pos, 32]
  void *kcache, *vcache;         // this layer: [slots, nkv*64] 16-bit
  int nh, nkv;
};

// act_f16: A, W and the 16-bit outputs are fp16 instead of bf16 (tcgen05 kind::f16 needs both operands alike)
// a_split: A is [M, 2K] bf16 = [hi | lo] halves of the activations (BF16_SPLIT mode; both halves meet the same W, fp32
// accumulate); out_split: a 16-bit store epilogue writes its result in that form, [M, 2N], for the GEMM that follows.
int gemm_tc_launch(const void* A, const void* W, const float* bias, void* out, int M, int N, int K,
                   int epilogue, const LmHeadArgs* lmh, cudaStream_t st, int act_f16 = 0, const RopeEpiArgs* rope = nullptr,
                   int a_split = 0, int out_split = 0, int out_f16 = -1 /* 16-bit outputs as fp16; -1 = as the operands */,
                   int act_kind = 0 /* EPI_GELU_BF16: B2S_ACT_* applied in the epilogue */);
int gemm_f32_launch(const float* A, const float* W, const float* bias, float* out, int M, int N, int K,
                    int epilogue, cudaStream_t st);

#ifdef __CUDACC__
// ---- device math -------------------------------------------------------------------------------
__device__ __forceinline__ float gelu_erf(float x) {
  // HF "gelu" = x * 0.5 * (1 + erf(x / sqrt(2)))   (HF falcon:537, gpt_neox:43)
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
// GELU for the tensor-core epilogues: x * Phi(x) with Phi(-|x|) = 0.5 erfc(|x|/sqrt 2) from Abramowitz-Stegun 7.1.26
// (|error| < 5e-7 absolute, three orders below the 16-bit rounding that follows; 16 FP instructions instead of 25 for
// erff -- epilogue instructions cost clock under the power cap, profiles/r1_gemm_power.md).  The fp32 check mode
// keeps erff.
__device__ __forceinline__ float gelu_fast(float x) {
  const float ax = fabsf(x) * 0.70710678118654752f;
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, ax, 1.0f)));
  float q = fmaf(1.061405429f, t, -1.453152027f);
  q = fmaf(q, t, 1.421413741f); q = fmaf(q, t, -0.284496736f); q = fmaf(q, t, 0.254829592f); q *= t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-ax * ax * 1.4426950408889634f));
  const float h = 0.5f * q * e;
  return x * (x >= 0.f ? 1.0f - h : h);
}
// MLP activation of the fused "activation + 16-bit store" epilogue: erf-GELU (HF falcon / gpt_neox), ReLU (HF opt), tanh-GELU
// (HF bloom bloom_gelu_forward)
__device__ __forceinline__ float mlp_act(float x, int kind) {
  if (kind == B2S_ACT_RELU) return fmaxf(x, 0.f);
  if (kind == B2S_ACT_GELU_TANH) return x * 0.5f * (1.0f + tanhf(0.79788456f * x * (1.0f + 0.044715f * x * x)));
  return gelu_fast(x);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// two fp32 -> packed 16-bit pair in the activation format of the run (bf16 or fp16)
__device__ __forceinline__ uint32_t pack_act16x2(float lo, float hi, int f16) { return f16 ? pack_f16x2(lo, hi) : pack_bf16x2(lo, hi); }
template <bool F16> __device__ __forceinline__ uint32_t pack16x2(float lo, float hi) {
  if constexpr (F16) return pack_f16x2(lo, hi); else return pack_bf16x2(lo, hi);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// eight consecutive elements (16 bytes of a 16-bit type, 32 bytes of fp32) as floats
__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const __half* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
}

__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void store8(__half* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_f16x2(v[0], v[1]), pack_f16x2(v[2], v[3]), pack_f16x2(v[4], v[5]), pack_f16x2(v[6], v[7]));
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}

// insert v into a descending top-5 list
__device__ __forceinline__ void top5_insert(float (&t)[5], float v) {
  if (v > t[4]) {
    if (v > t[0]) { t[4] = t[3]; t[3] = t[2]; t[2] = t[1]; t[1] = t[0]; t[0] = v; }
    else if (v > t[1]) { t[4] = t[3]; t[3] = t[2]; t[2] = t[1]; t[1] = v; }
    else if (v > t[2]) { t[4] = t[3]; t[3] = t[2]; t[2] = v; }
    else if (v > t[3]) { t[4] = t[3]; t[3] = v; }
    else t[4] = v;
  }
}
#endif

}  // namespace b2s
