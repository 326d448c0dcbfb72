// tcgen05 varlen causal attention for the prompt pass, second generation: P stays in tensor memory.
//
// Same contract as attn_tc_kernel's prefill mode (b2s_attn_tc.cu; HF:models/falcon/modeling_falcon.py:349-362):
// head_dim 64, <= 256 keys per prompt, one CTA per (prompt, 128-query tile, KV head) looping over the query heads
// that share the KV head.  What changed, and why (profiles/r1_attn_ncu.md): the first kernel spent its time in the
// 256 softmax threads (two per scheduler, latency-bound) and in moving P through shared memory.  Here
//   * P is written back over S in tensor memory (tcgen05.st, two 16-bit probabilities per column) and P.V reads its
//     A operand from tensor memory (tcgen05.mma [d], [a_tmem], b_desc): no shared-memory P tile, no proxy fence;
//   * O overlays the dead upper half of S, so an item needs max(keys, 128) columns and FOUR items are in flight
//     for <= 128 keys (two for <= 256): 16 softmax warps, four per scheduler;
//   * a chunk of 32 keys that every row of the warp may see skips the causal compare; chunks no row may see are
//     not computed (their P columns are zero-filled);
//   * the row sum is taken in fp32 before rounding; the 16-bit format is a template parameter, not a branch.
// The issuing thread polls: S of the next item when its lane is free and its Q tile has landed, P.V of the oldest
// item whose P is written.
#include "b2s_tc.cuh"
#include "b2s_decoder.cuh"

namespace b2s {

static constexpr int A2_HD = 64;
static constexpr int A2_BM = 128;
static constexpr int A2_MAXKV = 256;
static constexpr int A2_KV_BYTES = A2_MAXKV * A2_HD * 2;     // 32 KB per K or V
static constexpr int A2_Q_BYTES = A2_BM * A2_HD * 2;         // 16 KB per Q tile
static constexpr int A2_NQ = 6;                              // Q ring depth
static constexpr int A2_LANES = 4;
// head_dim 128 (StableLM-alpha, RedPajama-7B, Mistral, BLOOM ...): K, V and Q are staged as two 64-column sub-tiles each (a
// 128B-swizzled K-major tile is 64 elements wide), S accumulates over both, O = P.V is two N = 64 products side by side.
// K + V take 128 KB, so the Q ring is 3 deep; an item needs 256 tensor-memory columns (P over S[0,128), O in [128,256)):
// two items in flight whatever the key count.
template <int HD> constexpr int a2_nq() { return HD == 64 ? A2_NQ : 3; }
template <int HD> constexpr int a2_q_bytes() { return A2_BM * HD * 2; }
template <int HD> constexpr int a2_kv_bytes() { return A2_MAXKV * HD * 2; }
template <int HD> constexpr int a2_smem() { return 2 * a2_kv_bytes<HD>() + a2_nq<HD>() * a2_q_bytes<HD>() + 1024 + 256; }
static constexpr int A2_THREADS = A2_LANES * 128 + 32;       // 16 softmax warps + the issuer warp
static constexpr uint32_t A2_TMEM_COLS = 512;

struct Attn2Params {
  BatchMeta meta;
  uint16_t* out;          // [T, nh*64] 16-bit ([T, 2*nh*64] = hi | lo when split)
  int split;              // BF16_SPLIT: also store the bf16 rounding remainders nh*64 elements after the values
  int nh, nkv, rep;
  int q_base, q_hstride;
  float scale_log2e;
  const float* alibi;     // per-head slopes (ALIBI kernels), or null
};

__device__ __forceinline__ uint32_t a2_idesc(int m, int n, int b_mn, bool f16) {
  const uint32_t fmt = f16 ? 0u : 1u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}
__device__ __forceinline__ uint64_t a2_vdesc(uint32_t smem_addr) {   // V [keys, 64] as the MN-major B operand
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
__device__ __forceinline__ float a2_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void a2_mma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 16 (or 8) consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void a2_tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void a2_tmem_st8(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
      : "memory");
}
__device__ __forceinline__ void a2_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// wait without reading the clock every spin (try_wait itself suspends the thread for a while)
__device__ __forceinline__ void a2_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 27)) {
      printf("b2s attn_tc2: mbarrier wait timed out (block %d,%d thread %d bar %u parity %u)\n", blockIdx.x, blockIdx.y,
             threadIdx.x, bar, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ bool a2_test(uint32_t bar, uint32_t parity) {   // non-blocking probe
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

// One item for one thread (== tile row): wait for S, two passes over this row's S in tensor memory (max, then
// exp / pack / store P over S), hand P to the issuer, wait for O, normalise and write 64 outputs.
// a row of 64 outputs (two 32-column chunks of O) normalised and stored in the run's 16-bit format, or, with lo_off != 0, as
// bf16 [hi | lo] halves (BF16_SPLIT: the dense projection reads this row that way whatever format Q/K/V/P had)
template <bool F16>
__device__ __forceinline__ void a2_store64(const float (&o0)[32], const float (&o1)[32], float inv, uint16_t* out_row, int lo_off) {
  uint4* dst = reinterpret_cast<uint4*>(out_row);
  if (lo_off) {
    uint4* dlo = reinterpret_cast<uint4*>(out_row + lo_off);
    auto hi = [inv](float x) { return x * inv; };
    auto rem = [inv](float x) { const float y = x * inv; return y - __bfloat162float(__float2bfloat16_rn(y)); };
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      dst[q] = make_uint4(pack_bf16x2(hi(o0[8 * q]), hi(o0[8 * q + 1])), pack_bf16x2(hi(o0[8 * q + 2]), hi(o0[8 * q + 3])),
                          pack_bf16x2(hi(o0[8 * q + 4]), hi(o0[8 * q + 5])), pack_bf16x2(hi(o0[8 * q + 6]), hi(o0[8 * q + 7])));
      dlo[q] = make_uint4(pack_bf16x2(rem(o0[8 * q]), rem(o0[8 * q + 1])), pack_bf16x2(rem(o0[8 * q + 2]), rem(o0[8 * q + 3])),
                          pack_bf16x2(rem(o0[8 * q + 4]), rem(o0[8 * q + 5])), pack_bf16x2(rem(o0[8 * q + 6]), rem(o0[8 * q + 7])));
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      dst[4 + q] = make_uint4(pack_bf16x2(hi(o1[8 * q]), hi(o1[8 * q + 1])), pack_bf16x2(hi(o1[8 * q + 2]), hi(o1[8 * q + 3])),
                              pack_bf16x2(hi(o1[8 * q + 4]), hi(o1[8 * q + 5])), pack_bf16x2(hi(o1[8 * q + 6]), hi(o1[8 * q + 7])));
      dlo[4 + q] = make_uint4(pack_bf16x2(rem(o1[8 * q]), rem(o1[8 * q + 1])), pack_bf16x2(rem(o1[8 * q + 2]), rem(o1[8 * q + 3])),
                              pack_bf16x2(rem(o1[8 * q + 4]), rem(o1[8 * q + 5])), pack_bf16x2(rem(o1[8 * q + 6]), rem(o1[8 * q + 7])));
    }
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      dst[q] = make_uint4(pack16x2<F16>(o0[8 * q] * inv, o0[8 * q + 1] * inv), pack16x2<F16>(o0[8 * q + 2] * inv, o0[8 * q + 3] * inv),
                          pack16x2<F16>(o0[8 * q + 4] * inv, o0[8 * q + 5] * inv), pack16x2<F16>(o0[8 * q + 6] * inv, o0[8 * q + 7] * inv));
#pragma unroll
    for (int q = 0; q < 4; ++q)
      dst[4 + q] = make_uint4(pack16x2<F16>(o1[8 * q] * inv, o1[8 * q + 1] * inv), pack16x2<F16>(o1[8 * q + 2] * inv, o1[8 * q + 3] * inv),
                              pack16x2<F16>(o1[8 * q + 4] * inv, o1[8 * q + 5] * inv), pack16x2<F16>(o1[8 * q + 6] * inv, o1[8 * q + 7] * inv));
  }
}

// the first 16 of a chunk's 32 columns (columns 64..79 of an 80-wide head)
template <bool F16>
__device__ __forceinline__ void a2_store16(const float (&o)[32], float inv, uint16_t* out_row, int lo_off) {
  uint4* dst = reinterpret_cast<uint4*>(out_row);
  if (lo_off) {
    uint4* dlo = reinterpret_cast<uint4*>(out_row + lo_off);
    auto rem = [inv](float x) { const float y = x * inv; return y - __bfloat162float(__float2bfloat16_rn(y)); };
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      dst[q] = make_uint4(pack_bf16x2(o[8 * q] * inv, o[8 * q + 1] * inv), pack_bf16x2(o[8 * q + 2] * inv, o[8 * q + 3] * inv),
                          pack_bf16x2(o[8 * q + 4] * inv, o[8 * q + 5] * inv), pack_bf16x2(o[8 * q + 6] * inv, o[8 * q + 7] * inv));
      dlo[q] = make_uint4(pack_bf16x2(rem(o[8 * q]), rem(o[8 * q + 1])), pack_bf16x2(rem(o[8 * q + 2]), rem(o[8 * q + 3])),
                          pack_bf16x2(rem(o[8 * q + 4]), rem(o[8 * q + 5])), pack_bf16x2(rem(o[8 * q + 6]), rem(o[8 * q + 7])));
    }
  } else {
#pragma unroll
    for (int q = 0; q < 2; ++q)
      dst[q] = make_uint4(pack16x2<F16>(o[8 * q] * inv, o[8 * q + 1] * inv), pack16x2<F16>(o[8 * q + 2] * inv, o[8 * q + 3] * inv),
                          pack16x2<F16>(o[8 * q + 4] * inv, o[8 * q + 5] * inv), pack16x2<F16>(o[8 * q + 6] * inv, o[8 * q + 7] * inv));
  }
}

// One item for one thread (== tile row): wait for S, two passes over this row's S in tensor memory (max, then
// exp / pack / store P over S), hand P to the issuer, wait for O, normalise and write HD outputs.
// ALIBI: the score of key j is s_j * scale + slope * j (HF bloom build_alibi_tensor); `ab` = slope * log2(e).
// TABLE: the score of key j is s_j * scale + tb[j] (HF t5 position_bias: `tb` points at this row's slice of the per-head bias
// table laid out by key - query distance).
// HDR: the model's head dimension when it is staged in a wider layout (80 columns in the two 64-wide sub-tiles of HD = 128: the
// upper 48 columns of Q, K and V are TMA zero fill and only 16 columns of the second O slice are stored).
template <bool F16, int HD = 64, bool ALIBI = false, bool TABLE = false, int HDR = HD>
__device__ __forceinline__ void a2_item(uint32_t t_lane, uint32_t o_off, int n_live, int n_full, int n_p, int kvp, int limit, float sc,
                                        uint32_t bar_s, uint32_t bar_p, uint32_t bar_o, uint32_t bar_e, uint32_t ph, bool valid,
                                        uint16_t* out_row, int lo_off = 0, float ab = 0.f, const float* tb = nullptr) {
  constexpr float L2E = 1.4426950408889634f;
  const uint32_t t_o = t_lane + o_off;
  a2_wait(bar_s, ph);
  tcgen05_fence_after();
  // pass 1: row maximum over the visible keys (ALIBI: of the biased, scaled scores)
  float m = -INFINITY;
  for (int c = 0; c < n_live; ++c) {
    float v[32];
    tmem_ld32(t_lane + static_cast<uint32_t>(c * 32), v);
    if constexpr (TABLE) {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (c * 32 + j <= limit) m = fmaxf(m, fmaf(v[j], sc, __ldg(tb + c * 32 + j) * L2E));
    } else if constexpr (ALIBI) {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (c * 32 + j <= limit) m = fmaxf(m, fmaf(v[j], sc, ab * static_cast<float>(c * 32 + j)));
    } else if (c < n_full) {
#pragma unroll
      for (int j = 0; j < 32; ++j) m = fmaxf(m, v[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (c * 32 + j <= limit) m = fmaxf(m, v[j]);
    }
  }
  const float ms = (ALIBI || TABLE) ? m : m * sc;
  // pass 2: p = 2^(s*scale*log2e - max), rounded to 16 bits, two per column over the first half of S
  float lsum = 0.f;
  for (int c = 0; c < n_p; ++c) {
    uint32_t pk[16];
    if (c < n_live) {
      float v[32];
      tmem_ld32(t_lane + static_cast<uint32_t>(c * 32), v);
      if constexpr (TABLE) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const bool in0 = c * 32 + j <= limit, in1 = c * 32 + j + 1 <= limit;
          const float p0 = in0 ? a2_ex2(fmaf(v[j], sc, fmaf(__ldg(tb + c * 32 + j), L2E, -ms))) : 0.f;
          const float p1 = in1 ? a2_ex2(fmaf(v[j + 1], sc, fmaf(__ldg(tb + c * 32 + j + 1), L2E, -ms))) : 0.f;
          lsum += p0 + p1;
          pk[j >> 1] = pack16x2<F16>(p0, p1);
        }
      } else if constexpr (ALIBI) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float b0 = ab * static_cast<float>(c * 32 + j) - ms, b1 = ab * static_cast<float>(c * 32 + j + 1) - ms;
          const float p0 = (c * 32 + j <= limit) ? a2_ex2(fmaf(v[j], sc, b0)) : 0.f;
          const float p1 = (c * 32 + j + 1 <= limit) ? a2_ex2(fmaf(v[j + 1], sc, b1)) : 0.f;
          lsum += p0 + p1;
          pk[j >> 1] = pack16x2<F16>(p0, p1);
        }
      } else if (c < n_full) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float p0 = a2_ex2(fmaf(v[j], sc, -ms)), p1 = a2_ex2(fmaf(v[j + 1], sc, -ms));
          lsum += p0 + p1;
          pk[j >> 1] = pack16x2<F16>(p0, p1);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float p0 = (c * 32 + j <= limit) ? a2_ex2(fmaf(v[j], sc, -ms)) : 0.f;
          const float p1 = (c * 32 + j + 1 <= limit) ? a2_ex2(fmaf(v[j + 1], sc, -ms)) : 0.f;
          lsum += p0 + p1;
          pk[j >> 1] = pack16x2<F16>(p0, p1);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) pk[j] = 0u;
    }
    if (c * 32 + 32 <= kvp) a2_tmem_st16(t_lane + static_cast<uint32_t>(c * 16), pk);
    else a2_tmem_st8(t_lane + static_cast<uint32_t>(c * 16), pk);
  }
  a2_tmem_st_wait();
  tcgen05_fence_before();
  mbar_arrive(bar_p);
  // epilogue of this item: O / rowsum, 64 output columns at a time
  a2_wait(bar_o, ph);
  tcgen05_fence_after();
  const float inv = 1.f / lsum;
#pragma unroll
  for (int h = 0; h < HD / 64; ++h) {
    float o0[32], o1[32];
    tmem_ld32(t_o + static_cast<uint32_t>(64 * h), o0);
    tmem_ld32(t_o + static_cast<uint32_t>(64 * h + 32), o1);
    if (h == HD / 64 - 1) {
      tcgen05_fence_before();
      mbar_arrive(bar_e);                 // the lane's columns may be overwritten by the next S
    }
    if (HDR == 80 && h == 1) {
      if (valid) a2_store16<F16>(o0, inv, out_row + 64, lo_off);
    } else if (valid) {
      a2_store64<F16>(o0, o1, inv, out_row + 64 * h, lo_off);
    }
  }
}

template <bool F16, int HD, bool ALIBI, int HDR = HD>
__global__ void __launch_bounds__(A2_THREADS, 1)
attn_tc2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const Attn2Params p) {
  const int seq = blockIdx.x, mt = blockIdx.y, g = blockIdx.z;
  const int ql = p.meta.q_len[seq];
  if (mt * A2_BM >= ql) return;                        // uniform per CTA: nothing allocated yet
  const int q0 = p.meta.q_start[seq] + mt * A2_BM;
  const int kv0 = p.meta.kv_start[seq];
  const int kl = p.meta.kv_len[seq];
  const int n_items = p.rep;
  const int rows_here = min(A2_BM, ql - mt * A2_BM);
  const int base_lim = kl - ql + mt * A2_BM;           // last key visible to row 0 of the tile (causal)
  const int kv_tile = min(kl, base_lim + rows_here);   // keys visible to the last data row
  const int kvp = (kv_tile + 15) & ~15;                // MMA granularity (N % 16, K % 16)
  constexpr int NQ = a2_nq<HD>(), Q_BYTES = a2_q_bytes<HD>(), KV_BYTES = a2_kv_bytes<HD>(), NSUB = HD / 64;
  const int nl = (HD == 64 && kvp <= 128) ? 4 : 2;     // items in flight
  const uint32_t lane_cols = 512u / static_cast<uint32_t>(nl);
  const uint32_t o_off = nl == 4 ? 64u : 128u;         // O overlays S columns that P does not cover

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + KV_BYTES, sQ = base + 2 * KV_BYTES;
  const uint32_t bars = sQ + NQ * Q_BYTES;
  // barriers (8 B each): kv, q[6], s[4], p[4], o[4], e[4]; then the TMEM slot
  const uint32_t bar_kv = bars;
  auto bar_q = [&](int b) { return bars + 8u * (1 + b); };
  auto bar_s = [&](int l) { return bars + 8u * (7 + l); };    // S(i) accumulated            (tensor core -> softmax)
  auto bar_p = [&](int l) { return bars + 8u * (11 + l); };   // P(i) written, S(i) consumed (softmax -> issuer)
  auto bar_o = [&](int l) { return bars + 8u * (15 + l); };   // O(i) accumulated            (tensor core -> softmax)
  auto bar_e = [&](int l) { return bars + 8u * (19 + l); };   // O(i) read: lane free        (softmax -> issuer)
  const uint32_t tmem_slot = bars + 8u * 23;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == A2_LANES * 128) {
    mbar_init(bar_kv, 1);
    for (int b = 0; b < NQ; ++b) mbar_init(bar_q(b), 1);
    for (int l = 0; l < A2_LANES; ++l) {
      mbar_init(bar_s(l), 1); mbar_init(bar_p(l), 128); mbar_init(bar_o(l), 1); mbar_init(bar_e(l), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(A2_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  const int h_first = g * p.rep;

  if (warp == A2_LANES * 4) {
    if (lane == 0) {
      // ---------------- TMA + MMA issuer ----------------
      auto load_q = [&](int i) {
        const uint32_t bq = bar_q(i % NQ);
        mbar_arrive_expect_tx(bq, Q_BYTES);
#pragma unroll
        for (int sb = 0; sb < NSUB; ++sb) {
          if constexpr (HDR == 80)      // 3D map [rows][segments of 80][80]: the box's columns past 80 arrive as zeros
            tma_load_3d(sQ + (i % NQ) * Q_BYTES + sb * A2_Q_BYTES, &tmQ, bq, 64 * sb, (p.q_base + (h_first + i) * p.q_hstride) / 80, q0);
          else
            tma_load_2d(sQ + (i % NQ) * Q_BYTES + sb * A2_Q_BYTES, &tmQ, bq, p.q_base + (h_first + i) * p.q_hstride + 64 * sb, q0);
        }
      };
      mbar_arrive_expect_tx(bar_kv, 2 * KV_BYTES);
#pragma unroll
      for (int sb = 0; sb < NSUB; ++sb) {
        if constexpr (HDR == 80) {
          tma_load_3d(sK + sb * A2_KV_BYTES, &tmK, bar_kv, 64 * sb, g, kv0);
          tma_load_3d(sV + sb * A2_KV_BYTES, &tmV, bar_kv, 64 * sb, g, kv0);
        } else {
          tma_load_2d(sK + sb * A2_KV_BYTES, &tmK, bar_kv, g * HD + 64 * sb, kv0);
          tma_load_2d(sV + sb * A2_KV_BYTES, &tmV, bar_kv, g * HD + 64 * sb, kv0);
        }
      }
      for (int i = 0; i < NQ && i < n_items; ++i) load_q(i);
      a2_wait(bar_kv, 0);
      const uint32_t idesc_s = a2_idesc(A2_BM, kvp, 0, F16);
      const uint32_t idesc_o = a2_idesc(A2_BM, 64, 1, F16);      // one 64-column slice of O per product
      int is = 0, ipv = 0;
      uint32_t idle = 0;
      while (ipv < n_items) {
        bool did = false;
        if (is < n_items) {
          // S(is): its lane is free once the previous item of the lane has had its O read
          const int l = is & (nl - 1);
          const int qb = is % NQ;
          const bool free_lane = is < nl || a2_test(bar_e(l), static_cast<uint32_t>((is / nl) - 1) & 1u);
          if (free_lane && a2_test(bar_q(qb), static_cast<uint32_t>(is / NQ) & 1u)) {
            tcgen05_fence_after();
            const uint32_t t_s = tmem_base + static_cast<uint32_t>(l) * lane_cols;
#pragma unroll
            for (int sb = 0; sb < NSUB; ++sb) {      // S accumulates over the 64-wide slices of the head dimension
              const uint64_t qdesc = make_kmajor_sw128_desc(sQ + qb * Q_BYTES + sb * A2_Q_BYTES);
              const uint64_t kdesc = make_kmajor_sw128_desc(sK + sb * A2_KV_BYTES);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tcgen05_mma_bf16(t_s, qdesc + 2u * k, kdesc + 2u * k, idesc_s, (sb | k) != 0 ? 1u : 0u);
            }
            tcgen05_commit(bar_s(l));
            ++is; did = true;
          }
        }
        if (ipv < is) {
          const int l = ipv & (nl - 1);
          if (a2_test(bar_p(l), static_cast<uint32_t>(ipv / nl) & 1u)) {   // P(ipv) in tensor memory, S(ipv) consumed
            tcgen05_fence_after();
            const uint32_t t_p = tmem_base + static_cast<uint32_t>(l) * lane_cols;
            const uint32_t t_o = t_p + o_off;
#pragma unroll
            for (int sb = 0; sb < NSUB; ++sb) {
              const uint64_t vdesc = a2_vdesc(sV + sb * A2_KV_BYTES);
              for (int ks = 0; ks < kvp / 16; ++ks)   // 16 keys = 8 columns of packed P, 16 rows (2048 B) of V
                a2_mma_ts(t_o + 64u * sb, t_p + 8u * ks, vdesc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_o, ks != 0 ? 1u : 0u);
            }
            tcgen05_commit(bar_o(l));
            if (ipv + NQ < n_items) load_q(ipv + NQ);   // S(ipv) is complete, so its Q buffer is free
            ++ipv; did = true;
          }
        }
        if (did) idle = 0;
        else if (++idle > (1u << 28)) {
          printf("b2s attn_tc2: issuer stalled (block %d,%d S %d PV %d of %d)\n", blockIdx.x, blockIdx.y, is, ipv, n_items);
          __trap();
        }
      }
    }
  } else if (warp < 4 * nl) {
    // ---------------- softmax + epilogue: thread == tile row ----------------
    const int l = warp >> 2;                   // lane (item slot)
    const int quad = warp & 3;                 // TMEM lane quadrant of this warp
    const int row = quad * 32 + lane;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(l) * lane_cols;
    const bool valid = row < rows_here;
    const int limit = min(kl - 1, base_lim + row);              // last key this row may see
    const int lim_lo = min(kl - 1, base_lim + quad * 32);       // ... the first row of the warp
    const int lim_hi = min(kl - 1, base_lim + quad * 32 + 31);  // ... the last row of the warp
    const int n_live = lim_hi / 32 + 1;        // 32-key chunks some row of the warp sees
    const int n_full = (lim_lo + 1) / 32;      // chunks every row of the warp sees entirely (no compare needed)
    const int n_p = (kvp + 31) / 32;           // chunks of P the tensor core will read
    const float sc = p.scale_log2e;
    for (int i = l; i < n_items; i += nl) {
      const uint32_t ph = static_cast<uint32_t>(i / nl) & 1u;
      const size_t owidth = static_cast<size_t>(p.nh) * HDR;
      uint16_t* dst = p.out + static_cast<size_t>(q0 + row) * (owidth << p.split) + static_cast<size_t>(h_first + i) * HDR;
      float ab = 0.f;
      if constexpr (ALIBI) ab = p.alibi[h_first + i] * 1.4426950408889634f;
      a2_item<F16, HD, ALIBI, false, HDR>(t_lane, o_off, n_live, n_full, n_p, kvp, limit, sc, bar_s(l), bar_p(l), bar_o(l), bar_e(l), ph, valid, dst,
                              p.split ? static_cast<int>(owidth) : 0, ab);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(A2_TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Multi-head bidirectional attention with a relative-position bias table: the T5 encoder (HF t5:153-345; every head has
// its own K and V, every query row sees every key, score = q.k + bias[head][key - query], no 1/sqrt(d)).  One CTA per
// (prompt, 128-query tile) walks the heads; an item (= head) brings its own Q, K and V tiles, straight from the fused
// q|k|v projection output (no cache copy), into a ring of NS stages; stage s and tensor-memory lane s belong together.
// KVBOX 128: four 48 KB stages / four items in flight; KVBOX 256: two 80 KB stages / two items.  Per item the tensor core
// does 128 x keys x 64 twice and the 128 softmax threads of the lane do two passes over their row: the kernel is bound by
// the K/V bytes it streams (every q|k|v row is read once per query tile of its prompt).
// ---------------------------------------------------------------------------------------------
extern int g_opt_attn_ts;
struct AttnEncParams {
  BatchMeta meta;
  uint16_t* out;          // [T, nh*64] 16-bit ([T, 2*nh*64] = hi | lo when split)
  int split;
  int nh, inner;          // inner = nh*64: K at column inner, V at column 2*inner of the projection output
  float scale_log2e;
  const float* table;     // [nh, 2*R + 1]
  int R;
};
template <int KVBOX> constexpr int enc_stage_bytes() { return A2_Q_BYTES + 2 * KVBOX * A2_HD * 2; }
template <int KVBOX> constexpr int enc_stages() { return KVBOX == 128 ? 4 : 2; }
template <int KVBOX> constexpr int enc_smem() { return enc_stages<KVBOX>() * enc_stage_bytes<KVBOX>() + 1024 + 256; }

template <bool F16, int KVBOX>
__global__ void __launch_bounds__(A2_THREADS, 1)
attn_enc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnEncParams p) {
  constexpr int NS = enc_stages<KVBOX>();
  constexpr uint32_t STAGE = enc_stage_bytes<KVBOX>(), KV_BYTES = KVBOX * A2_HD * 2;
  const int mt = blockIdx.x, seq = blockIdx.y;      // query tiles of a prompt are neighbours in launch order: the second one
  const int kl = p.meta.q_len[seq];                 // finds the prompt's K and V in L2
  if (mt * A2_BM >= kl) return;
  const int row0 = p.meta.q_start[seq];
  const int q0 = row0 + mt * A2_BM;
  const int rows_here = min(A2_BM, kl - mt * A2_BM);
  const int kvp = (kl + 15) & ~15;
  const int n_items = p.nh;
  constexpr uint32_t lane_cols = 512u / NS;
  constexpr uint32_t o_off = NS == 4 ? 64u : 128u;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  auto sQ = [&](int s) { return base + static_cast<uint32_t>(s) * STAGE; };
  auto sK = [&](int s) { return sQ(s) + A2_Q_BYTES; };
  auto sV = [&](int s) { return sK(s) + KV_BYTES; };
  const uint32_t bars = base + NS * STAGE;
  auto bar_f = [&](int l) { return bars + 8u * l; };            // Q, K, V of the stage landed      (TMA -> issuer)
  auto bar_s = [&](int l) { return bars + 8u * (4 + l); };
  auto bar_p = [&](int l) { return bars + 8u * (8 + l); };
  auto bar_o = [&](int l) { return bars + 8u * (12 + l); };
  auto bar_e = [&](int l) { return bars + 8u * (16 + l); };
  const uint32_t tmem_slot = bars + 8u * 20;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == A2_LANES * 128) {
    for (int l = 0; l < NS; ++l) {
      mbar_init(bar_f(l), 1); mbar_init(bar_s(l), 1); mbar_init(bar_p(l), 128); mbar_init(bar_o(l), 1); mbar_init(bar_e(l), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(A2_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == A2_LANES * 4) {
    if (lane == 0) {
      const uint32_t idesc_s = a2_idesc(A2_BM, kvp, 0, F16);
      const uint32_t idesc_o = a2_idesc(A2_BM, A2_HD, 1, F16);
      int il = 0, is = 0, ipv = 0;
      uint32_t idle = 0;
      while (ipv < n_items) {
        bool did = false;
        if (il < n_items) {
          // the stage is free once P.V of the item that used it last has completed (S before it, in issue order)
          const int l = il % NS;
          if (il < NS || a2_test(bar_o(l), static_cast<uint32_t>((il / NS) - 1) & 1u)) {
            // 64-row boxes (8 KB, one after the other: the layout of a single tall box), only as many as hold rows of this
            // prompt: a 170-token prompt moves 192 key rows and, for its second query tile, 64 query rows instead of 256 / 128.
            // Rows of a stage that are not loaded keep older finite data: keys past kvp are not multiplied, P is zero
            // for keys past kl, and query rows past rows_here are never stored.
            const uint32_t nqb = static_cast<uint32_t>((rows_here + 63) >> 6), nkb = static_cast<uint32_t>((kl + 63) >> 6);
            mbar_arrive_expect_tx(bar_f(l), (nqb + 2 * nkb) * 8192u);
            for (uint32_t b = 0; b < nqb; ++b) tma_load_2d(sQ(l) + b * 8192u, &tmQ, bar_f(l), il * A2_HD, q0 + 64 * b);
            for (uint32_t b = 0; b < nkb; ++b) {
              tma_load_2d(sK(l) + b * 8192u, &tmKV, bar_f(l), p.inner + il * A2_HD, row0 + 64 * b);
              tma_load_2d(sV(l) + b * 8192u, &tmKV, bar_f(l), 2 * p.inner + il * A2_HD, row0 + 64 * b);
            }
            ++il; did = true;
          }
        }
        if (is < il) {
          const int l = is % NS;
          const bool free_lane = is < NS || a2_test(bar_e(l), static_cast<uint32_t>((is / NS) - 1) & 1u);
          if (free_lane && a2_test(bar_f(l), static_cast<uint32_t>(is / NS) & 1u)) {
            tcgen05_fence_after();
            const uint32_t t_s = tmem_base + static_cast<uint32_t>(l) * lane_cols;
            const uint64_t qdesc = make_kmajor_sw128_desc(sQ(l));
            const uint64_t kdesc = make_kmajor_sw128_desc(sK(l));
#pragma unroll
            for (int k = 0; k < A2_HD / 16; ++k) tcgen05_mma_bf16(t_s, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
            tcgen05_commit(bar_s(l));
            ++is; did = true;
          }
        }
        if (ipv < is) {
          const int l = ipv % NS;
          if (a2_test(bar_p(l), static_cast<uint32_t>(ipv / NS) & 1u)) {
            tcgen05_fence_after();
            const uint32_t t_p = tmem_base + static_cast<uint32_t>(l) * lane_cols;
            const uint64_t vdesc = a2_vdesc(sV(l));
            for (int ks = 0; ks < kvp / 16; ++ks)
              a2_mma_ts(t_p + o_off, t_p + 8u * ks, vdesc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_o, ks != 0 ? 1u : 0u);
            tcgen05_commit(bar_o(l));
            ++ipv; did = true;
          }
        }
        if (did) idle = 0;
        else if (++idle > (1u << 28)) {
          printf("b2s attn_enc: issuer stalled (block %d,%d load %d S %d PV %d of %d)\n", blockIdx.x, blockIdx.y, il, is, ipv, n_items);
          __trap();
        }
      }
    }
  } else if (warp < 4 * NS) {
    const int l = warp >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(l) * lane_cols;
    const bool valid = row < rows_here;
    const int limit = kl - 1;                  // every query row sees every key
    const int n_live = limit / 32 + 1;
    const int n_p = (kvp + 31) / 32;
    const size_t owidth = static_cast<size_t>(p.inner);
    const size_t tw = 2 * static_cast<size_t>(p.R) + 1;
    const int qpos = mt * A2_BM + row;         // (rows past the prompt's end compute on, unused: R >= 255 keeps them in the table)
    for (int i = l; i < n_items; i += NS) {
      const uint32_t ph = static_cast<uint32_t>(i / NS) & 1u;
      uint16_t* dst = p.out + static_cast<size_t>(q0 + row) * (owidth << p.split) + static_cast<size_t>(i) * A2_HD;
      const float* tb = p.table + static_cast<size_t>(i) * tw + (p.R - qpos);
      a2_item<F16, 64, false, true>(t_lane, o_off, n_live, 0, n_p, kvp, limit, p.scale_log2e, bar_s(l), bar_p(l), bar_o(l), bar_e(l), ph,
                                    valid, dst, p.split ? static_cast<int>(owidth) : 0, 0.f, tb);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(A2_TMEM_COLS) : "memory");
  }
}

template <bool F16, int KVBOX>
static int launch_enc_t(const AttnArgs& a, int n_seqs, int total_rows, int max_qlen, cudaStream_t st) {
  B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_enc_kernel<F16, KVBOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, enc_smem<KVBOX>()));
  CUtensorMap tq, tkv;      // one map over the projection output, 64 x 64 boxes, used for Q, K and V
  B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows), static_cast<uint64_t>(a.qkv_dim),
                            static_cast<uint64_t>(a.qkv_dim) * 2, 64, 64, a.f16));
  tkv = tq;
  AttnEncParams p;
  p.meta = a.meta;
  p.out = reinterpret_cast<uint16_t*>(a.out);
  p.split = a.split ? 1 : 0;
  p.nh = a.nh; p.inner = a.nh * A2_HD;
  p.scale_log2e = a.scale * 1.4426950408889634f;
  p.table = a.relbias; p.R = a.rel_R;
  dim3 grid(static_cast<unsigned>(ceil_div(max_qlen, A2_BM)), n_seqs);
  attn_enc_kernel<F16, KVBOX><<<grid, A2_THREADS, enc_smem<KVBOX>(), st>>>(tq, tkv, p);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

// q|k|v of every head side by side in one [rows, 3*nh*64] 16-bit buffer (a.qkv; a.kcache / a.vcache are not used)
bool attn_enc_supported(const AttnArgs& a) {
  return g_opt_attn_ts != 0 && a.hd == A2_HD && a.nkv == a.nh && a.noncausal && a.relbias != nullptr && a.rel_R >= A2_MAXKV - 1 &&
         a.max_kv <= A2_MAXKV && a.qkv_dim == 3 * a.nh * A2_HD && a.q_base == 0 && a.q_hstride == A2_HD && a.alibi == nullptr;
}

int launch_attn_enc(const AttnArgs& a, int n_seqs, int total_rows, int max_qlen, cudaStream_t st) {
  B2S_REQUIRE(attn_enc_supported(a), "attn_enc: shape not supported (head_dim 64, <= 256 keys, fused q|k|v rows, bias table of >= 511 entries)");
  if (a.max_kv <= 128) return a.f16 ? launch_enc_t<true, 128>(a, n_seqs, total_rows, max_qlen, st)
                                    : launch_enc_t<false, 128>(a, n_seqs, total_rows, max_qlen, st);
  return a.f16 ? launch_enc_t<true, 256>(a, n_seqs, total_rows, max_qlen, st) : launch_enc_t<false, 256>(a, n_seqs, total_rows, max_qlen, st);
}

// ---------------------------------------------------------------------------------------------
// One-token (look-ahead) steps: one CTA per (prompt, KV head); the rows of the tile are the query HEADS of the single
// new token (the qkv row is read as a [heads, 64] matrix), no causal mask.  One item per CTA for <= 128 heads per KV
// head, so the only parallelism is between CTAs: the CTA is made small (160 threads, K/V boxes of KVBOX keys, KVBOX
// tensor-memory columns) so that 2 (KVBOX 256) or 4 (KVBOX 128) CTAs share an SM and hide each other's TMA / MMA /
// barrier latencies.  The first-generation kernel ran this with one 209 KB CTA per SM.
// ---------------------------------------------------------------------------------------------
struct AttnDec2Params {
  BatchMeta meta;
  uint16_t* out;
  int split;              // as in Attn2Params
  int nh, nkv, rep;
  int q_base;             // column of head 0 in the qkv row (multiple of 64)
  int qrows_per_token;    // qkv_dim / 64
  float scale_log2e;
};

template <int KVBOX> constexpr int dec2_smem() { return 2 * KVBOX * A2_HD * 2 + A2_Q_BYTES + 1024 + 128; }

template <bool F16, int KVBOX>
__global__ void __launch_bounds__(160, KVBOX == 128 ? 4 : 2)
attn_dec2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const AttnDec2Params p) {
  constexpr uint32_t KV_BYTES = KVBOX * A2_HD * 2;
  constexpr uint32_t O_OFF = KVBOX / 2;                 // P covers [0, KVBOX/2), O the 64 columns after it
  const int seq = blockIdx.x, g = blockIdx.z;
  const int q0 = p.meta.q_start[seq];
  const int kv0 = p.meta.kv_start[seq];
  const int kl = p.meta.kv_len[seq];
  const int n_items = (p.rep + A2_BM - 1) / A2_BM;
  const int kvp = (kl + 15) & ~15;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + KV_BYTES, sQ = base + 2 * KV_BYTES;
  const uint32_t bars = sQ + A2_Q_BYTES;
  const uint32_t bar_kv = bars, bar_q = bars + 8, bar_s = bars + 16, bar_p = bars + 24, bar_o = bars + 32, bar_e = bars + 40;
  const uint32_t tmem_slot = bars + 48;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 128) {
    mbar_init(bar_kv, 1); mbar_init(bar_q, 1); mbar_init(bar_s, 1); mbar_init(bar_p, 128); mbar_init(bar_o, 1); mbar_init(bar_e, 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(static_cast<uint32_t>(KVBOX)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  const int h_first = g * p.rep;
  if (warp == 4) {
    if (lane == 0) {
      auto load_q = [&](int i) {
        mbar_arrive_expect_tx(bar_q, A2_Q_BYTES);
        tma_load_2d(sQ, &tmQ, bar_q, 0, q0 * p.qrows_per_token + p.q_base / A2_HD + h_first + i * A2_BM);
      };
      mbar_arrive_expect_tx(bar_kv, 2 * KV_BYTES);
      tma_load_2d(sK, &tmK, bar_kv, g * A2_HD, kv0);
      tma_load_2d(sV, &tmV, bar_kv, g * A2_HD, kv0);
      load_q(0);
      a2_wait(bar_kv, 0);
      const uint32_t idesc_s = a2_idesc(A2_BM, kvp, 0, F16);
      const uint32_t idesc_o = a2_idesc(A2_BM, A2_HD, 1, F16);
      const uint64_t kdesc = make_kmajor_sw128_desc(sK);
      const uint64_t qdesc = make_kmajor_sw128_desc(sQ);
      const uint64_t vdesc = a2_vdesc(sV);
      for (int i = 0; i < n_items; ++i) {
        const uint32_t ph = static_cast<uint32_t>(i) & 1u;
        if (i > 0) a2_wait(bar_e, ph ^ 1u);             // O(i-1) read: the columns are free
        a2_wait(bar_q, ph);
        tcgen05_fence_after();
#pragma unroll
        for (int k = 0; k < A2_HD / 16; ++k)
          tcgen05_mma_bf16(tmem_base, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
        tcgen05_commit(bar_s);
        a2_wait(bar_p, ph);                             // P(i) in tensor memory, S(i) (and so Q(i)) consumed
        tcgen05_fence_after();
        if (i + 1 < n_items) load_q(i + 1);
        for (int ks = 0; ks < kvp / 16; ++ks)
          a2_mma_ts(tmem_base + O_OFF, tmem_base + 8u * ks, vdesc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_o, ks != 0 ? 1u : 0u);
        tcgen05_commit(bar_o);
      }
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const int limit = kl - 1;                  // the new token sees every cached key, its own included
    const int n_live = limit / 32 + 1;
    const int n_full = kl / 32;
    const int n_p = (kvp + 31) / 32;
    for (int i = 0; i < n_items; ++i) {
      const bool valid = row < min(A2_BM, p.rep - i * A2_BM);
      const size_t owidth = static_cast<size_t>(p.nh) * A2_HD;
      uint16_t* dst = p.out + static_cast<size_t>(q0) * (owidth << p.split) + static_cast<size_t>(h_first + i * A2_BM + row) * A2_HD;
      a2_item<F16>(t_lane, O_OFF, n_live, n_full, n_p, kvp, limit, p.scale_log2e, bar_s, bar_p, bar_o, bar_e,
                   static_cast<uint32_t>(i) & 1u, valid, dst, p.split ? static_cast<int>(owidth) : 0);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(static_cast<uint32_t>(KVBOX)) : "memory");
  }
}

template <bool F16, int KVBOX>
static int launch_dec2(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, cudaStream_t st) {
  // per-device function attribute, and cheap: set on every launch rather than once per process
  B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_dec2_kernel<F16, KVBOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, dec2_smem<KVBOX>()));
  CUtensorMap tq, tk, tv;
  const uint64_t rpt = static_cast<uint64_t>(a.qkv_dim / A2_HD);   // qkv viewed as [total_rows * qkv_dim/64, 64]
  B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows) * rpt, A2_HD, A2_HD * 2, A2_BM, A2_HD, a.f16));
  const uint64_t kvw = static_cast<uint64_t>(a.nkv) * A2_HD;
  B2S_TRY(make_tmap_bf16_2d(&tk, a.kcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, KVBOX, A2_HD, a.f16));
  B2S_TRY(make_tmap_bf16_2d(&tv, a.vcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, KVBOX, A2_HD, a.f16));
  AttnDec2Params p;
  p.meta = a.meta;
  p.out = reinterpret_cast<uint16_t*>(a.out);
  p.split = a.split ? 1 : 0;
  p.nh = a.nh; p.nkv = a.nkv; p.rep = a.rep;
  p.q_base = a.q_base;
  p.qrows_per_token = a.qkv_dim / A2_HD;
  p.scale_log2e = a.scale * 1.4426950408889634f;
  dim3 grid(n_seqs, 1, a.nkv);
  attn_dec2_kernel<F16, KVBOX><<<grid, 160, dec2_smem<KVBOX>(), st>>>(tq, tk, tv, p);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int launch_attn_dec2(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, cudaStream_t st) {
  B2S_REQUIRE(attn_tc_decode_supported(a.hd, a.max_kv, a.qkv_dim, a.q_base, a.q_hstride),
              "attn_dec2: needs head_dim 64, <= 256 keys and contiguous 64-wide query heads");
  if (a.max_kv <= 128) return a.f16 ? launch_dec2<true, 128>(a, n_seqs, total_rows, cache_slots, st)
                                    : launch_dec2<false, 128>(a, n_seqs, total_rows, cache_slots, st);
  return a.f16 ? launch_dec2<true, 256>(a, n_seqs, total_rows, cache_slots, st)
               : launch_dec2<false, 256>(a, n_seqs, total_rows, cache_slots, st);
}

int g_opt_attn_ts = 1;   // b2s_set_option("attn_ts"): 1 = this kernel for the prompt pass, 0 = attn_tc_kernel

template <bool F16, int HD, bool ALIBI, int HDR = HD>
static int launch_tc2_t(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st) {
  // per-device function attribute, and cheap: set on every launch rather than once per process
  B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_tc2_kernel<F16, HD, ALIBI, HDR>, cudaFuncAttributeMaxDynamicSharedMemorySize, a2_smem<HD>()));
  CUtensorMap tq, tk, tv;      // boxes are 64 columns wide whatever the head dimension (one swizzle atom)
  const uint64_t kvw = static_cast<uint64_t>(a.nkv) * HDR;
  if constexpr (HDR == 80) {
    B2S_REQUIRE(a.qkv_dim % 80 == 0 && a.q_base % 80 == 0 && a.q_hstride % 80 == 0, "attn_tc2: 80-wide heads must start at multiples of 80");
    B2S_TRY(make_tmap_16bit_3d(&tq, a.qkv, static_cast<uint64_t>(total_rows), static_cast<uint64_t>(a.qkv_dim / 80), 80,
                               static_cast<uint64_t>(a.qkv_dim) * 2, A2_BM, 64, a.f16));
    B2S_TRY(make_tmap_16bit_3d(&tk, a.kcache, static_cast<uint64_t>(cache_slots), static_cast<uint64_t>(a.nkv), 80, kvw * 2, A2_MAXKV, 64, a.f16));
    B2S_TRY(make_tmap_16bit_3d(&tv, a.vcache, static_cast<uint64_t>(cache_slots), static_cast<uint64_t>(a.nkv), 80, kvw * 2, A2_MAXKV, 64, a.f16));
  } else {
    B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows), static_cast<uint64_t>(a.qkv_dim),
                              static_cast<uint64_t>(a.qkv_dim) * 2, A2_BM, 64, a.f16));
    B2S_TRY(make_tmap_bf16_2d(&tk, a.kcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, A2_MAXKV, 64, a.f16));
    B2S_TRY(make_tmap_bf16_2d(&tv, a.vcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, A2_MAXKV, 64, a.f16));
  }
  Attn2Params p;
  p.meta = a.meta;
  p.out = reinterpret_cast<uint16_t*>(a.out);
  p.split = a.split ? 1 : 0;
  p.nh = a.nh; p.nkv = a.nkv; p.rep = a.rep;
  p.q_base = a.q_base; p.q_hstride = a.q_hstride;
  p.scale_log2e = a.scale * 1.4426950408889634f;
  p.alibi = a.alibi;
  dim3 grid(n_seqs, static_cast<unsigned>(ceil_div(max_qlen, A2_BM)), a.nkv);
  attn_tc2_kernel<F16, HD, ALIBI, HDR><<<grid, A2_THREADS, a2_smem<HD>(), st>>>(tq, tk, tv, p);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int launch_attn_tc2(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st) {
  B2S_REQUIRE(attn_tc_supported(a.hd, a.max_kv), "attn_tc2: head_dim %d / %d keys not supported", a.hd, a.max_kv);
  B2S_REQUIRE(a.q_base % 8 == 0 && a.q_hstride % 8 == 0, "attn_tc2: query heads must start at 16-byte boundaries");
#define B2S_TC2(F, H, AL) return launch_tc2_t<F, H, AL>(a, n_seqs, total_rows, cache_slots, max_qlen, st)
  if (a.hd == 64) {
    if (a.alibi) { if (a.f16) B2S_TC2(true, 64, true); else B2S_TC2(false, 64, true); }
    if (a.f16) B2S_TC2(true, 64, false); else B2S_TC2(false, 64, false);
  }
  if (a.hd == 80) {      // staged as 128 with TMA zero fill (no ALiBi model has this head size)
    B2S_REQUIRE(a.alibi == nullptr, "attn_tc2: head_dim 80 with ALiBi is not built");
    if (a.f16) return launch_tc2_t<true, 128, false, 80>(a, n_seqs, total_rows, cache_slots, max_qlen, st);
    return launch_tc2_t<false, 128, false, 80>(a, n_seqs, total_rows, cache_slots, max_qlen, st);
  }
  if (a.alibi) { if (a.f16) B2S_TC2(true, 128, true); else B2S_TC2(false, 128, true); }
  if (a.f16) B2S_TC2(true, 128, false); else B2S_TC2(false, 128, false);
#undef B2S_TC2
}

}  // namespace b2s
