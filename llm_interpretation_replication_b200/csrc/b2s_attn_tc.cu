// tcgen05 varlen causal attention, head_dim 64, bf16 (Falcon MQA; also GQA/MHA with head_dim 64).
//
// Replaces F.scaled_dot_product_attention(q, k, v, is_causal=True) of the reference's decoder forward
// (HF:models/falcon/modeling_falcon.py:349-362) for a whole varlen-packed batch, in both phases of the scorer:
//   prefill    one CTA per (prompt, 128-query tile, KV head); rows of the MMA tile are query positions and the
//              CTA loops over the query heads that share the KV head (71 for Falcon's multi-query attention);
//   one-token  (look-ahead steps) one CTA per (prompt, KV head); rows of the tile are the query HEADS of the
//              single new token (the qkv row is read as a [heads, 64] matrix), one iteration, no causal mask.
// K and V of the prompt (<= 256 keys; the reference's prompts have <= 190 tokens) are TMA-staged
// ONCE into 128B-swizzled shared memory; Q tiles stream through a 3-deep TMA ring.  Per item:
//     S = Q.K^T      tcgen05.mma 128 x kvp x 64   (A, B K-major)            -> TMEM
//     softmax        thread == tile row: tcgen05.ld, mask, ex2, bf16 P      -> swizzled smem (A operand of P.V)
//     O = P.V        tcgen05.mma 128 x 64 x kvp   (A K-major, B = V MN-major) -> TMEM
//     epilogue       tcgen05.ld, 1/rowsum, bf16, 128 B per thread to global
// With <= 192 keys two items are in flight at once: warps 0-3 ("lane" 0) and 4-7 (lane 1) each own one S/O
// accumulator pair and one P buffer, and the single issuing thread (warp 8) interleaves S(i), P.V(i-1), so the
// tensor core works on one lane while the other lane's threads are in softmax.  Longer prompts use lane 0 only.
#include "b2s_tc.cuh"
#include "b2s_decoder.cuh"

namespace b2s {

extern int g_opt_attn_ts;
static inline bool g_opt_attn_ts_on() { return g_opt_attn_ts != 0; }

static constexpr int AT_HD = 64;
static constexpr int AT_BM = 128;          // tile rows
static constexpr int AT_MAXKV = 256;       // keys per prompt supported by this kernel
static constexpr int AT_KV2 = 192;         // keys up to which both lanes are used
static constexpr int AT_KV_BYTES = AT_MAXKV * AT_HD * 2;    // 32 KB per K or V
static constexpr int AT_Q_BYTES = AT_BM * AT_HD * 2;        // 16 KB per Q tile
static constexpr int AT_NQ = 3;                             // Q ring depth
static constexpr int AT_P_ATOM = AT_BM * 128;               // one 64-key swizzle atom of P: 16 KB
static constexpr int AT_P_LANE = (AT_KV2 / 64) * AT_P_ATOM; // 48 KB per lane (lane 0 may use 64 KB when alone)
static constexpr int AT_SMEM = 2 * AT_KV_BYTES + AT_NQ * AT_Q_BYTES + 2 * AT_P_LANE + 1024 + 128;
static constexpr int AT_THREADS = 288;     // 8 softmax warps + 1 issuer warp
static constexpr uint32_t AT_TMEM_COLS = 512;
static constexpr uint32_t AT_O_COL = 384;  // O accumulators: lane l at columns [384 + 64 l, +64)

struct AttnTcParams {
  BatchMeta meta;
  uint16_t* out;          // [T, nh*64] bf16 or fp16 bits
  int f16;                // Q/K/V/P/out are fp16 instead of bf16
  int nh, nkv, rep;
  int q_base, q_hstride;  // column of head h in the qkv row = q_base + h*q_hstride
  int decode;             // 0: rows are query positions (prefill); 1: rows are heads of one token
  int qrows_per_token;    // decode: qkv_dim / 64 (rows of the [*, 64] view per token)
  float scale_log2e;      // (1/sqrt(hd)) * log2(e)
};

// instruction descriptor: fp32 accumulate, bf16 A/B, A K-major, B K-major (b_mn = 0) or MN-major (b_mn = 1)
__device__ __forceinline__ uint32_t idesc_16(int m, int n, int b_mn, int f16) {
  const uint32_t fmt = f16 ? 0u : 1u;    // F16 = 0, BF16 = 1 for both operands
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(b_mn) << 16) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}
// V tile [keys, 64 dims] with 128-byte rows, 128B swizzle, used as the MN-major B operand (N = dims, K = keys):
// one 64-element swizzle atom along N (LBO unused), 8-key groups 1024 B apart (SBO).
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool HOLD_S>   // keep S in registers between the max and the exp pass when it fits (<= 128 keys)
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const AttnTcParams p) {
  const int seq = blockIdx.x, mt = blockIdx.y, g = blockIdx.z;
  const int ql = p.meta.q_len[seq];
  if (!p.decode && mt * AT_BM >= ql) return;          // uniform per CTA: nothing allocated yet
  const int q0 = p.meta.q_start[seq] + (p.decode ? 0 : mt * AT_BM);
  const int kv0 = p.meta.kv_start[seq];
  const int kl = p.meta.kv_len[seq];
  // rows of the tile that carry data, keys visible to the last of them, number of items
  int rows_here, kv_tile, n_items;
  if (p.decode) {
    n_items = (p.rep + AT_BM - 1) / AT_BM;
    rows_here = 0;                                     // per item, see below
    kv_tile = kl;
  } else {
    n_items = p.rep;
    rows_here = min(AT_BM, ql - mt * AT_BM);
    kv_tile = min(kl, kl - ql + mt * AT_BM + rows_here);
  }
  const int kvp = (kv_tile + 15) & ~15;               // MMA granularity (N % 16, K % 16)
  const bool two = (kvp <= AT_KV2) && (n_items > 1);
  const int nl = two ? 2 : 1;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + AT_KV_BYTES, sQ = base + 2 * AT_KV_BYTES;
  const uint32_t sP = sQ + AT_NQ * AT_Q_BYTES;
  const uint32_t bars = sP + 2 * AT_P_LANE;
  // barriers (8 B each): kv, q[3], s[2], p[2], o[2]; then the TMEM slot
  const uint32_t bar_kv = bars;
  auto bar_q = [&](int b) { return bars + 8u * (1 + b); };
  auto bar_s = [&](int l) { return bars + 8u * (4 + l); };
  auto bar_p = [&](int l) { return bars + 8u * (6 + l); };
  auto bar_o = [&](int l) { return bars + 8u * (8 + l); };
  const uint32_t tmem_slot = bars + 8u * 10;
  uint8_t* sP_gen = smem_raw + (sP - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 256) {
    mbar_init(bar_kv, 1);
    for (int b = 0; b < AT_NQ; ++b) mbar_init(bar_q(b), 1);
    for (int l = 0; l < 2; ++l) { mbar_init(bar_s(l), 1); mbar_init(bar_p(l), 128); mbar_init(bar_o(l), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(AT_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  const int h_first = g * p.rep;

  if (warp == 8) {
    if (lane == 0) {
      // ---------------- TMA + MMA issuer ----------------
      auto load_q = [&](int i) {
        const uint32_t bq = bar_q(i % AT_NQ);
        mbar_arrive_expect_tx(bq, AT_Q_BYTES);
        if (p.decode)
          tma_load_2d(sQ + (i % AT_NQ) * AT_Q_BYTES, &tmQ, bq, 0,
                      q0 * p.qrows_per_token + p.q_base / AT_HD + h_first + i * AT_BM);
        else
          tma_load_2d(sQ + (i % AT_NQ) * AT_Q_BYTES, &tmQ, bq, p.q_base + (h_first + i) * p.q_hstride, q0);
      };
      mbar_arrive_expect_tx(bar_kv, 2 * AT_KV_BYTES);
      tma_load_2d(sK, &tmK, bar_kv, g * AT_HD, kv0);
      tma_load_2d(sV, &tmV, bar_kv, g * AT_HD, kv0);
      for (int i = 0; i < AT_NQ && i < n_items; ++i) load_q(i);
      mbar_wait(bar_kv, 0);
      const uint32_t idesc_s = idesc_16(AT_BM, kvp, 0, p.f16);
      const uint32_t idesc_o = idesc_16(AT_BM, AT_HD, 1, p.f16);
      const uint64_t kdesc = make_kmajor_sw128_desc(sK);
      const uint64_t vdesc = make_mnmajor_sw128_desc(sV);
      const int skew = two ? 1 : 0;
      for (int i = 0; i < n_items + skew; ++i) {
        if (i < n_items) {
          // S(i): its lane's S accumulator was last read by item i - nl, whose P arrival was waited below
          const int l = two ? (i & 1) : 0;
          mbar_wait(bar_q(i % AT_NQ), (i / AT_NQ) & 1);
          tcgen05_fence_after();
          const uint64_t qdesc = make_kmajor_sw128_desc(sQ + (i % AT_NQ) * AT_Q_BYTES);
          const uint32_t t_s = tmem_base + static_cast<uint32_t>(l * AT_KV2);
#pragma unroll
          for (int k = 0; k < AT_HD / 16; ++k)
            tcgen05_mma_bf16(t_s, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
          tcgen05_commit(bar_s(l));
        }
        const int j = i - skew;
        if (j >= 0) {
          const int l = two ? (j & 1) : 0;
          mbar_wait(bar_p(l), (j / nl) & 1);     // P(j) written; S(j) and O(j - nl) fully read
          tcgen05_fence_after();
          const uint32_t pbase = sP + l * AT_P_LANE;
          const uint32_t t_o = tmem_base + AT_O_COL + static_cast<uint32_t>(l * AT_HD);
          for (int ks = 0; ks < kvp / 16; ++ks) {
            const uint64_t pdesc = make_kmajor_sw128_desc(pbase + (ks >> 2) * AT_P_ATOM) + 2u * (ks & 3);
            tcgen05_mma_bf16(t_o, pdesc, vdesc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_o, ks != 0 ? 1u : 0u);
          }
          tcgen05_commit(bar_o(l));
          if (j + AT_NQ < n_items) load_q(j + AT_NQ);   // S(j) is complete, so its Q buffer is free
        }
      }
    }
  } else if (warp < 4 * nl) {
    // ---------------- softmax + epilogue: thread == tile row ----------------
    const int l = warp >> 2;                   // lane
    const int quad = warp & 3;                 // TMEM lane quadrant
    const int row = quad * 32 + lane;
    const uint32_t t_s = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(l * AT_KV2);
    const uint32_t t_o = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + AT_O_COL + static_cast<uint32_t>(l * AT_HD);
    uint8_t* pl = sP_gen + l * AT_P_LANE;
    const int nchunks = (kvp + 31) / 32;
    for (int i = l; i < n_items; i += nl) {
      int limit, warp_limit;
      bool valid;
      uint16_t* optr;
      if (p.decode) {
        const int rows_i = min(AT_BM, p.rep - i * AT_BM);
        valid = row < rows_i;
        limit = kl - 1;
        warp_limit = kl - 1;
        optr = p.out + static_cast<size_t>(q0) * (static_cast<size_t>(p.nh) * AT_HD) +
               static_cast<size_t>(h_first + i * AT_BM + row) * AT_HD;
      } else {
        valid = row < rows_here;
        limit = kl - ql + mt * AT_BM + row;   // last visible key (causal)
        warp_limit = kl - ql + mt * AT_BM + min(quad * 32 + 31, rows_here - 1);
        optr = p.out + static_cast<size_t>(q0 + row) * (static_cast<size_t>(p.nh) * AT_HD) +
               static_cast<size_t>(h_first + i) * AT_HD;
      }
      const uint32_t ph = static_cast<uint32_t>((i / nl) & 1);
      mbar_wait(bar_s(l), ph);
      tcgen05_fence_after();
      float lsum = 0.f;
      // p = 2^(s*scale*log2e - max) for 32 keys, rounded to the 16-bit format and written to the swizzled A tile
      auto emit_chunk = [&](int c, const float (&v)[32], bool live, float ms) {
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float p0 = 0.f, p1 = 0.f;
          if (live && valid) {
            if (c * 32 + j <= limit) p0 = ex2_approx(fmaf(v[j], p.scale_log2e, -ms));
            if (c * 32 + j + 1 <= limit) p1 = ex2_approx(fmaf(v[j + 1], p.scale_log2e, -ms));
          }
          if (p.f16) {
            const __half2 h2 = __floats2half2_rn(p0, p1);
            lsum += __low2float(h2) + __high2float(h2);   // normalise by what the MMA actually multiplies
            pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
          } else {
            const __nv_bfloat162 b2 = __floats2bfloat162_rn(p0, p1);
            lsum += __low2float(b2) + __high2float(b2);
            pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&b2);
          }
        }
        // 32 keys = 64 B = four 16-byte chunks of this row inside atom (c/2); swizzle: chunk ^= row % 8
        uint8_t* prow = pl + (c >> 1) * AT_P_ATOM + row * 128;
        const int cbase = (c & 1) * 4;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int chunk = (cbase + q) ^ (row & 7);
          *reinterpret_cast<uint4*>(prow + chunk * 16) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
      };
      if (HOLD_S && nchunks <= 4) {
        // <= 128 keys (most tiles): S is read from TMEM ONCE and kept in registers between the max and the exp pass
        // (TMEM reads, not the math, bound this kernel)
        float v0[32], v1[32], v2[32], v3[32];
        const bool l0 = true, l1 = nchunks > 1 && 32 <= warp_limit, l2 = nchunks > 2 && 64 <= warp_limit,
                   l3 = nchunks > 3 && 96 <= warp_limit;        // warp-uniform
        float m = -INFINITY;
        tmem_ld32(t_s, v0);
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j <= limit) m = fmaxf(m, v0[j]);
        if (l1) {
          tmem_ld32(t_s + 32u, v1);
#pragma unroll
          for (int j = 0; j < 32; ++j) if (32 + j <= limit) m = fmaxf(m, v1[j]);
        }
        if (l2) {
          tmem_ld32(t_s + 64u, v2);
#pragma unroll
          for (int j = 0; j < 32; ++j) if (64 + j <= limit) m = fmaxf(m, v2[j]);
        }
        if (l3) {
          tmem_ld32(t_s + 96u, v3);
#pragma unroll
          for (int j = 0; j < 32; ++j) if (96 + j <= limit) m = fmaxf(m, v3[j]);
        }
        const float ms = m * p.scale_log2e;
        emit_chunk(0, v0, l0, ms);
        if (nchunks > 1) emit_chunk(1, v1, l1, ms);
        if (nchunks > 2) emit_chunk(2, v2, l2, ms);
        if (nchunks > 3) emit_chunk(3, v3, l3, ms);
      } else {
        // pass 1: row maximum over the visible keys
        float m = -INFINITY;
        for (int c = 0; c < nchunks; ++c) {
          if (c * 32 > warp_limit) break;                 // warp-uniform
          float v[32];
          tmem_ld32(t_s + static_cast<uint32_t>(c * 32), v);
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c * 32 + j <= limit) m = fmaxf(m, v[j]);
        }
        const float ms = m * p.scale_log2e;
        // pass 2: re-read S chunk by chunk
        for (int c = 0; c < nchunks; ++c) {
          const bool live = c * 32 <= warp_limit;         // warp-uniform
          float v[32];
          if (live) tmem_ld32(t_s + static_cast<uint32_t>(c * 32), v);
          emit_chunk(c, v, live, ms);
        }
      }
      // make the generic-proxy smem writes visible to the tensor core (async proxy), then signal
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tcgen05_fence_before();
      mbar_arrive(bar_p(l));
      // epilogue of this item
      mbar_wait(bar_o(l), ph);
      tcgen05_fence_after();
      float o0[32], o1[32];
      tmem_ld32(t_o, o0);
      tmem_ld32(t_o + 32, o1);
      if (valid) {
        const float inv = 1.f / lsum;
        uint4* dst = reinterpret_cast<uint4*>(optr);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          dst[q] = make_uint4(pack_act16x2(o0[8 * q] * inv, o0[8 * q + 1] * inv, p.f16), pack_act16x2(o0[8 * q + 2] * inv, o0[8 * q + 3] * inv, p.f16),
                              pack_act16x2(o0[8 * q + 4] * inv, o0[8 * q + 5] * inv, p.f16), pack_act16x2(o0[8 * q + 6] * inv, o0[8 * q + 7] * inv, p.f16));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          dst[4 + q] = make_uint4(pack_act16x2(o1[8 * q] * inv, o1[8 * q + 1] * inv, p.f16), pack_act16x2(o1[8 * q + 2] * inv, o1[8 * q + 3] * inv, p.f16),
                                  pack_act16x2(o1[8 * q + 4] * inv, o1[8 * q + 5] * inv, p.f16), pack_act16x2(o1[8 * q + 6] * inv, o1[8 * q + 7] * inv, p.f16));
      }
      tcgen05_fence_before();   // O and S reads are ordered before this thread's next bar_p arrival
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(AT_TMEM_COLS) : "memory");
  }
}

int g_opt_attn_hold_s = 0;   // b2s_set_option("attn_hold_s")

// The caller decides which kernel to use from attn_tc_supported(); there is no silent fallback here.
// prompt pass: head_dim 64 or 128 (second-generation kernel; the first-generation one below knows 64 only), <= 256 keys
bool attn_tc_supported(int hd, int max_kv) { return (hd == AT_HD || ((hd == 128 || hd == 80) && g_opt_attn_ts_on())) && max_kv <= AT_MAXKV; }
// one-token steps additionally need contiguous 64-wide query heads (the qkv row is read as a [heads, 64] matrix)
bool attn_tc_decode_supported(int hd, int max_kv, int qkv_dim, int q_base, int q_hstride) {
  return hd == AT_HD && max_kv <= AT_MAXKV && q_hstride == AT_HD && qkv_dim % AT_HD == 0 && q_base % AT_HD == 0;
}

extern int g_opt_attn_ts;
int launch_attn_tc2(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st);
int launch_attn_dec2(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, cudaStream_t st);

int launch_attn_tc(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st) {
  B2S_REQUIRE(attn_tc_supported(a.hd, a.max_kv), "attn_tc: head_dim %d / %d keys not supported", a.hd, a.max_kv);
  const bool decode = max_qlen == 1;
  if (!decode && g_opt_attn_ts) return launch_attn_tc2(a, n_seqs, total_rows, cache_slots, max_qlen, st);   // prompt pass: P-in-TMEM kernel
  if (decode && g_opt_attn_ts && attn_tc_decode_supported(a.hd, a.max_kv, a.qkv_dim, a.q_base, a.q_hstride))
    return launch_attn_dec2(a, n_seqs, total_rows, cache_slots, st);                                          // one-token steps: small CTAs
  B2S_REQUIRE(!decode || attn_tc_decode_supported(a.hd, a.max_kv, a.qkv_dim, a.q_base, a.q_hstride),
              "attn_tc: one-token mode needs contiguous 64-wide query heads");
  B2S_REQUIRE(!a.split, "attn_tc: the first-generation kernel (attn_ts = 0) does not write split outputs");
  // per-device function attribute, and cheap: set on every launch rather than once per process
  B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
  B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
  CUtensorMap tq, tk, tv;
  if (decode) {   // qkv viewed as [total_rows * qkv_dim/64, 64]
    const uint64_t rpt = static_cast<uint64_t>(a.qkv_dim / AT_HD);
    B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows) * rpt, AT_HD, AT_HD * 2, AT_BM, AT_HD, a.f16));
  } else {
    B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows), static_cast<uint64_t>(a.qkv_dim),
                              static_cast<uint64_t>(a.qkv_dim) * 2, AT_BM, AT_HD, a.f16));
  }
  const uint64_t kvw = static_cast<uint64_t>(a.nkv) * AT_HD;
  B2S_TRY(make_tmap_bf16_2d(&tk, a.kcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, AT_MAXKV, AT_HD, a.f16));
  B2S_TRY(make_tmap_bf16_2d(&tv, a.vcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, AT_MAXKV, AT_HD, a.f16));
  AttnTcParams p;
  p.meta = a.meta;
  p.out = reinterpret_cast<uint16_t*>(a.out);
  p.f16 = a.f16;
  p.nh = a.nh; p.nkv = a.nkv; p.rep = a.rep;
  p.q_base = a.q_base; p.q_hstride = a.q_hstride;
  p.decode = decode ? 1 : 0;
  p.qrows_per_token = a.qkv_dim / AT_HD;
  p.scale_log2e = a.scale * 1.4426950408889634f;
  dim3 grid(n_seqs, decode ? 1u : static_cast<unsigned>(ceil_div(max_qlen, AT_BM)), a.nkv);
  if (g_opt_attn_hold_s) attn_tc_kernel<true><<<grid, AT_THREADS, AT_SMEM, st>>>(tq, tk, tv, p);
  else attn_tc_kernel<false><<<grid, AT_THREADS, AT_SMEM, st>>>(tq, tk, tv, p);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // namespace b2s
