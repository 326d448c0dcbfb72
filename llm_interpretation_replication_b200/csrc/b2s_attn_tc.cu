// tcgen05 varlen causal prefill attention, head_dim 64, bf16 (Falcon MQA; also GQA/MHA with head_dim 64).
//
// Replaces F.scaled_dot_product_attention(q, k, v, is_causal=True) of the reference's decoder forward
// (HF:models/falcon/modeling_falcon.py:349-362) for a whole varlen-packed batch.
//
// One CTA per (sequence, 128-query tile, KV head).  K and V of the sequence (<= 256 keys; the reference's
// prompts have <= 190 tokens) are TMA-staged ONCE into 128B-swizzled shared memory and reused
// by every query head that shares the KV head (71 heads for Falcon's multi-query attention); Q tiles are
// streamed with double-buffered TMA loads.  Per head:
//     S = Q.K^T      tcgen05.mma 128 x kvp x 64   (A, B K-major)          -> TMEM columns [0, kvp)
//     softmax        thread == query row: tcgen05.ld, causal mask, exp2, bf16 P -> swizzled smem (A operand)
//     O = P.V        tcgen05.mma 128 x 64 x kvp   (A K-major, B = V MN-major) -> TMEM columns [256, 320)
//     epilogue       tcgen05.ld, 1/rowsum, bf16, 128 B per thread to global
// Warps 0-3 are the softmax/epilogue threads (TMEM lane quadrant == warp), warp 4 lane 0 issues TMA and MMA.
#include "b2s_tc.cuh"
#include "b2s_decoder.cuh"

namespace b2s {

static constexpr int AT_HD = 64;
static constexpr int AT_BM = 128;          // queries per tile
static constexpr int AT_MAXKV = 256;       // keys per sequence supported by this kernel
static constexpr int AT_KV_BYTES = AT_MAXKV * AT_HD * 2;    // 32 KB per K or V
static constexpr int AT_Q_BYTES = AT_BM * AT_HD * 2;        // 16 KB per Q tile
static constexpr int AT_P_ATOM = AT_BM * 128;               // one 64-key swizzle atom of P: 16 KB
static constexpr int AT_P_BYTES = (AT_MAXKV / 64) * AT_P_ATOM;
static constexpr int AT_SMEM = 2 * AT_KV_BYTES + 2 * AT_Q_BYTES + AT_P_BYTES + 1024 + 128;
static constexpr int AT_THREADS = 160;
static constexpr uint32_t AT_TMEM_COLS = 512;
static constexpr uint32_t AT_O_COL = 256;

struct AttnTcParams {
  BatchMeta meta;
  __nv_bfloat16* out;     // [T, nh*64]
  int nh, nkv, rep;
  int q_base, q_hstride;  // column of head h in the qkv row = q_base + h*q_hstride
  float scale_log2e;      // (1/sqrt(hd)) * log2(e)
};

// instruction descriptor: fp32 accumulate, bf16 A/B, A K-major, B K-major (b_mn = 0) or MN-major (b_mn = 1)
__device__ __forceinline__ uint32_t idesc_bf16(int m, int n, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(b_mn) << 16) |
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

__global__ void __launch_bounds__(AT_THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const AttnTcParams p) {
  const int seq = blockIdx.x, mt = blockIdx.y, g = blockIdx.z;
  const int ql = p.meta.q_len[seq];
  if (mt * AT_BM >= ql) return;                       // uniform per CTA: nothing allocated yet
  const int q0 = p.meta.q_start[seq] + mt * AT_BM;
  const int kv0 = p.meta.kv_start[seq];
  const int kl = p.meta.kv_len[seq];
  const int rows_here = min(AT_BM, ql - mt * AT_BM);
  // keys visible to the last query row of this tile; padded to the MMA granularity (N % 16, K % 16)
  const int kv_tile = min(kl, kl - ql + mt * AT_BM + rows_here);
  const int kvp = (kv_tile + 15) & ~15;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + AT_KV_BYTES, sQ = base + 2 * AT_KV_BYTES;
  const uint32_t sP = sQ + 2 * AT_Q_BYTES;
  const uint32_t bars = sP + AT_P_BYTES;
  const uint32_t bar_kv = bars, bar_q0 = bars + 8, bar_q1 = bars + 16, bar_s = bars + 24, bar_p = bars + 32,
                 bar_o = bars + 40, tmem_slot = bars + 48;
  uint8_t* sP_gen = smem_raw + (sP - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 128) {
    mbar_init(bar_kv, 1); mbar_init(bar_q0, 1); mbar_init(bar_q1, 1);
    mbar_init(bar_s, 1); mbar_init(bar_p, 128); mbar_init(bar_o, 1);
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
  const int H = p.rep;

  if (warp == 4) {
    if (lane == 0) {
      // ---- TMA + MMA issuer ----
      mbar_arrive_expect_tx(bar_kv, 2 * AT_KV_BYTES);
      tma_load_2d(sK, &tmK, bar_kv, g * AT_HD, kv0);
      tma_load_2d(sV, &tmV, bar_kv, g * AT_HD, kv0);
      mbar_arrive_expect_tx(bar_q0, AT_Q_BYTES);
      tma_load_2d(sQ, &tmQ, bar_q0, p.q_base + h_first * p.q_hstride, q0);
      mbar_wait(bar_kv, 0);
      const uint32_t idesc_s = idesc_bf16(AT_BM, kvp, 0);
      const uint32_t idesc_o = idesc_bf16(AT_BM, AT_HD, 1);
      const uint64_t kdesc = make_kmajor_sw128_desc(sK);
      const uint64_t vdesc = make_mnmajor_sw128_desc(sV);
      for (int i = 0; i < H; ++i) {
        const int b = i & 1;
        if (i + 1 < H) {   // prefetch the next head's Q tile into the other buffer (its S MMA is long finished)
          const uint32_t bq = b ? bar_q0 : bar_q1;
          mbar_arrive_expect_tx(bq, AT_Q_BYTES);
          tma_load_2d(sQ + (b ^ 1) * AT_Q_BYTES, &tmQ, bq, p.q_base + (h_first + i + 1) * p.q_hstride, q0);
        }
        mbar_wait(b ? bar_q1 : bar_q0, (i >> 1) & 1);
        tcgen05_fence_after();
        const uint64_t qdesc = make_kmajor_sw128_desc(sQ + b * AT_Q_BYTES);
#pragma unroll
        for (int k = 0; k < AT_HD / 16; ++k)
          tcgen05_mma_bf16(tmem_base, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
        tcgen05_commit(bar_s);
        // P written by the 128 softmax threads (they have also finished reading O of the previous head)
        mbar_wait(bar_p, i & 1);
        tcgen05_fence_after();
        for (int ks = 0; ks < kvp / 16; ++ks) {
          const uint64_t pdesc = make_kmajor_sw128_desc(sP + (ks >> 2) * AT_P_ATOM) + 2u * (ks & 3);
          tcgen05_mma_bf16(tmem_base + AT_O_COL, pdesc, vdesc + static_cast<uint64_t>(ks) * (2048 >> 4), idesc_o,
                           ks != 0 ? 1u : 0u);
        }
        tcgen05_commit(bar_o);
      }
    }
  } else {
    // ---- softmax + epilogue: thread == query row ----
    const int row = warp * 32 + lane;
    const bool valid = row < rows_here;
    const int limit = kl - ql + mt * AT_BM + row;       // last visible key index (causal)
    const uint32_t t_s = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
    const uint32_t t_o = t_s + AT_O_COL;
    // keys the whole warp can skip: chunks starting beyond the warp's largest limit are fully masked
    const int warp_limit = kl - ql + mt * AT_BM + min(warp * 32 + 31, rows_here - 1);
    const int nchunks = kvp / 32 + ((kvp & 31) ? 1 : 0);
    __nv_bfloat16* orow = p.out + static_cast<size_t>(q0 + row) * (static_cast<size_t>(p.nh) * AT_HD);
    for (int i = 0; i < H; ++i) {
      mbar_wait(bar_s, i & 1);
      tcgen05_fence_after();
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
      // pass 2: p = exp2(s*scale*log2e - max), rounded to bf16, written to the swizzled A-operand tile
      float l = 0.f;
      for (int c = 0; c < nchunks; ++c) {
        const bool live = c * 32 <= warp_limit;         // warp-uniform
        float v[32];
        if (live) tmem_ld32(t_s + static_cast<uint32_t>(c * 32), v);
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float p0 = 0.f, p1 = 0.f;
          if (live && valid) {
            if (c * 32 + j <= limit) p0 = exp2f(fmaf(v[j], p.scale_log2e, -ms));
            if (c * 32 + j + 1 <= limit) p1 = exp2f(fmaf(v[j + 1], p.scale_log2e, -ms));
          }
          const __nv_bfloat162 b2 = __floats2bfloat162_rn(p0, p1);
          l += __low2float(b2) + __high2float(b2);      // normalise by what the MMA actually multiplies
          pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&b2);
        }
        // 32 keys = 64 B = four 16-byte chunks of this row inside atom (c/2); swizzle: chunk ^= row % 8
        uint8_t* prow = sP_gen + (c >> 1) * AT_P_ATOM + row * 128;
        const int cbase = (c & 1) * 4;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int chunk = (cbase + q) ^ (row & 7);
          *reinterpret_cast<uint4*>(prow + chunk * 16) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
      }
      // make the generic-proxy smem writes visible to the tensor core (async proxy), then signal
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tcgen05_fence_before();
      mbar_arrive(bar_p);
      // epilogue of this head
      mbar_wait(bar_o, i & 1);
      tcgen05_fence_after();
      float o0[32], o1[32];
      tmem_ld32(t_o, o0);
      tmem_ld32(t_o + 32, o1);
      if (valid) {
        const float inv = 1.f / l;
        uint4* dst = reinterpret_cast<uint4*>(orow + static_cast<size_t>(h_first + i) * AT_HD);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          dst[q] = make_uint4(pack_bf16x2(o0[8 * q] * inv, o0[8 * q + 1] * inv), pack_bf16x2(o0[8 * q + 2] * inv, o0[8 * q + 3] * inv),
                              pack_bf16x2(o0[8 * q + 4] * inv, o0[8 * q + 5] * inv), pack_bf16x2(o0[8 * q + 6] * inv, o0[8 * q + 7] * inv));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          dst[4 + q] = make_uint4(pack_bf16x2(o1[8 * q] * inv, o1[8 * q + 1] * inv), pack_bf16x2(o1[8 * q + 2] * inv, o1[8 * q + 3] * inv),
                                  pack_bf16x2(o1[8 * q + 4] * inv, o1[8 * q + 5] * inv), pack_bf16x2(o1[8 * q + 6] * inv, o1[8 * q + 7] * inv));
      }
      tcgen05_fence_before();   // O and S reads done before the next head's MMAs are released by bar_p
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(AT_TMEM_COLS) : "memory");
  }
}

// Host launcher.  Falls back (returns B2S_EUNSUPPORTED without setting an error the caller must report) is NOT
// done here: the caller decides which kernel to use from attn_tc_supported().
bool attn_tc_supported(int hd, int max_kv) { return hd == AT_HD && max_kv <= AT_MAXKV; }

int launch_attn_tc(const AttnArgs& a, int n_seqs, int total_rows, int cache_slots, int max_qlen, cudaStream_t st) {
  B2S_REQUIRE(attn_tc_supported(a.hd, a.max_kv), "attn_tc: head_dim %d / %d keys not supported", a.hd, a.max_kv);
  static bool attr_set = false;
  if (!attr_set) {
    B2S_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
    attr_set = true;
  }
  CUtensorMap tq, tk, tv;
  B2S_TRY(make_tmap_bf16_2d(&tq, a.qkv, static_cast<uint64_t>(total_rows), static_cast<uint64_t>(a.qkv_dim),
                            static_cast<uint64_t>(a.qkv_dim) * 2, AT_BM, AT_HD));
  const uint64_t kvw = static_cast<uint64_t>(a.nkv) * AT_HD;
  B2S_TRY(make_tmap_bf16_2d(&tk, a.kcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, AT_MAXKV, AT_HD));
  B2S_TRY(make_tmap_bf16_2d(&tv, a.vcache, static_cast<uint64_t>(cache_slots), kvw, kvw * 2, AT_MAXKV, AT_HD));
  AttnTcParams p;
  p.meta = a.meta;
  p.out = reinterpret_cast<__nv_bfloat16*>(a.out);
  p.nh = a.nh; p.nkv = a.nkv; p.rep = a.rep;
  p.q_base = a.q_base; p.q_hstride = a.q_hstride;
  p.scale_log2e = a.scale * 1.4426950408889634f;
  dim3 grid(n_seqs, static_cast<unsigned>(ceil_div(max_qlen, AT_BM)), a.nkv);
  attn_tc_kernel<<<grid, AT_THREADS, AT_SMEM, st>>>(tq, tk, tv, p);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // namespace b2s
