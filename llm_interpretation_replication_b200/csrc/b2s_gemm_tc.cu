// tcgen05 + TMA bf16 GEMM for sm_100a:  C[M,N] = A[M,K] . W[N,K]^T  (both operands K-major).
//
// Replaces the FalconLinear / nn.Linear calls of the reference's decoder forward
// (HF:models/falcon/modeling_falcon.py:60-65, HF:models/gpt_neox/modeling_gpt_neox.py:41-42,200-201)
// and, with the LM-head epilogue, lm_head + float() + softmax + topk of the scoring step
// (HF falcon:905-906; reference analysis/run_base_vs_instruct_100q.py:351-356).
//
// Structure: persistent CTAs (one per SM), 6 warps:
//   warp 0      TMA producer   cp.async.bulk.tensor 2D, 128B-swizzled boxes -> smem ring (STAGES deep)
//   warp 1      MMA issuer     one lane issues tcgen05.mma 128x256x16 (kind::f16, bf16 in / fp32 acc in TMEM)
//   warps 2..5  epilogue       tcgen05.ld 32x32b.x32 (thread == output row), fused epilogue, global stores
// TMEM holds two 128x256 fp32 accumulators (all 512 columns) so the epilogue of tile i overlaps the
// MMAs of tile i+1.  Tiles are rasterised in groups of GROUP_M row-blocks so that the CTAs running at
// the same time share A/W boxes in L2.
#include "b2s_tc.cuh"
#include <atomic>
#include <mutex>
#include <unordered_map>
#include <vector>
#include <stdarg.h>

namespace b2s {

static constexpr int BM = 128, BN = 256, BK = 64, UMMA_K = 16;
static constexpr unsigned long long HINT_EVICT_FIRST = 0x12F0000000000000ull, HINT_EVICT_LAST = 0x14F0000000000000ull,
                                    HINT_EVICT_NORMAL = 0x1000000000000000ull;
static constexpr int STAGES = 4;
static constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KB
static constexpr int B_STAGE_BYTES = BN * BK * 2;   // 32 KB
static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
static constexpr int EPI_STAGE_BYTES = 4096;   // per epilogue warp: 32 rows x 128 B, 128B-swizzled
static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 1024 /*barriers*/ + 8 * EPI_STAGE_BYTES;
static constexpr int NUM_THREADS = 192;
static constexpr uint32_t TMEM_COLS = 512;

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, num_kb;
  int prefetch;          // the producer asks the TMA unit to pull the boxes of the k-block this many steps ahead into L2 (0 = off)
  int band_n;            // CTA-pair 256x256 kernel: > 0 = column bands of this many tiles, all row blocks walked inside a band
  int group_m;           // row-blocks (of 128 rows) per rasterisation group
  int* wave_sync;        // experiment: global counters, one per synchronised wave (null = off)
  int wave_sync_every;   // pair kernels: all producers meet every this many tiles
  int pair_stride;       // experiment: CTA pair p works on the tile sequence of pair (p * pair_stride) % num_pairs
  int serp;              // serpentine order: odd groups walk the column tiles backwards; K walked backwards by odd column tiles (1) or odd waves (2)
  unsigned long long hint_a, hint_b;   // L2 cache-policy words for the TMA loads (0 = none)
  int tma_store;         // epilogue through shared memory + TMA bulk stores / reduce-adds
  int act_f16;           // both operands and the 16-bit outputs are fp16 instead of bf16 (idesc a/b_format F16)
  int act_kind;          // EPI_GELU_BF16: B2S_ACT_GELU (erf), B2S_ACT_RELU or B2S_ACT_GELU_TANH applied to acc + bias before the store
  int out_f16;           // the 16-bit outputs are fp16 (normally = act_f16; BF16_SPLIT writes fp16 q/k/v from bf16 operands)
  int k_wrap;            // BF16_SPLIT: A is [M, 2K] = [hi | lo] and num_kb counts its k-blocks; the W box of k-block kb is
                         // kb - k_wrap once kb >= k_wrap (0 = A and W walk K together)
  int out_split;         // BF16_SPLIT, 16-bit store epilogues: the output is [M, 2N]; hi goes to column c, lo to out_split + c
  void* out;             // bf16 or f32 [M,N]
  const float* bias;     // [N] or null
  LmHeadArgs lmh;
  RopeEpiArgs rope;
};

__device__ __forceinline__ void tma_load_2d_h(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar, int32_t c0, int32_t c1,
                                              unsigned long long hint) {
  if (hint == 0) { tma_load_2d(smem_dst, map, bar, c0, c1); return; }
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}

// L2 prefetch of one operand box (no shared-memory destination, no barrier): turns the DRAM miss of a later TMA load into an
// L2 hit.  The operand ring holds ~2 us of MMA work; a load that misses L2 under full DRAM load can take longer than that.
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* map, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1) : "memory");
}

__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int group_m, int serp, int& mb, int& nb) {
  const int per_group = group_m * tiles_n;
  const int g = tile / per_group;
  const int first_m = g * group_m;
  const int gm = min(group_m, tiles_m - first_m);
  const int within = tile - g * per_group;
  mb = first_m + within % gm;
  nb = within / gm;
  if (serp && (g & 1)) nb = tiles_n - 1 - nb;
}

// ---------------------------------------------------------------------------------------------
// Epilogues.  One thread owns one output row (TMEM lane) and sees 32 consecutive columns at a time.
// ---------------------------------------------------------------------------------------------
template <int EPI>
struct RowState {};
template <>
struct RowState<EPI_LMHEAD> {
  float m, s, top[5], best;
  int arg;
  float la, lb;
  bool has_a, has_b;
};

template <int EPI>
__device__ __forceinline__ void epi_chunk(const GemmParams& p, int row, int col0, float (&v)[32],
                                          RowState<EPI>& st, int ca, int cb, const float* resid) {
  const int N = p.N;
  if (col0 >= N) return;
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      if (col0 + j < N) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    }
  }
  if constexpr (EPI == EPI_STORE_BF16 || EPI == EPI_GELU_BF16) {
    if constexpr (EPI == EPI_GELU_BF16) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = mlp_act(v[j], p.act_kind);
    }
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + static_cast<size_t>(row) * N + col0;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      if (col0 + j < N) {   // N % 8 == 0 is required by the launcher
        uint4 w;
        w.x = pack_act16x2(v[j], v[j + 1], p.out_f16);
        w.y = pack_act16x2(v[j + 2], v[j + 3], p.out_f16);
        w.z = pack_act16x2(v[j + 4], v[j + 5], p.out_f16);
        w.w = pack_act16x2(v[j + 6], v[j + 7], p.out_f16);
        *reinterpret_cast<uint4*>(o + j) = w;
      }
    }
  } else if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_STORE_F32) {
    float* o = reinterpret_cast<float*>(p.out) + static_cast<size_t>(row) * N + col0;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      if (col0 + j < N) {
        float4 w = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        if constexpr (EPI == EPI_RESID_F32) {
          w.x += resid[j]; w.y += resid[j + 1]; w.z += resid[j + 2]; w.w += resid[j + 3];
        }
        *reinterpret_cast<float4*>(o + j) = w;
      }
    }
  } else if constexpr (EPI == EPI_LMHEAD) {
    float cm = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (col0 + j >= N) v[j] = -INFINITY;
      cm = fmaxf(cm, v[j]);
    }
    if (cm > st.m) {
      st.s *= __expf(st.m - cm);   // exp(-inf) = 0 on the first chunk
      st.m = cm;
    }
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float x = v[j];
      acc += __expf(x - st.m);
      if (x > st.best) { st.best = x; st.arg = col0 + j; }
      top5_insert(st.top, x);
      if (col0 + j == ca) { st.la = x; st.has_a = true; }
      if (col0 + j == cb) { st.lb = x; st.has_b = true; }
    }
    st.s += acc;
  }
}

__device__ __forceinline__ void load_resid_chunk(const GemmParams& p, int row, int col0, float (&r)[32]) {
  if (row < p.M) {
    const float* o = reinterpret_cast<const float*>(p.out) + static_cast<size_t>(row) * p.N + col0;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      if (col0 + j < p.N) {
        const float4 t = *reinterpret_cast<const float4*>(o + j);
        r[j] = t.x; r[j + 1] = t.y; r[j + 2] = t.z; r[j + 3] = t.w;
      }
    }
  }
}
// Pull this thread's 1 KB slice of the residual tile into L2 while the MMAs of the tile are still running.
__device__ __forceinline__ void prefetch_resid_tile(const GemmParams& p, int row, int nb) {
  if (row < p.M) {
    const float* o = reinterpret_cast<const float*>(p.out) + static_cast<size_t>(row) * p.N + nb * BN;
#pragma unroll
    for (int j = 0; j < BN; j += 32)
      if (nb * BN + j < p.N) asm volatile("prefetch.global.L2 [%0];" ::"l"(o + j));
  }
}

// One epilogue warp drains its 32 rows x 256 columns of an accumulator, then releases it (one arrival per warp on
// `release_bar`, which may live in the peer CTA of a pair: shared::cluster address).
template <int EPI>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, int row, int nb, uint32_t taddr, int lane,
                                              uint32_t release_bar) {
  RowState<EPI> st;
  int ca = -1, cb = -1;
  if constexpr (EPI == EPI_LMHEAD) {
    st.m = -INFINITY; st.s = 0.f; st.best = -INFINITY; st.arg = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) st.top[j] = -INFINITY;
    st.la = 0.f; st.lb = 0.f; st.has_a = false; st.has_b = false;
    if (row < p.M) { ca = p.lmh.cand_ids[2 * row]; cb = p.lmh.cand_ids[2 * row + 1]; }
  }
  if constexpr (EPI == EPI_QKV_ROPE) {
    // one 64-wide head = two 32-column chunks (x1 | x2 of rotate_half); cos/sin of this row's position once per tile
    int slot = 0;
    const float4 *c4 = nullptr, *s4 = nullptr;       // cos/sin rows of this token's position (L1/L2 resident tables)
    if (row < p.M) {
      const int pos = p.rope.pos[row];
      slot = p.rope.slot[row];
      c4 = reinterpret_cast<const float4*>(p.rope.cos + static_cast<size_t>(pos) * 32);
      s4 = reinterpret_cast<const float4*>(p.rope.sin + static_cast<size_t>(pos) * 32);
    }
#pragma unroll 1
    for (int hh = 0; hh < BN / 64; ++hh) {
      const int col = nb * BN + hh * 64;
      float x1[32], x2[32];
      tmem_ld32(taddr + static_cast<uint32_t>(hh * 64), x1);
      tmem_ld32(taddr + static_cast<uint32_t>(hh * 64 + 32), x2);
      if (row >= p.M || col >= p.N) continue;          // whole heads only: N is a multiple of 64
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j) { x1[j] += __ldg(p.bias + col + j); x2[j] += __ldg(p.bias + col + 32 + j); }
      }
      const int head = col >> 6;
      uint16_t* dst;
      if (head < p.rope.nh) dst = reinterpret_cast<uint16_t*>(p.out) + static_cast<size_t>(row) * p.N + col;
      else if (head < p.rope.nh + p.rope.nkv)
        dst = reinterpret_cast<uint16_t*>(p.rope.kcache) + (static_cast<size_t>(slot) * p.rope.nkv + (head - p.rope.nh)) * 64;
      else
        dst = reinterpret_cast<uint16_t*>(p.rope.vcache) + (static_cast<size_t>(slot) * p.rope.nkv + (head - p.rope.nh - p.rope.nkv)) * 64;
      if (head < p.rope.nh + p.rope.nkv) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 c = __ldg(c4 + j), sn = __ldg(s4 + j);
          const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const float a = x1[4 * j + u], b = x2[4 * j + u];
            x1[4 * j + u] = a * cc[u] - b * ss[u];
            x2[4 * j + u] = b * cc[u] + a * ss[u];
          }
        }
      }
      uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        d4[q] = make_uint4(pack_act16x2(x1[8 * q], x1[8 * q + 1], p.out_f16), pack_act16x2(x1[8 * q + 2], x1[8 * q + 3], p.out_f16),
                           pack_act16x2(x1[8 * q + 4], x1[8 * q + 5], p.out_f16), pack_act16x2(x1[8 * q + 6], x1[8 * q + 7], p.out_f16));
#pragma unroll
      for (int q = 0; q < 4; ++q)
        d4[4 + q] = make_uint4(pack_act16x2(x2[8 * q], x2[8 * q + 1], p.out_f16), pack_act16x2(x2[8 * q + 2], x2[8 * q + 3], p.out_f16),
                               pack_act16x2(x2[8 * q + 4], x2[8 * q + 5], p.out_f16), pack_act16x2(x2[8 * q + 6], x2[8 * q + 7], p.out_f16));
    }
  } else if constexpr (EPI == EPI_RESID_F32) {
    // read-modify-write of the fp32 residual: the loads of chunk c+1 are in flight while chunk c is processed
    float rbuf[2][32];
    load_resid_chunk(p, row, nb * BN, rbuf[0]);
#pragma unroll
    for (int c = 0; c < BN / 32; ++c) {
      if (c + 1 < BN / 32) load_resid_chunk(p, row, nb * BN + (c + 1) * 32, rbuf[(c + 1) & 1]);
      float v[32];
      tmem_ld32(taddr + static_cast<uint32_t>(c * 32), v);
      if (row < p.M) epi_chunk<EPI>(p, row, nb * BN + c * 32, v, st, ca, cb, rbuf[c & 1]);
    }
  } else {
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      float v[32];
      tmem_ld32(taddr + static_cast<uint32_t>(c * 32), v);   // warp-collective: all lanes participate
      if (row < p.M) epi_chunk<EPI>(p, row, nb * BN + c * 32, v, st, ca, cb, nullptr);
    }
  }
  // all TMEM reads of this accumulator are done: hand it back to the MMA warp
  tcgen05_fence_before();
  __syncwarp();
  if (lane == 0) asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(release_bar) : "memory");
  if constexpr (EPI == EPI_LMHEAD) {
    if (row < p.M) {
      float* o = p.lmh.partials + (static_cast<size_t>(row) * p.tiles_n + nb) * LMH_PART;
      o[0] = st.m; o[1] = st.s;
#pragma unroll
      for (int j = 0; j < 5; ++j) o[2 + j] = st.top[j];
      o[7] = __int_as_float(st.arg);
      if (st.has_a) p.lmh.cand_logits[2 * row] = st.la;
      if (st.has_b) p.lmh.cand_logits[2 * row + 1] = st.lb;
    }
  }
}

// one thread's 64 values (two 32-column chunks) -> its 128-byte line of the swizzled staging buffer
template <bool F16>
__device__ __forceinline__ void stage_store16(uint32_t my_line, uint32_t sw, const float (&x1)[32], const float (&x2)[32]) {
#pragma unroll
  for (uint32_t q = 0; q < 4; ++q) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_line + ((q ^ sw) << 4)),
                 "r"(pack16x2<F16>(x1[8 * q], x1[8 * q + 1])), "r"(pack16x2<F16>(x1[8 * q + 2], x1[8 * q + 3])),
                 "r"(pack16x2<F16>(x1[8 * q + 4], x1[8 * q + 5])), "r"(pack16x2<F16>(x1[8 * q + 6], x1[8 * q + 7])) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(my_line + (((q + 4) ^ sw) << 4)),
                 "r"(pack16x2<F16>(x2[8 * q], x2[8 * q + 1])), "r"(pack16x2<F16>(x2[8 * q + 2], x2[8 * q + 3])),
                 "r"(pack16x2<F16>(x2[8 * q + 4], x2[8 * q + 5])), "r"(pack16x2<F16>(x2[8 * q + 6], x2[8 * q + 7])) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Epilogue through shared memory + TMA (default): the warp's 32 x 64 (16-bit) or 32 x 32 (fp32) sub-tile is written
// to a 4 KB 128B-swizzled staging buffer and leaves as ONE bulk tensor store -- or, for the fp32 residual, one bulk
// tensor reduce-add, so the residual is never read back by the SM.  Stores are fully coalesced, need no bounds
// checks (the TMA unit clips at the tensor edge) and cost the LSU nothing.
//   row0  first row of this warp's 32-row slab;  stage  shared address of the warp's staging buffer
// ---------------------------------------------------------------------------------------------
template <int EPI, int NBUF>   // NBUF staging buffers of 4 KB per warp (2: a store may still be draining while the next is staged)
__device__ __forceinline__ void epilogue_tile_tma(const GemmParams& p, const CUtensorMap* tmC, int row0, int nb,
                                                  uint32_t taddr, int lane, uint32_t release_bar, uint32_t stage0,
                                                  uint32_t& flip /* which buffer comes next; lives across tiles */) {
  const int row = row0 + lane;
  const uint32_t sw = static_cast<uint32_t>(lane & 7);
  auto next_buffer = [&](uint32_t& stage, uint32_t& my_line) {
    stage = stage0 + (NBUF == 2 ? flip * EPI_STAGE_BYTES : 0u);
    my_line = stage + static_cast<uint32_t>(lane) * 128u;
    flip ^= 1u;
    if (lane == 0) { if constexpr (NBUF == 2) tma_store_wait_read1(); else tma_store_wait_read(); }
    __syncwarp();
  };
  uint32_t stage = stage0, my_line = stage0 + static_cast<uint32_t>(lane) * 128u;
  if constexpr (EPI == EPI_RESID_F32 || EPI == EPI_STORE_F32) {
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      const int col0 = nb * BN + c * 32;
      if (col0 >= p.N) break;
      float v[32];
      tmem_ld32(taddr + static_cast<uint32_t>(c * 32), v);
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (col0 + j < p.N) v[j] += __ldg(p.bias + col0 + j);
      }
      next_buffer(stage, my_line);                 // the store that used this buffer has finished reading it
#pragma unroll
      for (uint32_t q = 0; q < 8; ++q)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(my_line + ((q ^ sw) << 4)), "f"(v[4 * q]), "f"(v[4 * q + 1]),
                     "f"(v[4 * q + 2]), "f"(v[4 * q + 3]) : "memory");
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if constexpr (EPI == EPI_RESID_F32) tma_reduce_add_2d(tmC, stage, col0, row0);
        else tma_store_2d(tmC, stage, col0, row0);
        tma_store_commit();
      }
    }
  } else if constexpr (EPI == EPI_SWIGLU) {
    // Llama-family MLP (HF llama LlamaMLP: down(silu(gate(x)) * up(x))): the rows of W alternate gate_i, up_i, so a thread
    // holds both halves of every output element; 128 accumulator columns make one 64-column, 16-bit bulk store.
    // act_kind B2S_ACT_GELU_TANH: the gate is tanh-GELU (HF t5 T5DenseGatedActDense, "gelu_new") instead of SiLU
    const bool gelu_gate = p.act_kind == B2S_ACT_GELU_TANH;
    auto glu32 = [gelu_gate](const float (&a)[32], const float (&b)[32], float (&o)[32]) {
      if (gelu_gate) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          o[j] = mlp_act(a[2 * j], B2S_ACT_GELU_TANH) * a[2 * j + 1];
          o[16 + j] = mlp_act(b[2 * j], B2S_ACT_GELU_TANH) * b[2 * j + 1];
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          o[j] = __fdividef(a[2 * j], 1.f + __expf(-a[2 * j])) * a[2 * j + 1];
          o[16 + j] = __fdividef(b[2 * j], 1.f + __expf(-b[2 * j])) * b[2 * j + 1];
        }
      }
    };
#pragma unroll 1
    for (int g = 0; g < BN / 128; ++g) {
      const int col0 = nb * BN + g * 128;          // accumulator column; the output column is col0 / 2
      if (col0 >= p.N) break;
      float o1[32], o2[32];
      {
        float a[32], b[32];
        tmem_ld32(taddr + static_cast<uint32_t>(g * 128), a);
        tmem_ld32(taddr + static_cast<uint32_t>(g * 128 + 32), b);
        glu32(a, b, o1);
        tmem_ld32(taddr + static_cast<uint32_t>(g * 128 + 64), a);
        tmem_ld32(taddr + static_cast<uint32_t>(g * 128 + 96), b);
        glu32(a, b, o2);
      }
      next_buffer(stage, my_line);
      if (p.out_f16) stage_store16<true>(my_line, sw, o1, o2); else stage_store16<false>(my_line, sw, o1, o2);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { tma_store_2d(tmC, stage, col0 / 2, row0); tma_store_commit(); }
      if (p.out_split) {   // BF16_SPLIT: [hi | lo] halves for the down projection
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          o1[j] -= __bfloat162float(__float2bfloat16_rn(o1[j]));
          o2[j] -= __bfloat162float(__float2bfloat16_rn(o2[j]));
        }
        next_buffer(stage, my_line);
        stage_store16<false>(my_line, sw, o1, o2);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) { tma_store_2d(tmC, stage, p.out_split + col0 / 2, row0); tma_store_commit(); }
      }
    }
  } else {
    // 16-bit outputs: 64 columns (two TMEM chunks) per bulk store
    int slot = 0;
    const float4 *c4 = nullptr, *s4 = nullptr;
    if constexpr (EPI == EPI_QKV_ROPE) {
      if (row < p.M) {
        const int pos = p.rope.pos[row];
        slot = p.rope.slot[row];
        c4 = reinterpret_cast<const float4*>(p.rope.cos + static_cast<size_t>(pos) * 32);
        s4 = reinterpret_cast<const float4*>(p.rope.sin + static_cast<size_t>(pos) * 32);
      }
    }
#pragma unroll 1
    for (int g = 0; g < BN / 64; ++g) {
      const int col0 = nb * BN + g * 64;
      if (col0 >= p.N) break;
      float x1[32], x2[32];
      tmem_ld32(taddr + static_cast<uint32_t>(g * 64), x1);
      tmem_ld32(taddr + static_cast<uint32_t>(g * 64 + 32), x2);
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (col0 + j < p.N) x1[j] += __ldg(p.bias + col0 + j);
          if (col0 + 32 + j < p.N) x2[j] += __ldg(p.bias + col0 + 32 + j);
        }
      }
      if constexpr (EPI == EPI_GELU_BF16) {
        if (p.act_kind == B2S_ACT_GELU) {      // the bench family: no per-element branch
#pragma unroll
          for (int j = 0; j < 32; ++j) { x1[j] = gelu_fast(x1[j]); x2[j] = gelu_fast(x2[j]); }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) { x1[j] = mlp_act(x1[j], p.act_kind); x2[j] = mlp_act(x2[j], p.act_kind); }
        }
      }
      bool to_cache = false;
      if constexpr (EPI == EPI_QKV_ROPE) {
        const int head = col0 >> 6;
        if (head < p.rope.nh + p.rope.nkv && row < p.M) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 c = __ldg(c4 + j), sn = __ldg(s4 + j);
            const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float a = x1[4 * j + u], b = x2[4 * j + u];
              x1[4 * j + u] = a * cc[u] - b * ss[u];
              x2[4 * j + u] = b * cc[u] + a * ss[u];
            }
          }
        }
        if (head >= p.rope.nh) {   // key / value head: straight into this row's cache slot (rows map to arbitrary slots)
          to_cache = true;
          if (row < p.M) {
            uint16_t* dst = head < p.rope.nh + p.rope.nkv
                ? reinterpret_cast<uint16_t*>(p.rope.kcache) + (static_cast<size_t>(slot) * p.rope.nkv + (head - p.rope.nh)) * 64
                : reinterpret_cast<uint16_t*>(p.rope.vcache) + (static_cast<size_t>(slot) * p.rope.nkv + (head - p.rope.nh - p.rope.nkv)) * 64;
            uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              d4[q] = make_uint4(pack_act16x2(x1[8 * q], x1[8 * q + 1], p.out_f16), pack_act16x2(x1[8 * q + 2], x1[8 * q + 3], p.out_f16),
                                 pack_act16x2(x1[8 * q + 4], x1[8 * q + 5], p.out_f16), pack_act16x2(x1[8 * q + 6], x1[8 * q + 7], p.out_f16));
              d4[4 + q] = make_uint4(pack_act16x2(x2[8 * q], x2[8 * q + 1], p.out_f16), pack_act16x2(x2[8 * q + 2], x2[8 * q + 3], p.out_f16),
                                     pack_act16x2(x2[8 * q + 4], x2[8 * q + 5], p.out_f16), pack_act16x2(x2[8 * q + 6], x2[8 * q + 7], p.out_f16));
            }
          }
        }
      }
      if (to_cache) continue;
      next_buffer(stage, my_line);
      if (p.out_f16) stage_store16<true>(my_line, sw, x1, x2); else stage_store16<false>(my_line, sw, x1, x2);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { tma_store_2d(tmC, stage, col0, row0); tma_store_commit(); }
      if constexpr (EPI == EPI_STORE_BF16 || EPI == EPI_GELU_BF16) {
        if (p.out_split) {   // BF16_SPLIT: the next GEMM reads this buffer as [hi | lo]; lo = bf16(x - bf16(x))
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            x1[j] -= __bfloat162float(__float2bfloat16_rn(x1[j]));
            x2[j] -= __bfloat162float(__float2bfloat16_rn(x2[j]));
          }
          next_buffer(stage, my_line);
          stage_store16<false>(my_line, sw, x1, x2);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) { tma_store_2d(tmC, stage, p.out_split + col0, row0); tma_store_commit(); }
        }
      }
    }
  }
  // all TMEM reads of this accumulator are done: hand it back to the MMA warp
  tcgen05_fence_before();
  __syncwarp();
  if (lane == 0) asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(release_bar) : "memory");
}

// ---------------------------------------------------------------------------------------------
// Kernel
// ---------------------------------------------------------------------------------------------
// Wave meeting (b2s_set_option "gemm_wave_sync", default every 4 tiles).  CTAs that share an A or W panel only hit in
// L2 while they walk K together; nothing keeps persistent CTAs together, and once they drift the panels are re-read
// from DRAM (measured on the CTA-pair kernels: 63 -> 24 GB per launch and 1052 -> 1244 TFLOP/s sustained on the
// down projection, profiles/r1_gemm_wave_sync.txt).  Every `every` tiles the producer threads of all CTAs meet at a
// global counter.  Best effort: a CTA that waits longer than 4 ms (another grid is holding SMs) stops meeting, so the
// kernel can never dead-lock on residency; results do not depend on the meeting at all.
__device__ __forceinline__ void wave_meet(const GemmParams& p, int wave, int expect, bool& enabled) {
  if (!enabled || p.wave_sync == nullptr || wave == 0 || wave % p.wave_sync_every != 0) return;
  const int slot = wave / p.wave_sync_every;
  atomicAdd(p.wave_sync + slot, 1);
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while (*reinterpret_cast<volatile int*>(p.wave_sync + slot) < expect) {
    __nanosleep(64);
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > 4000000ull) { enabled = false; return; }
  }
}

template <int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES * STAGE_BYTES;
  // barrier layout (8 B each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr (4 B)
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // whole warp: allocate all 512 TMEM columns (1 CTA/SM, so no contention)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      bool meet = true;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int mb, nb;
        tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, p.serp, mb, nb);
        {
          const int wave = (tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x);
          wave_meet(p, wave, min(static_cast<int>(gridDim.x), num_tiles - wave * static_cast<int>(gridDim.x)), meet);
        }
        const bool rev = p.serp == 2 ? ((((tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x)) & 1) != 0) : (p.serp == 1 && (nb & 1));
        for (int kq = 0; kq < p.num_kb; ++kq) {
          const int kb = rev ? p.num_kb - 1 - kq : kq;   // the MMA side only counts k-blocks, so the order is free
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_STAGE_BYTES;
          mbar_arrive_expect_tx(full_bar(stage), STAGE_BYTES);
          const int kw = (p.k_wrap != 0 && kb >= p.k_wrap) ? kb - p.k_wrap : kb;   // lo half of A meets the same W block
          tma_load_2d_h(sa, &tmA, full_bar(stage), kb * BK, mb * BM, p.hint_a);
          tma_load_2d_h(sb, &tmB, full_bar(stage), kw * BK, nb * BN, p.hint_b);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      // a_format (bits 7..9) and b_format (bits 10..12): BF16 = 1, F16 = 0 -- the hardware wants both alike
      const uint32_t idesc = make_idesc_bf16(BM, BN) & ~(p.act_f16 ? ((7u << 7) | (7u << 10)) : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);   // epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t sb = sa + A_STAGE_BYTES;
          const uint64_t adesc = make_kmajor_sw128_desc(sa);
          const uint64_t bdesc = make_kmajor_sw128_desc(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 16 elements (32 B) along K inside the 128B swizzle atom: +2 in the (>>4) address field
            tcgen05_mma_bf16(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tcgen05_commit(empty_bar(stage));   // frees the smem slot when these MMAs have read it
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        tcgen05_commit(tfull_bar(acc));       // accumulator complete -> epilogue
      }
    }
  } else {
    // ===================== epilogue warps (2..5) =====================
    const int quad = warp & 3;                      // TMEM lane quadrant this warp may read
    uint32_t epi_flip = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      int mb, nb;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, p.serp, mb, nb);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int row = mb * BM + quad * 32 + lane;
      if constexpr (EPI == EPI_RESID_F32) prefetch_resid_tile(p, row, nb);
      mbar_wait_warp(tfull_bar(acc), acc_phase, lane);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(acc * BN);
      if constexpr (EPI != EPI_LMHEAD) {
        if (p.tma_store) {
          epilogue_tile_tma<EPI, 2>(p, &tmC, mb * BM + quad * 32, nb, taddr, lane, tempty_bar(acc),
                                    bars + 1024u + static_cast<uint32_t>((warp - 2) * 2 * EPI_STAGE_BYTES), epi_flip);
          continue;
        }
      }
      epilogue_tile<EPI>(p, row, nb, taddr, lane, tempty_bar(acc));
    }
    if (p.tma_store && lane == 0) tma_store_wait_all();   // bulk stores must have left shared memory before the CTA exits
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// CTA-pair kernel (cta_group::2): the two CTAs of a cluster compute one 256 x 256 tile.  Each CTA stages its own
// 128 rows of A and its own 128 rows (N) of W per k-block (32 KB per stage instead of 48 KB, so the ring is
// 6 deep and every W box is fetched once per pair); the leader CTA issues tcgen05.mma.cta_group::2 (M = 256)
// for both, and each CTA's TMEM receives its own 128 accumulator rows.  Barrier protocol:
//   full[s]    in the LEADER's smem; both CTAs' TMA loads complete_tx on it (64 KB per stage)
//   empty[s]   in each CTA's smem; released by the leader's multicast tcgen05.commit
//   tfull[a]   in each CTA's smem; multicast commit when an accumulator is complete
//   tempty[a]  in the LEADER's smem; the 4 epilogue warps of BOTH CTAs arrive (count 8)
// ---------------------------------------------------------------------------------------------
static constexpr int STAGES2 = 6;
static constexpr int STAGE2_BYTES = 2 * A_STAGE_BYTES;   // 16 KB A + 16 KB W
static constexpr int SMEM2_BYTES = STAGES2 * STAGE2_BYTES + 1024 + 1024 + 8 * EPI_STAGE_BYTES;
static constexpr int BM2 = 256;

__device__ __forceinline__ void tma_load_2d_cg2(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar, int32_t c0, int32_t c1,
                                                unsigned long long hint) {
  if (hint != 0) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(hint)
        : "memory");
    return;
  }
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tcgen05_commit_cg2_mc(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(static_cast<uint16_t>(3)) : "memory");
}
__device__ __forceinline__ void tcgen05_mma_bf16_cg2(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                     uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Column-band order: the W rows of one band (band_n x 256 rows of W, sized to stay L2-resident) are met by every row block
// before the next band starts, so W is read from DRAM once and A once per band.  For launches whose whole W is L2-sized
// but does not survive next to the A and residual streams (QKV, dense: 42 MB), the row-group order re-read W every wave.
__device__ __forceinline__ void tile_coords2_band(int tile, int tiles_m, int tiles_n, int band_n, int& mb, int& nb) {
  const int per_band = band_n * tiles_m;
  const int band = tile / per_band;
  const int first_n = band * band_n;
  const int bn = min(band_n, tiles_n - first_n);
  const int within = tile - band * per_band;
  nb = first_n + within % bn;
  mb = within / bn;
}
__device__ __forceinline__ void tile_coords2(int tile, int tiles_m, int tiles_n, int group_m, int serp, int& mb, int& nb) {
  const int GM = max(1, group_m / 2);    // groups of group_m x 128 rows = GM x 256 rows
  const int per_group = GM * tiles_n;
  const int g = tile / per_group;
  const int first_m = g * GM;
  const int gm = min(GM, tiles_m - first_m);
  const int within = tile - g * per_group;
  mb = first_m + within % gm;
  nb = within / gm;
  if (serp && (g & 1)) nb = tiles_n - 1 - nb;
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES2 * STAGE2_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES2 + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES2 + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES2 + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES2 + 4);
  constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's leader CTA

  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const bool leader = rank == 0;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m2 = (p.M + BM2 - 1) / BM2;
  const int num_tiles = tiles_m2 * p.tiles_n;
  const int num_pairs = gridDim.x >> 1;
  const int pair = static_cast<int>((static_cast<long long>(blockIdx.x >> 1) * p.pair_stride) % num_pairs);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES2; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();          // barrier inits and TMEM allocation of both CTAs are visible to both
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      bool meet = true;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int mb, nb;
        if (p.band_n > 0) tile_coords2_band(tile, tiles_m2, p.tiles_n, p.band_n, mb, nb);
        else tile_coords2(tile, tiles_m2, p.tiles_n, p.group_m, p.serp, mb, nb);
        wave_meet(p, (tile - pair) / num_pairs, 2 * min(num_pairs, num_tiles - ((tile - pair) / num_pairs) * num_pairs), meet);
        const bool rev = p.serp == 2 ? ((((tile - pair) / num_pairs) & 1) != 0) : (p.serp == 1 && (nb & 1));
        for (int kq = 0; kq < p.num_kb; ++kq) {
          const int kb = rev ? p.num_kb - 1 - kq : kq;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * STAGE2_BYTES;
          const uint32_t sb = sa + A_STAGE_BYTES;
          const uint32_t fb = full_bar(stage) & PEER_MASK;
          if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * STAGE2_BYTES);
          const int kw = (p.k_wrap != 0 && kb >= p.k_wrap) ? kb - p.k_wrap : kb;
          if (p.prefetch > 0 && !rev) {     // boxes of k-block kq + prefetch: this tile's, or the first ones of this pair's next tile
            int kp = kq + p.prefetch, pmb = mb, pnb = nb;
            bool ok = true;
            if (kp >= p.num_kb) {
              kp -= p.num_kb;
              const int nt = tile + num_pairs;
              ok = nt < num_tiles && kp < p.num_kb;
              if (ok) {
                if (p.band_n > 0) tile_coords2_band(nt, tiles_m2, p.tiles_n, p.band_n, pmb, pnb);
                else tile_coords2(nt, tiles_m2, p.tiles_n, p.group_m, p.serp, pmb, pnb);
              }
            }
            if (ok) {
              const int kpw = (p.k_wrap != 0 && kp >= p.k_wrap) ? kp - p.k_wrap : kp;
              tma_prefetch_2d(&tmA, kp * BK, pmb * BM2 + static_cast<int>(rank) * BM);
              if (kp < p.k_wrap || p.k_wrap == 0) tma_prefetch_2d(&tmB, kpw * BK, pnb * BN + static_cast<int>(rank) * (BN / 2));
            }
          }
          tma_load_2d_cg2(sa, &tmA, fb, kb * BK, mb * BM2 + static_cast<int>(rank) * BM, p.hint_a);
          tma_load_2d_cg2(sb, &tmB, fb, kw * BK, nb * BN + static_cast<int>(rank) * (BN / 2), p.hint_b);
          if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      const uint32_t idesc = make_idesc_bf16(BM2, BN) & ~(p.act_f16 ? ((7u << 7) | (7u << 10)) : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * STAGE2_BYTES;
          const uint32_t sb = sa + A_STAGE_BYTES;
          const uint64_t adesc = make_kmajor_sw128_desc(sa);
          const uint64_t bdesc = make_kmajor_sw128_desc(sb);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            tcgen05_mma_bf16_cg2(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          tcgen05_commit_cg2_mc(empty_bar(stage));
          if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
        }
        tcgen05_commit_cg2_mc(tfull_bar(acc));
      }
    }
  } else {
    // ===================== epilogue warps (2..5), both CTAs =====================
    const int quad = warp & 3;
    uint32_t epi_flip = 0;
    int it = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
      int mb, nb;
      if (p.band_n > 0) tile_coords2_band(tile, tiles_m2, p.tiles_n, p.band_n, mb, nb);
      else tile_coords2(tile, tiles_m2, p.tiles_n, p.group_m, p.serp, mb, nb);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int row = mb * BM2 + static_cast<int>(rank) * BM + quad * 32 + lane;
      if constexpr (EPI == EPI_RESID_F32) prefetch_resid_tile(p, row, nb);
      mbar_wait_warp(tfull_bar(acc), acc_phase, lane);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(acc * BN);
      if constexpr (EPI != EPI_LMHEAD) {
        if (p.tma_store) {
          epilogue_tile_tma<EPI, 2>(p, &tmC, mb * BM2 + static_cast<int>(rank) * BM + quad * 32, nb, taddr, lane,
                                    tempty_bar(acc) & PEER_MASK, bars + 1024u + static_cast<uint32_t>((warp - 2) * 2 * EPI_STAGE_BYTES), epi_flip);
          continue;
        }
      }
      epilogue_tile<EPI>(p, row, nb, taddr, lane, tempty_bar(acc) & PEER_MASK);
    }
    if (p.tma_store && lane == 0) tma_store_wait_all();   // bulk stores must have left shared memory before the CTA exits
  }

  tcgen05_fence_before();
  cluster_sync_all();          // neither CTA may exit (or free TMEM) while its peer can still signal it
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Big-tile CTA-pair kernel: the pair computes a 512 x 256 tile, each CTA 256 rows as two 128-row accumulators that
// fill all 512 TMEM columns.  Per k-block a CTA stages 256 rows of A (32 KB) and its 128 rows of W (16 KB) and the
// leader issues two M=256 MMAs (upper and lower row halves) against the same W stage, so 25 % fewer bytes cross
// L2 -> SM per FLOP than with 256 x 256 tiles and the W boxes are read half as often.  There is no second
// accumulator set: the epilogue (8 warps, 4 per accumulator) is exposed between tiles, which costs a few percent of
// tensor time but draws less power -- under the 1 kW cap that is the better trade (profiles/r1_gemm_power.md).
// ---------------------------------------------------------------------------------------------
static constexpr int STAGES3 = 4;
static constexpr int A3_STAGE_BYTES = 2 * A_STAGE_BYTES;            // 256 x 64 bf16
static constexpr int STAGE3_BYTES = A3_STAGE_BYTES + A_STAGE_BYTES; // + 128 x 64 of W
static constexpr int SMEM3_BYTES = STAGES3 * STAGE3_BYTES + 1024 + 1024 + 8 * EPI_STAGE_BYTES;
static constexpr int BM3 = 512;
static constexpr int NUM_THREADS3 = 320;   // producer, MMA, 8 epilogue warps

__device__ __forceinline__ void tile_coords3(int tile, int tiles_m, int tiles_n, int group_m, int serp, int& mb, int& nb) {
  const int GM = max(1, group_m / 4);    // groups of group_m x 128 rows = GM x 512 rows
  const int per_group = GM * tiles_n;
  const int g = tile / per_group;
  const int first_m = g * GM;
  const int gm = min(GM, tiles_m - first_m);
  const int within = tile - g * per_group;
  mb = first_m + within % gm;
  nb = within / gm;
  if (serp && (g & 1)) nb = tiles_n - 1 - nb;
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS3, 1)
gemm_tc3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES3 * STAGE3_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES3 + s); };
  const uint32_t tfull_bar = bars + 8u * (2 * STAGES3);
  const uint32_t tempty_bar = bars + 8u * (2 * STAGES3 + 1);
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES3 + 2);
  constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;

  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const bool leader = rank == 0;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m3 = (p.M + BM3 - 1) / BM3;
  const int num_tiles = tiles_m3 * p.tiles_n;
  const int num_pairs = gridDim.x >> 1;
  const int pair = static_cast<int>((static_cast<long long>(blockIdx.x >> 1) * p.pair_stride) % num_pairs);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES3; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    mbar_init(tempty_bar, 16);     // 8 epilogue warps in each CTA of the pair
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      bool meet = true;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int mb, nb;
        tile_coords3(tile, tiles_m3, p.tiles_n, p.group_m, p.serp, mb, nb);
        wave_meet(p, (tile - pair) / num_pairs, 2 * min(num_pairs, num_tiles - ((tile - pair) / num_pairs) * num_pairs), meet);
        const bool rev = p.serp == 2 ? ((((tile - pair) / num_pairs) & 1) != 0) : (p.serp == 1 && (nb & 1));
        for (int kq = 0; kq < p.num_kb; ++kq) {
          const int kb = rev ? p.num_kb - 1 - kq : kq;
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * STAGE3_BYTES;
          const uint32_t sb = sa + A3_STAGE_BYTES;
          const uint32_t fb = full_bar(stage) & PEER_MASK;
          if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * STAGE3_BYTES);
          const int kw = (p.k_wrap != 0 && kb >= p.k_wrap) ? kb - p.k_wrap : kb;
          if (p.prefetch > 0 && !rev) {
            int kp = kq + p.prefetch, pmb = mb, pnb = nb;
            bool ok = true;
            if (kp >= p.num_kb) {
              kp -= p.num_kb;
              const int nt = tile + num_pairs;
              ok = nt < num_tiles && kp < p.num_kb;
              if (ok) tile_coords3(nt, tiles_m3, p.tiles_n, p.group_m, p.serp, pmb, pnb);
            }
            if (ok) {
              const int kpw = (p.k_wrap != 0 && kp >= p.k_wrap) ? kp - p.k_wrap : kp;
              tma_prefetch_2d(&tmA, kp * BK, pmb * BM3 + static_cast<int>(rank) * 256);
              if (kp < p.k_wrap || p.k_wrap == 0) tma_prefetch_2d(&tmB, kpw * BK, pnb * BN + static_cast<int>(rank) * (BN / 2));
            }
          }
          tma_load_2d_cg2(sa, &tmA, fb, kb * BK, mb * BM3 + static_cast<int>(rank) * 256, p.hint_a);
          tma_load_2d_cg2(sb, &tmB, fb, kw * BK, nb * BN + static_cast<int>(rank) * (BN / 2), p.hint_b);
          if (++stage == STAGES3) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      const uint32_t idesc = make_idesc_bf16(256, BN) & ~(p.act_f16 ? ((7u << 7) | (7u << 10)) : 0u);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
        mbar_wait(tempty_bar, (static_cast<uint32_t>(it) & 1u) ^ 1u);   // both accumulators drained by all 16 warps
        tcgen05_fence_after();
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + stage * STAGE3_BYTES;
          const uint64_t adesc0 = make_kmajor_sw128_desc(sa);
          const uint64_t adesc1 = make_kmajor_sw128_desc(sa + A_STAGE_BYTES);      // rows 128..255 of this CTA's A box
          const uint64_t bdesc = make_kmajor_sw128_desc(sa + A3_STAGE_BYTES);
          const uint32_t accf = kb != 0 ? 1u : 0u;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            tcgen05_mma_bf16_cg2(tmem_base, adesc0 + 2u * k, bdesc + 2u * k, idesc, (accf | k) != 0 ? 1u : 0u);
            tcgen05_mma_bf16_cg2(tmem_base + static_cast<uint32_t>(BN), adesc1 + 2u * k, bdesc + 2u * k, idesc,
                                 (accf | k) != 0 ? 1u : 0u);
          }
          tcgen05_commit_cg2_mc(empty_bar(stage));
          if (++stage == STAGES3) { stage = 0; phase ^= 1u; }
        }
        tcgen05_commit_cg2_mc(tfull_bar);
      }
    }
  } else {
    // ===================== epilogue warps 2..9: warps 2-5 drain accumulator 0, warps 6-9 accumulator 1 =====================
    const int quad = warp & 3;
    const int acc = (warp - 2) >> 2;
    uint32_t epi_flip = 0;
    int it = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs, ++it) {
      int mb, nb;
      tile_coords3(tile, tiles_m3, p.tiles_n, p.group_m, p.serp, mb, nb);
      const int row = mb * BM3 + static_cast<int>(rank) * 256 + acc * BM + quad * 32 + lane;
      if constexpr (EPI == EPI_RESID_F32) prefetch_resid_tile(p, row, nb);
      mbar_wait_warp(tfull_bar, static_cast<uint32_t>(it) & 1u, lane);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(acc * BN);
      if constexpr (EPI != EPI_LMHEAD) {
        if (p.tma_store) {
          epilogue_tile_tma<EPI, 1>(p, &tmC, mb * BM3 + static_cast<int>(rank) * 256 + acc * BM + quad * 32, nb, taddr, lane,
                                 tempty_bar & PEER_MASK, bars + 1024u + static_cast<uint32_t>((warp - 2) * EPI_STAGE_BYTES), epi_flip);
          continue;
        }
      }
      epilogue_tile<EPI>(p, row, nb, taddr, lane, tempty_bar & PEER_MASK);
    }
    if (p.tma_store && lane == 0) tma_store_wait_all();   // bulk stores must have left shared memory before the CTA exits
  }

  tcgen05_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static int get_encode_fn() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
  });
  if (g_encode == nullptr) {
    set_error("cuTensorMapEncodeTiled not available from the CUDA driver");
    return B2S_ECUDA;
  }
  return B2S_OK;
}

// Encoded maps are pure functions of their arguments (no device state), and a scoring call asks for the same few
// hundred of them over and over (32 layers x 4 weights, a handful of activation buffers, a few row counts), so they are
// memoised: cuTensorMapEncodeTiled costs about as much as a kernel launch, and a one-token step is launch-bound.
struct TmapKey {
  const void* base; uint64_t rows, cols, pitch; uint32_t box_rows, box_cols; int esize, f16;
  bool operator==(const TmapKey& o) const {
    return base == o.base && rows == o.rows && cols == o.cols && pitch == o.pitch && box_rows == o.box_rows &&
           box_cols == o.box_cols && esize == o.esize && f16 == o.f16;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.base) * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.cols * 1315423911ull + k.pitch + (h << 6) + (h >> 2));
    h ^= (static_cast<uint64_t>(k.box_rows) << 32 | k.box_cols) + static_cast<uint64_t>(k.esize * 2 + k.f16) + (h << 6) + (h >> 2);
    return static_cast<size_t>(h);
  }
};
static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;
static std::mutex g_tmap_mutex;

int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch_bytes,
                 uint32_t box_rows, uint32_t box_cols, int esize, int f16) {
  B2S_TRY(get_encode_fn());
  const TmapKey key{base, rows, cols, pitch_bytes, box_rows, box_cols, esize, f16};
  {
    std::lock_guard<std::mutex> g(g_tmap_mutex);
    auto it = g_tmap_cache.find(key);
    if (it != g_tmap_cache.end()) { *map = it->second; return B2S_OK; }
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = esize == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                            : (f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  CUresult r = g_encode(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu pitch=%llu base=%p)", static_cast<int>(r),
              static_cast<unsigned long long>(rows), static_cast<unsigned long long>(cols),
              static_cast<unsigned long long>(pitch_bytes), base);
    return B2S_ECUDA;
  }
  {
    std::lock_guard<std::mutex> g(g_tmap_mutex);
    if (g_tmap_cache.size() > 16384) g_tmap_cache.clear();   // callers that keep changing buffers must not grow it forever
    g_tmap_cache.emplace(key, *map);
  }
  return B2S_OK;
}

int make_tmap_16bit_3d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t segs, uint64_t seg_cols, uint64_t pitch_bytes,
                       uint32_t box_rows, uint32_t box_cols, int f16) {
  B2S_TRY(get_encode_fn());
  cuuint64_t dims[3] = {seg_cols, segs, rows};
  cuuint64_t strides[2] = {seg_cols * 2, pitch_bytes};
  cuuint32_t box[3] = {box_cols, 1, box_rows};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims,
                        strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (3D) failed with CUresult %d (rows=%llu segs=%llu seg_cols=%llu pitch=%llu base=%p)", static_cast<int>(r),
              static_cast<unsigned long long>(rows), static_cast<unsigned long long>(segs), static_cast<unsigned long long>(seg_cols),
              static_cast<unsigned long long>(pitch_bytes), base);
    return B2S_ECUDA;
  }
  return B2S_OK;
}

int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch_bytes,
                      uint32_t box_rows, uint32_t box_cols, int f16) {
  return make_tmap_2d(map, base, rows, cols, pitch_bytes, box_rows, box_cols, 2, f16);
}

// row-major [rows, cols] bf16 matrix, box = [box_rows, 64 cols]
static int make_tmap_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows, int f16 = 0) {
  return make_tmap_bf16_2d(map, base, rows, cols, cols * 2, box_rows, BK, f16);
}

// Per-device facts (function attributes and the SM count belong to a device, and one process may drive several).
static constexpr int MAX_DEVICES = 64;
static int g_num_sms_dev[MAX_DEVICES] = {0};
static int current_device_sms(int* dev_out, int* sms_out) {
  int dev = 0;
  B2S_CHECK_CUDA(cudaGetDevice(&dev));
  B2S_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "gemm: device ordinal %d outside [0, %d)", dev, MAX_DEVICES);
  if (g_num_sms_dev[dev] == 0) B2S_CHECK_CUDA(cudaDeviceGetAttribute(&g_num_sms_dev[dev], cudaDevAttrMultiProcessorCount, dev));
  *dev_out = dev; *sms_out = g_num_sms_dev[dev];
  return B2S_OK;
}
extern int g_opt_attn_hold_s;
extern int g_opt_attn_ts;
extern int g_opt_compact;
int g_opt_tma_store = 1;   // 1: epilogue via shared memory + TMA bulk store / reduce-add; 0: per-thread global stores
int g_opt_hint = 0;        // TMA L2 cache hints: 0 none, 1 A evict_first + W evict_last, 2 A evict_last + W evict_first, 3 both evict_last
int g_opt_small_w_group = 8;   // rasterisation group when W is L2-sized (0 = rule off; b2s_set_option "gemm_small_w_group")
// b2s_set_option "gemm_serpentine": 0 (default) plain order.  2 = odd waves walk K backwards: +1 % end to end (each wave
// starts on the panel bytes the previous one touched last) but a row's fp32 summation order then depends on which wave
// its tile falls in, i.e. on the batch around it -- sharded and unsharded runs stop being bit-identical.  1 = direction by
// column tile: batch-invariant, but CTAs that share an A row-block no longer walk K together: -1 %.
int g_opt_prefetch = 0;      // b2s_set_option "gemm_prefetch": L2 prefetch distance of the CTA-pair producers in k-blocks (0 = off)
int g_opt_band_n = -1;       // b2s_set_option "gemm_band_n": 0 = row-group order; -1 (default) = column bands when W is L2-sized; k > 0 = bands of k tiles
int g_opt_policy = 1;        // b2s_set_option "gemm_policy" (see launch_epi)
int g_opt_wave_sync = 4;     // b2s_set_option "gemm_wave_sync": all producers meet every N tiles (0 = never)
// Meeting counters: one table per device (a __device__ array exists once on every GPU that loads the library).  Each
// launch takes its own slice at a rotating offset, so GEMMs that overlap on one device (several streams or host
// threads) count in different slots; the table holds the slices of some hundreds of launches before it wraps.
static constexpr int WAVE_TABLE = 65536;
static __device__ int g_wave_counters[WAVE_TABLE];
static std::atomic<unsigned> g_wave_cursor{0};
int g_opt_pair_stride = 1;   // b2s_set_option "gemm_pair_stride" (experiment; must be coprime to the number of CTA pairs)
int g_opt_serp = 0;
int g_opt_group_m = 0;     // 0: heuristic below; otherwise forced (b2s_set_option "gemm_group_m")
int g_opt_cta_group = 0;   // 0: per-launch policy (default); 1: one CTA per 128x256 tile; 2: CTA pairs on 256x256; 3: CTA pairs on 512x256 (b2s_set_option "gemm_cta_group")

// ---- optional per-launch timing (bench.py roofline): CUDA events on the launching stream ----------
struct GemmProfile {
  bool on = false;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;       // events handed out since the last read (pairs: start, stop)
  double flops = 0.0;
};
static GemmProfile g_prof;

static cudaEvent_t prof_event() {
  if (g_prof.used == g_prof.pool.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    g_prof.pool.push_back(e);
  }
  return g_prof.pool[g_prof.used++];
}

static int arm_wave_sync(GemmParams& q, int tiles, int pairs, cudaStream_t st) {
  if (g_opt_wave_sync <= 0) return B2S_OK;
  const int waves = static_cast<int>(ceil_div(tiles, pairs));
  const int need = waves / g_opt_wave_sync + 1;
  if (waves <= g_opt_wave_sync || need > WAVE_TABLE / 4) return B2S_OK;   // no meeting would happen / table too small
  int* ctr = nullptr;
  B2S_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ctr), g_wave_counters));   // address on the current device
  unsigned at = g_wave_cursor.fetch_add(static_cast<unsigned>(need)) % WAVE_TABLE;
  if (at + need > WAVE_TABLE) at = 0;                                      // slices never straddle the end of the table
  B2S_CHECK_CUDA(cudaMemsetAsync(ctr + at, 0, sizeof(int) * need, st));
  q.wave_sync = ctr + at; q.wave_sync_every = g_opt_wave_sync;
  return B2S_OK;
}

template <int EPI>
static int launch_epi(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tb2, const CUtensorMap& ta3,
                      const CUtensorMap& tc, const GemmParams& p, cudaStream_t st) {
  int dev = 0, num_sms = 0;
  B2S_TRY(current_device_sms(&dev, &num_sms));
  static bool attr_set[MAX_DEVICES] = {false};   // the dynamic shared memory opt-in is a per-device function attribute
  if (!attr_set[dev]) {
    B2S_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    B2S_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc2_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2_BYTES));
    B2S_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc3_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM3_BYTES));
    attr_set[dev] = true;
  }
  const int tiles = p.tiles_m * p.tiles_n;
  const int grid = tiles < num_sms ? tiles : num_sms;
  if (g_prof.on) cudaEventRecord(prof_event(), st);
  // Tiling policy (gemm_cta_group 0).  Under the 1 kW power cap the variants trade clock for tensor-pipe utilisation
  // (profiles/r1_gemm_power.md).  Since the wave meeting keeps the CTA-pair kernels' panels in L2 they win everywhere:
  // 512x256 pair tiles when K is long enough to amortise their exposed epilogue and there are enough 512-row blocks
  // (down projection), 256x256 pair tiles otherwise (QKV, dense, up + GELU, LM head, one-token steps).
  int cg = g_opt_cta_group;
  if (cg == 0) {
    if (g_opt_policy == 1)   // with the wave meeting the CTA-pair kernels keep their panels in L2: 512x256 pair tiles for long K
      cg = (p.M >= 4096 && p.K >= 8192) ? 3 : 2;
    else                     // round-1 policy before the wave meeting: single-CTA tiles for long-K and GELU launches
      cg = (p.M >= 4096 && (p.K >= 8192 || EPI == EPI_GELU_BF16)) ? 1 : 2;
  }
  const bool big = (cg == 3 && p.M >= 4096) || cg == 4;   // 4 = always (tests)
  if (big) {
    const int tiles3 = static_cast<int>(ceil_div(p.M, BM3)) * p.tiles_n;
    int pairs = num_sms / 2;
    if (tiles3 < pairs) pairs = tiles3;
    GemmParams q = p;
    if (pairs != num_sms / 2) q.pair_stride = 1;   // the permutation is only a bijection on the full grid
    B2S_TRY(arm_wave_sync(q, tiles3, pairs, st));
    gemm_tc3_kernel<EPI><<<2 * pairs, NUM_THREADS3, SMEM3_BYTES, st>>>(ta3, tb2, tc, q);
  } else if (cg >= 2) {
    const int tiles2 = static_cast<int>(ceil_div(p.M, BM2)) * p.tiles_n;
    int pairs = num_sms / 2;
    if (tiles2 < pairs) pairs = tiles2;
    GemmParams q = p;
    if (pairs != num_sms / 2) q.pair_stride = 1;
    B2S_TRY(arm_wave_sync(q, tiles2, pairs, st));
    gemm_tc2_kernel<EPI><<<2 * pairs, NUM_THREADS, SMEM2_BYTES, st>>>(ta, tb2, tc, q);
  } else {
    GemmParams q = p;
    B2S_TRY(arm_wave_sync(q, tiles, grid, st));
    gemm_tc_kernel<EPI><<<grid, NUM_THREADS, SMEM_BYTES, st>>>(ta, tb, tc, q);
  }
  if (g_prof.on) {
    cudaEventRecord(prof_event(), st);
    g_prof.flops += 2.0 * p.M * p.N * p.K;
  }
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int gemm_tc_launch(const void* A, const void* W, const float* bias, void* out, int M, int N, int K, int epilogue,
                   const LmHeadArgs* lmh, cudaStream_t st, int act_f16, const RopeEpiArgs* rope, int a_split, int out_split,
                   int out_f16, int act_kind) {
  B2S_REQUIRE(!a_split || (K % BK == 0 && !act_f16), "gemm_bf16: split activations need K %% 64 == 0 and bf16 operands (K=%d)", K);
  B2S_REQUIRE(!out_split || (N % (epilogue == EPI_SWIGLU ? 128 : 64) == 0 && g_opt_tma_store &&
                             (epilogue == EPI_STORE_BF16 || epilogue == EPI_GELU_BF16 || epilogue == EPI_SWIGLU)),
              "gemm_bf16: a split 16-bit output needs whole 64-column store groups, the TMA-store epilogue and a 16-bit store epilogue (N=%d)", N);
  B2S_REQUIRE(epilogue != EPI_SWIGLU || (g_opt_tma_store && N % 16 == 0), "gemm_bf16: the SwiGLU epilogue needs the TMA-store epilogue and N %% 16 == 0");
  B2S_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_bf16: empty problem M=%d N=%d K=%d", M, N, K);
  B2S_REQUIRE(K % 8 == 0, "gemm_bf16: K=%d must be a multiple of 8 (16-byte TMA row pitch)", K);
  B2S_REQUIRE(epilogue == EPI_LMHEAD || N % 8 == 0, "gemm_bf16: N=%d must be a multiple of 8", N);
  B2S_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "gemm_bf16: operands must be 16-byte aligned");
  int dev = 0, num_sms = 0;
  B2S_TRY(current_device_sms(&dev, &num_sms));
  CUtensorMap ta, tb, tb2, ta3;
  const uint64_t KA = static_cast<uint64_t>(K) * (a_split ? 2 : 1);   // split activations: A is [M, hi(K) | lo(K)]
  B2S_TRY(make_tmap_bf16(&ta, A, static_cast<uint64_t>(M), KA, BM, act_f16));
  B2S_TRY(make_tmap_bf16(&tb, W, static_cast<uint64_t>(N), static_cast<uint64_t>(K), BN, act_f16));
  B2S_TRY(make_tmap_bf16(&ta3, A, static_cast<uint64_t>(M), KA, 256, act_f16));   // big-tile kernel: 256-row A boxes
  B2S_TRY(make_tmap_bf16(&tb2, W, static_cast<uint64_t>(N), static_cast<uint64_t>(K), BN / 2, act_f16));   // per-CTA half of a W box
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = static_cast<int>(ceil_div(M, BM));
  p.tiles_n = static_cast<int>(ceil_div(N, BN));
  p.num_kb = static_cast<int>(ceil_div(K, BK)) * (a_split ? 2 : 1);
  p.k_wrap = a_split ? K / BK : 0;
  p.out_split = out_split ? (epilogue == EPI_SWIGLU ? N / 2 : N) : 0;
  // Rasterisation: the A panel of a group (group_m x 128 rows x K) is re-read by every wave of the group, so it
  // should stay L2-resident (<= ~40 MB); when K is so long that even 16 row-blocks exceed that, no cross-wave
  // reuse is possible and the near-square wave footprint (16 x 9 tiles) minimises the bytes each wave streams.
  {
    const double mb_bytes = 2.0 * BM * K;
    int gm = 64;
    while (gm > 8 && gm * mb_bytes > 40e6) gm >>= 1;      // K = 18176 (down projection): 8 row-blocks = 37 MB (measured: -0.9 % time, -9 % DRAM vs 16; profiles/r2_gemm_probe.json)
    // a W that can stay L2-resident by itself (qkv, dense: 42 MB) is better served by one-wave groups: the A panel
    // is then 9 MB and W is the only thing re-read (measured: profiles/r1_gemm_l2_hints.txt, DRAM 5.4 -> 4.1 GB on qkv)
    if (g_opt_small_w_group > 0 && 2.0 * N * K <= 48e6) gm = g_opt_small_w_group;
    p.group_m = g_opt_group_m > 0 ? g_opt_group_m : gm;
    p.prefetch = g_opt_prefetch;
    p.band_n = 0;
    if (g_opt_band_n > 0) p.band_n = g_opt_band_n;
    else if (g_opt_band_n < 0 && 2.0 * N * K <= 48e6 && 2.0 * N * K > 24e6 && p.serp == 0) {
      const int bands = static_cast<int>(ceil_div(static_cast<int64_t>(2.0 * N * K), 22000000));   // W bands of <= 22 MB
      p.band_n = static_cast<int>(ceil_div(p.tiles_n, bands));
    }
    p.serp = g_opt_serp;
    p.wave_sync = nullptr; p.wave_sync_every = 0;
    p.pair_stride = 1;
    {
      const int np = num_sms / 2;
      int a = g_opt_pair_stride, b = np;
      while (b) { const int t = a % b; a = b; b = t; }
      if (g_opt_pair_stride > 1 && a == 1) p.pair_stride = g_opt_pair_stride;
    }
  }
  p.out = out; p.bias = bias;
  p.act_f16 = act_f16 ? 1 : 0;
  p.out_f16 = out_f16 < 0 ? p.act_f16 : (out_f16 ? 1 : 0);
  p.act_kind = act_kind;
  {
    // 0 none | 1 A first, W last | 2 A last, W first | 3 both last | 4 both normal | 5 W last | 6 A first | 7 A normal, W last
    static const unsigned long long HA[8] = {0ull, HINT_EVICT_FIRST, HINT_EVICT_LAST, HINT_EVICT_LAST, HINT_EVICT_NORMAL, 0ull, HINT_EVICT_FIRST, HINT_EVICT_NORMAL};
    static const unsigned long long HB[8] = {0ull, HINT_EVICT_LAST, HINT_EVICT_FIRST, HINT_EVICT_LAST, HINT_EVICT_NORMAL, HINT_EVICT_LAST, 0ull, HINT_EVICT_LAST};
    p.hint_a = HA[g_opt_hint & 7];
    p.hint_b = HB[g_opt_hint & 7];
  }
  p.lmh = lmh ? *lmh : LmHeadArgs{nullptr, nullptr, nullptr};
  p.rope = rope ? *rope : RopeEpiArgs{};
  p.tma_store = (g_opt_tma_store && epilogue != EPI_LMHEAD) ? 1 : 0;
  CUtensorMap tc = ta;   // placeholder when the epilogue writes no [M,N] matrix
  if (p.tma_store) {
    const bool f32 = epilogue == EPI_RESID_F32 || epilogue == EPI_STORE_F32;
    const uint64_t NC = static_cast<uint64_t>(epilogue == EPI_SWIGLU ? N / 2 : N) * (out_split ? 2 : 1);
    B2S_TRY(make_tmap_2d(&tc, out, static_cast<uint64_t>(M), NC, NC * (f32 ? 4 : 2), 32, f32 ? 32 : 64, f32 ? 4 : 2, p.out_f16));
  }
  switch (epilogue) {
    case EPI_STORE_BF16: return launch_epi<EPI_STORE_BF16>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_GELU_BF16: return launch_epi<EPI_GELU_BF16>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_RESID_F32: return launch_epi<EPI_RESID_F32>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_STORE_F32: return launch_epi<EPI_STORE_F32>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_SWIGLU: return launch_epi<EPI_SWIGLU>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_QKV_ROPE:
      B2S_REQUIRE(rope && rope->pos && rope->slot && rope->cos && rope->sin && rope->kcache && rope->vcache,
                  "gemm_bf16: the QKV+RoPE epilogue needs its buffers");
      B2S_REQUIRE(N == (rope->nh + 2 * rope->nkv) * 64, "gemm_bf16: QKV+RoPE epilogue expects N = (nh + 2 nkv) * 64");
      return launch_epi<EPI_QKV_ROPE>(ta, tb, tb2, ta3, tc, p, st);
    case EPI_LMHEAD:
      B2S_REQUIRE(lmh && lmh->partials && lmh->cand_ids && lmh->cand_logits, "gemm_bf16: LM-head epilogue needs its buffers");
      return launch_epi<EPI_LMHEAD>(ta, tb, tb2, ta3, tc, p, st);
    default:
      set_error("gemm_bf16: unknown epilogue %d", epilogue);
      return B2S_EINVAL;
  }
}

}  // namespace b2s

extern "C" {

int b2s_set_option(const char* name, int32_t value) {
  B2S_REQUIRE(name != nullptr, "set_option: null name");
  if (std::string(name) == "gemm_cta_group") {
    B2S_REQUIRE(value >= 0 && value <= 4, "set_option: gemm_cta_group must be 0..4");
    b2s::g_opt_cta_group = value;
    return B2S_OK;
  }
  if (std::string(name) == "lookahead_compact") {
    b2s::g_opt_compact = value ? 1 : 0;
    return B2S_OK;
  }
  if (std::string(name) == "attn_step") {
    b2s::g_opt_attn_step = value ? 1 : 0;
    return B2S_OK;
  }
  if (std::string(name) == "int8_threshold_milli") {
    B2S_REQUIRE(value >= 0, "set_option: int8_threshold_milli must be >= 0");
    b2s::g_opt_int8_threshold = value / 1000.0f;
    return B2S_OK;
  }
  if (std::string(name) == "attn_ts") {
    b2s::g_opt_attn_ts = value ? 1 : 0;
    return B2S_OK;
  }
  if (std::string(name) == "attn_hold_s") {
    b2s::g_opt_attn_hold_s = value ? 1 : 0;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_tma_store") {
    b2s::g_opt_tma_store = value ? 1 : 0;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_l2_hint") {
    B2S_REQUIRE(value >= 0 && value <= 7, "set_option: gemm_l2_hint must be in [0, 7]");
    b2s::g_opt_hint = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_prefetch") {
    B2S_REQUIRE(value >= 0 && value <= 64, "set_option: gemm_prefetch must be in [0, 64]");
    b2s::g_opt_prefetch = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_band_n") {
    B2S_REQUIRE(value >= -1 && value <= 1024, "set_option: gemm_band_n must be in [-1, 1024]");
    b2s::g_opt_band_n = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_policy") {
    B2S_REQUIRE(value == 0 || value == 1, "set_option: gemm_policy must be 0 or 1");
    b2s::g_opt_policy = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_wave_sync") {
    B2S_REQUIRE(value >= 0 && value <= 1024, "set_option: gemm_wave_sync must be in [0, 1024]");
    b2s::g_opt_wave_sync = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_pair_stride") {
    B2S_REQUIRE(value >= 1 && value <= 1024, "set_option: gemm_pair_stride must be in [1, 1024]");
    b2s::g_opt_pair_stride = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_small_w_group") {
    B2S_REQUIRE(value >= 0 && value <= 1024, "set_option: gemm_small_w_group must be in [0, 1024]");
    b2s::g_opt_small_w_group = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_serpentine") {
    B2S_REQUIRE(value >= 0 && value <= 2, "set_option: gemm_serpentine must be 0, 1 or 2");
    b2s::g_opt_serp = value;
    return B2S_OK;
  }
  if (std::string(name) == "gemm_group_m") {
    B2S_REQUIRE(value >= 0 && value <= 1024, "set_option: gemm_group_m must be in [0, 1024]");
    b2s::g_opt_group_m = value;
    return B2S_OK;
  }
  b2s::set_error("set_option: unknown option '%s'", name);
  return B2S_EINVAL;
}

int b2s_profile_gemm(int32_t enable) {
  b2s::g_prof.on = enable != 0;
  return B2S_OK;
}

// Synchronises on the recorded events; returns the sums since the last read and clears them.
int b2s_profile_gemm_read(double* total_ms, int64_t* launches, double* flops) {
  using namespace b2s;
  double ms = 0.0;
  for (size_t i = 0; i + 1 < g_prof.used; i += 2) {
    B2S_CHECK_CUDA(cudaEventSynchronize(g_prof.pool[i + 1]));
    float t = 0.f;
    B2S_CHECK_CUDA(cudaEventElapsedTime(&t, g_prof.pool[i], g_prof.pool[i + 1]));
    ms += t;
  }
  if (total_ms) *total_ms = ms;
  if (launches) *launches = static_cast<int64_t>(g_prof.used / 2);
  if (flops) *flops = g_prof.flops;
  g_prof.used = 0;
  g_prof.flops = 0.0;
  return B2S_OK;
}

}  // extern "C"

