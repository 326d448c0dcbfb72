// Bootstrap kernels: NumPy-legacy MT19937 index stream (bit-exact), resample statistics, percentile CI.
//
// Replaces the `for _ in range(N_BOOTSTRAP)` loops of the reference:
//   analysis/run_base_vs_instruct_100q.py:655-658,692-695
//   survey_analysis/analyze_base_vs_instruct_mae_100q.py:310-315,380-388
//   analysis/analyze_results_base_versus_instruct.py:310-320,347-351
// NumPy legacy stream: np.random.seed(s) = Knuth init of 624 words;
// np.random.choice(n, size) = per draw, words masked to the smallest 2^b-1 >= n-1 and rejected while > n-1.
#include "b2s_common.cuh"
#include <mutex>

namespace b2s {

constexpr int MT_N = 624, MT_M = 397;

__global__ void mt_seed_kernel(uint32_t seed, uint32_t* __restrict__ state) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  uint32_t x = seed;
  state[0] = x;
  for (int i = 1; i < MT_N; ++i) {
    x = 1812433253u * (x ^ (x >> 30)) + static_cast<uint32_t>(i);
    state[i] = x;
  }
  state[MT_N] = MT_N;   // position: state exhausted, next draw twists first
}

__device__ __forceinline__ uint32_t mt_mix(uint32_t y) { return (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u); }
__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// Small draws (a few thousand, e.g. one np.random.choice(n, n) at a time): one CTA walks the stream 624 words at a time.  The twist is done in the four dependency phases of the
// in-place recurrence; acceptance flags are compacted with a block scan so that idx_out holds the accepted
// draws in stream order, exactly as the sequential reference consumes them.
__global__ void __launch_bounds__(640)
mt_choice_kernel(uint32_t* __restrict__ state, uint32_t rng, uint32_t mask, long long count, int32_t* __restrict__ idx_out) {
  __shared__ uint32_t mt[MT_N];
  __shared__ int warp_tot[20];
  __shared__ int s_pos, s_total, s_newpos;
  const int t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  if (t < MT_N) mt[t] = state[t];
  if (t == 0) s_pos = static_cast<int>(state[MT_N]);
  __syncthreads();
  long long produced = 0;
  while (produced < count) {
    int pos = s_pos;
    if (pos >= MT_N) {
      uint32_t y = 0;
      if (t < MT_N - 1) y = (mt[t] & 0x80000000u) | (mt[t + 1] & 0x7fffffffu);
      __syncthreads();
      if (t < MT_N - MT_M) mt[t] = mt[t + MT_M] ^ mt_mix(y);                       // kk in [0,227)
      __syncthreads();
      if (t >= MT_N - MT_M && t < 2 * (MT_N - MT_M)) mt[t] = mt[t - (MT_N - MT_M)] ^ mt_mix(y);   // [227,454)
      __syncthreads();
      if (t >= 2 * (MT_N - MT_M) && t < MT_N - 1) mt[t] = mt[t - (MT_N - MT_M)] ^ mt_mix(y);     // [454,623)
      __syncthreads();
      if (t == MT_N - 1) {
        const uint32_t yy = (mt[MT_N - 1] & 0x80000000u) | (mt[0] & 0x7fffffffu);
        mt[MT_N - 1] = mt[MT_M - 1] ^ mt_mix(yy);
      }
      pos = 0;
      __syncthreads();
    }
    const int w = pos + t;
    uint32_t v = 0;
    int acc = 0;
    if (w < MT_N) {
      v = mt_temper(mt[w]) & mask;
      acc = (v <= rng) ? 1 : 0;
    }
    // block exclusive scan of acc
    int incl = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int x = (lane < 20) ? warp_tot[lane] : 0;
      int xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, xi, o);
        if (lane >= o) xi += n;
      }
      if (lane < 20) warp_tot[lane] = xi - x;   // exclusive warp offsets
      if (lane == 19) s_total = xi;
    }
    __syncthreads();
    const int excl = warp_tot[warp] + incl - acc;
    const int total = s_total;
    const long long need = count - produced;
    if (acc && excl < need) idx_out[produced + excl] = static_cast<int32_t>(v);
    if (total < need) {
      if (t == 0) s_newpos = MT_N;          // whole block consumed, more draws needed
    } else if (acc && excl == need - 1) {
      s_newpos = w + 1;                     // the word after the last one consumed
    }
    __syncthreads();
    produced += (total < need) ? total : need;
    if (t == 0) s_pos = s_newpos;
    __syncthreads();
  }
  if (t < MT_N) state[t] = mt[t];
  if (t == 0) state[MT_N] = static_cast<uint32_t>(s_pos);
}

// ---- np.random.choice, device-wide ---------------------------------------------------------------------------
// The word stream x[0..623] = state, x[j] = x[j-227] ^ mix(x[j-624], x[j-623]) is the same sequence whether it is
// produced 624 words at a time in place (NumPy) or as a sliding recurrence.  Only the recurrence is serial, so it gets
// a kernel of its own that does nothing else (mt_gen_kernel: one CTA, 454 words per barrier); tempering, masking,
// rejection and the stream-order compaction of the accepted draws are data-parallel and run on the whole GPU
// (count -> scan -> scatter over 4096-word chunks).  The raw words are kept in scratch memory because they ARE the
// generator state: after the draw that consumes word e, NumPy's state array is the 624-word block containing e and its
// position is (e mod 624) + 1.
constexpr int MT_GEN_STEP = 454;      // words per barrier: j and j+227 by the same thread (x[j+227] needs x[j] from its own register)
constexpr int MT_RING = 2048;         // circular window of the stream in shared memory (>= 624 + 2*454, power of two)
constexpr int MT_CHUNK = 4096;        // words per CTA in the compaction kernels

__global__ void __launch_bounds__(256)
mt_gen_kernel(const uint32_t* __restrict__ state, uint32_t* __restrict__ raw, long long n_new /* multiple of MT_GEN_STEP */) {
  __shared__ uint32_t x[MT_RING];
  const int t = threadIdx.x;
  for (int i = t; i < MT_N; i += blockDim.x) { const uint32_t v = state[i]; x[i] = v; raw[i] = v; }
  __syncthreads();
  constexpr uint32_t RM = MT_RING - 1;
  uint32_t* out = raw + MT_N + t;                                   // word j of this thread; j + 227 is out[227]
  uint32_t r = static_cast<uint32_t>(MT_N + t);                    // j mod ring size (32-bit arithmetic in the serial loop)
  const long long steps = n_new / MT_GEN_STEP;
  for (long long sidx = 0; sidx < steps; ++sidx) {
    if (t < MT_N - MT_M) {   // 227 lanes
      const uint32_t a = x[(r - 624u) & RM], b = x[(r - 623u) & RM], c = x[(r - 227u) & RM];
      const uint32_t a2 = x[(r - 397u) & RM], b2 = x[(r - 396u) & RM];           // both older than this step
      const uint32_t v0 = c ^ mt_mix((a & 0x80000000u) | (b & 0x7fffffffu));
      const uint32_t v1 = v0 ^ mt_mix((a2 & 0x80000000u) | (b2 & 0x7fffffffu));
      x[r & RM] = v0; x[(r + 227u) & RM] = v1;
      out[0] = v0; out[227] = v1;
    }
    r += MT_GEN_STEP; out += MT_GEN_STEP;
    __syncthreads();
  }
}

// accepted draws of one chunk: word w is consumed iff w >= pos0 (the unconsumed tail of the current state block comes
// first) and its masked value is <= rng
__device__ __forceinline__ int mt_accept(const uint32_t* raw, long long w, long long n_words, long long pos0, uint32_t mask,
                                         uint32_t rng, uint32_t& v) {
  if (w >= n_words || w < pos0) return 0;
  v = mt_temper(raw[w]) & mask;
  return v <= rng ? 1 : 0;
}

__global__ void __launch_bounds__(256)
mt_count_kernel(const uint32_t* __restrict__ raw, const uint32_t* __restrict__ state, long long n_words, uint32_t mask, uint32_t rng,
                int* __restrict__ counts) {
  __shared__ int red[8];
  const long long pos0 = state[MT_N];
  const long long w0 = static_cast<long long>(blockIdx.x) * MT_CHUNK + threadIdx.x * 16;
  int c = 0;
  uint32_t v;
#pragma unroll
  for (int k = 0; k < 16; ++k) c += mt_accept(raw, w0 + k, n_words, pos0, mask, rng, v);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { int tsum = 0; for (int i = 0; i < 8; ++i) tsum += red[i]; counts[blockIdx.x] = tsum; }
}

// exclusive scan of the chunk counts (one CTA; 64-bit running offsets)
__global__ void __launch_bounds__(1024)
mt_scan_kernel(const int* __restrict__ counts, long long n_chunks, long long* __restrict__ offsets) {
  __shared__ long long wsum[32];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long c0 = 0; c0 < n_chunks; c0 += 1024) {
    const long long c = c0 + threadIdx.x;
    const long long v = c < n_chunks ? counts[c] : 0;
    long long incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const long long n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    long long before = carry;
    for (int k = 0; k < warp; ++k) before += wsum[k];
    if (c < n_chunks) offsets[c] = before + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n_chunks] = carry;     // total accepted in the generated words
}

__global__ void __launch_bounds__(256)
mt_scatter_kernel(const uint32_t* __restrict__ raw, const uint32_t* __restrict__ state, long long n_words, uint32_t mask, uint32_t rng,
                  const long long* __restrict__ offsets, long long count, int32_t* __restrict__ idx_out,
                  uint32_t* __restrict__ new_state) {
  __shared__ int wtot[8];
  __shared__ long long s_last;      // stream position of the last consumed word, if it lies in this chunk
  const long long pos0 = state[MT_N];
  const long long chunk_off = offsets[blockIdx.x];
  if (chunk_off >= count) return;                       // everything needed was produced by earlier chunks (uniform per CTA)
  if (threadIdx.x == 0) s_last = -1;
  const long long w0 = static_cast<long long>(blockIdx.x) * MT_CHUNK + threadIdx.x * 16;
  uint32_t v[16];
  int acc[16], c = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) { acc[k] = mt_accept(raw, w0 + k, n_words, pos0, mask, rng, v[k]); c += acc[k]; }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
  if (lane == 31) wtot[warp] = incl;
  __syncthreads();
  long long off = chunk_off + incl - c;
  for (int k = 0; k < warp; ++k) off += wtot[k];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    if (acc[k]) {
      if (off < count) idx_out[off] = static_cast<int32_t>(v[k]);
      if (off == count - 1) s_last = w0 + k;
      ++off;
    }
  }
  __syncthreads();
  const long long e = s_last;
  if (e >= 0) {          // this CTA saw the last draw: NumPy's state is now the block that contains word e.  It goes to a
    const long long blk = e / MT_N;   // staging copy: other CTAs are still reading the position out of `state`
    for (int i = threadIdx.x; i < MT_N; i += blockDim.x) new_state[i] = raw[blk * MT_N + i];
    if (threadIdx.x == 0) new_state[MT_N] = static_cast<uint32_t>(e - blk * MT_N + 1);
  }
}

__global__ void mt_commit_kernel(const uint32_t* __restrict__ new_state, uint32_t* __restrict__ state) {
  if (new_state[MT_N] == 0xFFFFFFFFu) {   // the generated words did not contain `count` accepted draws (a 40-sigma event)
    if (threadIdx.x == 0) printf("b2s mt19937_choice: word budget exhausted before the requested draws were produced\n");
    __trap();
  }
  for (int i = threadIdx.x; i <= MT_N; i += blockDim.x) state[i] = new_state[i];
}

// ---- Philox4x32-10 (counter-based stream for the non-NumPy mode) ----------------------------------
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

// One THREAD per resample b: its n indices are one contiguous row of the table, read with 16-byte loads, eight
// indices (two loads) in flight per step so that every 32-byte sector is consumed as soon as it arrives; vals sit in
// shared memory (<= 5 KB at the BASELINE sizes); fp64 accumulation in index order.  Without a table the indices come
// from the Philox4x32-10 counter (b, j/4).
__device__ __forceinline__ void boot_accum(const double* sv, int F, int split, int n, int id, double& lo, double& hi, double& ab) {
  for (int f = 0; f < F; ++f) {
    const double v = sv[f * n + id];
    if (f < split) lo += v; else hi += v;
    ab += fabs(v);
  }
}

__global__ void __launch_bounds__(256)
boot_resample_kernel(const double* __restrict__ vals, int F, int split, int n, long long B,
                     const int32_t* __restrict__ idx, unsigned long long seed, double* __restrict__ boot) {
  extern __shared__ double sv[];
  for (int i = threadIdx.x; i < F * n; i += blockDim.x) sv[i] = vals[i];
  __syncthreads();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  const bool vec = idx && (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(idx) & 15) == 0);
  for (long long b = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; b < B; b += stride) {
    double lo = 0.0, hi = 0.0, ab = 0.0;
    if (vec) {
      const int4* row = reinterpret_cast<const int4*>(idx + b * n);
      const int nv = n / 4;
      int j = 0;
      for (; j + 1 < nv; j += 2) {
        const int4 q0 = __ldg(row + j), q1 = __ldg(row + j + 1);
        boot_accum(sv, F, split, n, q0.x, lo, hi, ab); boot_accum(sv, F, split, n, q0.y, lo, hi, ab);
        boot_accum(sv, F, split, n, q0.z, lo, hi, ab); boot_accum(sv, F, split, n, q0.w, lo, hi, ab);
        boot_accum(sv, F, split, n, q1.x, lo, hi, ab); boot_accum(sv, F, split, n, q1.y, lo, hi, ab);
        boot_accum(sv, F, split, n, q1.z, lo, hi, ab); boot_accum(sv, F, split, n, q1.w, lo, hi, ab);
      }
      if (j < nv) {
        const int4 q0 = __ldg(row + j);
        boot_accum(sv, F, split, n, q0.x, lo, hi, ab); boot_accum(sv, F, split, n, q0.y, lo, hi, ab);
        boot_accum(sv, F, split, n, q0.z, lo, hi, ab); boot_accum(sv, F, split, n, q0.w, lo, hi, ab);
      }
    } else if (idx) {
      const int32_t* row = idx + b * n;
      for (int j = 0; j < n; ++j) boot_accum(sv, F, split, n, __ldg(row + j), lo, hi, ab);
    } else {
      for (int j0 = 0; j0 < n; j0 += 4) {
        uint32_t c[4] = {static_cast<uint32_t>(b), static_cast<uint32_t>(static_cast<unsigned long long>(b) >> 32),
                         static_cast<uint32_t>(j0 >> 2), 0u};
        philox4x32_10(c, k0, k1);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (j0 + u < n) boot_accum(sv, F, split, n, static_cast<int>(__umulhi(c[u], static_cast<uint32_t>(n))), lo, hi, ab);
      }
    }
    const double nlo = static_cast<double>(split) * n, nhi = static_cast<double>(F - split) * n;
    boot[b * 3 + 0] = split > 0 ? lo / nlo : nan("");
    boot[b * 3 + 1] = F > split ? hi / nhi : nan("");
    boot[b * 3 + 2] = ab / (static_cast<double>(F) * n);
  }
}

// Index-table fast path (n <= 256, n % 4 == 0; the BASELINE sizes).  The table is the only HBM stream (B * n * 4 bytes)
// and must be read at HBM speed, which the thread-per-row kernel above cannot do: a warp there touches 32 different rows
// per load.  Here a CTA stages the rows of 256 resamples with fully coalesced 16-byte loads, packs each index into one
// byte and stores the rows in shared memory with a stride of an odd number of words, so that afterwards thread t walks
// row t four indices per conflict-free 32-bit load.  The fp64 value gather is the other co-limiter (B * n * F shared
// loads): the [F, n] table is replicated 16 times, element e of copy c at word pair e * 16 + c, and thread t uses copy
// t mod 16 -- every lane of a half-warp then owns its bank pair and the gather never conflicts.  Each value row has its own
// running fp64 sum in index order; rows are combined at the end (differs from the general kernel's interleaved order in
// the last bits when several rows share a side).
constexpr int BT_ROWS = 256;
constexpr int BT_INFLIGHT = 7;     // 16-byte index loads a thread has in flight while staging (64 registers: four CTAs per SM)
template <int FT>    // FT = number of value rows when it is 1..4 or 6 (loops unroll, loads of four indices are in flight together); 0 = any F
__global__ void __launch_bounds__(BT_ROWS, 4)
boot_resample_tile_kernel(const double* __restrict__ vals, int F_rt, int split, int n, long long B, const int32_t* __restrict__ idx,
                          double* __restrict__ boot, int stride_words) {
  const int F = FT > 0 ? FT : F_rt;
  extern __shared__ double sv[];                                   // [F * n][16] replicated values
  uint32_t* sidx = reinterpret_cast<uint32_t*>(sv + static_cast<size_t>(F) * n * 16);   // [BT_ROWS][stride_words] packed indices
  for (int i = threadIdx.x; i < F * n * 16; i += BT_ROWS) sv[i] = vals[i >> 4];
  const int n4 = n >> 2;
  const long long tiles = (B + BT_ROWS - 1) / BT_ROWS;
  const double* tab = sv + (threadIdx.x & 15);                     // this thread's copy of the table
  const int r_first = static_cast<int>(threadIdx.x) / n4, c_first = static_cast<int>(threadIdx.x) - r_first * n4;   // int4 number tid of a tile
  const int dr = BT_ROWS / n4, dc = BT_ROWS - dr * n4;             // ... and what 256 int4 further on adds to (row, word)
  const int fstride = n * 16;                                      // doubles between value rows
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    __syncthreads();                                               // the previous tile's rows have been consumed (and sv is filled)
    const long long b0 = tile * BT_ROWS;
    const int R = static_cast<int>(B - b0 < BT_ROWS ? B - b0 : BT_ROWS);
    const int4* g = reinterpret_cast<const int4*>(idx + b0 * n);
    const int nq = R * n4;
    // The table streams at HBM speed only with enough bytes in flight: BT_INFLIGHT independent 16-byte loads per thread are
    // issued before the first one is consumed (28 KB per CTA, four CTAs per SM).  (row, word) of a staged int4 advance
    // incrementally -- an integer division per 16 bytes would cost more instructions than everything else in the tile.
    int r = r_first, c4 = c_first;
    for (int q0 = threadIdx.x; q0 < nq; q0 += BT_INFLIGHT * BT_ROWS) {
      int4 v[BT_INFLIGHT];
#pragma unroll
      for (int u = 0; u < BT_INFLIGHT; ++u) {
        const int q = q0 + u * BT_ROWS;
        if (q < nq)
          asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(g + q));
      }
#pragma unroll
      for (int u = 0; u < BT_INFLIGHT; ++u) {
        const int q = q0 + u * BT_ROWS;
        if (q < nq)
          sidx[r * stride_words + c4] = static_cast<uint32_t>(v[u].x) | (static_cast<uint32_t>(v[u].y) << 8) |
                                        (static_cast<uint32_t>(v[u].z) << 16) | (static_cast<uint32_t>(v[u].w) << 24);
        r += dr; c4 += dc;
        if (c4 >= n4) { c4 -= n4; ++r; }
      }
    }
    __syncthreads();
    if (threadIdx.x < R) {
      const uint32_t* row = sidx + threadIdx.x * stride_words;
      double lo = 0.0, hi = 0.0, ab0 = 0.0, ab1 = 0.0;
      if constexpr (FT > 0) {
        // one running sum per value row (no select on `split` inside the loop), four indices' loads in flight together
        double sum[FT];
#pragma unroll
        for (int f = 0; f < FT; ++f) sum[f] = 0.0;
        const uint32_t tab_addr = static_cast<uint32_t>(__cvta_generic_to_shared(tab));
        const uint32_t fbytes = static_cast<uint32_t>(fstride) * 8u;
        for (int j4 = 0; j4 < n4; ++j4) {
          const uint32_t w = row[j4];
          double x[4][FT];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            // byte u of w is the index; its table row starts 128 bytes * index after this thread's copy
            const uint32_t a = tab_addr + ((u == 0 ? (w << 7) : (u == 1 ? (w >> 1) : (u == 2 ? (w >> 9) : (w >> 17)))) & (255u << 7));
#pragma unroll
            for (int f = 0; f < FT; ++f)
              asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x[u][f]) : "r"(a + static_cast<uint32_t>(f) * fbytes));
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int f = 0; f < FT; ++f) {
              sum[f] += x[u][f];
              if (u & 1) ab1 += fabs(x[u][f]); else ab0 += fabs(x[u][f]);
            }
          }
        }
#pragma unroll
        for (int f = 0; f < FT; ++f) { if (f < split) lo += sum[f]; else hi += sum[f]; }
      } else {
        for (int j4 = 0; j4 < n4; ++j4) {
          const uint32_t w = row[j4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const double* p = tab + ((w >> (8 * u)) & 255u) * 16;
            for (int f = 0; f < F; ++f) {
              const double x = p[f * fstride];
              if (f < split) lo += x; else hi += x;
              ab0 += fabs(x);
            }
          }
        }
      }
      const long long b = b0 + threadIdx.x;
      const double nlo = static_cast<double>(split) * n, nhi = static_cast<double>(F - split) * n;
      boot[b * 3 + 0] = split > 0 ? lo / nlo : nan("");
      boot[b * 3 + 1] = F > split ? hi / nhi : nan("");
      boot[b * 3 + 2] = (ab0 + ab1) / (static_cast<double>(F) * n);
    }
  }
}

// The index table the Philox mode of the resample kernels draws, written out: idx[b, j] = floor(n * u / 2^32) with u the
// (j mod 4)-th word of Philox4x32-10(counter = (b, j / 4), key = seed).  For consumers that need the indices themselves
// (the correlation bootstraps) of the reference loops that draw from an unseeded generator.
__global__ void __launch_bounds__(256)
philox_indices_kernel(unsigned long long seed, int n, long long B, int32_t* __restrict__ idx) {
  const int n4 = (n + 3) / 4;
  const long long total = B * n4;
  const uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i / n4;
    const int q = static_cast<int>(i - b * n4);
    uint32_t c[4] = {static_cast<uint32_t>(b), static_cast<uint32_t>(static_cast<unsigned long long>(b) >> 32), static_cast<uint32_t>(q), 0u};
    philox4x32_10(c, k0, k1);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (4 * q + u < n) idx[b * n + 4 * q + u] = static_cast<int32_t>(__umulhi(c[u], static_cast<uint32_t>(n)));
  }
}

// Question-resampled model/human agreement metrics (survey_analysis/analyze_llm_agreement_simple_bootstrap.py:89-148): a
// resample is the SET of distinct sampled questions (the script matches model rows with `prompt in sampled_questions`), a
// question is used when the model has a value for it; MAE, MSE, MAPE over the finite |h - m| / |h| terms (in percent) and
// scipy's pearsonr of (human, model) when both have spread, else NaN.  Fewer than min_matched rows: all NaN (the script
// returns None).  One thread per resample; n <= 1024 questions.
__global__ void __launch_bounds__(128)
agreement_boot_kernel(const double* __restrict__ human, const double* __restrict__ model, int n, const int32_t* __restrict__ idx,
                      long long B, int min_matched, double* __restrict__ out) {
  const long long b = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (b >= B) return;
  uint32_t mask[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) mask[i] = 0u;
  const int32_t* row = idx + b * n;
  for (int j = 0; j < n; ++j) { const int q = row[j]; mask[q >> 5] |= 1u << (q & 31); }
  int cnt = 0, n_ape = 0;
  double sh = 0.0, sm = 0.0, sae = 0.0, sse = 0.0, sape = 0.0;
  for (int q = 0; q < n; ++q) {
    const double m = model[q];
    if (((mask[q >> 5] >> (q & 31)) & 1u) && m == m) {
      const double h = human[q], d = h - m;
      sh += h; sm += m; sae += fabs(d); sse += d * d; ++cnt;
      const double ape = fabs(d / h);
      if (isfinite(ape)) { sape += ape; ++n_ape; }
    }
  }
  double* o = out + b * 4;
  if (cnt < min_matched) { o[0] = o[1] = o[2] = o[3] = nan(""); return; }
  const double mh = sh / cnt, mm = sm / cnt;
  double vxx = 0.0, vyy = 0.0, vxy = 0.0;
  for (int q = 0; q < n; ++q) {
    const double m = model[q];
    if (((mask[q >> 5] >> (q & 31)) & 1u) && m == m) {
      const double dx = human[q] - mh, dy = m - mm;
      vxx += dx * dx; vyy += dy * dy; vxy += dx * dy;
    }
  }
  o[0] = sae / cnt;
  o[1] = sse / cnt;
  o[2] = n_ape > 0 ? sape / n_ape * 100.0 : nan("");
  o[3] = (cnt > 1 && vxx > 0.0 && vyy > 0.0) ? vxy / sqrt(vxx * vyy) : nan("");
}

// ---- summary: select statistic, sort, percentiles, tail fractions -----------------------------------
__device__ __forceinline__ double boot_stat(const double* boot, long long b, int stat) {
  if (stat == 0) return boot[b * 3];
  if (stat == 1) return boot[b * 3 + 1] - boot[b * 3];
  return boot[b * 3 + 2];
}

__global__ void boot_select_kernel(const double* __restrict__ boot, long long B, long long P, int stat,
                                   double* __restrict__ keys, double* __restrict__ stat_out) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < P;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    double v = INFINITY;
    if (i < B) { v = boot_stat(boot, i, stat); if (stat_out) stat_out[i] = v; }
    keys[i] = v;
  }
}

// global-memory bitonic compare-exchange step (used for strides that do not fit one CTA's shared memory)
__global__ void bitonic_global_step(double* __restrict__ a, long long P, long long j, long long k) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < P;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long l = i ^ j;
    if (l > i) {
      const double x = a[i], y = a[l];
      const bool up = (i & k) == 0;
      if ((x > y) == up) { a[i] = y; a[l] = x; }
    }
  }
}

constexpr int SORT_CHUNK = 4096;   // doubles per CTA in shared memory (32 KB)
// sorts each SORT_CHUNK block fully (k <= SORT_CHUNK), or finishes the strides j < SORT_CHUNK of a larger k
__global__ void __launch_bounds__(1024)
bitonic_shared_kernel(double* __restrict__ a, long long k_start, long long k_end) {
  __shared__ double s[SORT_CHUNK];
  const long long base = static_cast<long long>(blockIdx.x) * SORT_CHUNK;
  for (int i = threadIdx.x; i < SORT_CHUNK; i += blockDim.x) s[i] = a[base + i];
  __syncthreads();
  for (long long k = k_start; k <= k_end; k <<= 1) {
    for (long long j = (k >> 1) < SORT_CHUNK ? (k >> 1) : (SORT_CHUNK >> 1); j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < SORT_CHUNK; i += blockDim.x) {
        const int l = i ^ static_cast<int>(j);
        if (l > i) {
          const double x = s[i], y = s[l];
          const bool up = ((base + i) & k) == 0;
          if ((x > y) == up) { s[i] = y; s[l] = x; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < SORT_CHUNK; i += blockDim.x) a[base + i] = s[i];
}

__device__ __forceinline__ double np_percentile_sorted(const double* x, long long B, double q) {
  // numpy "linear": virtual index (B-1)*q/100, _lerp(a, b, t)
  const double h = (static_cast<double>(B) - 1.0) * (q / 100.0);
  long long lo = static_cast<long long>(floor(h));
  if (lo < 0) lo = 0;
  if (lo > B - 1) lo = B - 1;
  const long long hi = (lo + 1 < B) ? lo + 1 : B - 1;
  const double t = h - static_cast<double>(lo);
  const double a = x[lo], b = x[hi];
  const double diff = b - a;
  double r = a + diff * t;
  if (t >= 0.5) r = b - diff * (1.0 - t);
  return r;
}

__global__ void __launch_bounds__(1024)
boot_summary_kernel(const double* __restrict__ sorted, long long B, double q_lo, double q_hi, double observed,
                    double* __restrict__ summary) {
  __shared__ double red[4][32];
  double sum = 0.0, le0 = 0.0, ge0 = 0.0, geo = 0.0;
  for (long long i = threadIdx.x; i < B; i += blockDim.x) {
    const double v = sorted[i];
    sum += v; le0 += (v <= 0.0); ge0 += (v >= 0.0); geo += (v >= observed);
  }
  double vals[4] = {sum, le0, ge0, geo};
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vals[q] += __shfl_xor_sync(0xffffffffu, vals[q], o);
    if (lane == 0) red[q][warp] = vals[q];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot[4] = {0, 0, 0, 0};
    for (int q = 0; q < 4; ++q)
      for (int w = 0; w < static_cast<int>(blockDim.x >> 5); ++w) tot[q] += red[q][w];
    summary[0] = np_percentile_sorted(sorted, B, q_lo);
    summary[1] = np_percentile_sorted(sorted, B, q_hi);
    summary[2] = tot[1] / static_cast<double>(B);
    summary[3] = tot[2] / static_cast<double>(B);
    summary[4] = tot[0] / static_cast<double>(B);
    summary[5] = tot[3] / static_cast<double>(B);
    summary[6] = sorted[0];
    summary[7] = sorted[B - 1];
  }
}

// Scratch for b2s_mt19937_choice comes from a memory pool the library owns (one per device, created on first use):
// stream-ordered allocation without touching the application's default pool, and with a release threshold so that
// the bootstrap entry points, which call it back to back, reuse the same block instead of going to the driver each time.
static cudaMemPool_t g_boot_pool[64] = {nullptr};
static std::mutex g_boot_pool_mutex;
static int boot_pool(cudaMemPool_t* out) {
  int dev = 0;
  B2S_CHECK_CUDA(cudaGetDevice(&dev));
  B2S_REQUIRE(dev >= 0 && dev < 64, "mt19937_choice: device ordinal %d outside [0, 64)", dev);
  std::lock_guard<std::mutex> g(g_boot_pool_mutex);
  if (g_boot_pool[dev] == nullptr) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    B2S_CHECK_CUDA(cudaMemPoolCreate(&g_boot_pool[dev], &props));
    unsigned long long keep = 1ull << 30;     // blocks up to 1 GiB stay in the pool between calls
    B2S_CHECK_CUDA(cudaMemPoolSetAttribute(g_boot_pool[dev], cudaMemPoolAttrReleaseThreshold, &keep));
  }
  *out = g_boot_pool[dev];
  return B2S_OK;
}

static long long next_pow2(long long v) {
  long long p = SORT_CHUNK;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace b2s

using namespace b2s;

extern "C" {

int b2s_mt19937_seed(uint32_t seed, uint32_t* state, void* stream) {
  B2S_REQUIRE(state, "mt19937_seed: null state");
  mt_seed_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(seed, state);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int b2s_mt19937_choice(uint32_t* state, int32_t n, int64_t count, int32_t* idx_out, void* stream) {
  B2S_REQUIRE(state && idx_out, "mt19937_choice: null argument");
  B2S_REQUIRE(n >= 1, "mt19937_choice: n=%d must be >= 1", n);
  B2S_REQUIRE(count >= 0, "mt19937_choice: negative count");
  if (count == 0) return B2S_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n == 1) {   // numpy: rng == 0 draws nothing from the stream and returns 0
    B2S_CHECK_CUDA(cudaMemsetAsync(idx_out, 0, 4 * static_cast<size_t>(count), st));
    return B2S_OK;
  }
  const uint32_t rng = static_cast<uint32_t>(n - 1);
  uint32_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  if (count <= 16384) {   // one launch; the device-wide path below has a fixed cost of six launches and an allocation
    mt_choice_kernel<<<1, 640, 0, st>>>(state, rng, mask, static_cast<long long>(count), idx_out);
    B2S_CHECK_LAUNCH();
    return B2S_OK;
  }
  // Word budget of a segment of `c` draws: c / p expected (p = n / (mask + 1) >= 1/2), plus 40 standard deviations of the
  // negative-binomial spread, rounded up to whole generator steps.  mt_commit_kernel traps if that were ever not enough.
  const double pacc = static_cast<double>(n) / (static_cast<double>(mask) + 1.0);
  auto words_for = [&](long long c) {
    const double w = c / pacc + 40.0 * sqrt(static_cast<double>(c) * (1.0 - pacc)) / pacc + 1024.0;
    return static_cast<long long>(ceil_div(static_cast<int64_t>(w) + 1, MT_GEN_STEP)) * MT_GEN_STEP;
  };
  const long long SEG = 1LL << 26;                                   // draws per segment (bounds the scratch to ~0.6 GB)
  const long long seg_max = count < SEG ? count : SEG;
  const long long words_max = MT_N + words_for(seg_max);
  const long long chunks_max = ceil_div(words_max, MT_CHUNK);
  const size_t raw_bytes = align_up(4 * static_cast<size_t>(words_max), 256);
  const size_t cnt_bytes = align_up(4 * static_cast<size_t>(chunks_max), 256);
  const size_t off_bytes = align_up(8 * static_cast<size_t>(chunks_max + 1), 256);
  uint8_t* scratch = nullptr;
  cudaMemPool_t pool;
  B2S_TRY(boot_pool(&pool));
  B2S_CHECK_CUDA(cudaMallocFromPoolAsync(reinterpret_cast<void**>(&scratch), raw_bytes + cnt_bytes + off_bytes + 4096, pool, st));
  uint32_t* raw = reinterpret_cast<uint32_t*>(scratch);
  int* counts = reinterpret_cast<int*>(scratch + raw_bytes);
  long long* offsets = reinterpret_cast<long long*>(scratch + raw_bytes + cnt_bytes);
  uint32_t* new_state = reinterpret_cast<uint32_t*>(scratch + raw_bytes + cnt_bytes + off_bytes);
  int rc = B2S_OK;
  for (long long done = 0; done < count && rc == B2S_OK; done += SEG) {
    const long long c = count - done < SEG ? count - done : SEG;
    const long long n_new = words_for(c), n_words = MT_N + n_new;
    const unsigned n_chunks = static_cast<unsigned>(ceil_div(n_words, MT_CHUNK));
    cudaMemsetAsync(new_state + MT_N, 0xFF, 4, st);
    mt_gen_kernel<<<1, 256, 0, st>>>(state, raw, n_new);
    mt_count_kernel<<<n_chunks, 256, 0, st>>>(raw, state, n_words, mask, rng, counts);
    mt_scan_kernel<<<1, 1024, 0, st>>>(counts, n_chunks, offsets);
    mt_scatter_kernel<<<n_chunks, 256, 0, st>>>(raw, state, n_words, mask, rng, offsets, c, idx_out + done, new_state);
    mt_commit_kernel<<<1, 640, 0, st>>>(new_state, state);
    g_launch_count.fetch_add(4, std::memory_order_relaxed);
    if (cudaGetLastError() != cudaSuccess) { set_error("mt19937_choice: kernel launch failed"); rc = B2S_ECUDA; }
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
  }
  cudaFreeAsync(scratch, st);
  return rc;
}

int b2s_philox_indices(uint64_t seed, int32_t n, int64_t B, int32_t* idx_out, void* stream) {
  B2S_REQUIRE(idx_out, "philox_indices: null argument");
  B2S_REQUIRE(n >= 1 && B >= 1, "philox_indices: bad sizes n=%d B=%lld", n, static_cast<long long>(B));
  const long long total = static_cast<long long>(B) * ((n + 3) / 4);
  const long long blocks = ceil_div(total, 256);
  philox_indices_kernel<<<static_cast<unsigned>(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      seed, n, B, idx_out);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int b2s_agreement_bootstrap(const double* human, const double* model, int32_t n, const int32_t* idx, int64_t B,
                            int32_t min_matched, double* out, void* stream) {
  B2S_REQUIRE(human && model && idx && out, "agreement_bootstrap: null argument");
  B2S_REQUIRE(n >= 1 && n <= 1024 && B >= 1, "agreement_bootstrap: n=%d outside [1, 1024] or B=%lld < 1", n, static_cast<long long>(B));
  agreement_boot_kernel<<<static_cast<unsigned>(ceil_div(B, 128)), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      human, model, n, idx, B, min_matched, out);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

int b2s_bootstrap_resample(const double* vals, int32_t F, int32_t split, int32_t n, int64_t B, const int32_t* idx,
                           uint64_t philox_seed, double* boot, void* stream) {
  B2S_REQUIRE(vals && boot, "bootstrap_resample: null argument");
  B2S_REQUIRE(F >= 1 && n >= 1 && B >= 1 && split >= 0 && split <= F, "bootstrap_resample: bad sizes F=%d split=%d n=%d B=%lld", F,
              split, n, static_cast<long long>(B));
  const size_t smem = static_cast<size_t>(F) * n * sizeof(double);
  B2S_REQUIRE(smem <= 200 * 1024, "bootstrap_resample: F*n=%d values exceed the shared-memory table (25600)", F * n);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  {
    int dev = 0, sms = 0;
    B2S_CHECK_CUDA(cudaGetDevice(&dev));
    B2S_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int sw = (n + 3) / 4;
    if (sw % 2 == 0) ++sw;                                              // odd word stride: conflict-free row walks
    const size_t tile_smem = static_cast<size_t>(F) * n * 16 * sizeof(double) + static_cast<size_t>(BT_ROWS) * sw * 4;
    if (idx != nullptr && n % 4 == 0 && n <= 256 && (reinterpret_cast<uintptr_t>(idx) & 15) == 0 && tile_smem <= 110 * 1024) {
      const long long tiles = ceil_div(B, BT_ROWS);
      const int per_sm = static_cast<int>((220 * 1024) / (tile_smem + 1024));
      const long long cap = static_cast<long long>(sms) * (per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm));   // registers allow four CTAs per SM
      const unsigned grid = static_cast<unsigned>(tiles < cap ? tiles : cap);
#define B2S_BOOT_TILE(FT)                                                                                                     \
  do {                                                                                                                        \
    B2S_CHECK_CUDA(cudaFuncSetAttribute(boot_resample_tile_kernel<FT>, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                        static_cast<int>(tile_smem)));                                                        \
    B2S_CHECK_CUDA(cudaFuncSetAttribute(boot_resample_tile_kernel<FT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100)); \
    boot_resample_tile_kernel<FT><<<grid, BT_ROWS, tile_smem, st>>>(vals, F, split, n, B, idx, boot, sw);                     \
  } while (0)
      switch (F) {
        case 1: B2S_BOOT_TILE(1); break;
        case 2: B2S_BOOT_TILE(2); break;
        case 3: B2S_BOOT_TILE(3); break;
        case 4: B2S_BOOT_TILE(4); break;
        case 6: B2S_BOOT_TILE(6); break;
        default: B2S_BOOT_TILE(0); break;
      }
#undef B2S_BOOT_TILE
      B2S_CHECK_LAUNCH();
      return B2S_OK;
    }
  }
  if (smem > 48 * 1024)
    B2S_CHECK_CUDA(cudaFuncSetAttribute(boot_resample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  int dev = 0, sms = 0;
  B2S_CHECK_CUDA(cudaGetDevice(&dev));
  B2S_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  long long blocks = ceil_div(B, 256);
  const long long cap = static_cast<long long>(sms) * 8;   // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  boot_resample_kernel<<<static_cast<unsigned>(blocks), 256, smem, st>>>(vals, F, split, n, B, idx, philox_seed, boot);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

size_t b2s_bootstrap_scratch_bytes(int64_t B) {
  if (B <= 0) return 0;
  return static_cast<size_t>(next_pow2(B)) * sizeof(double);
}

int b2s_bootstrap_summary(const double* boot, int64_t B, int32_t stat, double q_lo, double q_hi, double observed,
                          double* summary, double* stat_out, void* scratch, size_t scratch_bytes, void* stream) {
  B2S_REQUIRE(boot && summary && scratch, "bootstrap_summary: null argument");
  B2S_REQUIRE(B >= 1 && stat >= 0 && stat <= 2, "bootstrap_summary: bad B=%lld or stat=%d", static_cast<long long>(B), stat);
  B2S_REQUIRE(q_lo >= 0 && q_lo <= 100 && q_hi >= 0 && q_hi <= 100, "bootstrap_summary: percentiles outside [0,100]");
  const long long P = next_pow2(B);
  B2S_REQUIRE(scratch_bytes >= static_cast<size_t>(P) * sizeof(double), "bootstrap_summary: scratch of %zu bytes is too small, %zu needed",
              scratch_bytes, static_cast<size_t>(P) * sizeof(double));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  double* keys = reinterpret_cast<double*>(scratch);
  const unsigned gs = static_cast<unsigned>(ceil_div(P, 256) < 1184 ? ceil_div(P, 256) : 1184);
  boot_select_kernel<<<gs, 256, 0, st>>>(boot, B, P, stat, keys, stat_out);
  B2S_CHECK_LAUNCH();
  const unsigned chunks = static_cast<unsigned>(P / SORT_CHUNK);
  bitonic_shared_kernel<<<chunks, 1024, 0, st>>>(keys, 2, SORT_CHUNK);
  B2S_CHECK_LAUNCH();
  for (long long k = 2LL * SORT_CHUNK; k <= P; k <<= 1) {
    for (long long j = k >> 1; j >= SORT_CHUNK; j >>= 1) {
      bitonic_global_step<<<gs, 256, 0, st>>>(keys, P, j, k);
      B2S_CHECK_LAUNCH();
    }
    bitonic_shared_kernel<<<chunks, 1024, 0, st>>>(keys, k, k);
    B2S_CHECK_LAUNCH();
  }
  boot_summary_kernel<<<1, 1024, 0, st>>>(keys, B, q_lo, q_hi, observed, summary);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // extern "C"
