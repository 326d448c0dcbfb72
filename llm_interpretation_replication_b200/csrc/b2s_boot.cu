// Bootstrap kernels: NumPy-legacy MT19937 index stream (bit-exact), resample statistics, percentile CI.
//
// Replaces the `for _ in range(N_BOOTSTRAP)` loops of the reference:
//   analysis/run_base_vs_instruct_100q.py:655-658,692-695
//   survey_analysis/analyze_base_vs_instruct_mae_100q.py:310-315,380-388
//   analysis/analyze_results_base_versus_instruct.py:310-320,347-351
// NumPy legacy stream: np.random.seed(s) = Knuth init of 624 words;
// np.random.choice(n, size) = per draw, words masked to the smallest 2^b-1 >= n-1 and rejected while > n-1.
#include "b2s_common.cuh"

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

// One CTA walks the stream 624 words at a time.  The twist is done in the four dependency phases of the
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
  mt_choice_kernel<<<1, 640, 0, st>>>(state, rng, mask, static_cast<long long>(count), idx_out);
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
