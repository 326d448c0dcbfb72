// Prompt-resampled inter-model correlation bootstrap.
//
// Replaces the body of the loop of reference analysis/model_comparison_graph.py:261-291
//     for i in range(n_bootstrap):
//         bootstrap_pivot = pivot_df.loc[unique_prompts[np.random.choice(n_prompts, n_prompts)]]
//         bootstrap_pivot.corr(method='pearson' | 'spearman')  -> values of the model pairs -> mean / median / std
// i.e. n_bootstrap pandas DataFrame.corr calls (pairwise-complete observations, min_periods = 1) on a
// [prompts, models] table.  The index stream comes from b2s_mt19937_choice (bit-exact NumPy legacy stream).
//   corr_pearson_kernel   thread per (resample, pair): two passes over the resampled rows both models answered
//                         (pandas libalgos.nancorr is a Welford pass; same value to ~1e-15)
//   corr_spearman_kernel  CTA per (resample, pair): compact the rows both models answered, average-tie ranks of
//                         either column (bitonic sort of (value, slot) in shared memory, pandas rank_1d 'average'),
//                         Pearson of the ranks (pandas libalgos.nancorr_spearman)
//   corr_stats_kernel     CTA per resample: mean / median / population std of the non-NaN pair correlations
//                         (np.mean / np.median / np.std of `bootstrap_correlations`)
#include "b2s_common.cuh"
#include <math_constants.h>

namespace b2s {

constexpr int CORR_MAX_P = 2048;      // Spearman: rows per resample that fit the shared-memory sort
constexpr int CORR_MAX_PAIRS = 2048;  // stats kernel: pairs per resample (the reference samples at most 1000)

// min_valid > 0: a column takes part in a resample only if at least min_valid of its resampled rows are not NaN (the
// reference's cross-prompt bootstrap drops respondents who answered fewer than 5 of the sampled questions,
// survey_analysis/survey_analysis_consolidated.py:520-523); its pairs are NaN for that resample.
__global__ void __launch_bounds__(128)
corr_pearson_kernel(const double* __restrict__ X, int P, int M, const int32_t* __restrict__ idx, int B,
                    const int32_t* __restrict__ pairs, int n_pairs, int min_valid, double* __restrict__ corr) {
  const int b = blockIdx.y;
  const int pr = blockIdx.x * blockDim.x + threadIdx.x;
  if (pr >= n_pairs) return;
  const int ci = pairs[2 * pr], cj = pairs[2 * pr + 1];
  const int32_t* ib = idx + static_cast<size_t>(b) * P;
  // Values are shifted by the first complete row so that a constant column gives EXACTLY zero variance (pandas'
  // Welford pass does: dx = vx - meanx is 0 at every step), whatever its value; NaN then, as in pandas.
  double x0 = 0.0, y0 = 0.0, sx = 0.0, sy = 0.0;
  int n = 0, nx = 0, ny = 0;
  for (int r = 0; r < P; ++r) {
    const size_t row = static_cast<size_t>(ib[r]) * M;
    const double x = X[row + ci], y = X[row + cj];
    nx += (x == x); ny += (y == y);
    if (x == x && y == y) {
      if (n == 0) { x0 = x; y0 = y; }
      sx += x - x0; sy += y - y0; ++n;
    }
  }
  if (nx < min_valid || ny < min_valid) n = 0;
  double out = CUDART_NAN;
  if (n >= 1) {
    const double mx = sx / n, my = sy / n;
    double sxx = 0.0, syy = 0.0, sxy = 0.0;
    for (int r = 0; r < P; ++r) {
      const size_t row = static_cast<size_t>(ib[r]) * M;
      const double x = X[row + ci], y = X[row + cj];
      if (x == x && y == y) {
        const double dx = (x - x0) - mx, dy = (y - y0) - my;
        sxx += dx * dx; syy += dy * dy; sxy += dx * dy;
      }
    }
    const double div = sqrt(sxx * syy);
    if (div != 0.0) out = sxy / div;
  }
  corr[static_cast<size_t>(b) * n_pairs + pr] = out;
}

// ascending bitonic sort of (key, slot) pairs; n2 = power of two >= count, padded with +inf keys
__device__ void bitonic_kv(double* key, int* slot, int n2) {
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < n2; t += blockDim.x) {
        const int p = t ^ j;
        if (p > t) {
          const bool up = (t & k) == 0;
          const double a = key[t], c = key[p];
          if ((a > c) == up) {
            key[t] = c; key[p] = a;
            const int s = slot[t]; slot[t] = slot[p]; slot[p] = s;
          }
        }
      }
      __syncthreads();
    }
  }
}

// average-tie ranks (1-based) of vals[0..n) into rank[0..n); key/slot are scratch of n2 entries
__device__ void rank_average(const double* vals, int n, int n2, double* key, int* slot, double* rank) {
  for (int t = threadIdx.x; t < n2; t += blockDim.x) {
    key[t] = t < n ? vals[t] : CUDART_INF;
    slot[t] = t;
  }
  __syncthreads();
  bitonic_kv(key, slot, n2);
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    if (t == 0 || key[t] != key[t - 1]) {           // first of a tie group: [t, e)
      int e = t + 1;
      while (e < n && key[e] == key[t]) ++e;
      const double rk = 0.5 * static_cast<double>(t + e - 1) + 1.0;
      for (int q = t; q < e; ++q) rank[slot[q]] = rk;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
corr_spearman_kernel(const double* __restrict__ X, int P, int M, const int32_t* __restrict__ idx, int B,
                     const int32_t* __restrict__ pairs, int n_pairs, int n2, double* __restrict__ corr) {
  extern __shared__ double sm[];
  double* xs = sm;                 // [n2]
  double* ys = xs + n2;            // [n2]
  double* key = ys + n2;           // [n2]
  double* rx = key + n2;           // [n2]
  double* ry = rx + n2;            // [n2]
  int* slot = reinterpret_cast<int*>(ry + n2);   // [n2]
  __shared__ int s_n;
  __shared__ double s_red[3][8];
  const int b = blockIdx.y, pr = blockIdx.x;
  const int ci = pairs[2 * pr], cj = pairs[2 * pr + 1];
  const int32_t* ib = idx + static_cast<size_t>(b) * P;
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  for (int r = threadIdx.x; r < P; r += blockDim.x) {
    const size_t row = static_cast<size_t>(ib[r]) * M;
    const double x = X[row + ci], y = X[row + cj];
    if (x == x && y == y) {
      const int s = atomicAdd(&s_n, 1);   // slot order is arbitrary; ranks and their exact sums do not depend on it
      xs[s] = x; ys[s] = y;
    }
  }
  __syncthreads();
  const int n = s_n;
  double out = CUDART_NAN;
  if (n >= 1) {
    int m2 = 1;
    while (m2 < n) m2 <<= 1;
    rank_average(xs, n, m2, key, slot, rx);
    rank_average(ys, n, m2, key, slot, ry);
    const double mean = 0.5 * (n + 1);
    double sxx = 0.0, syy = 0.0, sxy = 0.0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
      const double dx = rx[t] - mean, dy = ry[t] - mean;
      sxx += dx * dx; syy += dy * dy; sxy += dx * dy;    // multiples of 1/4 below 2^53: exact in any order
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sxx += __shfl_xor_sync(0xffffffffu, sxx, o);
      syy += __shfl_xor_sync(0xffffffffu, syy, o);
      sxy += __shfl_xor_sync(0xffffffffu, sxy, o);
    }
    if (l == 0) { s_red[0][w] = sxx; s_red[1][w] = syy; s_red[2][w] = sxy; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double a = 0.0, c = 0.0, d = 0.0;
      for (int q = 0; q < 8; ++q) { a += s_red[0][q]; c += s_red[1][q]; d += s_red[2][q]; }
      const double div = sqrt(a * c);
      if (div != 0.0) out = d / div;
    }
  }
  if (threadIdx.x == 0) corr[static_cast<size_t>(b) * n_pairs + pr] = out;
}

// per resample: [mean, median, std (ddof 0), count] of the non-NaN pair correlations
__global__ void __launch_bounds__(256)
corr_stats_kernel(const double* __restrict__ corr, int n_pairs, int n2, double* __restrict__ stats) {
  extern __shared__ double sm[];
  double* key = sm;                                  // [n2]
  int* slot = reinterpret_cast<int*>(key + n2);      // [n2] (unused payload of the shared sort)
  __shared__ int s_n;
  __shared__ double s_red[8];
  const int b = blockIdx.x;
  const double* c = corr + static_cast<size_t>(b) * n_pairs;
  if (threadIdx.x == 0) s_n = 0;
  for (int t = threadIdx.x; t < n2; t += blockDim.x) { key[t] = CUDART_INF; slot[t] = t; }
  __syncthreads();
  for (int t = threadIdx.x; t < n_pairs; t += blockDim.x) {
    const double v = c[t];
    if (v == v) key[atomicAdd(&s_n, 1)] = v;
  }
  __syncthreads();
  const int n = s_n;
  double* o = stats + static_cast<size_t>(b) * 4;
  if (n == 0) {
    if (threadIdx.x == 0) { o[0] = o[1] = o[2] = CUDART_NAN; o[3] = 0.0; }
    return;
  }
  bitonic_kv(key, slot, n2);                         // sorted: sums below run in a fixed order -> deterministic
  auto block_sum = [&](double v) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
    for (int q = 16; q > 0; q >>= 1) v += __shfl_xor_sync(0xffffffffu, v, q);
    __syncthreads();
    if (l == 0) s_red[w] = v;
    __syncthreads();
    double t = 0.0;
    for (int q = 0; q < 8; ++q) t += s_red[q];
    return t;
  };
  double s = 0.0;
  for (int t = threadIdx.x; t < n; t += blockDim.x) s += key[t];
  const double mean = block_sum(s) / n;
  double q = 0.0;
  for (int t = threadIdx.x; t < n; t += blockDim.x) { const double d = key[t] - mean; q += d * d; }
  const double var = block_sum(q) / n;
  if (threadIdx.x == 0) {
    o[0] = mean;
    o[1] = (n & 1) ? key[n >> 1] : 0.5 * (key[(n >> 1) - 1] + key[n >> 1]);
    o[2] = sqrt(var);
    o[3] = static_cast<double>(n);
  }
}

// more pairs than the shared-memory sort holds: mean, population std and count only (median = NaN); two passes, sums in
// a fixed order per thread and a fixed tree across threads -> deterministic
__global__ void __launch_bounds__(256)
corr_stats_big_kernel(const double* __restrict__ corr, int n_pairs, double* __restrict__ stats) {
  __shared__ double s_red[8];
  __shared__ double s_cnt[8];
  const int b = blockIdx.x;
  const double* c = corr + static_cast<size_t>(b) * n_pairs;
  auto block_sum = [&](double v, double* red) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
    for (int q = 16; q > 0; q >>= 1) v += __shfl_xor_sync(0xffffffffu, v, q);
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    double t = 0.0;
    for (int q = 0; q < 8; ++q) t += red[q];
    return t;
  };
  double sum = 0.0, cnt = 0.0;
  for (int t = threadIdx.x; t < n_pairs; t += blockDim.x) { const double v = c[t]; if (v == v) { sum += v; cnt += 1.0; } }
  const double n = block_sum(cnt, s_cnt);
  const double tot = block_sum(sum, s_red);
  double* o = stats + static_cast<size_t>(b) * 4;
  if (n == 0.0) { if (threadIdx.x == 0) { o[0] = o[1] = o[2] = CUDART_NAN; o[3] = 0.0; } return; }
  const double mean = tot / n;
  double q = 0.0;
  for (int t = threadIdx.x; t < n_pairs; t += blockDim.x) { const double v = c[t]; if (v == v) q += (v - mean) * (v - mean); }
  const double var = block_sum(q, s_red) / n;
  if (threadIdx.x == 0) { o[0] = mean; o[1] = CUDART_NAN; o[2] = sqrt(var); o[3] = n; }
}

}  // namespace b2s

extern "C" int b2s_corr_bootstrap_ex(const double* X, int32_t P, int32_t M, const int32_t* idx, int32_t B, const int32_t* pairs,
                                     int32_t n_pairs, int32_t method, int32_t min_valid, double* corr, double* stats, void* stream);

extern "C" int b2s_corr_bootstrap(const double* X, int32_t P, int32_t M, const int32_t* idx, int32_t B, const int32_t* pairs,
                                  int32_t n_pairs, int32_t method, double* corr, double* stats, void* stream) {
  return b2s_corr_bootstrap_ex(X, P, M, idx, B, pairs, n_pairs, method, 0, corr, stats, stream);
}

extern "C" int b2s_corr_bootstrap_ex(const double* X, int32_t P, int32_t M, const int32_t* idx, int32_t B, const int32_t* pairs,
                                     int32_t n_pairs, int32_t method, int32_t min_valid, double* corr, double* stats, void* stream) {
  using namespace b2s;
  B2S_REQUIRE(X && idx && pairs && corr, "corr_bootstrap: null argument");
  B2S_REQUIRE(P > 0 && M > 1 && B > 0 && n_pairs > 0, "corr_bootstrap: empty problem (P=%d M=%d B=%d pairs=%d)", P, M, B, n_pairs);
  B2S_REQUIRE(B <= 65535, "corr_bootstrap: at most 65535 resamples per call");
  B2S_REQUIRE(method == 0 || method == 1, "corr_bootstrap: method must be 0 (pearson) or 1 (spearman)");
  B2S_REQUIRE(min_valid >= 0 && (min_valid == 0 || method == 0), "corr_bootstrap: min_valid is supported for pearson only");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (method == 0) {
    dim3 grid(static_cast<unsigned>(ceil_div(n_pairs, 128)), static_cast<unsigned>(B));
    corr_pearson_kernel<<<grid, 128, 0, st>>>(X, P, M, idx, B, pairs, n_pairs, min_valid, corr);
    B2S_CHECK_LAUNCH();
  } else {
    B2S_REQUIRE(P <= CORR_MAX_P, "corr_bootstrap: spearman supports at most %d rows per resample (got %d)", CORR_MAX_P, P);
    int n2 = 1;
    while (n2 < P) n2 <<= 1;
    const size_t smem = static_cast<size_t>(n2) * (5 * sizeof(double) + sizeof(int));
    if (smem > 48 * 1024)   // per-device attribute: set whenever it is needed
      B2S_CHECK_CUDA(cudaFuncSetAttribute(corr_spearman_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    dim3 grid(static_cast<unsigned>(n_pairs), static_cast<unsigned>(B));
    corr_spearman_kernel<<<grid, 256, smem, st>>>(X, P, M, idx, B, pairs, n_pairs, n2, corr);
    B2S_CHECK_LAUNCH();
  }
  if (stats && n_pairs > CORR_MAX_PAIRS) {     // mean / std / count without the median (no shared-memory sort of that size)
    corr_stats_big_kernel<<<static_cast<unsigned>(B), 256, 0, st>>>(corr, n_pairs, stats);
    B2S_CHECK_LAUNCH();
  } else if (stats) {
    int n2 = 1;
    while (n2 < n_pairs) n2 <<= 1;
    const size_t smem = static_cast<size_t>(n2) * (sizeof(double) + sizeof(int));
    corr_stats_kernel<<<static_cast<unsigned>(B), 256, smem, st>>>(corr, n_pairs, n2, stats);
    B2S_CHECK_LAUNCH();
  }
  return B2S_OK;
}
