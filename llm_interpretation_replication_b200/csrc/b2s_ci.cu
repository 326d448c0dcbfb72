// Simulated-individual correlation bootstrap.
//
// Replaces the body of `for i in range(n_bootstrap): calculate_individual_correlations(..., n_samples=100, seed=i)`
// of reference survey_analysis/bootstrap_confidence_intervals.py:54-99,126-148 -- 1.8e7 scipy.pearsonr calls in
// pure Python there.  One thread per seed i runs the whole stream of that seed on a private NumPy-legacy MT19937:
//   np.random.seed(i); for each model (CSV order): n_samples x { group = randint(1, 6); for the group's 10
//   questions that have a prompt, statistics and a model answer: human = clip(normal(mean, std), 0, 1) };
//   pearsonr(human, model) when >= 8 questions and no NaN; mean of the correlations per model type.
// legacy normal = loc + scale * gauss, gauss = Marsaglia polar method with the cached second value
// (SURVEY.md Appendix B); randint(1, 6) = 1 + masked rejection with mask 7.
// The 624-word states of the CTA's 64 seeds live in shared memory, word-interleaved.
#include "b2s_common.cuh"

namespace b2s {

constexpr int CI_THREADS = 64;
constexpr int CI_N = 624, CI_M = 397;
constexpr int CI_G = 5, CI_Q = 10;

struct CiArgs {
  const uint8_t* avail;      // [n_models, 5, 10] question usable for this model
  const double* rel;         // [n_models, 5, 10] model relative probability (NaN allowed)
  const double* mean;        // [5, 10] human mean on the 0-1 scale
  const double* stdv;        // [5, 10]
  const int32_t* model_type; // [n_models] 0 = base, 1 = instruct
  int n_models, n_samples, n_seeds;
  uint32_t seed0;
  double* means_out;         // [n_seeds, 2] mean correlation of base / instruct models (NaN if none)
  double* corrs_out;         // [n_models, n_samples] per-sample correlations of the FIRST seed (NaN = skipped), or null
  uint32_t* final_state;     // [625] MT state after the LAST seed's run, or null
};

struct Rng {
  uint32_t* mt;   // this thread's words at mt[j * CI_THREADS]
  int pos;
  bool has_gauss;
  double gauss;
  __device__ __forceinline__ uint32_t& w(int j) { return mt[j * CI_THREADS]; }
  __device__ void seed(uint32_t s) {
    uint32_t x = s;
    w(0) = x;
    for (int i = 1; i < CI_N; ++i) { x = 1812433253u * (x ^ (x >> 30)) + static_cast<uint32_t>(i); w(i) = x; }
    pos = CI_N; has_gauss = false; gauss = 0.0;
  }
  __device__ void twist() {
    for (int kk = 0; kk < CI_N; ++kk) {
      const uint32_t y = (w(kk) & 0x80000000u) | (w((kk + 1) % CI_N) & 0x7fffffffu);
      w(kk) = w((kk + CI_M) % CI_N) ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    pos = 0;
  }
  __device__ __forceinline__ uint32_t next() {
    if (pos >= CI_N) twist();
    uint32_t y = w(pos++);
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
  }
  __device__ __forceinline__ double dbl() {
    const uint32_t a = next() >> 5, b = next() >> 6;
    return (a * 67108864.0 + b) / 9007199254740992.0;
  }
  __device__ double normal01() {
    if (has_gauss) { has_gauss = false; return gauss; }
    double x1, x2, r2;
    do {
      x1 = 2.0 * dbl() - 1.0;
      x2 = 2.0 * dbl() - 1.0;
      r2 = __dadd_rn(__dmul_rn(x1, x1), __dmul_rn(x2, x2));   // no FMA contraction: keeps the rejection decisions
    } while (r2 >= 1.0 || r2 == 0.0);
    const double f = sqrt(-2.0 * log(r2) / r2);
    gauss = f * x1; has_gauss = true;
    return f * x2;
  }
  __device__ __forceinline__ int bounded(uint32_t rng, uint32_t mask) {
    uint32_t v;
    do { v = next() & mask; } while (v > rng);
    return static_cast<int>(v);
  }
};

// numpy's pairwise summation for n <= 10: eight running sums for the first eight, then the tail in order
__device__ __forceinline__ double np_sum(const double* a, int n) {
  if (n < 8) { double s = 0.0; for (int i = 0; i < n; ++i) s += a[i]; return s; }
  double s = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
  for (int i = 8; i < n; ++i) s += a[i];
  return s;
}

// scipy.stats.pearsonr: r = dot(xm / ||xm||, ym / ||ym||) clipped to [-1, 1]; NaN for a constant input
__device__ double pearson(const double* x, const double* y, int n) {
  const double xm = np_sum(x, n) / n, ym = np_sum(y, n) / n;
  double dx[CI_Q], dy[CI_Q], sx = 0.0, sy = 0.0;
  bool cx = true, cy = true;
  for (int i = 0; i < n; ++i) {
    dx[i] = x[i] - xm; dy[i] = y[i] - ym;
    sx += dx[i] * dx[i]; sy += dy[i] * dy[i];
    cx = cx && (x[i] == x[0]); cy = cy && (y[i] == y[0]);
  }
  if (cx || cy) return nan("");
  const double nx = sqrt(sx), ny = sqrt(sy);
  double r = 0.0;
  for (int i = 0; i < n; ++i) r += (dx[i] / nx) * (dy[i] / ny);
  return fmax(fmin(r, 1.0), -1.0);
}

__global__ void __launch_bounds__(CI_THREADS)
ci_sim_kernel(CiArgs a) {
  extern __shared__ uint32_t smem_u[];
  uint32_t* states = smem_u;                                         // [624][64]
  double* tab = reinterpret_cast<double*>(states + CI_N * CI_THREADS);
  double* s_mean = tab;                                              // [50]
  double* s_std = tab + CI_G * CI_Q;                                 // [50]
  double* s_rel = tab + 2 * CI_G * CI_Q;                             // [n_models][50]
  uint8_t* s_avail = reinterpret_cast<uint8_t*>(s_rel + a.n_models * CI_G * CI_Q);
  for (int i = threadIdx.x; i < CI_G * CI_Q; i += blockDim.x) { s_mean[i] = a.mean[i]; s_std[i] = a.stdv[i]; }
  for (int i = threadIdx.x; i < a.n_models * CI_G * CI_Q; i += blockDim.x) { s_rel[i] = a.rel[i]; s_avail[i] = a.avail[i]; }
  __syncthreads();
  const int sidx = blockIdx.x * CI_THREADS + threadIdx.x;
  if (sidx >= a.n_seeds) return;
  Rng rng;
  rng.mt = states + threadIdx.x;
  rng.seed(a.seed0 + static_cast<uint32_t>(sidx));
  double sum[2] = {0.0, 0.0};
  long long cnt[2] = {0, 0};
  for (int m = 0; m < a.n_models; ++m) {
    const int ty = a.model_type[m];
    for (int s = 0; s < a.n_samples; ++s) {
      const int g = rng.bounded(4u, 7u);                             // np.random.randint(1, 6) - 1
      double hv[CI_Q], mv[CI_Q];
      int n = 0;
      bool has_nan = false;
      const int off = (m * CI_G + g) * CI_Q;
      for (int k = 0; k < CI_Q; ++k) {
        if (s_avail[off + k]) {
          const double x = __dadd_rn(s_mean[g * CI_Q + k], __dmul_rn(s_std[g * CI_Q + k], rng.normal01()));
          hv[n] = fmin(fmax(x, 0.0), 1.0);
          mv[n] = s_rel[off + k];
          has_nan = has_nan || isnan(mv[n]);
          ++n;
        }
      }
      double corr = nan("");
      if (n >= 8 && !has_nan) corr = pearson(hv, mv, n);
      if (!isnan(corr)) { sum[ty] += corr; cnt[ty]++; }
      if (a.corrs_out && sidx == 0) a.corrs_out[static_cast<size_t>(m) * a.n_samples + s] = corr;
    }
  }
  a.means_out[2 * sidx] = cnt[0] ? sum[0] / static_cast<double>(cnt[0]) : nan("");
  a.means_out[2 * sidx + 1] = cnt[1] ? sum[1] / static_cast<double>(cnt[1]) : nan("");
  if (a.final_state && sidx == a.n_seeds - 1) {
    for (int j = 0; j < CI_N; ++j) a.final_state[j] = rng.w(j);
    a.final_state[CI_N] = static_cast<uint32_t>(rng.pos);
  }
}

}  // namespace b2s

using namespace b2s;

extern "C" int b2s_ci_simulate(const uint8_t* avail, const double* rel, const double* mean, const double* stdv,
                               const int32_t* model_type, int32_t n_models, int32_t n_samples, uint32_t seed0, int32_t n_seeds,
                               double* means_out, double* corrs_out, uint32_t* final_state, void* stream) {
  B2S_REQUIRE(avail && rel && mean && stdv && model_type && means_out, "ci_simulate: null argument");
  B2S_REQUIRE(n_models >= 1 && n_samples >= 1 && n_seeds >= 1, "ci_simulate: bad sizes");
  const size_t smem = static_cast<size_t>(CI_N) * CI_THREADS * 4 + (2 + static_cast<size_t>(n_models)) * CI_G * CI_Q * 8 +
                      static_cast<size_t>(n_models) * CI_G * CI_Q + 64;
  B2S_REQUIRE(smem <= 227 * 1024, "ci_simulate: %d models need %zu bytes of shared memory (limit 227 KB)", n_models, smem);
  B2S_CHECK_CUDA(cudaFuncSetAttribute(ci_sim_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  CiArgs a{avail, rel, mean, stdv, model_type, n_models, n_samples, n_seeds, seed0, means_out, corrs_out, final_state};
  ci_sim_kernel<<<static_cast<unsigned>(ceil_div(n_seeds, CI_THREADS)), CI_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}
