// fp32 CUDA-core GEMM for the FP32_CHECK compute mode: C[M,N] = A[M,K] . W[N,K]^T (+bias), same
// epilogues as the tensor-core kernel.  This is the <=1e-5 parity mode named in BASELINE.json; it is not
// the throughput path.  64x64 tiles, BK=16, 256 threads, 4x4 register micro-tile, fp32 FMA accumulate.
#include "b2s_common.cuh"

namespace b2s {

static constexpr int FT = 64, FK = 16;

template <int EPI>
__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                float* __restrict__ out, int M, int N, int K) {
  __shared__ float sA[FK][FT + 4];
  __shared__ float sW[FK][FT + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * FT, n0 = blockIdx.x * FT;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += FK) {
    // each thread loads 4 elements of A and 4 of W: tile is 64 rows x 16 k
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int idx = threadIdx.x + r * 256;   // 0..1023
      const int row = idx >> 4, kk = idx & 15;
      const int gm = m0 + row, gn = n0 + row, gk = k0 + kk;
      sA[kk][row] = (gm < M && gk < K) ? A[static_cast<size_t>(gm) * K + gk] : 0.f;
      sW[kk][row] = (gn < N && gk < K) ? W[static_cast<size_t>(gn) * K + gk] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < FK; ++kk) {
      float a[4], w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = sA[kk][ty * 4 + i]; w[i] = sW[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j] + (bias ? bias[gn] : 0.f);
      float* o = out + static_cast<size_t>(gm) * N + gn;
      if constexpr (EPI == EPI_GELU_BF16) v = gelu_erf(v);
      if constexpr (EPI == EPI_RESID_F32) v += *o;
      *o = v;
    }
  }
}

int gemm_f32_launch(const float* A, const float* W, const float* bias, float* out, int M, int N, int K, int epilogue,
                    cudaStream_t st) {
  B2S_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_f32: empty problem M=%d N=%d K=%d", M, N, K);
  dim3 grid(static_cast<unsigned>(ceil_div(N, FT)), static_cast<unsigned>(ceil_div(M, FT)));
  switch (epilogue) {
    case EPI_STORE_BF16:   // in fp32 mode "store" means store fp32
    case EPI_STORE_F32: gemm_f32_kernel<EPI_STORE_F32><<<grid, 256, 0, st>>>(A, W, bias, out, M, N, K); break;
    case EPI_GELU_BF16: gemm_f32_kernel<EPI_GELU_BF16><<<grid, 256, 0, st>>>(A, W, bias, out, M, N, K); break;
    case EPI_RESID_F32: gemm_f32_kernel<EPI_RESID_F32><<<grid, 256, 0, st>>>(A, W, bias, out, M, N, K); break;
    default: set_error("gemm_f32: unknown epilogue %d", epilogue); return B2S_EINVAL;
  }
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // namespace b2s
