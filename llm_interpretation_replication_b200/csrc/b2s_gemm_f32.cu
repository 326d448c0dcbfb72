// fp32 CUDA-core GEMM for the FP32_CHECK compute mode: C[M,N] = A[M,K] . W[N,K]^T (+bias), same
// epilogues as the tensor-core kernel.  This is the <=1e-5 parity mode named in BASELINE.json; it is not
// the throughput path, but the full-depth parity runs (tests/test_gpu_full_depth.py, tools/parity_histogram.py)
// push whole BASELINE configs through it, so it is a register-tiled kernel rather than a naive one:
// 128x128 tiles, BK=16, 256 threads, 8x8 register micro-tile (two 4-wide strips per dimension so that the
// shared-memory reads are conflict-free float4s), next k-slab prefetched into registers while the current one
// is multiplied.  Every output element is one fmaf chain over k = 0..K-1 in ascending order, so the result does not
// depend on the tile shape or on M (bit-identical to the 64x64 kernel of round 1).
#include "b2s_common.cuh"

namespace b2s {

static constexpr int FT = 128, FK = 16;

template <int EPI, bool VEC>
__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                float* __restrict__ out, int M, int N, int K) {
  __shared__ __align__(16) float sA[FK][FT + 4];   // +4: the transposing stores of a warp spread over more banks
  __shared__ __align__(16) float sW[FK][FT + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * FT, n0 = blockIdx.x * FT;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  // global -> register staging: the 128 x 16 slab is 512 float4; thread t takes float4 t and t + 256
  float4 ra[2], rw[2];
  auto load_slab = [&](int k0) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int idx = threadIdx.x + r * 256;
      const int row = idx >> 2, kq = (idx & 3) * 4;
      const int gm = m0 + row, gn = n0 + row, gk = k0 + kq;
      if constexpr (VEC) {   // K % 4 == 0 and 16-byte aligned bases: whole float4s are in or out of range
        ra[r] = (gm < M && gk < K) ? *reinterpret_cast<const float4*>(A + static_cast<size_t>(gm) * K + gk) : make_float4(0.f, 0.f, 0.f, 0.f);
        rw[r] = (gn < N && gk < K) ? *reinterpret_cast<const float4*>(W + static_cast<size_t>(gn) * K + gk) : make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        float a[4], w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          a[u] = (gm < M && gk + u < K) ? A[static_cast<size_t>(gm) * K + gk + u] : 0.f;
          w[u] = (gn < N && gk + u < K) ? W[static_cast<size_t>(gn) * K + gk + u] : 0.f;
        }
        ra[r] = make_float4(a[0], a[1], a[2], a[3]);
        rw[r] = make_float4(w[0], w[1], w[2], w[3]);
      }
    }
  };
  auto store_slab = [&]() {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int idx = threadIdx.x + r * 256;
      const int row = idx >> 2, kq = (idx & 3) * 4;
      sA[kq][row] = ra[r].x; sA[kq + 1][row] = ra[r].y; sA[kq + 2][row] = ra[r].z; sA[kq + 3][row] = ra[r].w;
      sW[kq][row] = rw[r].x; sW[kq + 1][row] = rw[r].y; sW[kq + 2][row] = rw[r].z; sW[kq + 3][row] = rw[r].w;
    }
  };

  load_slab(0);
  for (int k0 = 0; k0 < K; k0 += FK) {
    store_slab();
    __syncthreads();
    if (k0 + FK < K) load_slab(k0 + FK);   // in flight while this slab is multiplied
#pragma unroll
    for (int kk = 0; kk < FK; ++kk) {
      // rows ty*4..+3 and 64+ty*4..+3; columns tx*4..+3 and 64+tx*4..+3
      const float4 a0 = *reinterpret_cast<const float4*>(&sA[kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[kk][64 + ty * 4]);
      const float4 w0 = *reinterpret_cast<const float4*>(&sW[kk][tx * 4]);
      const float4 w1 = *reinterpret_cast<const float4*>(&sW[kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int gn = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (gn >= N) continue;
      float v = acc[i][j] + (bias ? bias[gn] : 0.f);
      float* o = out + static_cast<size_t>(gm) * N + gn;
      if constexpr (EPI == EPI_GELU_BF16) v = gelu_erf(v);
      if constexpr (EPI == EPI_RESID_F32) v += *o;
      *o = v;
    }
  }
}

template <int EPI>
static void launch_f32(const float* A, const float* W, const float* bias, float* out, int M, int N, int K, dim3 grid,
                       cudaStream_t st) {
  const bool vec = K % 4 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0;
  if (vec) gemm_f32_kernel<EPI, true><<<grid, 256, 0, st>>>(A, W, bias, out, M, N, K);
  else gemm_f32_kernel<EPI, false><<<grid, 256, 0, st>>>(A, W, bias, out, M, N, K);
}

int gemm_f32_launch(const float* A, const float* W, const float* bias, float* out, int M, int N, int K, int epilogue,
                    cudaStream_t st) {
  B2S_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_f32: empty problem M=%d N=%d K=%d", M, N, K);
  dim3 grid(static_cast<unsigned>(ceil_div(N, FT)), static_cast<unsigned>(ceil_div(M, FT)));
  switch (epilogue) {
    case EPI_STORE_BF16:   // in fp32 mode "store" means store fp32
    case EPI_STORE_F32: launch_f32<EPI_STORE_F32>(A, W, bias, out, M, N, K, grid, st); break;
    case EPI_GELU_BF16: launch_f32<EPI_GELU_BF16>(A, W, bias, out, M, N, K, grid, st); break;
    case EPI_RESID_F32: launch_f32<EPI_RESID_F32>(A, W, bias, out, M, N, K, grid, st); break;
    default: set_error("gemm_f32: unknown epilogue %d", epilogue); return B2S_EINVAL;
  }
  B2S_CHECK_LAUNCH();
  return B2S_OK;
}

}  // namespace b2s
