// gemm_f32.cu -- SIMT fp32 GEMM  C[M,N] = A[M,K] @ W[N,K]^T  with the same fused gdEpilogue.
// Used for "fp32 mode" (1e-5 parity runs) and one-off setup products (timestep-embedding table,
// seed-pose embedding); the timed bf16 path uses gemm_tcgen05.cu.  64x64 tile, 16-deep K slab,
// 256 threads, 4x4 register micro-tile, fp32 FMA accumulation in ascending-k order.
#include "common.cuh"

namespace gd {

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16;

__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw, int M, int N, int K,
                const gdEpilogue ep) {
  __shared__ float sA[SG_BK][SG_BM + 4];
  __shared__ float sW[SG_BK][SG_BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;           // 16 x 16 threads; each 4 rows x 4 cols
  const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * SG_BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int lr = tid >> 2;            // 0..63 : row inside tile to load
  const int lk = (tid & 3) * 4;       // 0,4,8,12 : k offset to load (4 consecutive)
  for (int k0 = 0; k0 < K; k0 += SG_BK) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + lk + q;
      const int am = m0 + lr, wn = n0 + lr;
      sA[lk + q][lr] = (am < M && k < K) ? A[(size_t)am * lda + k] : 0.f;
      sW[lk + q][lr] = (wn < N && k < K) ? W[(size_t)wn * ldw + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sW[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    const int n = n0 + tx * 4;
    if (m < M && n < N) epilogue_row_chunk<4>(ep, m, n, N, acc[i]);
  }
}

}  // namespace gd

extern "C" int gd_gemm_f32(const float* A, int32_t lda, const float* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                           const gdEpilogue* ep, void* stream) {
  using namespace gd;
  GD_REQUIRE(A && W && ep, "gd_gemm_f32: null pointer");
  GD_REQUIRE(M > 0 && N > 0 && K > 0 && lda >= K && ldw >= K, "gd_gemm_f32: bad shape M=%d N=%d K=%d", M, N, K);
  int rc = validate_epilogue(ep, M, N);
  if (rc) return rc;
  dim3 grid(ceil_div(M, SG_BM), ceil_div(N, SG_BN));
  gemm_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, lda, W, ldw, M, N, K, *ep);
  GD_LAUNCH_CHECK();
  return GD_OK;
}
