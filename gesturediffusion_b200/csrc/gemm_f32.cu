// gemm_f32.cu -- SIMT fp32 GEMM  C[M,N] = A[M,K] @ W[N,K]^T  with the same fused gdEpilogue.
// Used for "fp32 mode" (1e-5 parity runs) and one-off setup products (timestep-embedding table,
// seed-pose embedding); the timed bf16 path uses gemm_tcgen05.cu.  64x64 tile, 16-deep K slab,
// 256 threads, 4x4 register micro-tile, fp32 FMA accumulation in ascending-k order.
// Few tiles and a long K (the seed-pose embedding: [B, 10 x njoints] x [512, 10 x njoints]^T, mdm.py:186-187 -- 16 tiles
// of 312 K slabs at batch 128) go to gemm_f32_splitk_kernel: a thread-block cluster of 2 / 4 / 8 CTAs shares one tile,
// CTA r sums K slice r (ascending k, next slab prefetched into registers), the partial tiles meet in CTA 0 through
// distributed shared memory and are added in rank order (deterministic), then the same epilogue.
#include <cooperative_groups.h>

#include "common.cuh"

namespace gd {

int num_sms();   // gemm_tcgen05.cu

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

constexpr int SG_MAX_SPLIT = 8;

__global__ void __launch_bounds__(256)
gemm_f32_splitk_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw, int M, int N, int K,
                       int k_per_cta, const gdEpilogue ep) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ float sA[SG_BK][SG_BM + 4];
  __shared__ float sW[SG_BK][SG_BN + 4];
  __shared__ float sP[SG_BM][SG_BN];                 // this CTA's partial tile (read by CTA 0 of the cluster)
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * SG_BN;
  const int rank = (int)cluster.block_rank(), nsplit = (int)cluster.num_blocks();
  const int kbeg = rank * k_per_cta, kend = min(K, kbeg + k_per_cta);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  const int am = m0 + lr, wn = n0 + lr;
  const float* ap = A + (size_t)min(am, M - 1) * lda;
  const float* wp = W + (size_t)min(wn, N - 1) * ldw;
  float ra[4], rw[4];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = k0 + lk + q;
      ra[q] = (am < M && k < kend) ? __ldg(ap + k) : 0.f;
      rw[q] = (wn < N && k < kend) ? __ldg(wp + k) : 0.f;
    }
  };
  if (kbeg < kend) fetch(kbeg);
  for (int k0 = kbeg; k0 < kend; k0 += SG_BK) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      sA[lk + q][lr] = ra[q];
      sW[lk + q][lr] = rw[q];
    }
    __syncthreads();
    if (k0 + SG_BK < kend) fetch(k0 + SG_BK);        // in flight while this slab is multiplied
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
  if (rank != 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<float4*>(&sP[ty * 4 + i][tx * 4]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
  cluster.sync();                                    // partial tiles are written
  if (rank == 0) {
    for (int r = 1; r < nsplit; ++r) {               // rank order: the sum does not depend on timing
      const float* rp = cluster.map_shared_rank(&sP[0][0], r);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 q = *reinterpret_cast<const float4*>(rp + (ty * 4 + i) * SG_BN + tx * 4);
        acc[i][0] += q.x; acc[i][1] += q.y; acc[i][2] += q.z; acc[i][3] += q.w;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + ty * 4 + i;
      const int n = n0 + tx * 4;
      if (m < M && n < N) epilogue_row_chunk<4>(ep, m, n, N, acc[i]);
    }
  }
  cluster.sync();                                    // no CTA exits while CTA 0 may still read its shared memory
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
  const int tiles = (int)(grid.x * grid.y), sms = num_sms();
  int split = 1;
  while (split < SG_MAX_SPLIT && tiles * split * 2 <= sms && K >= 512 * split) split *= 2;
  if (split > 1) {
    const int k_per_cta = ceil_div(ceil_div(K, SG_BK), split) * SG_BK;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid.x, grid.y, split);
    cfg.blockDim = dim3(256);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = split;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    GD_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_f32_splitk_kernel, A, lda, W, ldw, M, N, K, k_per_cta, *ep));
    GD_LAUNCH_CHECK();
    return GD_OK;
  }
  gemm_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, lda, W, ldw, M, N, K, *ep);
  GD_LAUNCH_CHECK();
  return GD_OK;
}
