// probe: which cp.async.bulk.tensor store forms work (fp32, no swizzle, 2D / 3D, negative coordinates)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
template <int RANK>
__global__ void k(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2) {
  extern __shared__ __align__(1024) float sm[];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (float)i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(sm);
    if (RANK == 3)
      asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(&tm), "r"(s),
                   "r"(c0), "r"(c1), "r"(c2) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(&tm), "r"(s),
                   "r"(c0), "r"(c1) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}
int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeFn enc = (EncodeFn)fn;
  struct Case { int T, N, Q, c0, c1, c2; const char* what; };
  Case cases[] = {{196, 263, 4, 4, 250, 1, "column box past N"}, {196, 263, 4, 180, 0, 1, "frame box past T"},
                  {32, 32, 6, 0, 0, 0, "box == dims"}, {20, 24, 6, 0, 0, 0, "box larger than dims"},
                  {196, 263, 4, 195, 0, 3, "one frame inside"}, {196, 263, 4, 196, 0, 1, "fully outside (frames)"}};
  for (auto& c : cases) {
    float* d;
    cudaMalloc(&d, sizeof(float) * c.T * c.N * c.Q);
    cudaMemset(d, 0, sizeof(float) * c.T * c.N * c.Q);
    CUtensorMap tm;
    cuuint64_t dims[3] = {(cuuint64_t)c.T, (cuuint64_t)c.N, (cuuint64_t)c.Q};
    cuuint64_t strides[2] = {(cuuint64_t)c.T * 4, (cuuint64_t)c.T * c.N * 4};
    cuuint32_t box[3] = {32, 32, 1}, es[3] = {1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    k<3><<<1, 128, 4096>>>(tm, c.c0, c.c1, c.c2);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%-26s T %d N %d: encode %d, run %s\n", c.what, c.T, c.N, (int)r, cudaGetErrorString(e));
    fflush(stdout);
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
