// noalloc.cu -- an entry point for callers that cannot allocate (the
// reference has no counterpart): the fp32-mode post-norm sub-layer tail
//   X = LayerNorm(residual + A[M,K] @ W[N,K]^T + bias) * gamma + beta          (mdm.py:90-96, norm_first=False)
// which the engine runs as gd_gemm_f32 (residual epilogue) -> Y32 scratch [M, N] fp32 -> gd_layernorm with the scratch
// in a caller-provided workspace buffer.  Here the scratch is a file-scope device array of a fixed maximum size, so the
// call needs no workspace and performs no cudaMalloc / cudaFree / allocating library call: usable under CUDA-graph
// capture, in tight serving loops and in harnesses that forbid allocation.
//
// Limits and sharing (also in include/gd_b200.h):
//   * M * N <= GD_NOALLOC_MAX_ELEMS (1 Mi fp32 = 4 MiB, e.g. 2048 rows x 512); larger problems are rejected with
//     GD_ERR_INVALID and a message -- use gd_gemm_f32 + gd_layernorm with your own buffer instead;
//   * every GPU that loads the library holds the 4 MiB until the process exits;
//   * ALL calls on a device share the one array, which keeps the previous call's contents.  Nothing has to start at
//     zero: the GEMM writes every element of the [M, N] region before the LayerNorm kernel reads it (same stream, so
//     ordered).  Calls that overlap on the GPU (different streams, or graphs replayed concurrently) race on it.
#include "common.cuh"

#define GD_NOALLOC_MAX_ELEMS (1 << 20)

namespace gd {

static __device__ __align__(256) float g_noalloc_scratch[GD_NOALLOC_MAX_ELEMS];

// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, not a device address: the address is
// queried once per device, with that device current, and cached.
static float* g_scratch_ptr[kMaxDevices] = {};

static int scratch_for_current_device(float** out) {
  int dev = -1;
  GD_CUDA_OK(cudaGetDevice(&dev));
  GD_REQUIRE(dev >= 0 && dev < kMaxDevices, "gd_linear_res_ln_f32_noalloc: device index %d out of range", dev);
  if (g_scratch_ptr[dev] == nullptr) {
    float* p = nullptr;
    GD_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    GD_REQUIRE(p != nullptr, "gd_linear_res_ln_f32_noalloc: cudaGetSymbolAddress returned NULL");
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return GD_OK;
}

}  // namespace gd

extern "C" int gd_noalloc_max_elems(void) { return GD_NOALLOC_MAX_ELEMS; }

extern "C" int gd_linear_res_ln_f32_noalloc(const float* A, int32_t lda, const float* W, int32_t ldw, int32_t M,
                                            int32_t N, int32_t K, const float* bias, const float* residual,
                                            int32_t ld_res, const float* gamma, const float* beta, float eps,
                                            float* out_f32, void* out_h16, int32_t h16_fmt, void* stream) {
  using namespace gd;
  GD_REQUIRE(A && W && residual && gamma && beta && (out_f32 || out_h16),
             "gd_linear_res_ln_f32_noalloc: null pointer");
  GD_REQUIRE(M > 0 && N > 0 && K > 0, "gd_linear_res_ln_f32_noalloc: bad shape M=%d N=%d K=%d", M, N, K);
  GD_REQUIRE((long long)M * (long long)N <= (long long)GD_NOALLOC_MAX_ELEMS,
             "gd_linear_res_ln_f32_noalloc: M*N = %lld exceeds the built-in scratch of %d fp32 elements (4 MiB); "
             "use gd_gemm_f32 + gd_layernorm with a caller-provided buffer for larger problems",
             (long long)M * (long long)N, (int)GD_NOALLOC_MAX_ELEMS);
  GD_REQUIRE(ld_res >= N, "gd_linear_res_ln_f32_noalloc: ld_res < N");
  float* y = nullptr;
  int rc = scratch_for_current_device(&y);
  if (rc) return rc;
  gdEpilogue ep = {};
  ep.bias = bias;
  ep.residual = residual;
  ep.ld_res = ld_res;
  ep.act = GD_ACT_NONE;
  ep.out_f32 = y;                                   // dense [M, N] in the scratch
  ep.ld_out = N;
  rc = gd_gemm_f32(A, lda, W, ldw, M, N, K, &ep, stream);
  if (rc) return rc;
  return gd_layernorm(y, gamma, beta, out_f32, out_h16, h16_fmt, M, N, eps, stream);
}
