// common.cuh -- shared device/host helpers for the gd_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/gd_b200.h"

namespace gd {

// ------------------------------------------------------------------ host-side error plumbing
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;
extern unsigned long long* g_ffn_trace;     // debug timeline buffer (gd_ffn_trace), NULL = off

#define GD_REQUIRE(cond, ...)                                 \
  do {                                                        \
    if (!(cond)) {                                            \
      gd::set_error(__VA_ARGS__);                             \
      return GD_ERR_INVALID;                                  \
    }                                                         \
  } while (0)

#define GD_CUDA_OK(expr)                                                              \
  do {                                                                                \
    cudaError_t e__ = (expr);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      gd::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return GD_ERR_CUDA;                                                             \
    }                                                                                 \
  } while (0)

// GD_PDL=0 turns programmatic dependent launch off (kernels then serialise the ordinary way).
// GD_PDL is a bit mask over kernel kinds: 1 plain GEMM with at least one tile per SM, 2 GEMM+LayerNorm, 4 attention,
// 8 plain GEMM with fewer tiles than SMs (small batches).  Default 14: measured on the cfg2a step (ms per denoise step,
// same box): none 1.698, all 1.681, large GEMM only 1.707, GEMM+LN and attention 1.671; on the batch-1 step (cfg1, every
// GEMM is "small"): 0.354 with the GEMMs included against 0.365 without and 0.394 with none (profiles/r02k_cfg1_pdl.txt).
enum { PDL_GEMM = 1, PDL_GEMM_LN = 2, PDL_ATTENTION = 4, PDL_GEMM_SMALL = 8 };
inline bool pdl_enabled(int kind) {
  static int mask = -1;
  if (mask < 0) {
    const char* e = getenv("GD_PDL");
    mask = e ? atoi(e) : 14;
  }
  return (mask & kind) != 0;
}
// Launch with the programmatic-stream-serialization attribute: the grid may be scheduled while the previous
// kernel of the stream drains; the kernel itself orders its global accesses with ptx::grid_dep_wait().
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int kind, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  int n = 0;
  if (pdl_enabled(kind)) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define GD_LAUNCH_CHECK()                                                             \
  do {                                                                                \
    gd::g_launches.fetch_add(1, std::memory_order_relaxed);                           \
    cudaError_t e__ = cudaGetLastError();                                             \
    if (e__ != cudaSuccess) {                                                         \
      gd::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
      return GD_ERR_CUDA;                                                             \
    }                                                                                 \
  } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Host-side one-off state (cudaFuncSetAttribute, scratch allocations) is per DEVICE: function attributes belong to the
// device's context, and one process may drive several GPUs (a model on cuda:1 while cuda:0 is busy elsewhere).
constexpr int kMaxDevices = 64;
static inline int current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kMaxDevices) d = 0;
  return d;
}

// ------------------------------------------------------------------ device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

// 16-bit packing in the runtime-selected format (GD_BF16 / GD_FP16); fp16 saturates instead of
// overflowing to inf.
__device__ __forceinline__ uint32_t pack_h16x2(float lo, float hi, int fmt) {
  uint32_t r;
  if (fmt == GD_FP16) {
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));   // one F2FP, clamps to +-65504
    return r;
  }
  return pack_bf16x2(lo, hi);
}
// inverse of pack_h16x2: (low half, high half) as floats
__device__ __forceinline__ float2 unpack_h16x2(uint32_t u, int fmt) {
  float2 f;
  if (fmt == GD_FP16) {
    asm("{\n\t.reg .f16 l, h;\n\tmov.b32 {l, h}, %2;\n\tcvt.f32.f16 %0, l;\n\tcvt.f32.f16 %1, h;\n\t}"
        : "=f"(f.x), "=f"(f.y) : "r"(u));
    return f;
  }
  f.x = __uint_as_float(u << 16);
  f.y = __uint_as_float(u & 0xFFFF0000u);
  return f;
}
__device__ __forceinline__ uint16_t cvt_h16(float v, int fmt) {
  return (uint16_t)(pack_h16x2(v, 0.f, fmt) & 0xFFFFu);
}

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float silu(float x) { return x / (1.0f + __expf(-x)); }

// erf-GELU for the 16-bit tensor-core epilogues (outputs are rounded to 16 bit afterwards) as
//   gelu(x) = 0.5 x (1 + tanh(x (a0 + a1 x^2 + a2 x^4)))
// with (a0, a1, a2) a minimax fit of atanh(erf(x / sqrt 2)) / x: max |error| against float64 erf-GELU over [-6, 6]
// 2.5e-5 in exact arithmetic (the textbook tanh-GELU constants give 4.7e-4), plus the MUFU.TANH error (2^-11
// relative on tanh: <= 2.4e-4 |x| absolute, about one extra fp16 rounding of the result).  x^2 is clamped to 25
// so that the polynomial never turns over: beyond |x| = 5 the argument is 1.5 x and tanh is +-1 to 6e-7.
// 9 instructions (FMUL, FMNMX, 2 FFMA, FMUL, MUFU.TANH, FMUL, FFMA + the bias FADD) instead of 17 for the previous
// erfc(z) = 2^q(z) form: the GELU epilogue is issue-bound (linear1 35.5 us against 21.0 us without epilogue
// arithmetic at the cfg2a shape), so the instruction count is what its speed is made of.
__device__ __forceinline__ float gelu_fast(float x) {
  const float x2 = fminf(x * x, 25.0f);
  float p = fmaf(x2, -3.5170863903880907e-4f, 3.7006827112020954e-2f);
  p = fmaf(x2, p, 7.975065863776912e-1f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(p * x));
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
}

__device__ __forceinline__ void f4_add(float4& a, const float4 b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}

// ------------------------------------------------------------------ fused GEMM epilogue
// Processes CH consecutive columns [n0, n0+CH) of logical row m.  See gdEpilogue in gd_b200.h.
template <int CH>
__device__ __forceinline__ void epilogue_row_chunk(const gdEpilogue& ep, int m, int n0, int N,
                                                   float (&v)[CH]) {
  const int nvalid = min(CH, N - n0);
  if (nvalid <= 0) return;
  if (ep.bias) {
#pragma unroll
    for (int i = 0; i < CH; ++i)
      if (i < nvalid) v[i] += __ldg(ep.bias + n0 + i);
  }
  if (ep.rowtab) {
    const float* tr = ep.rowtab + (size_t)((m % ep.tab_period) + ep.tab_offset) * ep.ld_tab + n0;
#pragma unroll
    for (int i = 0; i < CH; ++i)
      if (i < nvalid) v[i] += __ldg(tr + i);
  }
  const int ncopy = ep.dup > 1 ? ep.dup : 1;
  int base_row = m;
  if (ep.map_t > 0) base_row = (m / ep.map_t) * ep.map_s + ep.map_off + (m % ep.map_t);
  for (int c = 0; c < ncopy; ++c) {
    float w[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) w[i] = v[i];
    if (ep.seqbias) {
      const float* sb = ep.seqbias + (size_t)(c * ep.seq_count + m / ep.seq_rows) * ep.ld_seq + n0;
#pragma unroll
      for (int i = 0; i < CH; ++i)
        if (i < nvalid) w[i] += __ldg(sb + i);
    }
    const size_t r = (size_t)base_row + (size_t)c * ep.dup_row_stride;
    if (ep.residual) {
      const float* rr = ep.residual + r * ep.ld_res + n0;
      if (nvalid == CH && (CH % 4 == 0) && ((((size_t)rr) & 15) == 0)) {
#pragma unroll
        for (int i = 0; i < CH; i += 4) {
          float4 q = *reinterpret_cast<const float4*>(rr + i);
          w[i] += q.x; w[i + 1] += q.y; w[i + 2] += q.z; w[i + 3] += q.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (i < nvalid) w[i] += rr[i];
      }
    }
    if (ep.act == GD_ACT_GELU_ERF) {
#pragma unroll
      for (int i = 0; i < CH; ++i) w[i] = gelu_erf(w[i]);
    } else if (ep.act == GD_ACT_SILU) {
#pragma unroll
      for (int i = 0; i < CH; ++i) w[i] = silu(w[i]);
    }
    if (ep.out_t) {
      const int seq = m / ep.t_s, s = m % ep.t_s;
      if (s >= ep.t_skip) {
        float* o = ep.out_t + ((size_t)seq * N + n0) * ep.t_t + (s - ep.t_skip);
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (i < nvalid) o[(size_t)i * ep.t_t] = w[i];
      }
      continue;
    }
    if (ep.out_f32) {
      float* o = ep.out_f32 + r * ep.ld_out + n0;
      if (nvalid == CH && (CH % 4 == 0) && ((((size_t)o) & 15) == 0)) {
#pragma unroll
        for (int i = 0; i < CH; i += 4)
          *reinterpret_cast<float4*>(o + i) = make_float4(w[i], w[i + 1], w[i + 2], w[i + 3]);
      } else {
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (i < nvalid) o[i] = w[i];
      }
    }
    if (ep.out_h16) {
      uint16_t* o = reinterpret_cast<uint16_t*>(ep.out_h16) + r * ep.ld_out_h16 + n0;
      if (nvalid == CH && (CH % 8 == 0) && ((((size_t)o) & 15) == 0)) {
#pragma unroll
        for (int i = 0; i < CH; i += 8) {
          uint4 q;
          q.x = pack_h16x2(w[i], w[i + 1], ep.h16_fmt);
          q.y = pack_h16x2(w[i + 2], w[i + 3], ep.h16_fmt);
          q.z = pack_h16x2(w[i + 4], w[i + 5], ep.h16_fmt);
          q.w = pack_h16x2(w[i + 6], w[i + 7], ep.h16_fmt);
          *reinterpret_cast<uint4*>(o + i) = q;
        }
      } else {
#pragma unroll
        for (int i = 0; i < CH; ++i)
          if (i < nvalid) o[i] = cvt_h16(w[i], ep.h16_fmt);
      }
    }
  }
}

int validate_epilogue(const gdEpilogue* ep, int M, int N);

}  // namespace gd
