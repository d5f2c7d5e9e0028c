// layernorm.cu -- row LayerNorm (norm1/norm2 of nn.TransformerEncoderLayer, mdm.py:90-96), eps 1e-5.
// One warp per row, the row held in registers (d <= 1024, d % 128 == 0): 128-bit loads, two-pass
// mean / variance (sum of squared deviations -- matches the reference's numerics far inside 1e-5),
// shuffle reductions, fp32 and/or 16-bit (bf16/fp16) outputs.  HBM-bound: reads 4*d B/row, writes 4*d + 2*d B/row.
#include "common.cuh"

namespace gd {

template <int VEC>   // VEC float4 per lane: d = VEC * 128
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ in, const float* __restrict__ gamma, const float* __restrict__ beta,
                 float* __restrict__ out_f32, uint16_t* __restrict__ out_h16, int h16_fmt, int rows, float eps) {
  constexpr int D = VEC * 128;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float4* src = reinterpret_cast<const float4*>(in + (size_t)row * D);
  float4 x[VEC];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    x[i] = __ldcs(src + lane + 32 * i);
    s += (x[i].x + x[i].y) + (x[i].z + x[i].w);
  }
  const float mean = warp_sum(s) * (1.0f / D);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    x[i].x -= mean; x[i].y -= mean; x[i].z -= mean; x[i].w -= mean;
    q += (x[i].x * x[i].x + x[i].y * x[i].y) + (x[i].z * x[i].z + x[i].w * x[i].w);
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / D) + eps);
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * i);
    float4 y;
    y.x = x[i].x * rstd * g.x + b.x;
    y.y = x[i].y * rstd * g.y + b.y;
    y.z = x[i].z * rstd * g.z + b.z;
    y.w = x[i].w * rstd * g.w + b.w;
    if (out_f32) reinterpret_cast<float4*>(out_f32 + (size_t)row * D)[lane + 32 * i] = y;
    if (out_h16) {
      uint2 p;
      p.x = pack_h16x2(y.x, y.y, h16_fmt);
      p.y = pack_h16x2(y.z, y.w, h16_fmt);
      reinterpret_cast<uint2*>(out_h16 + (size_t)row * D)[lane + 32 * i] = p;
    }
  }
}

}  // namespace gd

extern "C" int gd_layernorm(const float* in, const float* gamma, const float* beta, float* out_f32, void* out_h16,
                            int32_t h16_fmt, int32_t rows, int32_t d, float eps, void* stream) {
  using namespace gd;
  GD_REQUIRE(in && gamma && beta && (out_f32 || out_h16), "gd_layernorm: null pointer");
  GD_REQUIRE(!out_h16 || h16_fmt == GD_BF16 || h16_fmt == GD_FP16, "gd_layernorm: bad h16_fmt");
  GD_REQUIRE(rows > 0 && d % 128 == 0 && d >= 128 && d <= 1024, "gd_layernorm: d=%d must be a multiple of 128 in [128,1024]", d);
  const int warps = 8;
  dim3 grid(ceil_div(rows, warps));
  cudaStream_t st = (cudaStream_t)stream;
  uint16_t* ob = reinterpret_cast<uint16_t*>(out_h16);
  switch (d / 128) {
#define GD_LN_CASE(V) case V: layernorm_kernel<V><<<grid, warps * 32, 0, st>>>(in, gamma, beta, out_f32, ob, h16_fmt, rows, eps); break;
    GD_LN_CASE(1) GD_LN_CASE(2) GD_LN_CASE(3) GD_LN_CASE(4) GD_LN_CASE(5) GD_LN_CASE(6) GD_LN_CASE(7) GD_LN_CASE(8)
#undef GD_LN_CASE
  }
  GD_LAUNCH_CHECK();
  return GD_OK;
}
