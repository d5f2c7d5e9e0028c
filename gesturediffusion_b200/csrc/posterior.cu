// posterior.cu -- the fused classifier-free-guidance combine + q_posterior mean + noise-add update
// (cfg_sampler.py:28; gaussian_diffusion.py:253-275, 307-311, 354-376, 532-547), plus the small
// sampler-side helpers (q_sample, Philox fill, InputProcess permute, loop state).
//
// HBM-bound elementwise kernel over [B, J, T] fp32.  One CTA owns a slab of 32 consecutive j rows of
// one sample (contiguous 32*T floats): 128-bit coalesced loads of cond/uncond/x_t/(noise), 128-bit
// stores of x_{t-1}; the slab is kept in shared memory so the NEXT step's GEMM operand
// a_next[(b*T+t), j] (the permute of InputProcess, mdm.py:350-352) is written from the same pass.
// Algorithmic bytes: 4 reads + 1 write = 20 B/element with injected noise, 16 B/element with the
// in-kernel Philox4x32-10 generator (+2 B/element for the bf16 operand copy).
#include "common.cuh"

namespace gd {

// ------------------------------------------------------------------ Philox4x32-10 + Box-Muller
struct Philox4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ Philox4 philox4x32_10(Philox4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c.x;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c.z;
    Philox4 n;
    n.x = (uint32_t)(p1 >> 32) ^ c.y ^ k0;
    n.y = (uint32_t)p1;
    n.z = (uint32_t)(p0 >> 32) ^ c.w ^ k1;
    n.w = (uint32_t)p0;
    c = n;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

// Four N(0,1) values for element group g (= flat index within the sample >> 2) of `sample` at `step`.
__device__ __forceinline__ float4 philox_normal4(uint32_t g, uint32_t step, uint64_t sample, uint64_t seed) {
  Philox4 c{g, step, (uint32_t)sample, (uint32_t)(sample >> 32)};
  const Philox4 r = philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  const float s = 2.3283064365386963e-10f;   // 2^-32
  const float u1 = ((float)r.x + 0.5f) * s, u2 = ((float)r.y + 0.5f) * s;
  const float u3 = ((float)r.z + 0.5f) * s, u4 = ((float)r.w + 0.5f) * s;
  // Box-Muller with the hardware approximations (log2 / sin / cos on the special-function unit, ~2^-21 absolute
  // error): this is a noise source, not a parity path -- parity runs inject their noise.
  const float ta = -2.0f * __logf(fminf(u1, 0.99999994f)), tb = -2.0f * __logf(fminf(u3, 0.99999994f));
  const float ra = ta * rsqrtf(ta), rb = tb * rsqrtf(tb);       // ta, tb >= 1.2e-7 > 0
  float sa, ca, sb, cb;
  __sincosf(6.283185307179586f * u2, &sa, &ca);
  __sincosf(6.283185307179586f * u4, &sb, &cb);
  return make_float4(ra * ca, ra * sa, rb * cb, rb * sb);
}

constexpr int PS_ROWS = 32;       // j rows per CTA
constexpr int PS_THREADS = 256;

__device__ __forceinline__ float post_one(float c, float u, float sc, bool has_u, float xt, float nz, float coef1,
                                          float coef2, float sigma, float c_recip, float c_recipm1, int eps_mode,
                                          int ddim, int clip, bool has_mask, bool msk, float inp, float* x0_out,
                                          float* mean_out) {
  float out = has_u ? (u + sc * (c - u)) : c;
  if (has_mask) out = msk ? inp : out;
  float x0 = eps_mode ? (c_recip * xt - c_recipm1 * out) : out;
  if (clip) x0 = fminf(fmaxf(x0, -1.0f), 1.0f);
  float mean;
  if (ddim) {
    // ddim_sample (gaussian_diffusion.py:761-782): eps re-derived from pred_xstart (:763), eq. 12 with
    // coef1 = sqrt(alpha_bar_prev), coef2 = sqrt(1 - alpha_bar_prev - sigma^2)
    const float eps = (c_recip * xt - x0) / c_recipm1;
    mean = x0 * coef1 + coef2 * eps;
  } else {
    mean = coef1 * x0 + coef2 * xt;
  }
  *x0_out = x0;
  *mean_out = mean;
  return mean + sigma * nz;
}

__global__ void __launch_bounds__(PS_THREADS, 5)
posterior_step_kernel(const gdPosteriorArgs a, uint32_t* block_counter) {
  extern __shared__ float tile[];                 // [PS_ROWS][T + pad] when a_next is requested
  const int b = blockIdx.y;
  const int j0 = blockIdx.x * PS_ROWS;
  const int rows = min(PS_ROWS, a.J - j0);
  const int T = a.T;
  const int tstride = T + 1 + (T & 1);            // odd stride -> conflict-free transposed reads
  const int step = *a.step_ptr;
  const float coef1 = a.coef[0 * a.n_steps + step];
  const float coef2 = a.coef[1 * a.n_steps + step];
  // row 2: log-variance of the ancestral step, or (ddim) sigma itself
  const float sigma = (step != 0) ? (a.ddim ? a.coef[2 * a.n_steps + step] : expf(0.5f * a.coef[2 * a.n_steps + step]))
                                  : 0.0f;
  const float c_recip = a.coef[3 * a.n_steps + step];
  const float c_recipm1 = a.coef[4 * a.n_steps + step];
  const float sc = a.uncond ? a.scale[b] : 1.0f;
  const bool has_u = a.uncond != nullptr;
  const bool has_mask = a.inpaint_mask != nullptr && a.inpainted != nullptr;
  const size_t sample_off = (size_t)b * a.J * T;
  const size_t slab_off = sample_off + (size_t)j0 * T;
  const size_t noise_off = (a.const_noise ? 0 : sample_off) + (size_t)j0 * T;
  const uint64_t rng_sample = a.const_noise ? (uint64_t)a.sample_base : (uint64_t)(a.sample_base + b);
  const uint64_t rng_seed = a.seed_dev ? *a.seed_dev : a.seed;
  const int n = rows * T;                          // floats in this slab
  // float4 / uchar4 access needs the element offsets AND every base pointer 16-byte (mask: 4-byte) aligned: the
  // guidance path passes uncond = OUT + B*J*T, which is only 4-byte aligned when B*J*T % 4 != 0 (odd T, B = 1 ...)
  const uintptr_t bases = (uintptr_t)a.cond | (uintptr_t)a.uncond | (uintptr_t)a.x_t | (uintptr_t)a.noise |
                          (uintptr_t)a.inpainted | (uintptr_t)a.x_prev | (uintptr_t)a.pred_xstart | (uintptr_t)a.mean;
  const bool vec_ok = ((slab_off & 3) == 0) && ((noise_off & 3) == 0) && ((((size_t)j0 * T) & 3) == 0) &&
                      ((bases & 15) == 0) && (((uintptr_t)a.inpaint_mask & 3) == 0);
  const int n4 = vec_ok ? (n >> 2) : 0;

  for (int i = threadIdx.x; i < n4; i += PS_THREADS) {
    const size_t e = slab_off + 4 * (size_t)i;
    const float4 c = __ldcs(reinterpret_cast<const float4*>(a.cond + e));
    const float4 u = has_u ? __ldcs(reinterpret_cast<const float4*>(a.uncond + e)) : c;
    const float4 x = __ldcs(reinterpret_cast<const float4*>(a.x_t + e));
    float4 z;
    if (a.noise) z = __ldcs(reinterpret_cast<const float4*>(a.noise + noise_off + 4 * (size_t)i));
    else z = philox_normal4((uint32_t)(((size_t)j0 * T >> 2) + i), (uint32_t)step, rng_sample, rng_seed);
    float4 inp = make_float4(0, 0, 0, 0);
    uchar4 mk = make_uchar4(0, 0, 0, 0);
    if (has_mask) {
      inp = *reinterpret_cast<const float4*>(a.inpainted + e);
      mk = *reinterpret_cast<const uchar4*>(a.inpaint_mask + e);
    }
    float4 x0, mean, o;
    o.x = post_one(c.x, u.x, sc, has_u, x.x, z.x, coef1, coef2, sigma, c_recip, c_recipm1, a.eps_mode, a.ddim, a.clip, has_mask, mk.x != 0, inp.x, &x0.x, &mean.x);
    o.y = post_one(c.y, u.y, sc, has_u, x.y, z.y, coef1, coef2, sigma, c_recip, c_recipm1, a.eps_mode, a.ddim, a.clip, has_mask, mk.y != 0, inp.y, &x0.y, &mean.y);
    o.z = post_one(c.z, u.z, sc, has_u, x.z, z.z, coef1, coef2, sigma, c_recip, c_recipm1, a.eps_mode, a.ddim, a.clip, has_mask, mk.z != 0, inp.z, &x0.z, &mean.z);
    o.w = post_one(c.w, u.w, sc, has_u, x.w, z.w, coef1, coef2, sigma, c_recip, c_recipm1, a.eps_mode, a.ddim, a.clip, has_mask, mk.w != 0, inp.w, &x0.w, &mean.w);
    if (a.x_prev) *reinterpret_cast<float4*>(a.x_prev + e) = o;
    if (a.pred_xstart) *reinterpret_cast<float4*>(a.pred_xstart + e) = x0;
    if (a.mean) *reinterpret_cast<float4*>(a.mean + e) = mean;
    if (a.a_next) {
      const int f = 4 * i;
      const float ov[4] = {o.x, o.y, o.z, o.w};
      int r = f / T, t = f - r * T;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        tile[r * tstride + t] = ov[q];
        if (++t == T) { t = 0; ++r; }
      }
    }
  }
  // scalar tail / unaligned fallback
  for (int f = 4 * n4 + threadIdx.x; f < n; f += PS_THREADS) {
    const size_t e = slab_off + f;
    const float c = a.cond[e];
    const float u = has_u ? a.uncond[e] : c;
    float z;
    if (a.noise) z = a.noise[noise_off + f];
    else {
      const size_t flat = (size_t)j0 * T + f;
      const float4 z4 = philox_normal4((uint32_t)(flat >> 2), (uint32_t)step, rng_sample, rng_seed);
      const int l = (int)(flat & 3);
      z = l == 0 ? z4.x : (l == 1 ? z4.y : (l == 2 ? z4.z : z4.w));
    }
    float x0, mean;
    const float o = post_one(c, u, sc, has_u, a.x_t[e], z, coef1, coef2, sigma, c_recip, c_recipm1, a.eps_mode,
                             a.ddim, a.clip, has_mask, has_mask ? a.inpaint_mask[e] != 0 : false,
                             has_mask ? a.inpainted[e] : 0.f, &x0, &mean);
    if (a.x_prev) a.x_prev[e] = o;
    if (a.pred_xstart) a.pred_xstart[e] = x0;
    if (a.mean) a.mean[e] = mean;
    if (a.a_next) {
      const int r = f / T, t = f - r * T;
      tile[r * tstride + t] = o;
    }
  }
  if (a.a_next) {
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane < rows) {
      for (int t = w; t < T; t += PS_THREADS / 32) {
        const float v = tile[lane * tstride + t];
        const size_t o = ((size_t)b * T + t) * a.ld_a + j0 + lane;
        if (a.a_fmt == GD_F32) reinterpret_cast<float*>(a.a_next)[o] = v;
        else reinterpret_cast<uint16_t*>(a.a_next)[o] = cvt_h16(v, a.a_fmt);
      }
    }
  }
  // Device-side loop state: the last CTA to finish advances step and the model timesteps
  // (respace.py:124-129 map), so a captured step graph can be replayed without host involvement.
  if (a.next_t != nullptr && block_counter != nullptr) {
    // No fence: the last CTA consumes nothing other CTAs wrote; it only has to overwrite *step_ptr after every CTA
    // has read it, and each CTA's reads happen before its own barrier below (kernel-boundary ordering does the rest).
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t total = gridDim.x * gridDim.y;
      last = (atomicAdd(block_counter, 1u) == total - 1);
    }
    __syncthreads();
    if (last) {
      const int ns = step > 0 ? step - 1 : 0;
      const long long tv = a.timestep_map ? a.timestep_map[ns] : (long long)ns;
      for (int i = threadIdx.x; i < a.B; i += PS_THREADS) a.next_t[i] = tv;
      if (threadIdx.x == 0) {
        *a.step_ptr = ns;
        *block_counter = 0;
      }
    }
  }
}

static uint32_t* g_block_counter_dev[gd::kMaxDevices] = {};        // one counter per device

}  // namespace gd

extern "C" int gd_posterior_step(const gdPosteriorArgs* args, void* stream) {
  using namespace gd;
  GD_REQUIRE(args, "gd_posterior_step: null args");
  const gdPosteriorArgs& a = *args;
  GD_REQUIRE(a.cond && a.x_t && a.coef && a.step_ptr, "gd_posterior_step: cond/x_t/coef/step_ptr required");
  GD_REQUIRE(a.B > 0 && a.J > 0 && a.T > 0 && a.n_steps > 0, "gd_posterior_step: bad shape");
  GD_REQUIRE(!a.uncond || a.scale, "gd_posterior_step: scale required with uncond");
  GD_REQUIRE(a.x_prev || a.pred_xstart || a.mean, "gd_posterior_step: no output");
  GD_REQUIRE(!a.a_next || (a.ld_a >= a.J && a.a_fmt >= GD_F32 && a.a_fmt <= GD_FP16), "gd_posterior_step: ld_a < J or bad a_fmt");
  GD_REQUIRE(a.B <= 65535, "gd_posterior_step: B too large for grid.y");
  uint32_t*& g_block_counter = g_block_counter_dev[current_device()];
  if (a.next_t && !g_block_counter) {
    GD_CUDA_OK(cudaMalloc(&g_block_counter, sizeof(uint32_t)));     // one-off per device, before any capture
    GD_CUDA_OK(cudaMemset(g_block_counter, 0, sizeof(uint32_t)));
  }
  const int tstride = a.T + 1 + (a.T & 1);
  const size_t smem = a.a_next ? (size_t)PS_ROWS * tstride * sizeof(float) : 0;
  GD_REQUIRE(smem <= 200 * 1024, "gd_posterior_step: T=%d too long for the transposing slab", a.T);
  static size_t smem_set_dev[kMaxDevices] = {};
  size_t& smem_set = smem_set_dev[current_device()];
  if (smem > 48 * 1024 && smem > smem_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(posterior_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  dim3 grid(ceil_div(a.J, PS_ROWS), a.B);
  posterior_step_kernel<<<grid, PS_THREADS, smem, (cudaStream_t)stream>>>(a, a.next_t ? g_block_counter : nullptr);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

// ------------------------------------------------------------------ helpers
namespace gd {

__global__ void philox_fill_kernel(float* out, int J, int T, uint64_t seed, long long sample_base, int step) {
  const int b = blockIdx.y;
  const size_t per = (size_t)J * T;
  const size_t groups = (per + 3) >> 2;
  for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < groups; g += (size_t)gridDim.x * blockDim.x) {
    const float4 z = philox_normal4((uint32_t)g, (uint32_t)step, (uint64_t)(sample_base + b), seed);
    const float zv[4] = {z.x, z.y, z.z, z.w};
    for (int q = 0; q < 4; ++q)
      if (4 * g + q < per) out[(size_t)b * per + 4 * g + q] = zv[q];
  }
}

__global__ void q_sample_kernel(const float* x0, const float* noise, const float* ta, const float* tb, int step,
                                float* out, long long n) {
  const float ca = ta[step], cb = tb[step];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = ca * x0[i] + cb * noise[i];
}

// src [B, C, T] fp32 -> dst[(b*T+t)*ld + col_off + c]; 32x32 smem transpose per (b, c-tile, t-tile).
__global__ void __launch_bounds__(256)
pack_rows_kernel(const float* __restrict__ src, void* __restrict__ dst, int C, int T, int ld, int col_off, int fmt) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 32, t0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int c = c0 + r, t = t0 + tx;
    tile[r][tx] = (c < C && t < T) ? src[((size_t)b * C + c) * T + t] : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int t = t0 + r, c = c0 + tx;
    if (t < T && c < C) {
      const size_t o = ((size_t)b * T + t) * ld + col_off + c;
      if (fmt == GD_F32) reinterpret_cast<float*>(dst)[o] = tile[tx][r];
      else reinterpret_cast<uint16_t*>(dst)[o] = cvt_h16(tile[tx][r], fmt);
    }
  }
}

__global__ void set_loop_state_kernel(int* step_ptr, long long* t, const long long* map, int step, int B) {
  const long long tv = map ? map[step] : (long long)step;
  for (int i = threadIdx.x; i < B; i += blockDim.x) t[i] = tv;
  if (threadIdx.x == 0) *step_ptr = step;
}

}  // namespace gd

extern "C" int gd_philox_normal(float* out, int32_t B, int32_t J, int32_t T, uint64_t seed, int64_t sample_base,
                                int32_t step, void* stream) {
  using namespace gd;
  GD_REQUIRE(out && B > 0 && J > 0 && T > 0 && B <= 65535, "gd_philox_normal: bad args");
  const size_t groups = ((size_t)J * T + 3) >> 2;
  dim3 grid((unsigned)((groups + 255) / 256 > 1024 ? 1024 : (groups + 255) / 256), B);
  philox_fill_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(out, J, T, seed, sample_base, step);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

extern "C" int gd_q_sample(const float* x0, const float* noise, const float* sqrt_acp, const float* sqrt_1macp,
                           int32_t step, float* out, int64_t n, void* stream) {
  using namespace gd;
  GD_REQUIRE(x0 && noise && sqrt_acp && sqrt_1macp && out && n > 0 && step >= 0, "gd_q_sample: bad args");
  const int blocks = (int)((n + 255) / 256 > 148 * 8 ? 148 * 8 : (n + 255) / 256);
  q_sample_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x0, noise, sqrt_acp, sqrt_1macp, step, out, n);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

extern "C" int gd_pack_rows(const float* src, void* dst, int32_t B, int32_t C, int32_t T, int32_t ld, int32_t col_off,
                            int32_t dst_fmt, void* stream) {
  using namespace gd;
  GD_REQUIRE(src && dst && B > 0 && C > 0 && T > 0 && ld >= col_off + C && B <= 65535, "gd_pack_rows: bad args");
  dim3 grid(ceil_div(T, 32), ceil_div(C, 32), B);
  GD_REQUIRE(dst_fmt >= GD_F32 && dst_fmt <= GD_FP16, "gd_pack_rows: bad dst_fmt");
  pack_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, C, T, ld, col_off, dst_fmt);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

namespace gd {
// One thread per (feature row, frame): out = x * std + mean, routed to the rotation (feature % 6 < 3) or position
// tensor of the take at frame offset t_off.
__global__ void unnormalize_split_kernel(const float* __restrict__ x, const float* __restrict__ mean,
                                         const float* __restrict__ stdv, float* __restrict__ pos,
                                         float* __restrict__ rot, int J, int T, int t_off, int T_total) {
  const int f = blockIdx.y, b = blockIdx.z;
  const int joint = f / 6, c = f - joint * 6;
  const float m = mean ? mean[f] : 0.f, sd = stdv ? stdv[f] : 1.f;
  float* dst = (c < 3 ? rot : pos);
  if (!dst) return;
  const size_t orow = ((size_t)b * (J / 6) + joint) * 3 + (c < 3 ? c : c - 3);
  const float* src = x + ((size_t)b * J + f) * T;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < T; t += gridDim.x * blockDim.x)
    dst[orow * T_total + t_off + t] = __fadd_rn(__fmul_rn(src[t], sd), m);   // two roundings, like data * std + mean
}
}  // namespace gd

extern "C" int gd_unnormalize_split(const float* x, const float* mean, const float* stdv, float* pos, float* rot,
                                    int32_t B, int32_t J, int32_t T, int32_t t_off, int32_t T_total, void* stream) {
  using namespace gd;
  GD_REQUIRE(x && (pos || rot), "gd_unnormalize_split: null pointer");
  GD_REQUIRE(B > 0 && B <= 65535 && J > 0 && J <= 65535 && J % 6 == 0 && T > 0 && t_off >= 0 && t_off + T <= T_total,
             "gd_unnormalize_split: bad shape B=%d J=%d T=%d t_off=%d T_total=%d (J must be a multiple of 6)", B, J, T,
             t_off, T_total);
  dim3 grid(ceil_div(T, 128), J, B);
  unnormalize_split_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(x, mean, stdv, pos, rot, J, T, t_off, T_total);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

extern "C" int gd_set_loop_state(int32_t* step_ptr, int64_t* t, const int64_t* timestep_map, int32_t step, int32_t B,
                                 void* stream) {
  using namespace gd;
  GD_REQUIRE(step_ptr && t && B > 0 && step >= 0, "gd_set_loop_state: bad args");
  set_loop_state_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(step_ptr, (long long*)t, (const long long*)timestep_map,
                                                            step, B);
  GD_LAUNCH_CHECK();
  return GD_OK;
}
