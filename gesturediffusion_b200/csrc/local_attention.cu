// local_attention.cu -- RoPE -> causal windowed LocalAttention -> RoPE of the V2 denoiser
// (mdm.py:176-213; local_attention.py:43-62 SinusoidalEmbeddings/rotate_half/apply_rotary_pos_emb,
//  :92-172 LocalAttention.forward with window 10, look_backward 1, look_forward 0, q = k = v),
// and the per-step conditioning token (mdm.py:130,154-160,197; mdm_old.py:94-99,109-112).
//
// One CTA per (sequence, head): the [T, e] head slab is read once from HBM (fp32), rotated with the
// cos/sin tables and kept in shared memory (row stride e+1 -> conflict-free row-parallel reads);
// each warp then owns query rows: lanes = keys (<= 2*window) for the scores, shuffle softmax,
// lanes = channels for the PV sum, second rotation at position t+1, coalesced fp32 + 16-bit stores
// straight into rows 1..T of the encoder input.  HBM-bound: 4 B/elem in, 4 (+2) B/elem out.
// Semantics kept from the reference: the query at in-window offset i of window w attends to all of
// window w-1 and to offsets 0..i of window w; window 0 has no look-back (its padding keys are masked,
// contributing exactly 0); scale = e^-0.5; V is the rotated tensor as well.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"

namespace gd {

constexpr int LA_THREADS = 256;

__global__ void __launch_bounds__(LA_THREADS)
local_attention_kernel(const float* __restrict__ z, const float* __restrict__ rcos, const float* __restrict__ rsin,
                       float* __restrict__ x_f32, uint16_t* __restrict__ x_h16, int h16_fmt, int T, int d,
                       int e, int window, float scale) {
  extern __shared__ float sm[];
  const int es = e + 1;
  float* slab = sm;                                   // [T][e+1] rotated q=k=v
  float* scratch = sm + (size_t)T * es;               // [warps][e]
  const int head = blockIdx.x, seq = blockIdx.y;
  const int half = e >> 1;
  const float* zs = z + (size_t)seq * T * d + head * e;
  for (int idx = threadIdx.x; idx < T * e; idx += LA_THREADS) {
    const int t = idx / e, i = idx - t * e;
    const float x = zs[(size_t)t * d + i];
    const float p = (i < half) ? -zs[(size_t)t * d + i + half] : zs[(size_t)t * d + i - half];
    slab[t * es + i] = x * rcos[t * e + i] + p * rsin[t * e + i];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* my = scratch + warp * e;
  const int S = T + 1;
  for (int t = warp; t < T; t += LA_THREADS / 32) {
    const int w = t / window;
    const int kbeg = (w > 0 ? (w - 1) * window : 0);
    const int nk = t - kbeg + 1;                       // <= 2*window
    const float* q = slab + t * es;
    float out_acc[4] = {0.f, 0.f, 0.f, 0.f};           // channels lane, lane+32, ... (e <= 128)
    float run_max = -INFINITY, run_sum = 0.f;
    for (int kb = 0; kb < nk; kb += 32) {              // normally a single pass (nk <= 20)
      const int k = kbeg + kb + lane;
      float sc = -INFINITY;
      if (kb + lane < nk) {
        const float* kr = slab + k * es;
        float acc = 0.f;
        for (int i = 0; i < e; ++i) acc = fmaf(q[i], kr[i], acc);
        sc = acc * scale;
      }
      const float bm = warp_max(sc);
      const float nm = fmaxf(run_max, bm);
      const float corr = expf(run_max - nm);           // 0 on the first pass
      const float p = (kb + lane < nk) ? expf(sc - nm) : 0.f;
      run_sum = run_sum * corr + warp_sum(p);
      run_max = nm;
#pragma unroll
      for (int c = 0; c < 4; ++c) out_acc[c] *= corr;
      const int cnt = min(32, nk - kb);
      for (int j = 0; j < cnt; ++j) {
        const float pj = __shfl_sync(0xffffffffu, p, j);
        const float* vr = slab + (kbeg + kb + j) * es;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int ch = lane + 32 * c;
          if (ch < e) out_acc[c] = fmaf(pj, vr[ch], out_acc[c]);
        }
      }
    }
    const float inv = 1.0f / run_sum;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int ch = lane + 32 * c;
      if (ch < e) my[ch] = out_acc[c] * inv;
    }
    __syncwarp();
    // second rotation at sequence position t+1 (token 0 is the conditioning token)
    const size_t orow = ((size_t)seq * S + 1 + t) * d + head * e;
    const float* c2 = rcos + (size_t)(t + 1) * e;
    const float* s2 = rsin + (size_t)(t + 1) * e;
    for (int ch = lane; ch < e; ch += 32) {
      const float x = my[ch];
      const float p = (ch < half) ? -my[ch + half] : my[ch - half];
      const float y = x * c2[ch] + p * s2[ch];
      if (x_f32) x_f32[orow + ch] = y;
      if (x_h16) x_h16[orow + ch] = cvt_h16(y, h16_fmt);
    }
    __syncwarp();
  }
}

// Fast path for the reference's configuration (head dim 64, window 10) on the tensor cores, one warp per window.  The rotated slab is kept as a
// 16-bit PAIR (hi = fp16(x), lo = fp16(x - hi): x to ~2^-22 relative), and every product runs as three
// mma.sync.m16n8k16 passes (hi*hi + hi*lo + lo*hi, fp32 accumulation), so the result keeps fp32-level accuracy while
// the 10 x 20 x 64 score and 10 x 64 x 20 output products of a window cost 36 + 48 tensor instructions instead of
// ~1300 FMA instructions (a SIMT one-warp-per-window kernel, removed, was issue-bound: 87 us for the cfg3 shape against
// 24 us of HBM time; this one takes 69 us, bound by the legacy HMMA issue rate).
//   scores  S[16 x 24] = Q[16 x 64] K[24 x 64]^T   (rows 10..15 and keys 20..23 are padding, masked)
//   softmax on the accumulator fragments (a row lives in the four lanes of a quad)
//   output  O[16 x 64] = P[16 x 32] V[32 x 64]     (P straight from the score fragments, V through ldmatrix.trans)
//   second RoPE at position t + 1: channels c and c + 32 sit in accumulator tiles nt and nt + 4 of the same thread.
constexpr int LM_THREADS = 192, LM_E = 64, LM_W = 10, LM_RS = LM_E + 8, LM_PAD = 12;   // row stride 72 halves = 144 B

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                          uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
               "{%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// (hi, lo) fp16 pair of two floats, packed for an A / B fragment register
__device__ __forceinline__ void split_h2(float x, float y, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(x, y);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn(x - hf.x, y - hf.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}

__global__ void __launch_bounds__(LM_THREADS)
local_attention_mma_kernel(const float* __restrict__ z, const float* __restrict__ rcos, const float* __restrict__ rsin,
                           float* __restrict__ x_f32, uint16_t* __restrict__ x_h16, int h16_fmt, int T, int d,
                           float scale_log2e) {
  extern __shared__ __align__(16) uint8_t smraw[];
  __half* shi = reinterpret_cast<__half*>(smraw);          // [T + LM_PAD][LM_RS] rotated q = k = v, hi halves
  __half* slo = shi + (size_t)(T + LM_PAD) * LM_RS;        // ... lo halves
  const int head = blockIdx.x, seq = blockIdx.y;
  constexpr int E = LM_E, HALF = E / 2;
  const float* zs = z + (size_t)seq * T * d + head * E;
  // Load + first rotation: 16 threads per row (4 channels each); the rotation partner of channel i is i +- 32, held by
  // lane ^ 8 of the same warp.  Four loads per thread are issued before the first is used (the phase is latency-bound: one
  // CTA reads 30 KB with 192 threads; batches of twelve, or the tables fetched after the batch, measured slower: 73.8 / 76.5 vs 68.9 us).
  constexpr int kLoadBatch = 4;
  const int n_items = (T + LM_PAD) * (E / 4);
  for (int base = 0; base < n_items; base += kLoadBatch * LM_THREADS) {
    float4 x[kLoadBatch], c[kLoadBatch], sn[kLoadBatch];
#pragma unroll
    for (int u = 0; u < kLoadBatch; ++u) {
      const int idx = base + u * LM_THREADS + (int)threadIdx.x;
      const int t = idx / (E / 4), i = (idx - t * (E / 4)) * 4;
      x[u] = make_float4(0.f, 0.f, 0.f, 0.f);              // padding rows: finite operands under zero probabilities
      c[u] = x[u];
      sn[u] = x[u];
      if (idx < n_items && t < T) {
        x[u] = *reinterpret_cast<const float4*>(zs + (size_t)t * d + i);
        c[u] = __ldg(reinterpret_cast<const float4*>(rcos + t * E + i));
        sn[u] = __ldg(reinterpret_cast<const float4*>(rsin + t * E + i));
      }
    }
#pragma unroll
    for (int u = 0; u < kLoadBatch; ++u) {
      const int idx = base + u * LM_THREADS + (int)threadIdx.x;
      const int t = idx / (E / 4), i = (idx - t * (E / 4)) * 4;
      float4 pr;                                           // whole warps run this (LM_THREADS and n_items are multiples of 32)
      pr.x = __shfl_xor_sync(0xffffffffu, x[u].x, 8);
      pr.y = __shfl_xor_sync(0xffffffffu, x[u].y, 8);
      pr.z = __shfl_xor_sync(0xffffffffu, x[u].z, 8);
      pr.w = __shfl_xor_sync(0xffffffffu, x[u].w, 8);
      if (idx >= n_items) continue;
      const float sg = (i < HALF) ? -1.f : 1.f;
      float4 o;
      o.x = x[u].x * c[u].x + sg * pr.x * sn[u].x;
      o.y = x[u].y * c[u].y + sg * pr.y * sn[u].y;
      o.z = x[u].z * c[u].z + sg * pr.z * sn[u].z;
      o.w = x[u].w * c[u].w + sg * pr.w * sn[u].w;
      uint2 h, l;
      split_h2(o.x, o.y, h.x, l.x);
      split_h2(o.z, o.w, h.y, l.y);
      *reinterpret_cast<uint2*>(shi + t * LM_RS + i) = h;
      *reinterpret_cast<uint2*>(slo + t * LM_RS + i) = l;
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  const uint32_t shi_s = (uint32_t)__cvta_generic_to_shared(shi), slo_s = (uint32_t)__cvta_generic_to_shared(slo);
  const int S = T + 1, n_win = T / LM_W;
  for (int w = warp; w < n_win; w += LM_THREADS / 32) {
    const int t0 = w * LM_W, kbeg = t0 - LM_W;
    // ---- scores: three 8-key accumulator tiles (keys kbeg .. kbeg + 23), K = 64 in four steps
    float sc[3][4];
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) sc[j][q] = 0.f;
    {
      // ldmatrix row addresses (byte offsets inside a slab).  Negative key rows (window 0's look-back) are clamped:
      // their probabilities are masked to zero below.
      const uint32_t qoff = (uint32_t)((t0 + (lane & 7) + ((lane >> 3) & 1) * 8) * LM_RS + (lane >> 4) * 8) * 2;
      const int krow01 = kbeg + (lane & 7) + (lane >> 4) * 8;
      const uint32_t koff01 = (uint32_t)((krow01 < 0 ? 0 : krow01) * LM_RS + ((lane >> 3) & 1) * 8) * 2;
      const int krow2 = kbeg + 16 + (lane & 7);
      const uint32_t koff2 = (uint32_t)(krow2 * LM_RS + ((lane >> 3) & 1) * 8) * 2;
#pragma unroll
      for (int ks = 0; ks < E / 16; ++ks) {
        uint32_t ah[4], al[4], bh[6], bl[6];
        ldsm_x4(shi_s + qoff + ks * 32, ah[0], ah[1], ah[2], ah[3]);
        ldsm_x4(slo_s + qoff + ks * 32, al[0], al[1], al[2], al[3]);
        ldsm_x4(shi_s + koff01 + ks * 32, bh[0], bh[1], bh[2], bh[3]);
        ldsm_x4(slo_s + koff01 + ks * 32, bl[0], bl[1], bl[2], bl[3]);
        ldsm_x2(shi_s + koff2 + ks * 32, bh[4], bh[5]);
        ldsm_x2(slo_s + koff2 + ks * 32, bl[4], bl[5]);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          mma_16816(sc[j], ah[0], ah[1], ah[2], ah[3], bh[2 * j], bh[2 * j + 1]);
          mma_16816(sc[j], ah[0], ah[1], ah[2], ah[3], bl[2 * j], bl[2 * j + 1]);
          mma_16816(sc[j], al[0], al[1], al[2], al[3], bh[2 * j], bh[2 * j + 1]);
        }
      }
    }
    // ---- causal window softmax on the fragments: this thread holds queries g and g + 8, keys 8 j + 2 tq + {0, 1}
    uint32_t ph[2][4], pl[2][4];                           // P as the A operand of two 16-key steps (hi, lo)
    float inv[2];
    {
      float p[2][8];                                       // [row half][key slot j * 2 + e], fourth key tile = 0
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int qi = g + 8 * r;
        float m = -INFINITY;
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
          for (int e2 = 0; e2 < 2; ++e2) {
            const int key = 8 * j + 2 * tq + e2;
            const bool ok = qi < LM_W && (key < LM_W ? (w > 0) : (key < 2 * LM_W && key - LM_W <= qi));
            const float v = ok ? sc[j][2 * r + e2] * scale_log2e : -INFINITY;
            p[r][2 * j + e2] = v;
            m = fmaxf(m, v);
          }
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        const float mm = (m == -INFINITY) ? 0.f : m;         // padding rows 10..15: everything masked
        float sum = 0.f;
#pragma unroll
        for (int q = 0; q < 6; ++q) {
          p[r][q] = exp2f(p[r][q] - mm);                     // exp2f(-inf) = 0 for the masked keys
          sum += p[r][q];
        }
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        inv[r] = sum > 0.f ? 1.0f / sum : 0.f;
        p[r][6] = 0.f;
        p[r][7] = 0.f;
      }
      // score tile j -> A registers: step s = keys 16 s .. 16 s + 15: a0 / a1 = tile 2 s (rows g / g + 8), a2 / a3 = tile 2 s + 1
#pragma unroll
      for (int s2 = 0; s2 < 2; ++s2) {
        split_h2(p[0][4 * s2], p[0][4 * s2 + 1], ph[s2][0], pl[s2][0]);
        split_h2(p[1][4 * s2], p[1][4 * s2 + 1], ph[s2][1], pl[s2][1]);
        split_h2(p[0][4 * s2 + 2], p[0][4 * s2 + 3], ph[s2][2], pl[s2][2]);
        split_h2(p[1][4 * s2 + 2], p[1][4 * s2 + 3], ph[s2][3], pl[s2][3]);
      }
    }
    // ---- O = P V: eight 8-channel accumulator tiles, keys in two steps of 16 (rows kbeg .. kbeg + 31 of the slab)
    float o[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int q = 0; q < 4; ++q) o[nt][q] = 0.f;
#pragma unroll
    for (int s2 = 0; s2 < 2; ++s2) {
      const int vrow = kbeg + 16 * s2 + (lane & 7) + ((lane >> 3) & 1) * 8;
      const uint32_t voff = (uint32_t)((vrow < 0 ? 0 : vrow) * LM_RS + (lane >> 4) * 8) * 2;
#pragma unroll
      for (int np = 0; np < 4; ++np) {                       // channel tiles 2 np, 2 np + 1
        uint32_t vh[4], vl[4];
        ldsm_x4_t(shi_s + voff + np * 32, vh[0], vh[1], vh[2], vh[3]);
        ldsm_x4_t(slo_s + voff + np * 32, vl[0], vl[1], vl[2], vl[3]);
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          float (&acc)[4] = o[2 * np + h2];
          mma_16816(acc, ph[s2][0], ph[s2][1], ph[s2][2], ph[s2][3], vh[2 * h2], vh[2 * h2 + 1]);
          mma_16816(acc, ph[s2][0], ph[s2][1], ph[s2][2], ph[s2][3], vl[2 * h2], vl[2 * h2 + 1]);
          mma_16816(acc, pl[s2][0], pl[s2][1], pl[s2][2], pl[s2][3], vh[2 * h2], vh[2 * h2 + 1]);
        }
      }
    }
    // ---- normalise, second rotation at sequence position t + 1 (token 0 is the conditioning token), stores
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int qi = g + 8 * r;
      if (qi >= LM_W) continue;
      const int t = t0 + qi;
      const float* c2 = rcos + (size_t)(t + 1) * E + 2 * tq;
      const float* s2p = rsin + (size_t)(t + 1) * E + 2 * tq;
      const size_t off0 = ((size_t)seq * S + 1 + t) * d + head * E + 2 * tq;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int pn = nt ^ 4;                               // channel c +- 32 lives in tile nt +- 4 of this thread
        const float sg = nt < 4 ? -1.f : 1.f;
        const float2 cc = __ldg(reinterpret_cast<const float2*>(c2 + 8 * nt));
        const float2 ss = __ldg(reinterpret_cast<const float2*>(s2p + 8 * nt));
        const float x0 = o[nt][2 * r] * inv[r], x1 = o[nt][2 * r + 1] * inv[r];
        const float q0 = o[pn][2 * r] * inv[r], q1 = o[pn][2 * r + 1] * inv[r];
        const float y0 = x0 * cc.x + sg * q0 * ss.x;
        const float y1 = x1 * cc.y + sg * q1 * ss.y;
        if (x_f32) *reinterpret_cast<float2*>(x_f32 + off0 + 8 * nt) = make_float2(y0, y1);
        if (x_h16) *reinterpret_cast<uint32_t*>(x_h16 + off0 + 8 * nt) = pack_h16x2(y0, y1, h16_fmt);
      }
    }
  }
}

__global__ void __launch_bounds__(128)
cond_tokens_kernel(const long long* __restrict__ t, const float* __restrict__ temb, const float* __restrict__ seed_emb,
                   const float* __restrict__ pe0, const float* __restrict__ coarse_tab,
                   const float* __restrict__ coarse_seed, float* __restrict__ coarse_out, float* __restrict__ x_f32,
                   uint16_t* __restrict__ x_h16, int h16_fmt, int B, int S, int d) {
  const int b = blockIdx.x, c = blockIdx.y;
  const long long tt = t[b];
  const size_t srow = (size_t)(c * B + b);
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    float v = temb[(size_t)tt * d + i] + seed_emb[srow * d + i];
    if (pe0) v += pe0[i];
    if (x_f32) x_f32[srow * S * d + i] = v;
    if (x_h16) x_h16[srow * S * d + i] = cvt_h16(v, h16_fmt);
    if (coarse_out) coarse_out[srow * d + i] = coarse_tab[(size_t)tt * d + i] + coarse_seed[srow * d + i];
  }
}

}  // namespace gd

extern "C" int gd_local_attention(const float* z, const float* rope_cos, const float* rope_sin, float* x_f32,
                                  void* x_h16, int32_t h16_fmt, int32_t n_seq, int32_t T, int32_t d, int32_t heads,
                                  int32_t window, void* stream) {
  using namespace gd;
  GD_REQUIRE(z && rope_cos && rope_sin && (x_f32 || x_h16), "gd_local_attention: null pointer");
  GD_REQUIRE(!x_h16 || h16_fmt == GD_BF16 || h16_fmt == GD_FP16, "gd_local_attention: bad h16_fmt");
  GD_REQUIRE(n_seq > 0 && n_seq <= 65535 && T > 0 && heads > 0 && d % heads == 0, "gd_local_attention: bad shape");
  GD_REQUIRE(window > 0 && T % window == 0,
             "sequence length must be divisible by window size for local attention (T=%d, window=%d)", T, window);
  const int e = d / heads;
  GD_REQUIRE(e % 2 == 0 && e <= 128, "gd_local_attention: head dim %d must be even and <= 128", e);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const bool fast_ok = e == LM_E && window == LM_W && d % 4 == 0 && al16(z) && al16(rope_cos) && al16(rope_sin) &&
                       (!x_f32 || al16(x_f32)) && (!x_h16 || al16(x_h16));
  const size_t smem_m = (size_t)(T + LM_PAD) * LM_RS * 2 * sizeof(uint16_t);
  if (fast_ok && smem_m <= 200 * 1024) {
    static size_t set_m_dev[kMaxDevices] = {};
    size_t& set_m = set_m_dev[current_device()];
    if (smem_m > 48 * 1024 && smem_m > set_m) {
      GD_CUDA_OK(cudaFuncSetAttribute(local_attention_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem_m));
      set_m = smem_m;
    }
    local_attention_mma_kernel<<<dim3(heads, n_seq), LM_THREADS, smem_m, (cudaStream_t)stream>>>(
        z, rope_cos, rope_sin, x_f32, (uint16_t*)x_h16, h16_fmt, T, d, 1.4426950408889634f / sqrtf((float)e));
    GD_LAUNCH_CHECK();
    return GD_OK;
  }
  const size_t smem = ((size_t)T * (e + 1) + (LA_THREADS / 32) * e) * sizeof(float);
  GD_REQUIRE(smem <= 200 * 1024, "gd_local_attention: T=%d too long for the shared-memory slab", T);
  static size_t set_dev[kMaxDevices] = {};
  size_t& set = set_dev[current_device()];
  if (smem > 48 * 1024 && smem > set) {
    GD_CUDA_OK(cudaFuncSetAttribute(local_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = smem;
  }
  dim3 grid(heads, n_seq);
  local_attention_kernel<<<grid, LA_THREADS, smem, (cudaStream_t)stream>>>(
      z, rope_cos, rope_sin, x_f32, (uint16_t*)x_h16, h16_fmt, T, d, e, window, 1.0f / sqrtf((float)e));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

extern "C" int gd_cond_tokens(const int64_t* t, const float* temb_table, const float* seed_emb, const float* pe0,
                              const float* coarse_tab, const float* coarse_seed, float* coarse_out, float* x_f32,
                              void* x_h16, int32_t h16_fmt, int32_t B, int32_t n_branch, int32_t S, int32_t d,
                              void* stream) {
  using namespace gd;
  GD_REQUIRE(t && temb_table && seed_emb && (x_f32 || x_h16 || coarse_out), "gd_cond_tokens: null pointer");
  GD_REQUIRE(!x_h16 || h16_fmt == GD_BF16 || h16_fmt == GD_FP16, "gd_cond_tokens: bad h16_fmt");
  GD_REQUIRE(B > 0 && n_branch > 0 && n_branch <= 65535 && S > 0 && d > 0, "gd_cond_tokens: bad shape");
  GD_REQUIRE(!coarse_out || (coarse_tab && coarse_seed), "gd_cond_tokens: coarse_out needs coarse_tab and coarse_seed");
  dim3 grid(B, n_branch);
  cond_tokens_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>((const long long*)t, temb_table, seed_emb, pe0, coarse_tab,
                                                            coarse_seed, coarse_out, x_f32, (uint16_t*)x_h16, h16_fmt,
                                                            B, S, d);
  GD_LAUNCH_CHECK();
  return GD_OK;
}
