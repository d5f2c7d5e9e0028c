// attention.cu -- full (unmasked) softmax self-attention of nn.MultiheadAttention inside
// nn.TransformerEncoderLayer (mdm.py:90-96,216): softmax(Q K^T / sqrt(dh)) V per (sequence, head).
// Input is the packed in_proj output qkv[n_seq*S, 3d] (q | k | v); output out[n_seq*S, d].
//
//  * 16-bit path (bf16 or fp16, templated): flash-style single pass, 64 queries per CTA (4 warps x 16 rows), 64-key blocks
//    staged in XOR-swizzled shared memory with cp.async (128-bit), QK^T and PV on mma.sync
//    m16n8k16 bf16 tensor-core tiles with fp32 accumulation, online softmax with quad shuffles.
//    (~5% of the step's flops; S <= a few hundred, so one CTA sees the whole K/V of its head.)
//  * fp32 path: SIMT reference-precision kernel for "fp32 mode" parity runs.
#include <stdlib.h>

#include "common.cuh"

namespace gd {

// ------------------------------------------------------------------ small PTX helpers
__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem, bool valid) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(s));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(s));
}
template <bool FP16>
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (FP16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

constexpr int AT_BQ = 64, AT_BK = 64, AT_THREADS = 128;

// element offset of 16-byte chunk `c` of row `r` in a [rows][DH] bf16 tile, XOR-swizzled
template <int DH>
__device__ __forceinline__ int swz(int r, int c) {
  constexpr int CPR = DH / 8;                       // chunks per row: 4, 8 or 16
  constexpr int SH = CPR >= 8 ? 0 : (CPR == 4 ? 1 : 2);
  constexpr int MASK = CPR >= 8 ? 7 : CPR - 1;
  return r * DH + ((c ^ ((r >> SH) & MASK)) << 3);
}

template <int DH>
__device__ __forceinline__ void load_tile(__nv_bfloat16* s, const __nv_bfloat16* g, int row0, int S, size_t ld) {
  constexpr int CPR = DH / 8;                          // 16B chunks per row
  for (int i = threadIdx.x; i < 64 * CPR; i += AT_THREADS) {
    const int r = i / CPR, c = i % CPR;
    const bool ok = (row0 + r) < S;
    const __nv_bfloat16* src = g + (size_t)(ok ? row0 + r : 0) * ld + c * 8;
    cp_async_16(s + swz<DH>(r, c), src, ok);
  }
}

template <int DH, bool FP16>
__global__ void __launch_bounds__(AT_THREADS)
attention_h16_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int S, int d,
                      float scale_log2e) {
  __shared__ __align__(128) __nv_bfloat16 sQ[AT_BQ * DH];
  __shared__ __align__(128) __nv_bfloat16 sK[AT_BK * DH];
  __shared__ __align__(128) __nv_bfloat16 sV[AT_BK * DH];
  const int seq = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * AT_BQ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t ld = (size_t)3 * d;
  const __nv_bfloat16* gq = qkv + (size_t)seq * S * ld + head * DH;
  const __nv_bfloat16* gk = gq + d;
  const __nv_bfloat16* gv = gq + 2 * d;

  load_tile<DH>(sQ, gq, q0, S, ld);
  cp_async_wait_all();
  __syncthreads();

  uint32_t qf[DH / 16][4];
#pragma unroll
  for (int kk = 0; kk < DH / 16; ++kk)
    ldsm_x4(qf[kk], sQ + swz<DH>(warp * 16 + (lane & 15), kk * 2 + (lane >> 4)));

  float o[DH / 8][4];
#pragma unroll
  for (int i = 0; i < DH / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};

  for (int k0 = 0; k0 < S; k0 += AT_BK) {
    __syncthreads();                                  // previous block's smem reads are done
    load_tile<DH>(sK, gk, k0, S, ld);
    load_tile<DH>(sV, gv, k0, S, ld);
    cp_async_wait_all();
    __syncthreads();

    float s[AT_BK / 8][4];
#pragma unroll
    for (int nt = 0; nt < AT_BK / 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < DH / 16; kk += 2) {
#pragma unroll
      for (int nt = 0; nt < AT_BK / 8; ++nt) {
        uint32_t b[4];
        ldsm_x4(b, sK + swz<DH>(nt * 8 + (lane & 7), kk * 2 + (lane >> 3)));
        mma_16816<FP16>(s[nt], qf[kk], b[0], b[1]);
        mma_16816<FP16>(s[nt], qf[kk + 1], b[2], b[3]);
      }
    }
    // mask keys beyond S, online softmax (rows lane/4 and lane/4+8; columns (lane%4)*2 + {0,1})
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < AT_BK / 8; ++nt) {
      const int key = k0 + nt * 8 + (lane & 3) * 2;
      if (key >= S) s[nt][0] = s[nt][2] = -INFINITY;
      if (key + 1 >= S) s[nt][1] = s[nt][3] = -INFINITY;
      mx[0] = fmaxf(mx[0], fmaxf(s[nt][0], s[nt][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[nt][2], s[nt][3]));
    }
    float corr[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float mnew = fmaxf(mrow[r], mx[r]);        // finite: every block has >= 1 valid key
      corr[r] = exp2f((mrow[r] - mnew) * scale_log2e);
      mrow[r] = mnew;
    }
    float ls[2] = {0.f, 0.f};
#pragma unroll
    for (int nt = 0; nt < AT_BK / 8; ++nt) {
      s[nt][0] = exp2f((s[nt][0] - mrow[0]) * scale_log2e);
      s[nt][1] = exp2f((s[nt][1] - mrow[0]) * scale_log2e);
      s[nt][2] = exp2f((s[nt][2] - mrow[1]) * scale_log2e);
      s[nt][3] = exp2f((s[nt][3] - mrow[1]) * scale_log2e);
      ls[0] += s[nt][0] + s[nt][1];
      ls[1] += s[nt][2] + s[nt][3];
    }
    lrow[0] = lrow[0] * corr[0] + ls[0];
    lrow[1] = lrow[1] * corr[1] + ls[1];
#pragma unroll
    for (int i = 0; i < DH / 8; ++i) {
      o[i][0] *= corr[0]; o[i][1] *= corr[0];
      o[i][2] *= corr[1]; o[i][3] *= corr[1];
    }
#pragma unroll
    for (int kk = 0; kk < AT_BK / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_h16x2(s[2 * kk][0], s[2 * kk][1], FP16 ? GD_FP16 : GD_BF16);
      pa[1] = pack_h16x2(s[2 * kk][2], s[2 * kk][3], FP16 ? GD_FP16 : GD_BF16);
      pa[2] = pack_h16x2(s[2 * kk + 1][0], s[2 * kk + 1][1], FP16 ? GD_FP16 : GD_BF16);
      pa[3] = pack_h16x2(s[2 * kk + 1][2], s[2 * kk + 1][3], FP16 ? GD_FP16 : GD_BF16);
#pragma unroll
      for (int nd = 0; nd < DH / 8; nd += 2) {
        uint32_t b[4];
        ldsm_x4_t(b, sV + swz<DH>(kk * 16 + (lane & 15), nd + (lane >> 4)));
        mma_16816<FP16>(o[nd], pa, b[0], b[1]);
        mma_16816<FP16>(o[nd + 1], pa, b[2], b[3]);
      }
    }
  }
  // finalize: quad-reduce the row sums, normalise, store bf16
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
    lrow[r] = 1.0f / lrow[r];
  }
  const int row_a = q0 + warp * 16 + (lane >> 2), row_b = row_a + 8;
  __nv_bfloat16* go = out + (size_t)seq * S * d + head * DH + (lane & 3) * 2;
#pragma unroll
  for (int nd = 0; nd < DH / 8; ++nd) {
    if (row_a < S) *reinterpret_cast<uint32_t*>(go + (size_t)row_a * d + nd * 8) = pack_h16x2(o[nd][0] * lrow[0], o[nd][1] * lrow[0], FP16 ? GD_FP16 : GD_BF16);
    if (row_b < S) *reinterpret_cast<uint32_t*>(go + (size_t)row_b * d + nd * 8) = pack_h16x2(o[nd][2] * lrow[1], o[nd][3] * lrow[1], FP16 ? GD_FP16 : GD_BF16);
  }
}

// ------------------------------------------------------------------ fp32 SIMT path
// One warp per query row; scores for all S keys kept in shared memory (two-pass softmax).
template <int DH>
__global__ void __launch_bounds__(256)
attention_f32_kernel(const float* __restrict__ qkv, float* __restrict__ out, int S, int d, float scale) {
  extern __shared__ float sm[];
  constexpr int WARPS = 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = blockIdx.z, head = blockIdx.y, row = blockIdx.x * WARPS + warp;
  float* sq = sm + warp * DH;
  float* sp = sm + WARPS * DH + warp * S;
  if (row >= S) return;
  const size_t ld = (size_t)3 * d;
  const float* base = qkv + (size_t)seq * S * ld + head * DH;
  for (int i = lane; i < DH; i += 32) sq[i] = base[(size_t)row * ld + i];
  __syncwarp();
  float mx = -INFINITY;
  for (int k = lane; k < S; k += 32) {
    const float4* kr = reinterpret_cast<const float4*>(base + d + (size_t)k * ld);
    float acc = 0.f;
#pragma unroll 8
    for (int i = 0; i < DH / 4; ++i) {
      const float4 kv = kr[i];
      acc = fmaf(sq[4 * i], kv.x, acc);
      acc = fmaf(sq[4 * i + 1], kv.y, acc);
      acc = fmaf(sq[4 * i + 2], kv.z, acc);
      acc = fmaf(sq[4 * i + 3], kv.w, acc);
    }
    acc *= scale;
    sp[k] = acc;
    mx = fmaxf(mx, acc);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int k = lane; k < S; k += 32) {
    const float p = expf(sp[k] - mx);
    sp[k] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  __syncwarp();
  const float inv = 1.0f / sum;
  for (int c = lane; c < DH; c += 32) {
    float acc = 0.f;
    const float* vc = base + 2 * d + c;
    for (int k = 0; k < S; ++k) acc = fmaf(sp[k], vc[(size_t)k * ld], acc);
    out[((size_t)seq * S + row) * d + head * DH + c] = acc * inv;
  }
}

bool attention_tc_supported(int S, int n_heads, int head_dim);
int launch_attention_tc(const void* qkv, void* out, int n_seq, int S, int n_heads, int fmt, cudaStream_t stream);

static bool attention_legacy_forced() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GD_ATTN_LEGACY");
    v = (e && e[0] == '1') ? 1 : 0;
  }
  return v == 1;
}

}  // namespace gd

extern "C" int gd_attention_h16(const void* qkv, void* out, int32_t n_seq, int32_t S, int32_t n_heads,
                                int32_t head_dim, int32_t fmt, void* stream) {
  using namespace gd;
  GD_REQUIRE(qkv && out && n_seq > 0 && S > 0 && n_heads > 0, "gd_attention_h16: bad args");
  GD_REQUIRE(n_seq <= 65535 && n_heads <= 65535, "gd_attention_h16: grid too large");
  GD_REQUIRE(fmt == GD_BF16 || fmt == GD_FP16, "gd_attention_h16: fmt must be GD_BF16 or GD_FP16");
  const int d = n_heads * head_dim;
  GD_REQUIRE(d % 8 == 0, "gd_attention_h16: d must be a multiple of 8");
  const float sl2 = (1.0f / sqrtf((float)head_dim)) * 1.4426950408889634f;
  dim3 grid(ceil_div(S, AT_BQ), n_heads, n_seq);
  cudaStream_t st = (cudaStream_t)stream;
  // tcgen05/TMEM kernel for the encoder shapes (head_dim 128, S <= 256); the mma.sync flash kernel below
  // covers other head dims and longer sequences (GD_ATTN_LEGACY=1 forces it, for A/B timing).
  if (attention_tc_supported(S, n_heads, head_dim) && (((uintptr_t)qkv) & 15) == 0 && (((uintptr_t)out) & 15) == 0 &&
      !attention_legacy_forced())
    return launch_attention_tc(qkv, out, n_seq, S, n_heads, fmt, st);
  const __nv_bfloat16* q = (const __nv_bfloat16*)qkv;
  __nv_bfloat16* o = (__nv_bfloat16*)out;
#define GD_ATH(DHV)                                                                              \
  if (fmt == GD_FP16) attention_h16_kernel<DHV, true><<<grid, AT_THREADS, 0, st>>>(q, o, S, d, sl2); \
  else attention_h16_kernel<DHV, false><<<grid, AT_THREADS, 0, st>>>(q, o, S, d, sl2);
  if (head_dim == 128) { GD_ATH(128) }
  else if (head_dim == 64) { GD_ATH(64) }
  else if (head_dim == 32) { GD_ATH(32) }
  else {
    set_error("gd_attention_h16: head_dim %d not in {32,64,128}", head_dim);
    return GD_ERR_UNSUPPORTED;
  }
#undef GD_ATH
  GD_LAUNCH_CHECK();
  return GD_OK;
}

extern "C" int gd_attention_f32(const float* qkv, float* out, int32_t n_seq, int32_t S, int32_t n_heads,
                                int32_t head_dim, void* stream) {
  using namespace gd;
  GD_REQUIRE(qkv && out && n_seq > 0 && S > 0 && n_heads > 0, "gd_attention_f32: bad args");
  GD_REQUIRE(n_seq <= 65535 && n_heads <= 65535, "gd_attention_f32: grid too large");
  const int d = n_heads * head_dim;
  GD_REQUIRE(d % 4 == 0, "gd_attention_f32: d must be a multiple of 4");
  const size_t smem = (size_t)(8 * head_dim + 8 * S) * sizeof(float);
  GD_REQUIRE(smem <= 160 * 1024, "gd_attention_f32: S=%d too long", S);
  const float scale = 1.0f / sqrtf((float)head_dim);
  dim3 grid(ceil_div(S, 8), n_heads, n_seq);
  cudaStream_t st = (cudaStream_t)stream;
#define GD_ATF(DHV)                                                                                              \
  {                                                                                                              \
    static size_t set_dev[kMaxDevices] = {}; size_t& set = set_dev[current_device()];                                                                                       \
    if (smem > 48 * 1024 && smem > set) {                                                                        \
      GD_CUDA_OK(cudaFuncSetAttribute(attention_f32_kernel<DHV>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                      (int)smem));                                                               \
      set = smem;                                                                                                \
    }                                                                                                            \
    attention_f32_kernel<DHV><<<grid, 256, smem, st>>>(qkv, out, S, d, scale);                                   \
  }
  if (head_dim == 128) GD_ATF(128)
  else if (head_dim == 64) GD_ATF(64)
  else if (head_dim == 32) GD_ATF(32)
  else {
    set_error("gd_attention_f32: head_dim %d not in {32,64,128}", head_dim);
    return GD_ERR_UNSUPPORTED;
  }
#undef GD_ATF
  GD_LAUNCH_CHECK();
  return GD_OK;
}
