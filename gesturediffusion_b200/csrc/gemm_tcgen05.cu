// gemm_tcgen05.cu -- persistent warp-specialised 16-bit (bf16 or fp16 operands) GEMM for sm_100a.
//   C[M,N] = A[M,K] @ W[N,K]^T  (+ fused gdEpilogue), fp32 accumulation in TMEM.
// Roles (320 threads): warp 0 = TMA producer, warp 1 = UMMA issuer (one elected lane),
// warps 2..9 = epilogue (TMEM -> registers -> fused epilogue -> global).  The accumulator is
// double buffered in TMEM (2 x BN fp32 columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1; A/W tiles stream through a kStages-deep TMA/mbarrier ring with 128B swizzle.
// Replaces every aten::addmm of the reference hot path (see gd_b200.h).
//
// Epilogues are compile-time specialised (template EPI) so that each instantiation stays small
// enough for the instruction cache (ncu on the first, fully runtime-dispatched version showed
// stall_no_inst as the top stall):
//   EPI_GENERIC            every gdEpilogue feature, one thread per accumulator row (used by the two
//                          GEMMs per denoise step that need row maps / duplication / transposed stores)
//   EPI_BIAS_H16           out_h16 = acc + bias                      (in_proj QKV)
//   EPI_BIAS_GELU_H16      out_h16 = gelu_erf(acc + bias)            (linear1)
//   EPI_BIAS_RES_F32       out_f32 = acc + bias + residual           (out_proj, linear2)
// The specialised ones stage each 32x32 accumulator block through shared memory (XOR-swizzled, conflict
// free) so that global stores are row-contiguous 64/128-byte segments, and fetch the residual one block
// ahead of the TMEM read.
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kBM = 128;          // UMMA M (cta_group::1)
constexpr int kBK = 64;           // 64 bf16 = 128 B = one swizzle row
constexpr int kUmmaK = 16;
constexpr int kEpiWarps = 8;
constexpr int kGemmThreads = (2 + kEpiWarps) * 32;

enum { EPI_GENERIC = 0, EPI_BIAS_H16 = 1, EPI_BIAS_GELU_H16 = 2, EPI_BIAS_RES_F32 = 3 };

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 256;
  static constexpr int kStagingBytes = kEpiWarps * 32 * 32 * 4;                 // one 32x32 fp32 block per epilogue warp
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarBytes + 1024;   // +1024 alignment slack
  static constexpr int kTmemCols = 2 * BN;                                      // 256 or 512 (power of 2)
};

// ---- specialised epilogue of one 32-row x 32-column accumulator block held as v[32] (row = lane) -------------
// stg: this warp's 4 KB staging block (shared-space address).  Block element (r, c) lives in 16-byte chunk
// (c / 4) ^ (r & 7) of row r: conflict-free for the row-wise writes and for both read-back patterns.
template <int EPI>
__device__ __forceinline__ void epi_block(const gdEpilogue& ep, uint32_t stg, int lane, int m_warp, int n0, int M,
                                          const float (&v)[32], const float4 (&pre)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    ptx::sts_v4(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  __syncwarp();
  if constexpr (EPI == EPI_BIAS_RES_F32) {
    // lane -> (row it*4 + lane/8, columns 4*(lane%8) .. +3): 8 lanes cover one 128-byte row segment
    const int c4 = lane & 7, rsub = lane >> 3, n = n0 + c4 * 4;
    const float4 b = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int rl = it * 4 + rsub, mm = m_warp + rl;
      float4 w = ptx::lds_v4(stg + rl * 128 + ((c4 ^ (rl & 7)) << 4));
      f4_add(w, b);
      f4_add(w, pre[it]);
      if (mm < M) *reinterpret_cast<float4*>(ep.out_f32 + (size_t)mm * ep.ld_out + n) = w;
    }
  } else {
    // 16-bit outputs: lane -> (row it*8 + lane/4, columns 8*(lane%4) .. +7): one 16-byte store per lane,
    // 4 lanes cover one 64-byte row segment
    const int c8 = lane & 3, rsub = lane >> 2, n = n0 + c8 * 8;
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(ep.bias + n + 4));
    uint16_t* out = reinterpret_cast<uint16_t*>(ep.out_h16);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int rl = it * 8 + rsub, mm = m_warp + rl;
      float4 w0 = ptx::lds_v4(stg + rl * 128 + (((2 * c8) ^ (rl & 7)) << 4));
      float4 w1 = ptx::lds_v4(stg + rl * 128 + (((2 * c8 + 1) ^ (rl & 7)) << 4));
      f4_add(w0, b0);
      f4_add(w1, b1);
      if constexpr (EPI == EPI_BIAS_GELU_H16) {
        w0.x = gelu_fast(w0.x); w0.y = gelu_fast(w0.y); w0.z = gelu_fast(w0.z); w0.w = gelu_fast(w0.w);
        w1.x = gelu_fast(w1.x); w1.y = gelu_fast(w1.y); w1.z = gelu_fast(w1.z); w1.w = gelu_fast(w1.w);
      }
      uint4 q;
      q.x = pack_h16x2(w0.x, w0.y, ep.h16_fmt);
      q.y = pack_h16x2(w0.z, w0.w, ep.h16_fmt);
      q.z = pack_h16x2(w1.x, w1.y, ep.h16_fmt);
      q.w = pack_h16x2(w1.z, w1.w, ep.h16_fmt);
      if (mm < M) *reinterpret_cast<uint4*>(out + (size_t)mm * ep.ld_out_h16 + n) = q;
    }
  }
  __syncwarp();
}

// residual block one step ahead: 8 independent 16-byte loads per lane in flight
__device__ __forceinline__ void prefetch_residual(const gdEpilogue& ep, int lane, int m_warp, int n0, int M,
                                                  float4 (&pre)[8]) {
  const int c4 = lane & 7, rsub = lane >> 3, n = n0 + c4 * 4;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int mm = m_warp + it * 4 + rsub;
    pre[it] = (mm < M) ? *reinterpret_cast<const float4*>(ep.residual + (size_t)mm * ep.ld_res + n)
                       : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_h16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                         int M, int N, int K, uint32_t idesc, const gdEpilogue ep) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
  uint8_t* bars = staging + Cfg::kStagingBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bars);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_tiles = (M + kBM - 1) / kBM;
  const int n_tiles = (N + BN - 1) / BN;
  const int total_tiles = m_tiles * n_tiles;
  const int k_blocks = (K + kBK - 1) / kBK;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(&tfull_bar[b], 1);
      ptx::mbar_init(&tempty_bar[b], kEpiWarps);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBK, m_blk * kBM);
          ptx::tma_load_2d(sb, &tmap_w, &full_bar[stage], kb * kBK, n_blk * BN);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);
        ptx::mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);     // epilogue drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(buf * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            const uint64_t da = ptx::umma_desc_k_sw128(sa + k * kUmmaK * 2);
            const uint64_t db = ptx::umma_desc_k_sw128(sb + k * kUmmaK * 2);
            ptx::umma_h16(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&empty_bar[stage]);               // smem slot reusable when MMAs retire
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        ptx::umma_commit(&tfull_bar[buf]);                   // accumulator complete
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;
    const int quarter = warp & 3;            // TMEM lane quarter this warp may access
    const int half = ew >> 2;                // which half of the BN columns
    constexpr int kColsPerWarp = BN / 2;
    const uint32_t stg = ptx::smem_u32(staging) + (uint32_t)ew * (32 * 32 * 4);
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      const int m_warp = m_blk * kBM + quarter * 32;
      const int n_base = n_blk * BN + half * kColsPerWarp;
      float4 pre[8];
      if constexpr (EPI == EPI_BIAS_RES_F32) {
        if (n_base < N) prefetch_residual(ep, lane, m_warp, n_base, M, pre);    // before the accumulator is ready
      }
      ptx::mbar_wait(&tfull_bar[buf], use & 1);
      ptx::tc_fence_after();
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
#pragma unroll 1
      for (int c = 0; c < kColsPerWarp; c += 32) {
        const int n0 = n_base + c;
        if (n0 >= N) break;                                  // warp-uniform
        float v[32];
        ptx::tmem_ld_32x32(taddr0 + c, v);
        if constexpr (EPI == EPI_GENERIC) {
          if (m_warp + lane < M) epilogue_row_chunk<32>(ep, m_warp + lane, n0, N, v);
        } else {
          float4 cur[8];
          if constexpr (EPI == EPI_BIAS_RES_F32) {
#pragma unroll
            for (int it = 0; it < 8; ++it) cur[it] = pre[it];
            if (c + 32 < kColsPerWarp && n0 + 32 < N) prefetch_residual(ep, lane, m_warp, n0 + 32, M, pre);
          }
          epi_block<EPI>(ep, stg, lane, m_warp, n0, M, v, cur);
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[buf]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess || p == nullptr) {
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// 2D bf16 row-major [rows, cols] with row stride ld elements; box = [box_rows, 64 cols], SW128.
int make_tmap_2d(CUtensorMap* out, const void* base, int rows, int cols, int ld, int box_rows, int fmt) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return GD_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, fmt == GD_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d)", (int)r, rows, cols, ld);
    return GD_ERR_CUDA;
  }
  return GD_OK;
}

static int g_num_sms = 0;

int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  }
  return g_num_sms;
}

// Which compile-time epilogue serves this gdEpilogue (EPI_GENERIC handles everything).
static int pick_epilogue(const gdEpilogue* ep, int N) {
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (N % 32 != 0 || !ep->bias || !al16(ep->bias) || ep->rowtab || ep->seqbias || ep->map_t > 0 || ep->dup > 1 ||
      ep->out_t)
    return EPI_GENERIC;
  if (ep->out_h16 && !ep->out_f32 && !ep->residual && al16(ep->out_h16) && ep->ld_out_h16 % 8 == 0) {
    if (ep->act == GD_ACT_NONE) return EPI_BIAS_H16;
    if (ep->act == GD_ACT_GELU_ERF) return EPI_BIAS_GELU_H16;
    return EPI_GENERIC;
  }
  if (ep->out_f32 && !ep->out_h16 && ep->residual && ep->act == GD_ACT_NONE && al16(ep->out_f32) &&
      al16(ep->residual) && ep->ld_out % 4 == 0 && ep->ld_res % 4 == 0)
    return EPI_BIAS_RES_F32;
  return EPI_GENERIC;
}

template <int BN, int EPI>
static int launch_gemm(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int fmt,
                       const gdEpilogue* ep, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(gemm_h16_tcgen05_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::kSmemBytes));
    attr_set = true;
  }
  CUtensorMap ta, tw;
  int rc = make_tmap_2d(&ta, A, M, K, lda, kBM, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw, W, N, K, ldw, BN, fmt);
  if (rc) return rc;
  const int tiles = ceil_div(M, kBM) * ceil_div(N, BN);
  const int grid = tiles < g_num_sms ? tiles : g_num_sms;
  gemm_h16_tcgen05_kernel<BN, EPI><<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(
      ta, tw, M, N, K, ptx::umma_idesc_h16_f32(kBM, BN, fmt == GD_FP16), *ep);
  GD_LAUNCH_CHECK();
  return GD_OK;
}

}  // namespace gd

extern "C" int gd_gemm_h16(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                           int32_t fmt, const gdEpilogue* ep, void* stream) {
  using namespace gd;
  GD_REQUIRE(A && W && ep, "gd_gemm_h16: null pointer");
  GD_REQUIRE(M > 0 && N > 0 && K > 0, "gd_gemm_h16: bad shape M=%d N=%d K=%d", M, N, K);
  GD_REQUIRE(fmt == GD_BF16 || fmt == GD_FP16, "gd_gemm_h16: fmt must be GD_BF16 or GD_FP16");
  GD_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K,
             "gd_gemm_h16: lda/ldw must be multiples of 8 and >= K (lda=%d ldw=%d K=%d)", lda, ldw, K);
  GD_REQUIRE((((uintptr_t)A) & 15) == 0 && (((uintptr_t)W) & 15) == 0, "gd_gemm_h16: operands must be 16B aligned");
  int rc = validate_epilogue(ep, M, N);
  if (rc) return rc;
  if (g_num_sms == 0) {
    int dev = 0;
    GD_CUDA_OK(cudaGetDevice(&dev));
    GD_CUDA_OK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  // Tile choice: BN=256 halves A re-reads and epilogue barrier traffic; BN=128 gives finer waves.
  const long tiles256 = (long)ceil_div(M, kBM) * ceil_div(N, 256);
  const bool use256 = (N >= 256) && (tiles256 >= 2L * g_num_sms);
  const int epi = pick_epilogue(ep, N);
  cudaStream_t st = (cudaStream_t)stream;
#define GD_GEMM_CASE(E)                                                                    \
  case E:                                                                                  \
    return use256 ? launch_gemm<256, E>(A, lda, W, ldw, M, N, K, fmt, ep, st)              \
                  : launch_gemm<128, E>(A, lda, W, ldw, M, N, K, fmt, ep, st);
  switch (epi) {
    GD_GEMM_CASE(EPI_BIAS_H16)
    GD_GEMM_CASE(EPI_BIAS_GELU_H16)
    GD_GEMM_CASE(EPI_BIAS_RES_F32)
    default: break;
  }
  return use256 ? launch_gemm<256, EPI_GENERIC>(A, lda, W, ldw, M, N, K, fmt, ep, st)
                : launch_gemm<128, EPI_GENERIC>(A, lda, W, ldw, M, N, K, fmt, ep, st);
#undef GD_GEMM_CASE
}
