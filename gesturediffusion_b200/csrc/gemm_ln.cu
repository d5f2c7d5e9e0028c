// gemm_ln.cu -- 16-bit tcgen05 GEMM with bias + residual + LayerNorm fused into the epilogue:
//   X = LayerNorm(residual + A @ W^T + bias) * gamma + beta        (post-norm sub-layer tail of
//   nn.TransformerEncoderLayer: out_proj + norm1 and linear2 + norm2, mdm.py:90-96)
// written as fp32 (the residual stream) and as the 16-bit operand of the next GEMM, so the pre-norm sum
// never goes to HBM (saves one fp32 write + one fp32 read of the whole activation per sub-layer and the
// separate LayerNorm launch).
//
// A LayerNorm row spans the full N = d columns; one CTA's TMEM holds two 128 x BN fp32 accumulators (double
// buffered), so a row is split over the TWO CTAs of a thread-block cluster (N = 2*BN, CTA rank r owns columns
// [r*BN, (r+1)*BN)).  Per 128-row tile each epilogue thread owns one row x BN/2 columns:
//   pass A  y = acc + bias + residual; the residual arrives as [32 x 32] fp32 TMA boxes in a per-warp ring, two
//           blocks ahead of the consumer; running (mean, M2) by Chan's update; y written back to TMEM in place;
//   exchange  the four partial (mean, M2) of a row -- 2 column halves x 2 CTAs -- go to both CTAs' shared memory
//           with st.async, which completes transaction bytes on the destination's mbarrier (one expect_tx arrival
//           per tile and CTA): data and signal travel together, no fence and no separate arrive;
//   pass B  z = (y - mean) * rstd * gamma + beta from TMEM into swizzled staging blocks (fp32 128B swizzle, 16-bit
//           64B swizzle), one TMA store per output and chunk (the hardware clips the ragged last tile).
// Statistics are two-pass quality (no E[x^2] - mean^2 cancellation).
//
// MODE selects how the fp32 residual stream travels through HBM (the kernel's memory phases run at HBM speed, so
// bytes are time):
//   0  fp32 residual in, fp32 and / or 16-bit out (gd_gemm_h16_ln)
//   1  fp32 residual in, split out: hi = round16(x), lo = round16(x - hi)  (gd_gemm_h16_ln_split, first sub-layer)
//   2  split residual in (x = hi + lo, exact to ~2^-22 relative), split out
// hi is the 16-bit operand the next GEMM needs anyway, so the split form costs 2 + 2 instead of 4 + 2 bytes per
// element on the way out and the residual read is two 16-bit boxes instead of one fp32 box.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kLnBM = 128, kLnBK = 64, kLnEpiWarps = 8, kLnThreads = 320;

// ALIAS = false: per epilogue warp two residual blocks in flight (also across tile boundaries) plus separate
// staging blocks.  ALIAS = true (one more operand stage for long K, main loop bound by the operand bytes in flight): the staging blocks alias the
// two residual blocks, so the first residual blocks of a tile are only requested when the previous tile has left.
template <int BN, int STAGES, bool ALIAS, int CL = 2>
struct LnCfg {
  static constexpr bool kAlias = ALIAS;
  static constexpr int kWarpRing = kAlias ? 2 * 4096 : 3 * 4096 + 2048;
  static constexpr int kABytes = kLnBM * kLnBK * 2;
  static constexpr int kBBytes = BN * kLnBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kRingBytes = kLnEpiWarps * kWarpRing;
  static constexpr int kVecBytes = 3 * BN * 4;                  // bias, gamma, beta slices of this CTA
  static constexpr int kParts = 2 * CL;                         // partial statistics per row: 2 column halves x CL CTAs
  static constexpr int kStatBytes = 2 * kParts * kLnBM * 8;     // [parity][part][row] (mean, M2)
  static constexpr int kBarBytes = 512;
  static constexpr int kSmemBytes = STAGES * kStageBytes + kRingBytes + kVecBytes + kStatBytes + kBarBytes;
  static constexpr int kTmemCols = 2 * BN;
};

struct LnParams {
  const float* bias; const float* gamma; const float* beta;
  const void* residual; int ld_res;   // fp32, or the 16-bit hi half of a split residual (MODE 2)
  void* out_f32; int ld_out;          // fp32 output (MODE 0) or the 16-bit lo output (MODE 1 / 2)
  void* out_h16; int ld_out_h16; int h16_fmt;
  float eps;
  int M, K;
  int tile_rows;   // rows per tile, in (96, 128]: the launcher balances the tiles over the clusters' rounds (see launch_gemm_ln)
  uint32_t idesc;
  int dbg;     // GD_LN_DBG diagnostics (results wrong): 1 no statistics exchange, 2 no global stores, 4 no residual, 8 no epilogue
};

template <int BN, int STAGES, bool ALIAS, int MODE, int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(kLnThreads, 1)
gemm_ln_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                       const __grid_constant__ CUtensorMap tmap_res, const __grid_constant__ CUtensorMap tmap_rlo,
                       const __grid_constant__ CUtensorMap tmap_o32, const __grid_constant__ CUtensorMap tmap_o16,
                       const __grid_constant__ CUtensorMap tmap_res_t, const __grid_constant__ CUtensorMap tmap_rlo_t,
                       const __grid_constant__ CUtensorMap tmap_o32_t, const __grid_constant__ CUtensorMap tmap_o16_t,
                       const LnParams p) {
  // the *_t maps are the same tensors with a box of (tile_rows - 96) rows: the fourth lane quarter of a tile shorter than
  // 128 rows loads / stores through them, so that it never touches the next tile's rows
  // MODE 0: tmap_res fp32 residual, tmap_o32 fp32 output, tmap_o16 16-bit output (tmap_rlo unused)
  // MODE 1: tmap_res fp32 residual, tmap_o32 = lo output, tmap_o16 = hi output
  // MODE 2: tmap_res / tmap_rlo = hi / lo residual (16-bit, 64B swizzle), outputs as MODE 1
  using Cfg = LnCfg<BN, STAGES, ALIAS, CL>;
  constexpr int kLnStages = STAGES;
  constexpr bool kAlias = Cfg::kAlias;
  constexpr int N = CL * BN;
  constexpr uint32_t P = (uint32_t)Cfg::kParts;
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* ring = smem + kLnStages * Cfg::kStageBytes;
  float* vecs = reinterpret_cast<float*>(ring + Cfg::kRingBytes);                // bias | gamma | beta
  uint8_t* stats = reinterpret_cast<uint8_t*>(vecs) + Cfg::kVecBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stats + Cfg::kStatBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = full_bar + kLnStages;
  uint64_t* tfull_bar = empty_bar + kLnStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* stat_bar = tempty_bar + 2;
  uint64_t* res_bar = stat_bar + 2;                          // [epilogue warp][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + 2 * kLnEpiWarps);

  // warp index through a shuffle: warp-uniform for the compiler, so the role branches stay uniform and the MMA warp's
  // descriptor arithmetic runs in the uniform datapath (no ELECT / R2UR waterfall around every UTCHMMA)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int n_clusters = (int)gridDim.x / CL, cluster_id = (int)blockIdx.x / CL;
  const int M = p.M;
  const int tile_rows = p.tile_rows;
  const int m_tiles = (M + tile_rows - 1) / tile_rows;
  const int k_blocks = (p.K + kLnBK - 1) / kLnBK;
  const int n_cta = (int)rank * BN;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_w);
    ptx::prefetch_tensormap(&tmap_res);
    if constexpr (MODE == 2) ptx::prefetch_tensormap(&tmap_rlo);
    ptx::prefetch_tensormap(&tmap_o32);
    ptx::prefetch_tensormap(&tmap_o16);
    if (tile_rows < kLnBM) {
      ptx::prefetch_tensormap(&tmap_res_t);
      if constexpr (MODE == 2) ptx::prefetch_tensormap(&tmap_rlo_t);
      ptx::prefetch_tensormap(&tmap_o32_t);
      ptx::prefetch_tensormap(&tmap_o16_t);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2 * kLnEpiWarps; ++i) ptx::mbar_init(&res_bar[i], 1);
    for (int s = 0; s < kLnStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(&tfull_bar[b], 1);
      ptx::mbar_init(&tempty_bar[b], kLnEpiWarps);
      ptx::mbar_init(&stat_bar[b], 1);                     // one expect_tx arrival per tile; the data are the transactions
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols);
  for (int i = threadIdx.x; i < BN; i += kLnThreads) {
    vecs[i] = p.bias[n_cta + i];
    vecs[BN + i] = p.gamma[n_cta + i];
    vecs[2 * BN + i] = p.beta[n_cta + i];
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();                                   // peer's barriers are initialised before any remote arrive
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::grid_dep_wait();                                      // prologue above overlaps the previous kernel's tail
  ptx::grid_dep_launch();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // (whole warp in uniform control flow, one elected lane issues)
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int m_blk = cluster_id; m_blk < m_tiles; m_blk += n_clusters) {
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          if (ptx::elect_one()) {
            // the A box has tile_rows rows; rows beyond keep stale bytes and feed accumulator lanes nobody stores
            ptx::mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(tile_rows * kLnBK * 2 + Cfg::kBBytes));
            ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kLnBK, m_blk * tile_rows);
            ptx::tma_load_2d(sb, &tmap_w, &full_bar[stage], kb * kLnBK, n_cta);
          }
          __syncwarp();
          if (++stage == kLnStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer
    // the whole warp runs the loop (uniform control flow, barrier waits by all lanes); one elected lane issues
    {
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t smem_u = __shfl_sync(0xffffffffu, ptx::smem_u32(smem), 0);
      for (int m_blk = cluster_id; m_blk < m_tiles; m_blk += n_clusters, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);
        ptx::mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_u + (uint32_t)(buf * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kLnBK / 16; ++k) {
              const uint64_t da = ptx::umma_desc_k_sw128(sa + k * 32);
              const uint64_t db = ptx::umma_desc_k_sw128(sb + k * 32);
              ptx::umma_h16(tmem_d, da, db, p.idesc, (kb | k) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kLnStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit(&tfull_bar[buf]);
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int half = ew >> 2;
    constexpr int CW = BN / 2;                               // columns per warp
    constexpr int NCH = CW / 32;
    // per-warp smem: blocks 0/1 receive the residual ([32 rows x 32 fp32] TMA boxes, 128B swizzle: 16-byte chunk
    // (c/4) ^ (r&7) of row r), two blocks in flight ahead of the consumer and across tile boundaries; block 2
    // (same layout) and a [32 x 32] 16-bit block (64B swizzle) stage the two outputs for the TMA stores of pass B.
    const uint32_t rbuf = ptx::smem_u32(ring) + (uint32_t)ew * (kAlias ? 2 * 4096 : 3 * 4096);
    const uint32_t stg = kAlias ? rbuf : rbuf + 2 * 4096;
    const uint32_t stg16 =
        kAlias ? rbuf + 4096 : ptx::smem_u32(ring) + (uint32_t)kLnEpiWarps * (3 * 4096) + (uint32_t)ew * 2048;
    const uint32_t rbar = ptx::smem_u32(&res_bar[2 * ew]);
    const uint32_t vec_s = ptx::smem_u32(vecs) + (uint32_t)(half * CW) * 4;      // this warp's bias slice
    const uint32_t stats_s = ptx::smem_u32(stats);
    uint32_t stats_cl[CL], sbar_cl[CL];
    const uint32_t sbar_s = ptx::smem_u32(stat_bar);
#pragma unroll
    for (int r = 0; r < CL; ++r) {
      stats_cl[r] = ptx::mapa(stats_s, (uint32_t)r);
      sbar_cl[r] = ptx::mapa(sbar_s, (uint32_t)r);
    }
    const uint32_t part = rank * 2u + (uint32_t)half;
    const int row_in_tile = quarter * 32 + lane;
    const int n_warp = n_cta + half * CW;
    const int my_tiles = (m_tiles - cluster_id + n_clusters - 1) / n_clusters;
    const int total_blocks = my_tiles * NCH;
    // rows of the tile in this warp's lane quarter: 32, or tile_rows - 96 in the fourth quarter of a shorter tile
    const int rows_q = min(32, tile_rows - quarter * 32);
    const bool tail_q = rows_q < 32;
    const CUtensorMap* map_res = tail_q ? &tmap_res_t : &tmap_res;
    const CUtensorMap* map_rlo = tail_q ? &tmap_rlo_t : &tmap_rlo;
    const CUtensorMap* map_o32 = tail_q ? &tmap_o32_t : &tmap_o32;
    const CUtensorMap* map_o16 = tail_q ? &tmap_o16_t : &tmap_o16;
    auto issue_residual = [&](int g) {                       // lane 0: residual block g of this warp's sequence
      const int tl = g / NCH, c = g - tl * NCH;
      const uint32_t bar = rbar + (uint32_t)(g & 1) * 8;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rows_q * 128) : "memory");
      const int row0 = (cluster_id + tl * n_clusters) * tile_rows + quarter * 32;
      const uint32_t dst = rbuf + (uint32_t)(g & 1) * 4096;
      ptx::tma_load_2d_s(dst, map_res, bar, n_warp + c * 32, row0);             // fp32 box, or the hi box (2 KB)
      if constexpr (MODE == 2) ptx::tma_load_2d_s(dst + 2048, map_rlo, bar, n_warp + c * 32, row0);
    };
    if (lane == 0 && !(p.dbg & 12)) {
      if (total_blocks > 0) issue_residual(0);
      if (total_blocks > 1) issue_residual(1);
    }
    int local = 0;
    for (int m_blk = cluster_id; m_blk < m_tiles; m_blk += n_clusters, ++local) {
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      const uint32_t par = (uint32_t)(local & 1);
      const int m_warp = m_blk * tile_rows + quarter * 32;
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * CW);
      // this tile's statistics: 4 partials x 128 rows x 8 bytes land on stat_bar[par] of each CTA
      if (ew == 0 && lane == 0 && !(p.dbg & 8)) ptx::mbar_arrive_expect_tx(&stat_bar[par], P * kLnBM * 8);
      ptx::mbar_wait(&tfull_bar[buf], use & 1);
      ptx::tc_fence_after();
      if (p.dbg & 8) {
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&tempty_bar[buf]);
        continue;
      }

      // ---- pass A: y = acc + bias + residual -> TMEM, running row statistics
      float r_mean = 0.f, r_m2 = 0.f;
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const int g = local * NCH + c;
        uint32_t vr[32];
        ptx::tmem_ld_32x32_nowait(taddr0 + c * 32, vr);
        if (!(p.dbg & 4)) {                                  // residual block g has landed
          const uint32_t bar = rbar + (uint32_t)(g & 1) * 8, parity = (uint32_t)(g >> 1) & 1u;
          uint32_t done;
          do {
            asm volatile(
                "{\n\t.reg .pred P1;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, P1;\n\t}\n"
                : "=r"(done) : "r"(bar), "r"(parity) : "memory");
          } while (!done);
        }
        float v[32];
        const uint32_t rb = rbuf + (uint32_t)(g & 1) * 4096;
        if constexpr (MODE == 2) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {                      // (hi + lo) + bias; 16-byte chunk j = columns 8j .. 8j+7
            const uint32_t off = (uint32_t)lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4);
            const uint4 h = ptx::lds_v4u(rb + off);
            const uint4 l = ptx::lds_v4u(rb + 2048 + off);
            const float4 b0 = ptx::lds_v4(vec_s + (c * 32 + j * 8) * 4);
            const float4 b1 = ptx::lds_v4(vec_s + (c * 32 + j * 8 + 4) * 4);
            const float2 h0 = unpack_h16x2(h.x, p.h16_fmt), h1 = unpack_h16x2(h.y, p.h16_fmt);
            const float2 h2 = unpack_h16x2(h.z, p.h16_fmt), h3 = unpack_h16x2(h.w, p.h16_fmt);
            const float2 l0 = unpack_h16x2(l.x, p.h16_fmt), l1 = unpack_h16x2(l.y, p.h16_fmt);
            const float2 l2 = unpack_h16x2(l.z, p.h16_fmt), l3 = unpack_h16x2(l.w, p.h16_fmt);
            v[8 * j] = (h0.x + l0.x) + b0.x;
            v[8 * j + 1] = (h0.y + l0.y) + b0.y;
            v[8 * j + 2] = (h1.x + l1.x) + b0.z;
            v[8 * j + 3] = (h1.y + l1.y) + b0.w;
            v[8 * j + 4] = (h2.x + l2.x) + b1.x;
            v[8 * j + 5] = (h2.y + l2.y) + b1.y;
            v[8 * j + 6] = (h3.x + l3.x) + b1.z;
            v[8 * j + 7] = (h3.y + l3.y) + b1.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {                      // residual + bias while the TMEM load is in flight
            const float4 r = ptx::lds_v4(rb + lane * 128 + ((j ^ (lane & 7)) << 4));
            const float4 b = ptx::lds_v4(vec_s + (c * 32 + j * 4) * 4);
            v[4 * j] = r.x + b.x;
            v[4 * j + 1] = r.y + b.y;
            v[4 * j + 2] = r.z + b.z;
            v[4 * j + 3] = r.w + b.w;
          }
        }
        ptx::fence_proxy_async();                            // our reads of the block precede the next TMA write
        __syncwarp();
        if (lane == 0 && g + 2 < total_blocks && !(p.dbg & 4) && (!kAlias || c + 2 < NCH)) issue_residual(g + 2);
        ptx::tmem_ld_wait_on(vr);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[4 * j] += __uint_as_float(vr[4 * j]);
          v[4 * j + 1] += __uint_as_float(vr[4 * j + 1]);
          v[4 * j + 2] += __uint_as_float(vr[4 * j + 2]);
          v[4 * j + 3] += __uint_as_float(vr[4 * j + 3]);
          s0 += v[4 * j];
          s1 += v[4 * j + 1];
          s2 += v[4 * j + 2];
          s3 += v[4 * j + 3];
        }
        ptx::tmem_st_32x32(taddr0 + c * 32, v);
        const float cm = ((s0 + s1) + (s2 + s3)) * (1.0f / 32.0f);
        float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d0 = v[4 * j] - cm, d1 = v[4 * j + 1] - cm, d2 = v[4 * j + 2] - cm, d3 = v[4 * j + 3] - cm;
          q0 = fmaf(d0, d0, q0);
          q1 = fmaf(d1, d1, q1);
          q2 = fmaf(d2, d2, q2);
          q3 = fmaf(d3, d3, q3);
        }
        const float cq = (q0 + q1) + (q2 + q3);
        // Chan: combine (c*32, r_mean, r_m2) with (32, cm, cq)
        const float na = (float)(c * 32), nb = 32.f, nt = na + nb;
        const float delta = cm - r_mean;
        r_mean += delta * (nb / nt);
        r_m2 += cq + delta * delta * (na * nb / nt);
      }
      ptx::tmem_st_wait();

      // ---- exchange: every thread sends its row's partial (mean, M2) to both CTAs of the cluster; the store
      // itself completes transaction bytes on the destination's barrier (no fence, no separate arrive)
      {
        const uint32_t off = ((par * P + part) * kLnBM + (uint32_t)row_in_tile) * 8u;
#pragma unroll
        for (int r = 0; r < CL; ++r) ptx::st_async_v2(stats_cl[r] + off, r_mean, r_m2, sbar_cl[r] + par * 8u);
      }
      if (!(p.dbg & 1)) ptx::mbar_wait(&stat_bar[par], use & 1);
      float mean = 0.f, m2 = 0.f;
      {
        float pm[P], pq[P];
#pragma unroll
        for (uint32_t q = 0; q < P; ++q) {
          const uint32_t a = stats_s + ((par * P + q) * kLnBM + (uint32_t)row_in_tile) * 8u;
          asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(pm[q]), "=f"(pq[q]) : "r"(a) : "memory");
          mean += pm[q];
        }
        mean *= 1.0f / (float)P;
#pragma unroll
        for (uint32_t q = 0; q < P; ++q) m2 += pq[q] + (float)CW * (pm[q] - mean) * (pm[q] - mean);
      }
      const float rstd = rsqrtf(m2 * (1.0f / (float)N) + p.eps);
      const float shift = -mean * rstd;

      // ---- pass B: normalise from TMEM into the swizzled staging blocks, one TMA store per output and chunk
      // Split output with aliased staging (hi + lo of a chunk = one 4 KB ring block), two ways to use the 2 blocks:
      //   default   block 1 stages every chunk, block 0 already receives the NEXT tile's first residual block
      //             (requested here, a whole pass B ahead of its use)
      //   dbg & 64  chunks alternate between the two blocks (a chunk never waits for the previous chunk's store)
      constexpr bool kSplitAlias = (MODE != 0) && kAlias;
      const bool dbl = kSplitAlias && (p.dbg & 64);
      if (kSplitAlias && !dbl && lane == 0 && !(p.dbg & 4)) {
        const int g = (local + 1) * NCH;                     // NCH is even: block g lives in ring slot 0
        if (g < total_blocks) issue_residual(g);
      }
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        float v[32];
        ptx::tmem_ld_32x32(taddr0 + c * 32, v);
        if (c == NCH - 1) {                                  // accumulator buffer fully consumed
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty_bar[buf]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 g = ptx::lds_v4(vec_s + (BN + c * 32 + j * 4) * 4);
          const float4 b = ptx::lds_v4(vec_s + (2 * BN + c * 32 + j * 4) * 4);
          v[4 * j] = fmaf(fmaf(v[4 * j], rstd, shift), g.x, b.x);
          v[4 * j + 1] = fmaf(fmaf(v[4 * j + 1], rstd, shift), g.y, b.y);
          v[4 * j + 2] = fmaf(fmaf(v[4 * j + 2], rstd, shift), g.z, b.z);
          v[4 * j + 3] = fmaf(fmaf(v[4 * j + 3], rstd, shift), g.w, b.w);
        }
        if (lane == 0) {                                     // the staging block's previous stores have read it
          if (dbl) ptx::bulk_wait_read_1();
          else ptx::bulk_wait_read_all();
        }
        __syncwarp();
        const uint32_t sth = kSplitAlias ? rbuf + (dbl ? (uint32_t)(c & 1) * 4096 : 4096u) : stg16;
        const uint32_t stl = kSplitAlias ? sth + 2048 : stg;
        if constexpr (MODE == 0) {
          if (p.out_f32) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              ptx::sts_v4(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2],
                          v[4 * j + 3]);
          }
          if (p.out_h16) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              ptx::sts_v4u(stg16 + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4),
                           pack_h16x2(v[8 * j], v[8 * j + 1], p.h16_fmt), pack_h16x2(v[8 * j + 2], v[8 * j + 3], p.h16_fmt),
                           pack_h16x2(v[8 * j + 4], v[8 * j + 5], p.h16_fmt),
                           pack_h16x2(v[8 * j + 6], v[8 * j + 7], p.h16_fmt));
          }
        } else {
          // split output: hi = round16(z), lo = round16(z - hi); both as [32 x 32] 16-bit blocks (64B swizzle)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t hp[4], lp[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float a = v[8 * j + 2 * t], b = v[8 * j + 2 * t + 1];
              hp[t] = pack_h16x2(a, b, p.h16_fmt);
              const float2 f = unpack_h16x2(hp[t], p.h16_fmt);
              lp[t] = pack_h16x2(a - f.x, b - f.y, p.h16_fmt);
            }
            const uint32_t off = (uint32_t)lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4);
            ptx::sts_v4u(sth + off, hp[0], hp[1], hp[2], hp[3]);
            if (p.out_f32) ptx::sts_v4u(stl + off, lp[0], lp[1], lp[2], lp[3]);   // out_f32 holds the lo pointer here
          }
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0 && !(p.dbg & 2)) {
          if (p.out_f32) ptx::tma_store_2d(map_o32, MODE == 0 ? stg : stl, n_warp + c * 32, m_warp);
          if (p.out_h16) ptx::tma_store_2d(map_o16, MODE == 0 ? stg16 : sth, n_warp + c * 32, m_warp);
          ptx::bulk_commit();
        }
      }
      if (kAlias && lane == 0 && !(p.dbg & 4)) {             // staging has been read out: request the next tile's blocks
        const int g = (local + 1) * NCH;
        ptx::bulk_wait_read_all();
        if (g < total_blocks && !(kSplitAlias && !dbl)) issue_residual(g);
        if (g + 1 < total_blocks) issue_residual(g + 1);
      }
    }
    if (lane == 0) ptx::bulk_wait_all();                     // global writes complete before the CTA retires
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();                                   // no CTA exits while its peer may still write its smem
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// MODE 0: p.residual / p.out_f32 are fp32.  MODE 1: p.residual fp32, p.out_f32 carries the lo output (16-bit,
// leading dimension p.ld_out).  MODE 2: p.residual carries the hi residual and res_lo the lo residual (16-bit,
// leading dimension p.ld_res), outputs as MODE 1.
template <int BN, int STAGES, bool ALIAS, int MODE, int CL = 2>
static int launch_gemm_ln(const void* A, int lda, const void* W, int ldw, int fmt, const LnParams& p_in,
                          const void* res_lo, cudaStream_t stream) {
  using Cfg = LnCfg<BN, STAGES, ALIAS, CL>;
  static_assert(Cfg::kSmemBytes <= 232448, "shared memory");
  static bool attr_set_dev[kMaxDevices] = {};
  static int max_clusters_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  int& max_clusters = max_clusters_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(gemm_ln_tcgen05_kernel<BN, STAGES, ALIAS, MODE, CL>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    max_clusters = num_sms() / CL;
    if (CL > 2) {                                            // clusters that fit the GPCs at once (a GPC of 18 SMs holds 4 of 4)
      cudaLaunchConfig_t qc = {};
      qc.gridDim = dim3(CL * (num_sms() / CL));
      qc.blockDim = dim3(kLnThreads);
      qc.dynamicSmemBytes = Cfg::kSmemBytes;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = CL;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      qc.attrs = qa;
      qc.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, gemm_ln_tcgen05_kernel<BN, STAGES, ALIAS, MODE, CL>, &qc) == cudaSuccess && n > 0)
        max_clusters = n < max_clusters ? n : max_clusters;
      else
        (void)cudaGetLastError();
    }
    attr_set = true;
  }
  // Balanced row tiles.  With 128-row tiles the clusters run ceil(tiles / clusters) rounds and the last one is partly empty
  // (25 216 rows: 197 tiles over 74 clusters = 2.66 -> 3 rounds).  The kernel is paced by the bytes of a tile, not by its
  // MMAs, so the same number of rounds over EQUAL tiles of ceil(M / (rounds x clusters)) rows (114 here) moves fewer bytes
  // per round.  Only when that is more than 96 rows: the fourth lane quarter then works through maps with a shorter box.
  LnParams p = p_in;
  const int sms = num_sms();
  GD_REQUIRE(sms >= 2, "gd_gemm_h16_ln: cannot query the SM count");
  p.tile_rows = kLnBM;
  {
    static int balance = -1;                                 // GD_LN_BALANCE=0: 128-row tiles always
    if (balance < 0) {
      const char* e = getenv("GD_LN_BALANCE");
      balance = e ? atoi(e) : 1;
    }
    const int t128 = ceil_div(p.M, kLnBM), cl = t128 < max_clusters ? t128 : max_clusters;
    const int rounds = ceil_div(t128, cl);
    const int even = ceil_div(p.M, rounds * cl);
    if (balance && rounds >= 2 && even > 96 && even < kLnBM) p.tile_rows = even;
  }
  const int tail_rows = p.tile_rows < kLnBM ? p.tile_rows - 96 : 32;
  CUtensorMap ta, tw;
  int rc = make_tmap_2d(&ta, A, p.M, p.K, lda, p.tile_rows, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw, W, CL * BN, p.K, ldw, BN, fmt);
  if (rc) return rc;
  CUtensorMap tr[2], trlo[2], to32[2], to16[2];              // [0] 32-row boxes, [1] tail_rows-row boxes
  for (int t = 0; t < 2; ++t) {
    const int br = t == 0 ? 32 : tail_rows;
    if (MODE == 2) rc = make_tmap_2d_h16_sw64(&tr[t], p.residual, p.M, CL * BN, p.ld_res, br, p.h16_fmt);
    else rc = make_tmap_2d_f32(&tr[t], p.residual, p.M, CL * BN, p.ld_res, br);
    if (rc) return rc;
    trlo[t] = tr[t];
    if (MODE == 2) {
      rc = make_tmap_2d_h16_sw64(&trlo[t], res_lo, p.M, CL * BN, p.ld_res, br, p.h16_fmt);
      if (rc) return rc;
    }
    to32[t] = tr[t];                                         // placeholders when an output is absent (never used)
    to16[t] = tr[t];
    if (p.out_f32) {
      if (MODE == 0) rc = make_tmap_2d_f32(&to32[t], p.out_f32, p.M, CL * BN, p.ld_out, br);
      else rc = make_tmap_2d_h16_sw64(&to32[t], p.out_f32, p.M, CL * BN, p.ld_out, br, p.h16_fmt);
      if (rc) return rc;
    }
    if (p.out_h16) {
      rc = make_tmap_2d_h16_sw64(&to16[t], p.out_h16, p.M, CL * BN, p.ld_out_h16, br, p.h16_fmt);
      if (rc) return rc;
    }
  }
  const int m_tiles = ceil_div(p.M, p.tile_rows);
  const int clusters = m_tiles < max_clusters ? m_tiles : max_clusters;
  GD_CUDA_OK(launch_pdl(PDL_GEMM_LN, gemm_ln_tcgen05_kernel<BN, STAGES, ALIAS, MODE, CL>, dim3(CL * clusters),
                        dim3(kLnThreads), Cfg::kSmemBytes, stream, ta, tw, tr[0], trlo[0], to32[0], to16[0], tr[1], trlo[1],
                        to32[1], to16[1], p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

}  // namespace gd

namespace gd {
int launch_gemm_ln_small(const void* A, int lda, const void* W, int ldw, int M, int K, int fmt, const float* bias,
                         const float* residual, int ld_res, const float* gamma, const float* beta, float eps,
                         float* out_f32, int ld_out, void* out_h16, int ld_out_h16, int h16_fmt, cudaStream_t stream);
int gemm_ln_small_max_rows();
}

extern "C" int gd_gemm_h16_ln(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                              int32_t fmt, const float* bias, const float* residual, int32_t ld_res,
                              const float* gamma, const float* beta, float eps, float* out_f32, int32_t ld_out,
                              void* out_h16, int32_t ld_out_h16, int32_t h16_fmt, void* stream) {
  using namespace gd;
  GD_REQUIRE(A && W && bias && residual && gamma && beta && (out_f32 || out_h16), "gd_gemm_h16_ln: null pointer");
  GD_REQUIRE(M > 0 && K > 0, "gd_gemm_h16_ln: bad shape M=%d K=%d", M, K);
  if (N != 256 && N != 512) {
    set_error("gd_gemm_h16_ln: N=%d not in {256, 512} (use gd_gemm_h16 + gd_layernorm)", N);
    return GD_ERR_UNSUPPORTED;
  }
  GD_REQUIRE(fmt == GD_BF16 || fmt == GD_FP16, "gd_gemm_h16_ln: fmt must be GD_BF16 or GD_FP16");
  GD_REQUIRE(!out_h16 || h16_fmt == GD_BF16 || h16_fmt == GD_FP16, "gd_gemm_h16_ln: bad h16_fmt");
  GD_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K,
             "gd_gemm_h16_ln: lda/ldw must be multiples of 8 and >= K (lda=%d ldw=%d K=%d)", lda, ldw, K);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  GD_REQUIRE(al16(A) && al16(W) && al16(residual) && (!out_f32 || al16(out_f32)) && (!out_h16 || al16(out_h16)),
             "gd_gemm_h16_ln: operands must be 16-byte aligned");
  GD_REQUIRE(ld_res >= N && ld_res % 4 == 0 && (!out_f32 || (ld_out >= N && ld_out % 4 == 0)) &&
                 (!out_h16 || (ld_out_h16 >= N && ld_out_h16 % 8 == 0)),
             "gd_gemm_h16_ln: leading dimensions must be >= N and multiples of 4 (16-bit output: 8)");
  {
    // few rows: an 8-CTA cluster per row tile (gemm_ln_small.cu), while all row tiles fit the GPU at once
    if (N == 512 && M <= gemm_ln_small_max_rows())
      return launch_gemm_ln_small(A, lda, W, ldw, M, K, fmt, bias, residual, ld_res, gamma, beta, eps, out_f32, ld_out,
                                  out_h16, ld_out_h16, h16_fmt, (cudaStream_t)stream);
  }
  LnParams p;
  p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.residual = residual; p.ld_res = ld_res;
  p.out_f32 = out_f32; p.ld_out = ld_out;
  p.out_h16 = out_h16; p.ld_out_h16 = ld_out_h16; p.h16_fmt = h16_fmt;
  p.eps = eps; p.M = M; p.K = K;
  {
    static int dbg = -1;
    if (dbg < 0) {
      const char* e = getenv("GD_LN_DBG");
      dbg = e ? atoi(e) : 0;
    }
    p.dbg = dbg;
  }
  p.idesc = ptx::umma_idesc_h16_f32(kLnBM, N / 2, fmt == GD_FP16);
  static int force_stages = -1;
  if (force_stages < 0) {
    const char* e = getenv("GD_LN_STAGES");
    force_stages = e ? atoi(e) : 0;
  }
  const bool deep = force_stages ? force_stages > 2 : true;
  if (N == 512)
    return deep ? launch_gemm_ln<256, 3, true, 0>(A, lda, W, ldw, fmt, p, nullptr, (cudaStream_t)stream)
                : launch_gemm_ln<256, 2, false, 0>(A, lda, W, ldw, fmt, p, nullptr, (cudaStream_t)stream);
  return launch_gemm_ln<128, 3, false, 0>(A, lda, W, ldw, fmt, p, nullptr, (cudaStream_t)stream);
}

extern "C" int gd_gemm_h16_ln_split(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N,
                                    int32_t K, int32_t fmt, const float* bias, const float* res_f32, int32_t ld_res,
                                    const void* res_hi, const void* res_lo, int32_t ld_res16, const float* gamma,
                                    const float* beta, float eps, void* out_hi, void* out_lo, int32_t ld_out16,
                                    int32_t h16_fmt, void* stream) {
  using namespace gd;
  const bool split_in = res_f32 == nullptr;
  GD_REQUIRE(A && W && bias && gamma && beta && out_hi, "gd_gemm_h16_ln_split: null pointer");
  GD_REQUIRE(split_in ? (res_hi && res_lo) : (!res_hi && !res_lo),
             "gd_gemm_h16_ln_split: give either res_f32 or both res_hi and res_lo");
  GD_REQUIRE(M > 0 && K > 0, "gd_gemm_h16_ln_split: bad shape M=%d K=%d", M, K);
  if (N != 256 && N != 512) {
    set_error("gd_gemm_h16_ln_split: N=%d not in {256, 512}", N);
    return GD_ERR_UNSUPPORTED;
  }
  GD_REQUIRE(fmt == GD_BF16 || fmt == GD_FP16, "gd_gemm_h16_ln_split: fmt must be GD_BF16 or GD_FP16");
  GD_REQUIRE(h16_fmt == GD_BF16 || h16_fmt == GD_FP16, "gd_gemm_h16_ln_split: bad h16_fmt");
  GD_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K,
             "gd_gemm_h16_ln_split: lda/ldw must be multiples of 8 and >= K (lda=%d ldw=%d K=%d)", lda, ldw, K);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  GD_REQUIRE(al16(A) && al16(W) && al16(res_f32) && al16(res_hi) && al16(res_lo) && al16(out_hi) && al16(out_lo),
             "gd_gemm_h16_ln_split: operands must be 16-byte aligned");
  GD_REQUIRE(ld_out16 >= N && ld_out16 % 8 == 0 &&
                 (split_in ? (ld_res16 >= N && ld_res16 % 8 == 0) : (ld_res >= N && ld_res % 4 == 0)),
             "gd_gemm_h16_ln_split: leading dimensions must be >= N and multiples of 8 (fp32 residual: 4)");
  LnParams p;
  p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.residual = split_in ? res_hi : (const void*)res_f32;
  p.ld_res = split_in ? ld_res16 : ld_res;
  p.out_f32 = out_lo; p.ld_out = ld_out16;
  p.out_h16 = out_hi; p.ld_out_h16 = ld_out16; p.h16_fmt = h16_fmt;
  p.eps = eps; p.M = M; p.K = K;
  {
    static int dbg = -1;
    if (dbg < 0) {
      const char* e = getenv("GD_LN_DBG");
      dbg = e ? atoi(e) : 0;
    }
    p.dbg = dbg;
  }
  p.idesc = ptx::umma_idesc_h16_f32(kLnBM, N / 2, fmt == GD_FP16);
  cudaStream_t st = (cudaStream_t)stream;
  static int shallow_k = -1;                                 // GD_LN_SHALLOW_K=k: K <= k uses 2 stages + separate staging
  if (shallow_k < 0) {
    const char* e = getenv("GD_LN_SHALLOW_K");
    shallow_k = e ? atoi(e) : 0;
  }
  if (N == 512 && K <= shallow_k)
    return split_in ? launch_gemm_ln<256, 2, false, 2>(A, lda, W, ldw, fmt, p, res_lo, st)
                    : launch_gemm_ln<256, 2, false, 1>(A, lda, W, ldw, fmt, p, nullptr, st);
  if (N == 512)
    return split_in ? launch_gemm_ln<256, 3, true, 2>(A, lda, W, ldw, fmt, p, res_lo, st)
                    : launch_gemm_ln<256, 3, true, 1>(A, lda, W, ldw, fmt, p, nullptr, st);
  return split_in ? launch_gemm_ln<128, 3, false, 2>(A, lda, W, ldw, fmt, p, res_lo, st)
                  : launch_gemm_ln<128, 3, false, 1>(A, lda, W, ldw, fmt, p, nullptr, st);
}
