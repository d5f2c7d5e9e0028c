// ffn_ln_ws.cu -- the feed-forward sub-layer kernel of ffn_ln.cu with the epilogue split over TWO sets of warps:
//   X' = LayerNorm(X + gelu(X_hi @ W1^T + b1) @ W2^T + b2) * gamma + beta          (mdm.py:90-96, post-norm)
// Same tiling, operands, TMEM map, H exchange and arithmetic as ffn_ln.cu (a 128-row tile per 2-CTA cluster, CTA r owns
// output columns [256 r, +256); H in chunks of 128 hidden columns, chunk j by CTA j % 2; residual by identity MMAs).
// What ffn_ln.cu's timeline showed: the tile is paced by its dependency chain, and ~10 k of its ~33 k
// clocks are the LayerNorm tail, during which the eight epilogue warps cannot produce the next tile's H chunks and the
// single-buffered output accumulator keeps GEMM2 of the next tile from starting.  Here
//   warps 0..7   (GELU warpgroups) only turn GEMM1 accumulators into H chunks: 32 rows x 64 hidden columns per warp
//                and chunk (one 4 KB slab of one K block) -> own slot + one bulk copy into the peer's slot.  They never
//                wait for a tail: the next tile's chunk 0 is ready long before GEMM2 of that tile may start.
//   warps 8..15  (tail warpgroups) only run LayerNorm tails: pass A moves the warp's 32 rows x 128 output columns from
//                TMEM into REGISTERS (y = acc + b2, Chan statistics) and releases the accumulator at once -- the residual
//                MMAs and GEMM2 of the next tile start ~2 k clocks after the last GEMM2 of this one -- then exchange,
//                normalise from the registers, hi then lo -> a dedicated 2 KB staging block per warp -> TMA stores.
//   warp 16      TMA producer (ring of four 32 KB stages)
//   warp 17      UMMA issuer; the next tile's first two GEMM1 chunks are issued behind this tile's last two GEMM2 pairs
//   warps 18, 19 idle (they fill the control warpgroup)
// Registers: 640 threads compile to 96 per thread; the control warpgroup gives its registers back (setmaxnreg.dec 24),
// the GELU warpgroups go down to 72 and the two tail warpgroups take 152 (setmaxnreg.inc) for the 128 held output values
// (the pool is what the CTA was launched with: 640 x 96 = 61 440 registers >= 256 x 72 + 128 x 24 + 256 x 152).
// b2 / gamma / beta are read through the read-only path (warp-uniform addresses): their 3 KB pay for the fourth stage.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kWsBM = 128, kWsBK = 64, kWsThreads = 640;
constexpr int kWsGeluWarps = 8, kWsTailWarps = 8;
constexpr int kWsProducerWarp = 16, kWsMmaWarp = 17;
constexpr int kWsD = 512, kWsBN = 256, kWsCH = 128;
constexpr int kWsStages = 4;
constexpr int kWsStageBytes = 32768;
constexpr int kWsSlotBytes = 32768;            // one H chunk: two [128 rows x 64] 128B-swizzled K blocks
constexpr int kWsRingOff = 0;
constexpr int kWsHbufOff = kWsStages * kWsStageBytes;
constexpr int kWsStgOff = kWsHbufOff + 2 * kWsSlotBytes;           // 2 KB staging block per tail warp
constexpr int kWsI64Off = kWsStgOff + kWsTailWarps * 2048;         // 64 x 64 identity, K-major 128B swizzle (8 KB)
constexpr int kWsStatOff = kWsI64Off + 8192;                       // [parity][part][row] (mean, M2)
constexpr int kWsBarOff = kWsStatOff + 2 * 4 * kWsBM * 8;
constexpr int kWsSmemBytes = kWsBarOff + 512;
static_assert(kWsSmemBytes <= 232448, "shared memory");

struct WsParams {
  const float* b1; const float* b2; const float* gamma; const float* beta;
  int has_lo;        // the lo half of the result is wanted (not the last sub-layer)
  int h16_fmt;
  float eps;
  int M, F;
  uint32_t idesc1, idesc2, idesc3;
  unsigned long long* trace;   // optional timeline buffer (gd_ffn_trace): 4 roles x 4096 entries of (tag << 48 | clock)
  int dbg;           // GD_FFN_DBG diagnostics (results wrong): 1 no GELU arithmetic, 2 no global stores, 8 no operand loads
};

template <uint32_t R>
__device__ __forceinline__ void ws_reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(R)); }
template <uint32_t R>
__device__ __forceinline__ void ws_reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R)); }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kWsThreads, 1)
ffn_ln_ws_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                         const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_xlo,
                         const __grid_constant__ CUtensorMap tmap_ohi, const __grid_constant__ CUtensorMap tmap_olo,
                         const WsParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* ring = smem + kWsRingOff;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWsBarOff);
  uint64_t* full_bar = bars;                                 // [stages]
  uint64_t* empty_bar = full_bar + kWsStages;                // [stages]
  uint64_t* hacc_full = empty_bar + kWsStages;               // [2]  GEMM1 chunk accumulator complete
  uint64_t* hacc_empty = hacc_full + 2;                      // [2]  ... drained by the GELU warps (4 arrivals)
  uint64_t* hfull_loc = hacc_empty + 2;                      // locally produced H chunk written (4 warp arrivals)
  uint64_t* hfull_rem = hfull_loc + 1;                       // peer's H chunk has landed (transaction bytes)
  uint64_t* hempty = hfull_rem + 1;                          // [2]  slot's GEMM2 retired in BOTH CTAs (2 commits)
  uint64_t* y_full = hempty + 2;
  uint64_t* y_empty = y_full + 1;                            // the tail warps hold the tile's output in registers (8)
  uint64_t* stat_bar = y_empty + 1;                          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  const uint32_t rank = ptx::cluster_ctarank();
  // timeline of CTA 0: role 0 = GELU warp 0, role 1 = MMA warp, role 2 = producer, role 3 = tail warp 8 (lane 0 each)
  int tr_n = 0;
  unsigned long long* tr_buf = nullptr;
  if (p.trace != nullptr && blockIdx.x == 0 && lane == 0) {
    if (warp == 0) tr_buf = p.trace;
    else if (warp == kWsMmaWarp) tr_buf = p.trace + 4096;
    else if (warp == kWsProducerWarp) tr_buf = p.trace + 8192;
    else if (warp == kWsGeluWarps) tr_buf = p.trace + 12288;
  }
  auto TR = [&](unsigned tag) {
    if (tr_buf != nullptr && tr_n < 4096) tr_buf[tr_n++] = ((unsigned long long)tag << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull);
  };
  const int n_clusters = (int)gridDim.x >> 1, cluster_id = (int)blockIdx.x >> 1;
  const int M = p.M;
  const int m_tiles = (M + kWsBM - 1) / kWsBM;
  const int NC = p.F / (2 * kWsCH);                          // chunks this CTA produces per tile (>= 2)
  const int n_cta = (int)rank * kWsBN;
  constexpr int kKB1 = kWsD / kWsBK;                         // 8 K blocks per GEMM1 chunk
  constexpr int kKB2 = kWsCH / kWsBK;                        // 2 K blocks per GEMM2 chunk

  if (warp == kWsProducerWarp && lane == 0) {
    ptx::prefetch_tensormap(&tmap_x);
    ptx::prefetch_tensormap(&tmap_w1);
    ptx::prefetch_tensormap(&tmap_w2);
    ptx::prefetch_tensormap(&tmap_xlo);
    ptx::prefetch_tensormap(&tmap_ohi);
    ptx::prefetch_tensormap(&tmap_olo);
  }
  if (warp == kWsMmaWarp && lane == 0) {
    for (int s = 0; s < kWsStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(&hacc_full[b], 1);
      ptx::mbar_init(&hacc_empty[b], kWsGeluWarps);
      ptx::mbar_init(&hempty[b], 2);
      ptx::mbar_init(&stat_bar[b], 1);
    }
    ptx::mbar_init(hfull_loc, kWsGeluWarps);
    ptx::mbar_init(hfull_rem, 1);
    ptx::mbar_init(y_full, 1);
    ptx::mbar_init(y_empty, kWsTailWarps);
    ptx::fence_barrier_init();
  }
  if (warp == 0) ptx::tmem_alloc(tmem_slot, 512);
  {
    // I64[n][k] = (n == k) as a K-major 128B-swizzled UMMA B operand (ffn_ln.cu): "Y += X I^T" adds the residual on
    // the tensor pipe (exact: products with 1.0).
    const uint32_t one = (p.h16_fmt == GD_FP16) ? 0x3C00u : 0x3F80u;
    const uint32_t i64_s = ptx::smem_u32(smem + kWsI64Off);
    for (int idx = threadIdx.x; idx < 64 * 8; idx += kWsThreads) {
      const int n = idx >> 3, c = idx & 7;
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      if (c == (n >> 3)) w[(n & 7) >> 1] = one << (16 * (n & 1));
      ptx::sts_v4u(i64_s + n * 128 + ((c ^ (n & 7)) << 4), w[0], w[1], w[2], w[3]);
    }
    ptx::fence_proxy_async();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();                                   // peer's barriers are initialised before any remote arrive
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::grid_dep_wait();                                      // prologue above overlaps the previous kernel's tail
  ptx::grid_dep_launch();
  auto blk_of = [&](int mi) { return mi; };

  if (warp >= kWsProducerWarp) {
    // ================================================================== control warpgroup (warps 16..19)
    ws_reg_dec<24>();
    if (warp == kWsProducerWarp) {
      // ---------------------------------------------------------------- TMA producer
      if (lane == 0) {
        int stage = 0;
        uint32_t phase = 0;
        auto load_g1 = [&](int m_blk, int i) {               // own chunk i of the tile: global chunk 2 i + rank
          const int j = 2 * i + (int)rank;
          TR(1);
          for (int kb = 0; kb < kKB1; ++kb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kWsStageBytes;
            if (p.dbg & 8) {                                 // diagnostics: no operand traffic at all
              ptx::mbar_arrive(&full_bar[stage]);
            } else {
              ptx::mbar_arrive_expect_tx(&full_bar[stage], kWsStageBytes);
              ptx::tma_load_2d(sa, &tmap_x, &full_bar[stage], kb * kWsBK, m_blk * kWsBM);
              ptx::tma_load_2d(sa + 16384, &tmap_w1, &full_bar[stage], kb * kWsBK, j * kWsCH);
            }
            if (++stage == kWsStages) { stage = 0; phase ^= 1; }
          }
        };
        auto load_g2 = [&](int j) {                          // W2[n_cta .. +256, chunk j's 128 hidden columns]
          TR(2);
          for (int kb = 0; kb < kKB2; ++kb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kWsStageBytes;
            if (p.dbg & 8) {
              ptx::mbar_arrive(&full_bar[stage]);
            } else {
              ptx::mbar_arrive_expect_tx(&full_bar[stage], kWsStageBytes);
              ptx::tma_load_2d(sa, &tmap_w2, &full_bar[stage], j * kWsCH + kb * kWsBK, n_cta);
            }
            if (++stage == kWsStages) { stage = 0; phase ^= 1; }
          }
        };
        auto load_res = [&](int m_blk) {                     // [X_hi | X_lo] blocks of this CTA's 256 output columns
          for (int sb = 0; sb < kWsBN / kWsBK; ++sb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * kWsStageBytes;
            ptx::mbar_arrive_expect_tx(&full_bar[stage], kWsStageBytes);
            ptx::tma_load_2d(sa, &tmap_x, &full_bar[stage], n_cta + sb * kWsBK, m_blk * kWsBM);
            ptx::tma_load_2d(sa + 16384, &tmap_xlo, &full_bar[stage], n_cta + sb * kWsBK, m_blk * kWsBM);
            if (++stage == kWsStages) { stage = 0; phase ^= 1; }
          }
        };
        if (cluster_id < m_tiles) {
          load_g1(blk_of(cluster_id), 0);
          load_g1(blk_of(cluster_id), 1);
        }
        for (int mi = cluster_id; mi < m_tiles; mi += n_clusters) {
          const int m_blk = blk_of(mi);
          const bool has_next = mi + n_clusters < m_tiles;
          load_res(m_blk);
          for (int i = 0; i < NC; ++i) {                     // same order as the MMA warp
            load_g2(2 * i + (int)rank);
            load_g2(2 * i + 1 - (int)rank);
            if (i + 2 < NC) load_g1(m_blk, i + 2);
            else if (has_next) load_g1(blk_of(mi + n_clusters), i + 2 - NC);   // the next tile's first two chunks
          }
        }
      }
    } else if (warp == kWsMmaWarp) {
      // ---------------------------------------------------------------- UMMA issuer
      // the whole warp runs the schedule (uniform control flow, barrier waits by all lanes); one elected lane issues
      int stage = 0;
      uint32_t phase = 0;
      uint32_t g1n = 0;                                      // GEMM1 chunks issued so far (all tiles)
      uint32_t pairn = 0;                                    // chunk pairs consumed so far (all tiles)
      const uint32_t smem_u = __shfl_sync(0xffffffffu, ptx::smem_u32(smem), 0);
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t hbuf_s = smem_u + kWsHbufOff;
      auto g1 = [&]() {
        const uint32_t buf = g1n & 1u, use = g1n >> 1;
        TR(10);
        ptx::mbar_wait(&hacc_empty[buf], (use & 1u) ^ 1u);   // the GELU warps have drained this accumulator
        ptx::tc_fence_after();
        TR(11);
        const uint32_t tmem_d = tmem_u + 256u + buf * 128u;
        for (int kb = 0; kb < kKB1; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(kWsRingOff + stage * kWsStageBytes);
          const uint32_t sb = sa + 16384;
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kWsBK / 16; ++k)
              ptx::umma_h16(tmem_d, ptx::umma_desc_k_sw128(sa + k * 32), ptx::umma_desc_k_sw128(sb + k * 32), p.idesc1,
                            (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kWsStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit(&hacc_full[buf]);
        __syncwarp();
        TR(12);
        ++g1n;
      };
      auto g2 = [&](int j) {
        const uint32_t slot = (uint32_t)j & 1u;
        TR(20);
        if (slot == rank) {
          ptx::mbar_wait(hfull_loc, pairn & 1u);
        } else {
          if (ptx::elect_one()) ptx::mbar_arrive_expect_tx(hfull_rem, kWsSlotBytes);   // the peer's 8 slabs are the transactions
          __syncwarp();
          ptx::mbar_wait(hfull_rem, pairn & 1u);
        }
        ptx::tc_fence_after();
        TR(21);
        for (int kb = 0; kb < kKB2; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = hbuf_s + slot * kWsSlotBytes + (uint32_t)kb * 16384u;
          const uint32_t sb = smem_u + (uint32_t)(kWsRingOff + stage * kWsStageBytes);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kWsBK / 16; ++k)
              ptx::umma_h16(tmem_u, ptx::umma_desc_k_sw128(sa + k * 32), ptx::umma_desc_k_sw128(sb + k * 32), p.idesc2, 1u);
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kWsStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit_mc(&hempty[slot], 3);   // this CTA is done with the slot: tell both CTAs
        __syncwarp();
        TR(22);
      };
      const uint32_t i64_s = smem_u + kWsI64Off;
      auto gres = [&]() {                                    // Y[:, 64 sb ..] = X_hi I^T + X_lo I^T: the residual opens the tile
        for (int sb = 0; sb < kWsBN / kWsBK; ++sb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(kWsRingOff + stage * kWsStageBytes);
          if (ptx::elect_one()) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
              for (int k = 0; k < kWsBK / 16; ++k)
                ptx::umma_h16(tmem_u + (uint32_t)(sb * kWsBK), ptx::umma_desc_k_sw128(sa + h * 16384 + k * 32),
                              ptx::umma_desc_k_sw128(i64_s + k * 32), p.idesc3, (h | k) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kWsStages) { stage = 0; phase ^= 1; }
        }
      };
      uint32_t t = 0;
      if (cluster_id < m_tiles) {
        g1();
        g1();
      }
      for (int mi = cluster_id; mi < m_tiles; mi += n_clusters, ++t) {
        const bool has_next = mi + n_clusters < m_tiles;
        ptx::mbar_wait(y_empty, (t & 1u) ^ 1u);              // the previous tile's output sits in the tail warps' registers
        ptx::tc_fence_after();
        gres();
        for (int i = 0; i < NC; ++i) {
          g2(2 * i + (int)rank);                             // the locally produced chunk first (no DSMEM hop)
          g2(2 * i + 1 - (int)rank);
          ++pairn;
          if (i == NC - 1) {
            if (ptx::elect_one()) ptx::umma_commit(y_full);
            __syncwarp();
            TR(30);
          }
          if (i + 2 < NC || has_next) g1();                  // chunk i + 2, or chunk i + 2 - NC of the NEXT tile
        }
      }
    }
  } else if (warp < kWsGeluWarps) {
    // ================================================================== GELU warpgroups (warps 0..7)
    // H chunk: TMEM -> + b1 -> GELU -> 16 bit -> own slot + peer's slot.  This warp owns rows [32 q, 32 q + 32) of K block
    // kbk (64 hidden columns) of every chunk: one 4 KB slab.
    ws_reg_dec<72>();
    const int quarter = warp & 3, kbk = warp >> 2;
    const uint32_t slab = ptx::smem_u32(smem + kWsHbufOff) + rank * kWsSlotBytes + (uint32_t)kbk * 16384u +
                          (uint32_t)quarter * 4096u;
    const uint32_t slab_peer = ptx::mapa(slab, rank ^ 1u);
    const uint32_t hfull_rem_peer = ptx::mapa(ptx::smem_u32(hfull_rem), rank ^ 1u);
    uint32_t e1n = 0;                                        // own chunks produced so far (all tiles)
    for (int mi = cluster_id; mi < m_tiles; mi += n_clusters) {
#pragma unroll 1
      for (int i = 0; i < NC; ++i, ++e1n) {
        const int j = 2 * i + (int)rank;
        const uint32_t buf = e1n & 1u, use = e1n >> 1;
        TR(40);
        ptx::mbar_wait(&hacc_full[buf], use & 1u);
        ptx::tc_fence_after();
        TR(41);
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + 256u + buf * 128u + (uint32_t)kbk * 64u;
        const float* b1p = p.b1 + j * kWsCH + kbk * 64;
        uint32_t v[32];
        uint32_t pk[16];                                     // gelu(acc + b1) as 16-bit pairs: 32 hidden columns
        auto half_chunk = [&](int h) {                       // hidden columns [32 h, 32 h + 32) of the K block
#pragma unroll
          for (int q = 0; q < 4; ++q) {                      // 16-byte piece: 8 hidden columns
            const float4 c0 = __ldg(reinterpret_cast<const float4*>(b1p + 32 * h + 8 * q));   // warp-uniform, L1-resident
            const float4 c1 = __ldg(reinterpret_cast<const float4*>(b1p + 32 * h + 8 * q + 4));
            float w[8];
            w[0] = __uint_as_float(v[8 * q]) + c0.x;     w[1] = __uint_as_float(v[8 * q + 1]) + c0.y;
            w[2] = __uint_as_float(v[8 * q + 2]) + c0.z; w[3] = __uint_as_float(v[8 * q + 3]) + c0.w;
            w[4] = __uint_as_float(v[8 * q + 4]) + c1.x; w[5] = __uint_as_float(v[8 * q + 5]) + c1.y;
            w[6] = __uint_as_float(v[8 * q + 6]) + c1.z; w[7] = __uint_as_float(v[8 * q + 7]) + c1.w;
            if (!(p.dbg & 1)) {
#pragma unroll
              for (int e = 0; e < 8; ++e) w[e] = gelu_fast(w[e]);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) pk[4 * q + e] = pack_h16x2(w[2 * e], w[2 * e + 1], p.h16_fmt);
          }
        };
        auto store_half = [&](int h) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            ptx::sts_v4u(slab + lane * 128 + (((4 * h + q) ^ (lane & 7)) << 4), pk[4 * q], pk[4 * q + 1], pk[4 * q + 2],
                         pk[4 * q + 3]);
        };
        ptx::tmem_ld_32x32_nowait(taddr, v);
        ptx::tmem_ld_wait_on(v);
        half_chunk(0);
        ptx::tmem_ld_32x32_nowait(taddr + 32, v);
        // the slot is free when both CTAs' GEMM2 of the previous own chunk have retired (the arithmetic above overlaps
        // that turnaround)
        TR(43);
        ptx::mbar_wait(&hempty[rank], (e1n & 1u) ^ 1u);
        TR(44);
        store_half(0);
        ptx::tmem_ld_wait_on(v);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&hacc_empty[buf]);   // GEMM1 of chunk i + 2 may overwrite the accumulator
        half_chunk(1);
        store_half(1);
        ptx::fence_proxy_async();                            // generic-proxy writes -> visible to UMMA / bulk copy
        __syncwarp();
        if (lane == 0) {
          ptx::mbar_arrive(hfull_loc);
          ptx::bulk_copy_to_cluster(slab_peer, slab, 4096, hfull_rem_peer);
        }
        TR(45);
      }
    }
  } else {
    // ================================================================== tail warpgroups (warps 8..15)
    ws_reg_inc<152>();
    const int tw = warp - kWsGeluWarps;
    const int quarter = warp & 3;                            // TMEM lane quarter this warp may access
    const int half = tw >> 2;
    constexpr int CW = kWsBN / 2;                            // output columns per warp
    constexpr int NCH = CW / 32;                             // 4 chunks of 32 columns
    const int row_in_tile = quarter * 32 + lane;
    const uint32_t stg = ptx::smem_u32(smem + kWsStgOff) + (uint32_t)tw * 2048u;
    const uint32_t stats_s = ptx::smem_u32(smem + kWsStatOff);
    const uint32_t stats_cl[2] = {ptx::mapa(stats_s, rank), ptx::mapa(stats_s, rank ^ 1u)};
    const uint32_t sbar_s = ptx::smem_u32(stat_bar);
    const uint32_t sbar_cl[2] = {ptx::mapa(sbar_s, rank), ptx::mapa(sbar_s, rank ^ 1u)};
    const uint32_t part = rank * 2u + (uint32_t)half;
    const int n_warp = n_cta + half * CW;
    const float* b2p = p.b2 + n_warp;                        // warp-uniform addresses: read-only path, L1-resident
    const float* gap = p.gamma + n_warp;
    const float* bep = p.beta + n_warp;
    uint32_t t = 0;
    for (int mi = cluster_id; mi < m_tiles; mi += n_clusters, ++t) {
      const int m_blk = blk_of(mi);
      // The accumulator already holds x + linear2(h) (the residual was added by the identity MMAs).  Pass A moves the
      // warp's 128 columns into REGISTERS (y = acc + b2, running (mean, M2) by Chan's update) and releases the
      // accumulator at once; pass B works from the registers.
      const uint32_t par = t & 1u, use2 = t >> 1;
      const int m_warp = m_blk * kWsBM + quarter * 32;
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * CW);
      if (tw == 0 && lane == 0) ptx::mbar_arrive_expect_tx(&stat_bar[par], 4 * kWsBM * 8);
      TR(50);
      ptx::mbar_wait(y_full, t & 1u);
      ptx::tc_fence_after();
      TR(51);
      float y[CW];                                           // this thread's row, the warp's 128 columns
      float r_mean = 0.f, r_m2 = 0.f;
      {
        uint32_t va[32], vb[32];
        ptx::tmem_ld_32x32_nowait(taddr0, va);
        auto pass_a = [&](int c, uint32_t (&vr)[32]) {
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(b2p + c * 32 + q * 4));
            y[c * 32 + 4 * q] = __uint_as_float(vr[4 * q]) + b.x;
            y[c * 32 + 4 * q + 1] = __uint_as_float(vr[4 * q + 1]) + b.y;
            y[c * 32 + 4 * q + 2] = __uint_as_float(vr[4 * q + 2]) + b.z;
            y[c * 32 + 4 * q + 3] = __uint_as_float(vr[4 * q + 3]) + b.w;
            s0 += y[c * 32 + 4 * q];
            s1 += y[c * 32 + 4 * q + 1];
            s2 += y[c * 32 + 4 * q + 2];
            s3 += y[c * 32 + 4 * q + 3];
          }
          const float cm = ((s0 + s1) + (s2 + s3)) * (1.0f / 32.0f);
          float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float d0 = y[c * 32 + 4 * q] - cm, d1 = y[c * 32 + 4 * q + 1] - cm;
            const float d2 = y[c * 32 + 4 * q + 2] - cm, d3 = y[c * 32 + 4 * q + 3] - cm;
            q0 = fmaf(d0, d0, q0);
            q1 = fmaf(d1, d1, q1);
            q2 = fmaf(d2, d2, q2);
            q3 = fmaf(d3, d3, q3);
          }
          const float cq = (q0 + q1) + (q2 + q3);
          const float na = (float)(c * 32), nb = 32.f, nt = na + nb;
          const float delta = cm - r_mean;
          r_mean += delta * (nb / nt);
          r_m2 += cq + delta * delta * (na * nb / nt);
        };
        // TMEM loads run one chunk ahead of the arithmetic (NCH = 4)
        ptx::tmem_ld_wait_on(va);
        ptx::tmem_ld_32x32_nowait(taddr0 + 32, vb);
        pass_a(0, va);
        ptx::tmem_ld_wait_on(vb);
        ptx::tmem_ld_32x32_nowait(taddr0 + 64, va);
        pass_a(1, vb);
        ptx::tmem_ld_wait_on(va);
        ptx::tmem_ld_32x32_nowait(taddr0 + 96, vb);
        pass_a(2, va);
        ptx::tmem_ld_wait_on(vb);
        ptx::tc_fence_before();                              // the output accumulator is fully consumed: the next tile's
        __syncwarp();                                        // residual MMAs and GEMM2 may overwrite it
        if (lane == 0) ptx::mbar_arrive(y_empty);
        pass_a(3, vb);
      }
      TR(52);
      // ---- exchange of the four partial (mean, M2) of every row (2 column halves x 2 CTAs)
      {
        const uint32_t off = ((par * 4u + part) * kWsBM + (uint32_t)row_in_tile) * 8u;
        ptx::st_async_v2(stats_cl[0] + off, r_mean, r_m2, sbar_cl[0] + par * 8u);
        ptx::st_async_v2(stats_cl[1] + off, r_mean, r_m2, sbar_cl[1] + par * 8u);
      }
      ptx::mbar_wait(&stat_bar[par], use2 & 1u);
      TR(53);
      float mean = 0.f, m2 = 0.f;
      {
        float pm[4], pq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t a = stats_s + ((par * 4u + (uint32_t)q) * kWsBM + (uint32_t)row_in_tile) * 8u;
          asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(pm[q]), "=f"(pq[q]) : "r"(a) : "memory");
          mean += pm[q];
        }
        mean *= 0.25f;
#pragma unroll
        for (int q = 0; q < 4; ++q) m2 += pq[q] + (float)CW * (pm[q] - mean) * (pm[q] - mean);
      }
      const float rstd = rsqrtf(m2 * (1.0f / (float)kWsD) + p.eps);
      const float shift = -mean * rstd;

      // ---- pass B from the registers: normalise, hi = round16(z), lo = round16(z - hi); hi, then lo, go through the
      // warp's own 2 KB staging block (one TMA store each).
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 g = __ldg(reinterpret_cast<const float4*>(gap + c * 32 + q * 4));
          const float4 b = __ldg(reinterpret_cast<const float4*>(bep + c * 32 + q * 4));
          y[c * 32 + 4 * q] = fmaf(fmaf(y[c * 32 + 4 * q], rstd, shift), g.x, b.x);
          y[c * 32 + 4 * q + 1] = fmaf(fmaf(y[c * 32 + 4 * q + 1], rstd, shift), g.y, b.y);
          y[c * 32 + 4 * q + 2] = fmaf(fmaf(y[c * 32 + 4 * q + 2], rstd, shift), g.z, b.z);
          y[c * 32 + 4 * q + 3] = fmaf(fmaf(y[c * 32 + 4 * q + 3], rstd, shift), g.w, b.w);
        }
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {               // 0: hi, 1: lo
          if (pass == 1 && !p.has_lo) break;
          if (lane == 0) ptx::bulk_wait_read_all();          // the block's previous store has read it
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float a = y[c * 32 + 8 * q + 2 * e], b = y[c * 32 + 8 * q + 2 * e + 1];
              const uint32_t hp = pack_h16x2(a, b, p.h16_fmt);
              if (pass == 0) {
                w[e] = hp;
              } else {
                const float2 f = unpack_h16x2(hp, p.h16_fmt);
                w[e] = pack_h16x2(a - f.x, b - f.y, p.h16_fmt);
              }
            }
            ptx::sts_v4u(stg + (uint32_t)lane * 64 + ((uint32_t)(q ^ ((lane >> 1) & 3)) << 4), w[0], w[1], w[2], w[3]);
          }
          ptx::fence_proxy_async();
          __syncwarp();
          if (lane == 0 && !(p.dbg & 2)) {
            ptx::tma_store_2d(pass == 0 ? &tmap_ohi : &tmap_olo, stg, n_warp + c * 32, m_warp);
            ptx::bulk_commit();
          }
        }
      }
      TR(54);
    }
    if (lane == 0) ptx::bulk_wait_all();                     // global writes complete before the CTA retires
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();                                   // no CTA exits while its peer may still write its smem
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// Called by gd_ffn_h16_ln_split (ffn_ln.cu), which has validated the arguments.
int launch_ffn_ln_ws(const void* x_hi, const void* x_lo, int ld_x, const void* W1, int ldw1, const float* b1,
                     const void* W2, int ldw2, const float* b2, int M, int N, int F, int fmt, const float* gamma,
                     const float* beta, float eps, void* out_hi, void* out_lo, int ld_out, int dbg, cudaStream_t stream) {
  WsParams p;
  p.b1 = b1; p.b2 = b2; p.gamma = gamma; p.beta = beta;
  p.has_lo = out_lo != nullptr;
  p.h16_fmt = fmt;
  p.eps = eps; p.M = M; p.F = F;
  p.trace = g_ffn_trace;
  p.dbg = dbg;
  p.idesc1 = ptx::umma_idesc_h16_f32(kWsBM, kWsCH, fmt == GD_FP16);
  p.idesc2 = ptx::umma_idesc_h16_f32(kWsBM, kWsBN, fmt == GD_FP16);
  p.idesc3 = ptx::umma_idesc_h16_f32(kWsBM, kWsBK, fmt == GD_FP16);
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(ffn_ln_ws_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWsSmemBytes));
    attr_set = true;
  }
  CUtensorMap tx, tw1, tw2, txl, toh, tol;
  int rc = make_tmap_2d(&tx, x_hi, M, N, ld_x, kWsBM, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw1, W1, F, N, ldw1, kWsCH, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw2, W2, N, F, ldw2, kWsBN, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&txl, x_lo, M, N, ld_x, kWsBM, fmt);
  if (rc) return rc;
  rc = make_tmap_2d_h16_sw64(&toh, out_hi, M, N, ld_out, 32, fmt);
  if (rc) return rc;
  tol = toh;
  if (out_lo) {
    rc = make_tmap_2d_h16_sw64(&tol, out_lo, M, N, ld_out, 32, fmt);
    if (rc) return rc;
  }
  const int sms = num_sms();
  GD_REQUIRE(sms >= 2, "gd_ffn_h16_ln_split: cannot query the SM count");
  const int m_tiles = ceil_div(M, kWsBM);
  const int clusters = m_tiles < sms / 2 ? m_tiles : sms / 2;
  GD_CUDA_OK(launch_pdl(PDL_GEMM_LN, ffn_ln_ws_tcgen05_kernel, dim3(2 * clusters), dim3(kWsThreads), kWsSmemBytes, stream, tx,
                        tw1, tw2, txl, toh, tol, p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

}  // namespace gd
