// ffn_ln.cu -- the whole feed-forward sub-layer of nn.TransformerEncoderLayer (post-norm, mdm.py:90-96) in ONE
// tcgen05 kernel:
//   X' = LayerNorm(X + gelu(X_hi @ W1^T + b1) @ W2^T + b2) * gamma + beta
// with the residual stream carried as the 16-bit pair X = X_hi + X_lo (gemm_ln.cu, MODE 2) and X_hi doubling as
// the GEMM operand.  The hidden activation H = gelu(linear1) [M, F] never goes to HBM (it was one 16-bit write
// and one 16-bit read of M x F per layer: 103 MB of the 258 MB the two separate kernels moved at the cfg2a shapes).
//
// A LayerNorm row spans N = d = 512 columns = two 256-column fp32 accumulators, so a 128-row tile is owned by the
// TWO CTAs of a cluster (as in gemm_ln.cu): CTA r accumulates output columns [256 r, 256 r + 256) in TMEM columns
// 0..255 and needs the tile's complete H for it.  H is produced in chunks of 128 hidden columns, chunk j by CTA
// j % 2: GEMM1 (M 128 x N 128 x K 512) into one of two 128-column TMEM accumulators -> the epilogue warps add b1,
// apply erf-GELU, round to 16 bit and write the chunk as the K-major 128B-swizzled A operand of GEMM2 into their
// own shared memory -> one bulk copy per warp (cp.async.bulk shared::cta -> shared::cluster, 4 KB slabs) puts the
// same bytes into the peer's shared memory and completes transaction bytes on the peer's barrier.  GEMM2
// (M 128 x N 256 x K 128 per chunk) then runs in both CTAs from shared memory.  A chunk slot is recycled when the
// GEMM2 MMAs of BOTH CTAs have retired: each CTA's tcgen05.commit multicasts its arrival to the pair.
//
// The residual is added on the tensor pipe: at the start of a tile "Y[:, 64 s ..] = X_hi I^T + X_lo I^T"
// with a 64 x 64 identity as the B operand (products with 1.0 and fp32 accumulation: the same sum the explicit add
// makes), so the LayerNorm tail only READS the accumulator in pass A and needs no residual loads.
//
// Roles (320 threads): warps 0..7 = epilogue (H chunks during the tile, then the LayerNorm tail of gemm_ln.cu:
// pass A statistics, exchange by st.async, pass B normalise -> swizzled staging -> TMA stores of hi / lo; the staging
// block of a warp is its own slab of the locally produced H slot), warp 8 =
// TMA producer (one ring of four 32 KB stages: [X 128x64 | W1 128x64] for GEMM1, [W2 256x64] for GEMM2,
// [X_hi 128x64 | X_lo 128x64] for the residual, in exactly the order the MMA thread consumes them), warp 9 = UMMA
// issuer.
//
// Status: parity-green and the default feed-forward path (GD_FUSE_FFN=0 selects the two
// kernels).  At d = 512, F = 1024, M = 25 216 it takes 69.5 us
// against 79.5 us for gd_gemm_h16 (linear1 + GELU) + gd_gemm_h16_ln_split in the loop benchmark and makes the
// cfg2a step 2-3 % shorter (1.497-1.529 vs 1.539-1.571 ms on the same boxes).  The in-kernel timeline
// (gd_ffn_trace, tools/trace_ffn.py) shows the limits: TMEM (Y 256 columns + two 128-column chunk accumulators = all
// 512 columns) forces N = 128 for GEMM1 (three quarters of the N = 256 rate), four 32 KB operand stages deliver
// ~57 B/clk, and the GELU chunks + LayerNorm tail keep the eight epilogue warps busy ~20 k of a tile's ~38 k clocks.
// Round 2 measurements (profiles/r02c_ffn_dbg_knobs.txt, r02f_trace_ffn_stg_passB.txt): without ANY operand loads the
// kernel still takes 64 of 71 us and without loads, GELU arithmetic, stores and H exchange 56 us, so the tile is paced
// by its dependency chain (GEMM1 -> GELU -> exchange -> GEMM2 -> LayerNorm tail on ONE set of epilogue warps with a
// single-buffered output accumulator), not by the feed.  The GELU arithmetic is issue-bound: the 9-instruction
// gelu_fast took the kernel from 71 to 66 us.  Replacing pass B's TMA stores by coalesced generic stores was slower
// (pass B 8.2 k vs 7.7 k clocks, kernel 71 vs 66 us) and was dropped; so was a ring of three 2 KB staging pieces with
// three TMA stores in flight per warp (pass B 6.2-6.8 k clocks, kernel 68.3 us: no gain, profiles/r02k_trace_ffn_ring.txt)
// -- pass B costs about 5.7 k clocks without any store (GD_FFN_DBG=2): 4 x (TMEM load, 24 LDS.128, ~150 FP / convert
// instructions, fence.proxy.async) on eight warps is latency-bound, and the single-buffered output accumulator keeps it on
// the tile's critical path.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kFfBM = 128, kFfBK = 64, kFfThreads = 320, kFfEpiWarps = 8;
// warps 0..7 = epilogue; the two single-lane control warps get the HIGHEST warp ids of their scheduler partitions
// (the arbiter favours them over the GELU-heavy epilogue warps, gemm_tcgen05.cu)
constexpr int kFfProducerWarp = kFfEpiWarps, kFfMmaWarp = kFfEpiWarps + 1;
constexpr int kFfD = 512;                      // model width: K of GEMM1, N of GEMM2 (two CTAs x 256 columns)
constexpr int kFfBN = 256;                     // output columns per CTA
constexpr int kFfCH = 128;                     // hidden columns per chunk
// Shared memory buys either a dedicated second staging block per epilogue warp (pass B never waits for the previous
// chunk's store) or a fourth operand stage; the operand feed paces the tile, so it gets the memory:
// 75.3 -> 69.5 us.
constexpr bool kFfDedicatedLn = false;
constexpr int kFfStages = kFfDedicatedLn ? 3 : 4;
constexpr int kFfStageBytes = 32768;
constexpr int kFfSlotBytes = 32768;            // one H chunk: two [128 rows x 64] 128B-swizzled K blocks
constexpr int kFfRingOff = 0;
constexpr int kFfHbufOff = kFfStages * kFfStageBytes;
constexpr int kFfLnOff = kFfHbufOff + 2 * kFfSlotBytes;            // dedicated 4 KB block per epilogue warp
constexpr int kFfI64Off = kFfLnOff + (kFfDedicatedLn ? kFfEpiWarps * 4096 : 0);           // 64 x 64 identity, K-major 128B swizzle (8 KB)
constexpr int kFfVecOff = kFfI64Off + 8192;                        // b2 | gamma | beta slices of this CTA
constexpr int kFfStatOff = kFfVecOff + 3 * kFfBN * 4;              // [parity][part][row] (mean, M2)
constexpr int kFfBarOff = kFfStatOff + 2 * 4 * kFfBM * 8;
constexpr int kFfSmemBytes = kFfBarOff + 512;

struct FfParams {
  const float* b1; const float* b2; const float* gamma; const float* beta;
  int has_lo;        // the lo half of the result is wanted (not the last sub-layer)
  int h16_fmt;
  float eps;
  int M, F;
  uint32_t idesc1, idesc2, idesc3;
  unsigned long long* trace;   // optional timeline buffer (gd_ffn_trace): 3 roles x 4096 entries of (tag << 48 | clock)
  int dbg;           // GD_FFN_DBG diagnostics (results wrong): 1 no GELU arithmetic, 2 no global stores, 4 no X loads,
                     // 8 no operand loads, 16 no H exchange between the CTAs, 32 no MMAs; 256 (results right): row tiles
                     // in descending order (measured: 0.4 % slower step, the producer's last rows are not what L2 keeps)
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFfThreads, 1)
ffn_ln_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                      const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_xlo,
                      const __grid_constant__ CUtensorMap tmap_ohi, const __grid_constant__ CUtensorMap tmap_olo,
                      const FfParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* ring = smem + kFfRingOff;
  float* vecs = reinterpret_cast<float*>(smem + kFfVecOff);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kFfBarOff);
  uint64_t* full_bar = bars;                                 // [stages]
  uint64_t* empty_bar = full_bar + kFfStages;                // [stages]
  uint64_t* hacc_full = empty_bar + kFfStages;               // [2]  GEMM1 chunk accumulator complete
  uint64_t* hacc_empty = hacc_full + 2;                      // [2]  ... drained by the epilogue warps
  uint64_t* hfull_loc = hacc_empty + 2;                      // locally produced H chunk written (8 warp arrivals)
  uint64_t* hfull_rem = hfull_loc + 1;                       // peer's H chunk has landed (transaction bytes)
  uint64_t* hempty = hfull_rem + 1;                          // [2]  slot's GEMM2 retired in BOTH CTAs (2 commits)
  uint64_t* y_full = hempty + 2;
  uint64_t* y_empty = y_full + 1;
  uint64_t* stat_bar = y_empty + 1;                          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + 2);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  const uint32_t rank = ptx::cluster_ctarank();
  // timeline of CTA 0: role 0 = epilogue warp 2 lane 0, role 1 = MMA thread, role 2 = TMA producer
  int tr_n = 0;
  unsigned long long* tr_buf = nullptr;
  if (p.trace != nullptr && blockIdx.x == 0 && lane == 0 && (warp == 0 || warp >= kFfEpiWarps))
    tr_buf = p.trace + (warp == 0 ? 0 : (warp == kFfMmaWarp ? 4096 : 8192));
  auto TR = [&](unsigned tag) {
    if (tr_buf != nullptr && tr_n < 4096) tr_buf[tr_n++] = ((unsigned long long)tag << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull);
  };
  const int n_clusters = (int)gridDim.x >> 1, cluster_id = (int)blockIdx.x >> 1;
  const int M = p.M;
  const int m_tiles = (M + kFfBM - 1) / kFfBM;
  const int NC = p.F / (2 * kFfCH);                          // chunks this CTA produces per tile (>= 2)
  const int n_cta = (int)rank * kFfBN;
  constexpr int kKB1 = kFfD / kFfBK;                         // 8 K blocks per GEMM1 chunk
  constexpr int kKB2 = kFfCH / kFfBK;                        // 2 K blocks per GEMM2 chunk

  if (warp == kFfProducerWarp && lane == 0) {
    ptx::prefetch_tensormap(&tmap_x);
    ptx::prefetch_tensormap(&tmap_w1);
    ptx::prefetch_tensormap(&tmap_w2);
    ptx::prefetch_tensormap(&tmap_xlo);
    ptx::prefetch_tensormap(&tmap_ohi);
    ptx::prefetch_tensormap(&tmap_olo);
  }
  if (warp == kFfMmaWarp && lane == 0) {
    for (int s = 0; s < kFfStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(&hacc_full[b], 1);
      ptx::mbar_init(&hacc_empty[b], kFfEpiWarps);
      ptx::mbar_init(&hempty[b], 2);
      ptx::mbar_init(&stat_bar[b], 1);
    }
    ptx::mbar_init(hfull_loc, kFfEpiWarps);
    ptx::mbar_init(hfull_rem, 1);
    ptx::mbar_init(y_full, 1);
    ptx::mbar_init(y_empty, kFfEpiWarps);
    ptx::fence_barrier_init();
  }
  if (warp == 0) ptx::tmem_alloc(tmem_slot, 512);
  for (int i = threadIdx.x; i < kFfBN; i += kFfThreads) {
    vecs[i] = p.b2[n_cta + i];
    vecs[kFfBN + i] = p.gamma[n_cta + i];
    vecs[2 * kFfBN + i] = p.beta[n_cta + i];
  }
  {
    // I64[n][k] = (n == k) as a K-major 128B-swizzled UMMA B operand: row n = 128 bytes, 16-byte chunk c (k = 8c..8c+7)
    // at position c ^ (n & 7).  "Y += X I^T" adds the residual on the tensor pipe (exact: products with 1.0).
    const uint32_t one = (p.h16_fmt == GD_FP16) ? 0x3C00u : 0x3F80u;
    const uint32_t i64_s = ptx::smem_u32(smem + kFfI64Off);
    for (int idx = threadIdx.x; idx < 64 * 8; idx += kFfThreads) {
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

  if (warp == kFfProducerWarp) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      auto load_g1 = [&](int m_blk, int i) {                 // own chunk i of the tile: global chunk 2 i + rank
        const int j = 2 * i + (int)rank;
        TR(1);
        for (int kb = 0; kb < kKB1; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * kFfStageBytes;
          if (p.dbg & 8) {                                   // diagnostics: no operand traffic at all
            ptx::mbar_arrive(&full_bar[stage]);
          } else if (p.dbg & 4) {                            // diagnostics: GEMM1 without the X block
            ptx::mbar_arrive_expect_tx(&full_bar[stage], 16384);
            ptx::tma_load_2d(sa + 16384, &tmap_w1, &full_bar[stage], kb * kFfBK, j * kFfCH);
          } else {
            ptx::mbar_arrive_expect_tx(&full_bar[stage], kFfStageBytes);
            ptx::tma_load_2d(sa, &tmap_x, &full_bar[stage], kb * kFfBK, m_blk * kFfBM);
            ptx::tma_load_2d(sa + 16384, &tmap_w1, &full_bar[stage], kb * kFfBK, j * kFfCH);
          }
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
      };
      auto load_g2 = [&](int j) {                            // W2[n_cta .. +256, chunk j's 128 hidden columns]
        TR(2);
        for (int kb = 0; kb < kKB2; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * kFfStageBytes;
          if (p.dbg & 8) {
            ptx::mbar_arrive(&full_bar[stage]);
          } else {
            ptx::mbar_arrive_expect_tx(&full_bar[stage], kFfStageBytes);
            ptx::tma_load_2d(sa, &tmap_w2, &full_bar[stage], j * kFfCH + kb * kFfBK, n_cta);
          }
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
      };
      auto load_res = [&](int m_blk) {                       // [X_hi | X_lo] blocks of this CTA's 256 output columns
        for (int sb = 0; sb < kFfBN / kFfBK; ++sb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * kFfStageBytes;
          ptx::mbar_arrive_expect_tx(&full_bar[stage], kFfStageBytes);
          ptx::tma_load_2d(sa, &tmap_x, &full_bar[stage], n_cta + sb * kFfBK, m_blk * kFfBM);
          ptx::tma_load_2d(sa + 16384, &tmap_xlo, &full_bar[stage], n_cta + sb * kFfBK, m_blk * kFfBM);
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
      };
      for (int mi = cluster_id; mi < m_tiles; mi += n_clusters) {
        const int m_blk = (p.dbg & 256) ? m_tiles - 1 - mi : mi;
        load_g1(m_blk, 0);
        load_g1(m_blk, 1);
        load_res(m_blk);
        for (int i = 0; i < NC; ++i) {                       // same order as the MMA thread
          load_g2(2 * i + (int)rank);
          load_g2(2 * i + 1 - (int)rank);
          if (i + 2 < NC) load_g1(m_blk, i + 2);
        }
      }
    }
  } else if (warp == kFfMmaWarp) {
    // ------------------------------------------------------------------ UMMA issuer
    // the whole warp runs the schedule (uniform control flow, barrier waits by all lanes); one elected lane issues
    {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t g1n = 0;                                      // GEMM1 chunks issued so far (all tiles)
      uint32_t pairn = 0;                                    // chunk pairs consumed so far (all tiles)
      const uint32_t smem_u = __shfl_sync(0xffffffffu, ptx::smem_u32(smem), 0);
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t hbuf_s = smem_u + kFfHbufOff;
      auto g1 = [&]() {
        const uint32_t buf = g1n & 1u, use = g1n >> 1;
        TR(10);
        ptx::mbar_wait(&hacc_empty[buf], (use & 1u) ^ 1u);   // the epilogue has drained this accumulator
        ptx::tc_fence_after();
        TR(11);
        const uint32_t tmem_d = tmem_u + 256u + buf * 128u;
        for (int kb = 0; kb < kKB1; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(kFfRingOff + stage * kFfStageBytes);
          const uint32_t sb = sa + 16384;
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kFfBK / 16; ++k)
              if (!(p.dbg & 32)) ptx::umma_h16(tmem_d, ptx::umma_desc_k_sw128(sa + k * 32), ptx::umma_desc_k_sw128(sb + k * 32), p.idesc1,
                            (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit(&hacc_full[buf]);
        __syncwarp();
        TR(12);
        ++g1n;
      };
      auto g2 = [&](int j, bool first) {
        const uint32_t slot = (uint32_t)j & 1u;
        TR(20);
        if (slot == rank) {
          ptx::mbar_wait(hfull_loc, pairn & 1u);
        } else {
          if (!(p.dbg & 16)) {
            if (ptx::elect_one()) ptx::mbar_arrive_expect_tx(hfull_rem, kFfSlotBytes);   // the peer's 8 slabs are the transactions
            __syncwarp();
            ptx::mbar_wait(hfull_rem, pairn & 1u);
          }
        }
        ptx::tc_fence_after();
        TR(21);
        for (int kb = 0; kb < kKB2; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = hbuf_s + slot * kFfSlotBytes + (uint32_t)kb * 16384u;
          const uint32_t sb = smem_u + (uint32_t)(kFfRingOff + stage * kFfStageBytes);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kFfBK / 16; ++k)
              if (!(p.dbg & 32)) ptx::umma_h16(tmem_u, ptx::umma_desc_k_sw128(sa + k * 32), ptx::umma_desc_k_sw128(sb + k * 32), p.idesc2,
                            (first && kb == 0 && k == 0) ? 0u : 1u);
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit_mc(&hempty[slot], 3);   // this CTA is done with the slot: tell both CTAs
        __syncwarp();
        TR(22);
      };
      const uint32_t i64_s = smem_u + kFfI64Off;
      auto gres = [&]() {                                    // Y[:, 64 sb ..] = X_hi I^T + X_lo I^T: the residual opens the tile
        for (int sb = 0; sb < kFfBN / kFfBK; ++sb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(kFfRingOff + stage * kFfStageBytes);
          if (ptx::elect_one()) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
              for (int k = 0; k < kFfBK / 16; ++k)
                ptx::umma_h16(tmem_u + (uint32_t)(sb * kFfBK), ptx::umma_desc_k_sw128(sa + h * 16384 + k * 32),
                              ptx::umma_desc_k_sw128(i64_s + k * 32), p.idesc3, (h | k) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == kFfStages) { stage = 0; phase ^= 1; }
        }
      };
      uint32_t t = 0;
      for (int m_blk = cluster_id; m_blk < m_tiles; m_blk += n_clusters, ++t) {
        g1();
        g1();
        // The residual MMAs initialise the output accumulator while the epilogue warps work on the tile's first H
        // chunk (the tensor pipe would wait for it anyway); every GEMM2 MMA then accumulates.
        ptx::mbar_wait(y_empty, (t & 1u) ^ 1u);              // the LayerNorm tail of the previous tile has left TMEM
        ptx::tc_fence_after();
        gres();
        for (int i = 0; i < NC; ++i) {
          // the slot turnaround (GEMM2 retired in both CTAs -> next chunk may be written) is the critical chain:
          // GEMM2 goes first, the locally produced chunk (no DSMEM hop) before the peer's; GEMM1 of chunk i + 2
          // has slack (double-buffered accumulator) and fills the tensor pipe behind it
          g2(2 * i + (int)rank, false);
          g2(2 * i + 1 - (int)rank, false);
          ++pairn;
          if (i + 2 < NC) g1();
        }
        if (ptx::elect_one()) ptx::umma_commit(y_full);
        __syncwarp();
        TR(30);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp;
    const int quarter = warp & 3;                            // TMEM lane quarter this warp may access
    const int half = ew >> 2;
    constexpr int CW = kFfBN / 2;                            // output columns per warp
    constexpr int NCH = CW / 32;                             // 4 chunks of 32 columns
    const int row_in_tile = quarter * 32 + lane;
    // H chunk: this thread owns row row_in_tile, hidden columns [64 half, 64 half + 64) of the chunk = one 128-byte
    // row of K block `half`; the warp's 32 rows are one contiguous 4 KB slab.
    const uint32_t slab = ptx::smem_u32(smem + kFfHbufOff) + rank * kFfSlotBytes + (uint32_t)half * 16384u +
                          (uint32_t)quarter * 4096u;
    const uint32_t slab_peer = ptx::mapa(slab, rank ^ 1u);
    const uint32_t hfull_rem_peer = ptx::mapa(ptx::smem_u32(hfull_rem), rank ^ 1u);
    // LayerNorm tail: block 0 (dedicated) and block 1 (= the slab) stage the outputs of alternate column chunks
    const uint32_t blk0 = ptx::smem_u32(smem + kFfLnOff) + (uint32_t)ew * 4096u;
    auto blk = [&](int c) { return (!kFfDedicatedLn || (c & 1)) ? slab : blk0; };
    const uint32_t vec_s = ptx::smem_u32(vecs) + (uint32_t)(half * CW) * 4;
    const uint32_t stats_s = ptx::smem_u32(smem + kFfStatOff);
    const uint32_t stats_cl[2] = {ptx::mapa(stats_s, rank), ptx::mapa(stats_s, rank ^ 1u)};
    const uint32_t sbar_s = ptx::smem_u32(stat_bar);
    const uint32_t sbar_cl[2] = {ptx::mapa(sbar_s, rank), ptx::mapa(sbar_s, rank ^ 1u)};
    const uint32_t part = rank * 2u + (uint32_t)half;
    const int n_warp = n_cta + half * CW;
    uint32_t t = 0;
    for (int mi = cluster_id; mi < m_tiles; mi += n_clusters, ++t) {
      const int m_blk = (p.dbg & 256) ? m_tiles - 1 - mi : mi;
      // ================= H chunks of this tile: TMEM -> + b1 -> GELU -> 16 bit -> own slot + peer's slot
      for (int i = 0; i < NC; ++i) {
        const int j = 2 * i + (int)rank;
        const uint32_t e1n = t * (uint32_t)NC + (uint32_t)i;
        const uint32_t buf = e1n & 1u, use = e1n >> 1;
        float4 bq[16];                                       // b1 of this thread's 64 columns (warp-uniform loads)
        {
          const float* b1p = p.b1 + j * kFfCH + half * 64;
#pragma unroll
          for (int q = 0; q < 16; ++q) bq[q] = __ldg(reinterpret_cast<const float4*>(b1p + 4 * q));
        }
        TR(40);
        ptx::mbar_wait(&hacc_full[buf], use & 1u);
        ptx::tc_fence_after();
        TR(41);
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + 256u + buf * 128u + (uint32_t)half * 64u;
        uint32_t v0[32], v1[32];
        ptx::tmem_ld_32x32_nowait(taddr, v0);
        ptx::tmem_ld_32x32_nowait(taddr + 32, v1);
        ptx::tmem_ld_wait_on(v0);
        ptx::tmem_ld_wait_on(v1);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&hacc_empty[buf]);   // GEMM1 of chunk i + 2 may overwrite the accumulator
        TR(42);
        uint32_t pk[32];                                     // gelu(acc + b1) as 16-bit pairs: 64 hidden columns
#pragma unroll
        for (int q = 0; q < 8; ++q) {                        // 16-byte chunk q = hidden columns 8q .. 8q+7 of the 64
          const uint32_t* src = (q < 4) ? &v0[8 * q] : &v1[8 * (q - 4)];
          const float4 c0 = bq[2 * q], c1 = bq[2 * q + 1];
          float w[8];
          w[0] = __uint_as_float(src[0]) + c0.x; w[1] = __uint_as_float(src[1]) + c0.y;
          w[2] = __uint_as_float(src[2]) + c0.z; w[3] = __uint_as_float(src[3]) + c0.w;
          w[4] = __uint_as_float(src[4]) + c1.x; w[5] = __uint_as_float(src[5]) + c1.y;
          w[6] = __uint_as_float(src[6]) + c1.z; w[7] = __uint_as_float(src[7]) + c1.w;
          if (!(p.dbg & 1)) {
#pragma unroll
            for (int e = 0; e < 8; ++e) w[e] = gelu_fast(w[e]);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) pk[4 * q + e] = pack_h16x2(w[2 * e], w[2 * e + 1], p.h16_fmt);
        }
        // the slot is free when both CTAs' GEMM2 of the previous own chunk have retired (the arithmetic above
        // overlaps that turnaround)
        TR(43);
        ptx::mbar_wait(&hempty[rank], (e1n & 1u) ^ 1u);
        TR(44);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          ptx::sts_v4u(slab + lane * 128 + ((q ^ (lane & 7)) << 4), pk[4 * q], pk[4 * q + 1], pk[4 * q + 2],
                       pk[4 * q + 3]);
        ptx::fence_proxy_async();                            // generic-proxy writes -> visible to UMMA / bulk copy
        __syncwarp();
        if (lane == 0) {
          ptx::mbar_arrive(hfull_loc);
          if (!(p.dbg & 16)) ptx::bulk_copy_to_cluster(slab_peer, slab, 4096, hfull_rem_peer);
        }
        TR(45);
      }

      // ================= LayerNorm tail of the tile.  The accumulator already holds x + linear2(h) (the residual was
      // added by the identity MMAs), so pass A only reads: y = acc + b2, running (mean, M2) by Chan's update.
      const uint32_t par = t & 1u, use2 = t >> 1;
      const int m_warp = m_blk * kFfBM + quarter * 32;
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * CW);
      if (ew == 0 && lane == 0) ptx::mbar_arrive_expect_tx(&stat_bar[par], 4 * kFfBM * 8);
      TR(50);
      ptx::mbar_wait(y_full, t & 1u);
      ptx::tc_fence_after();
      TR(51);
      float r_mean = 0.f, r_m2 = 0.f;
      {
        uint32_t va[32], vb[32];
        ptx::tmem_ld_32x32_nowait(taddr0, va);
        auto pass_a = [&](int c, uint32_t (&vr)[32]) {
          float v[32];
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b = ptx::lds_v4(vec_s + (c * 32 + q * 4) * 4);
            v[4 * q] = __uint_as_float(vr[4 * q]) + b.x;
            v[4 * q + 1] = __uint_as_float(vr[4 * q + 1]) + b.y;
            v[4 * q + 2] = __uint_as_float(vr[4 * q + 2]) + b.z;
            v[4 * q + 3] = __uint_as_float(vr[4 * q + 3]) + b.w;
            s0 += v[4 * q];
            s1 += v[4 * q + 1];
            s2 += v[4 * q + 2];
            s3 += v[4 * q + 3];
          }
          const float cm = ((s0 + s1) + (s2 + s3)) * (1.0f / 32.0f);
          float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float d0 = v[4 * q] - cm, d1 = v[4 * q + 1] - cm, d2 = v[4 * q + 2] - cm, d3 = v[4 * q + 3] - cm;
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
        pass_a(3, vb);
      }

      // ---- exchange of the four partial (mean, M2) of every row (2 column halves x 2 CTAs)
      {
        const uint32_t off = ((par * 4u + part) * kFfBM + (uint32_t)row_in_tile) * 8u;
        ptx::st_async_v2(stats_cl[0] + off, r_mean, r_m2, sbar_cl[0] + par * 8u);
        ptx::st_async_v2(stats_cl[1] + off, r_mean, r_m2, sbar_cl[1] + par * 8u);
      }
      // block 1 is this warp's slab of the local H slot: its last chunk has been consumed by both CTAs' GEMM2 (which
      // also means the bulk copy out of it has completed)
      TR(52);
      ptx::mbar_wait(&hempty[rank], (t * (uint32_t)NC + (uint32_t)NC - 1u) & 1u);
      ptx::mbar_wait(&stat_bar[par], use2 & 1u);
      TR(53);
      float mean = 0.f, m2 = 0.f;
      {
        float pm[4], pq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t a = stats_s + ((par * 4u + (uint32_t)q) * kFfBM + (uint32_t)row_in_tile) * 8u;
          asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(pm[q]), "=f"(pq[q]) : "r"(a) : "memory");
          mean += pm[q];
        }
        mean *= 0.25f;
#pragma unroll
        for (int q = 0; q < 4; ++q) m2 += pq[q] + (float)CW * (pm[q] - mean) * (pm[q] - mean);
      }
      const float rstd = rsqrtf(m2 * (1.0f / (float)kFfD) + p.eps);
      const float shift = -mean * rstd;

      // ---- pass B: normalise from TMEM, hi = round16(z), lo = round16(z - hi), staged in block c & 1
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        float v[32];
        ptx::tmem_ld_32x32(taddr0 + c * 32, v);
        if (c == NCH - 1) {                                  // the output accumulator is fully consumed
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(y_empty);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b2 = ptx::lds_v4(vec_s + (c * 32 + q * 4) * 4);
          const float4 g = ptx::lds_v4(vec_s + (kFfBN + c * 32 + q * 4) * 4);
          const float4 b = ptx::lds_v4(vec_s + (2 * kFfBN + c * 32 + q * 4) * 4);
          v[4 * q] = fmaf(fmaf(v[4 * q] + b2.x, rstd, shift), g.x, b.x);
          v[4 * q + 1] = fmaf(fmaf(v[4 * q + 1] + b2.y, rstd, shift), g.y, b.y);
          v[4 * q + 2] = fmaf(fmaf(v[4 * q + 2] + b2.z, rstd, shift), g.z, b.z);
          v[4 * q + 3] = fmaf(fmaf(v[4 * q + 3] + b2.w, rstd, shift), g.w, b.w);
        }
        if (lane == 0) {                                     // the block's previous store has read it
          if (kFfDedicatedLn) ptx::bulk_wait_read_1();
          else ptx::bulk_wait_read_all();
        }
        __syncwarp();
        const uint32_t sth = blk(c), stl = sth + 2048;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t hp[4], lp[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float a = v[8 * q + 2 * e], b = v[8 * q + 2 * e + 1];
            hp[e] = pack_h16x2(a, b, p.h16_fmt);
            const float2 f = unpack_h16x2(hp[e], p.h16_fmt);
            lp[e] = pack_h16x2(a - f.x, b - f.y, p.h16_fmt);
          }
          const uint32_t off = (uint32_t)lane * 64 + ((uint32_t)(q ^ ((lane >> 1) & 3)) << 4);
          ptx::sts_v4u(sth + off, hp[0], hp[1], hp[2], hp[3]);
          if (p.has_lo) ptx::sts_v4u(stl + off, lp[0], lp[1], lp[2], lp[3]);
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0 && !(p.dbg & 2)) {
          ptx::tma_store_2d(&tmap_ohi, sth, n_warp + c * 32, m_warp);
          if (p.has_lo) ptx::tma_store_2d(&tmap_olo, stl, n_warp + c * 32, m_warp);
          ptx::bulk_commit();
        }
      }
      TR(54);
      if (lane == 0) ptx::bulk_wait_read_all();              // staging read out before the slab takes the next tile's chunk
      __syncwarp();
      TR(55);
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

}  // namespace gd

namespace gd { unsigned long long* g_ffn_trace = nullptr; }
using gd::g_ffn_trace;
/* Debug only: timeline buffer (device, >= 3 * 4096 uint64) filled by CTA 0 of every later gd_ffn_h16_ln_split launch;
 * NULL turns it off. */
extern "C" int gd_ffn_trace(void* buf) {
  g_ffn_trace = static_cast<unsigned long long*>(buf);
  return GD_OK;
}

extern "C" int gd_ffn_h16_ln_split(const void* x_hi, const void* x_lo, int32_t ld_x, const void* W1, int32_t ldw1,
                                   const float* b1, const void* W2, int32_t ldw2, const float* b2, int32_t M,
                                   int32_t N, int32_t F, int32_t fmt, const float* gamma, const float* beta, float eps,
                                   void* out_hi, void* out_lo, int32_t ld_out, void* stream) {
  using namespace gd;
  GD_REQUIRE(x_hi && x_lo && W1 && b1 && W2 && b2 && gamma && beta && out_hi, "gd_ffn_h16_ln_split: null pointer");
  GD_REQUIRE(M > 0, "gd_ffn_h16_ln_split: bad shape M=%d", M);
  if (N != kFfD || F < 512 || F % 256 != 0) {
    set_error("gd_ffn_h16_ln_split: N=%d F=%d unsupported (N = 512, F a multiple of 256 >= 512)", N, F);
    return GD_ERR_UNSUPPORTED;
  }
  GD_REQUIRE(fmt == GD_BF16 || fmt == GD_FP16, "gd_ffn_h16_ln_split: fmt must be GD_BF16 or GD_FP16");
  GD_REQUIRE(ld_x % 8 == 0 && ld_x >= N && ld_out % 8 == 0 && ld_out >= N && ldw1 % 8 == 0 && ldw1 >= N &&
                 ldw2 % 8 == 0 && ldw2 >= F,
             "gd_ffn_h16_ln_split: leading dimensions must be multiples of 8 and cover the rows "
             "(ld_x=%d ld_out=%d ldw1=%d ldw2=%d)", ld_x, ld_out, ldw1, ldw2);
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  GD_REQUIRE(al16(x_hi) && al16(x_lo) && al16(W1) && al16(W2) && al16(out_hi) && al16(out_lo) && al16(b1),
             "gd_ffn_h16_ln_split: operands must be 16-byte aligned");
  FfParams p;
  p.b1 = b1; p.b2 = b2; p.gamma = gamma; p.beta = beta;
  p.has_lo = out_lo != nullptr;
  p.h16_fmt = fmt;
  p.eps = eps; p.M = M; p.F = F;
  p.trace = g_ffn_trace;
  p.idesc1 = ptx::umma_idesc_h16_f32(kFfBM, kFfCH, fmt == GD_FP16);
  p.idesc2 = ptx::umma_idesc_h16_f32(kFfBM, kFfBN, fmt == GD_FP16);
  p.idesc3 = ptx::umma_idesc_h16_f32(kFfBM, kFfBK, fmt == GD_FP16);
  {
    static int dbg = -1;
    if (dbg < 0) {
      const char* e = getenv("GD_FFN_DBG");
      dbg = e ? atoi(e) : 0;
    }
    p.dbg = dbg;
  }
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(ffn_ln_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFfSmemBytes));
    attr_set = true;
  }
  CUtensorMap tx, tw1, tw2, txl, toh, tol;
  int rc = make_tmap_2d(&tx, x_hi, M, N, ld_x, kFfBM, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw1, W1, F, N, ldw1, kFfCH, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw2, W2, N, F, ldw2, kFfBN, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&txl, x_lo, M, N, ld_x, kFfBM, fmt);
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
  const int m_tiles = ceil_div(M, kFfBM);
  const int clusters = m_tiles < sms / 2 ? m_tiles : sms / 2;
  GD_CUDA_OK(launch_pdl(PDL_GEMM_LN, ffn_ln_tcgen05_kernel, dim3(2 * clusters), dim3(kFfThreads), kFfSmemBytes,
                        (cudaStream_t)stream, tx, tw1, tw2, txl, toh, tol, p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}
