// gemm_tcgen05.cu -- persistent warp-specialised 16-bit (bf16 or fp16 operands) GEMM for sm_100a.
//   C[M,N] = A[M,K] @ W[N,K]^T  (+ fused gdEpilogue), fp32 accumulation in TMEM.
// Roles (320 threads): warps 0..7 = epilogue, warp 8 = TMA producer, warp 9 = UMMA issuer (one elected lane);
// epilogue (TMEM -> registers -> fused epilogue -> global).  The accumulator is
// double buffered in TMEM (2 x BN fp32 columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1; A/W tiles stream through a kStages-deep TMA/mbarrier ring with 128B swizzle.
// Replaces every aten::addmm of the reference hot path (see gd_b200.h).
//
// Epilogues are compile-time specialised (template EPI) so that each instantiation stays small
// enough for the instruction cache (ncu on the first, fully runtime-dispatched version showed
// stall_no_inst as the top stall):
//   EPI_BIAS_H16           out_h16 = acc + bias                      (in_proj QKV)
//   EPI_BIAS_GELU_H16      out_h16 = gelu_erf(acc + bias)            (linear1)
//                          both: bias as warp-uniform register loads ahead of the accumulator wait, 16-bit pairs
//                          staged in 128B-swizzled smem, one TMA store per [32 rows x 64 columns] box
//   EPI_BIAS_RES_F32       out_f32 = acc + bias + residual           (out_proj / linear2 when not LayerNorm-fused)
//   EPI_QUAD               bias, row table, per-sequence bias, residual, row map, copies on coalesced quads
//                          (InputProcess of both denoisers)
//   EPI_OUT_T              bias + transposed fp32 output             (OutputProcess + output[1:] + permute)
//   EPI_GENERIC            every gdEpilogue feature, one thread per accumulator row (whatever is left)
// RES_F32 / QUAD stage each 32x32 accumulator block through shared memory (XOR-swizzled, conflict free) so that
// global accesses are row-contiguous 128-byte segments.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kBM = 128;          // UMMA M (cta_group::1)
constexpr int kBK = 64;           // 64 bf16 = 128 B = one swizzle row
constexpr int kUmmaK = 16;
constexpr int kEpiWarps = 8;
constexpr int kGemmThreads = (2 + kEpiWarps) * 32;
// Warps 0..7 = epilogue, warp 8 = TMA producer, warp 9 = UMMA issuer.  The two single-lane control warps get the
// HIGHEST warp ids of their scheduler partitions (wid % 4 = 0 and 1): the warp arbiter favours high warp ids, and
// with the control warps at ids 0/1 the busy epilogue warps sharing their schedulers delayed every TMA issue and
// MMA issue (tensor pipe ~50 % idle with all operands resident).
constexpr int kProducerWarp = kEpiWarps, kMmaWarp = kEpiWarps + 1;

enum { EPI_GENERIC = 0, EPI_BIAS_H16 = 1, EPI_BIAS_GELU_H16 = 2, EPI_BIAS_RES_F32 = 3, EPI_QUAD = 4, EPI_OUT_T = 5,
       EPI_QUAD_IN1 = 6, EPI_QUAD_IN2 = 7 };

// One CTA per 128 x BN tile.  (A cta_group::2 variant -- CTA pair per 256 x BN tile, W tile split between the pair --
// was built and parity-tested in round 1 and measured slower at every encoder shape, 42.1 vs 39.9 us for QKV; it was
// removed in round 2)
template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarBytes = 256;
  static constexpr int kStagingBytes = kEpiWarps * 32 * 32 * 4;                 // one 4 KB block per epilogue warp
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarBytes;
  static constexpr int kTmemCols = 2 * BN;                                      // 256 or 512 (power of 2)
};

// ---- specialised epilogue of one 32-row x 32-column accumulator block held as v[32] (row = lane) -------------
// stg: this warp's 4 KB staging block (shared-space address).  Block element (r, c) lives in 16-byte chunk
// (c / 4) ^ (r & 7) of row r: conflict-free for the row-wise writes and for both read-back patterns.
__device__ __forceinline__ void epi_stage(uint32_t stg, int lane, const uint32_t (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    ptx::sts_v4u(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  __syncwarp();
}

// fp32 output + residual: lane -> (row it*4 + lane/8, columns 4*(lane%8) .. +3): 8 lanes cover one 128-byte row
// segment.  b = bias of the lane's 4 columns, pre = the residual block (fetched one block ahead).
__device__ __forceinline__ void epi_finish_f32(const gdEpilogue& ep, uint32_t stg, int lane, int m_warp, int n0, int M,
                                               const float4 b, const float4 (&pre)[8]) {
  const int c4 = lane & 7, rsub = lane >> 3, n = n0 + c4 * 4;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int rl = it * 4 + rsub, mm = m_warp + rl;
    float4 w = ptx::lds_v4(stg + rl * 128 + ((c4 ^ (rl & 7)) << 4));
    f4_add(w, b);
    f4_add(w, pre[it]);
    if (mm < M) *reinterpret_cast<float4*>(ep.out_f32 + (size_t)mm * ep.ld_out + n) = w;
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
                         const __grid_constant__ CUtensorMap tmap_out, int M, int N, int K, uint32_t idesc,
                         const gdEpilogue ep, int dbg) {
  using Cfg = GemmCfg<BN>;
  const int n_workers = (int)gridDim.x, worker = (int)blockIdx.x;
  extern __shared__ __align__(1024) uint8_t smem[];          // 128B-swizzled TMA/UMMA tiles need 1024-byte alignment
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
  uint8_t* bars = staging + Cfg::kStagingBytes;
  if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bars);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  // warp index through a shuffle: the compiler then knows it is warp-uniform, keeps the role branches uniform and the
  // MMA warp's descriptor arithmetic in the uniform datapath (no ELECT / R2UR waterfall around every UTCHMMA)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int m_tiles = (M + kBM - 1) / kBM;
  const int n_tiles = (N + BN - 1) / BN;
  const int total_tiles = m_tiles * n_tiles;
  const int k_blocks = (K + kBK - 1) / kBK;

  if (warp == kProducerWarp && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_w);
    if constexpr (EPI == EPI_BIAS_H16 || EPI == EPI_BIAS_GELU_H16) ptx::prefetch_tensormap(&tmap_out);
  }
  if (warp == kMmaWarp && lane == 0) {
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
  if (warp == 0) ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::grid_dep_wait();                                      // prologue above overlaps the previous kernel's tail
  ptx::grid_dep_launch();

  if (warp == kProducerWarp) {
    // ------------------------------------------------------------------ TMA producer
    // (whole warp in uniform control flow, one elected lane issues: tensor-map / barrier operands stay uniform)
    {
      int stage = 0;
      uint32_t phase = 0;
      uint8_t* const smem_u = smem;
      for (int tile = worker; tile < total_tiles; tile += n_workers) {
        const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_u + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          if (ptx::elect_one()) {
            if (dbg & 4) {                                   // diagnostics: no operand traffic
              ptx::mbar_arrive(&full_bar[stage]);
            } else {
              ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
              ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBK, m_blk * kBM);
              ptx::tma_load_2d(sb, &tmap_w, &full_bar[stage], kb * kBK, n_blk * BN);
            }
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ------------------------------------------------------------------ UMMA issuer
    // The whole warp runs the loop (uniform control flow, barrier waits by all lanes); one elected lane issues.
    {
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t smem_u = __shfl_sync(0xffffffffu, ptx::smem_u32(smem), 0);
      for (int tile = worker; tile < total_tiles; tile += n_workers, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);
        ptx::mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);          // epilogue drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_u + (uint32_t)(buf * BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_u + (uint32_t)(stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kBK / kUmmaK; ++k) {
              if (dbg & 2) break;                            // diagnostics: no MMAs
              const uint64_t da = ptx::umma_desc_k_sw128(sa + k * kUmmaK * 2);
              const uint64_t db = ptx::umma_desc_k_sw128(sb + k * kUmmaK * 2);
              ptx::umma_h16(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);             // smem slot reusable when the MMAs retire
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit(&tfull_bar[buf]);     // accumulator complete
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp;
    const int quarter = warp & 3;            // TMEM lane quarter this warp may access
    const int half = ew >> 2;                // which half of the BN columns
    constexpr int kColsPerWarp = BN / 2;
    const uint32_t stg = ptx::smem_u32(staging) + (uint32_t)ew * (32 * 32 * 4);
    int local = 0;
    for (int tile = worker; tile < total_tiles; tile += n_workers, ++local) {
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      const int m_warp = m_blk * kBM + quarter * 32;
      const int n_base = n_blk * BN + half * kColsPerWarp;
      if (dbg & 1) {                                       // diagnostics: drain without the epilogue work
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
      } else if constexpr (EPI == EPI_GENERIC) {
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
#pragma unroll 1
        for (int c = 0; c < kColsPerWarp; c += 32) {
          const int n0 = n_base + c;
          if (n0 >= N) break;                                // warp-uniform
          float v[32];
          ptx::tmem_ld_32x32(taddr0 + c, v);
          if (m_warp + lane < M) epilogue_row_chunk<32>(ep, m_warp + lane, n0, N, v);
        }
      } else if constexpr (EPI == EPI_QUAD || EPI == EPI_QUAD_IN1 || EPI == EPI_QUAD_IN2) {
        // EPI_QUAD_IN1 / _IN2 are the same code with the feature tests of the two input projections folded at compile
        // time (the generic form runs ~3 k instructions per warp and tile, most of them tests and address arithmetic:
        // it was issue-bound at 3x the main loop's time):
        //   IN1 (V1: poseEmbedding + pe)   bias, row table, row map, 2 copies, fp32 + 16-bit out
        //   IN2 (V2: project_to_lat)       row table, per-sequence bias, 2 copies, fp32 out
        constexpr bool kSpec = EPI != EPI_QUAD;
        const bool has_bias = kSpec ? (EPI == EPI_QUAD_IN1) : ep.bias != nullptr;
        const bool has_tab = kSpec ? true : ep.rowtab != nullptr;
        const bool has_seq = kSpec ? (EPI == EPI_QUAD_IN2) : ep.seqbias != nullptr;
        const bool has_res = kSpec ? false : ep.residual != nullptr;
        const bool has_f32 = kSpec ? true : ep.out_f32 != nullptr;
        const bool has_h16 = kSpec ? (EPI == EPI_QUAD_IN1) : ep.out_h16 != nullptr;
        const bool has_map = kSpec ? (EPI == EPI_QUAD_IN1) : ep.map_t > 0;
        // Every linear gdEpilogue feature (bias, row table, per-sequence bias, residual, row map, copies; no
        // activation, no out_t) on row-contiguous quads: TMEM -> registers (row = lane) -> this warp's swizzled
        // staging block -> lane (it, c4) takes columns 4*c4 .. 4*c4+3 of row 4*it + lane/8, so table, residual and
        // output accesses are full 128-byte lines.
        const int c4 = lane & 7, rsub = lane >> 3;
        int tabrow[8], outrow[8], seqrow[8];                 // of this lane's eight quad rows (4*it + lane/8)
        {
          const int mm = m_warp + lane;                      // computed for row `lane`, handed over by shuffle
          const int tabrow_l = has_tab ? (mm % ep.tab_period) + ep.tab_offset : 0;
          const int seqrow_l = has_seq ? mm / ep.seq_rows : 0;
          const int mapped = has_map ? (mm / ep.map_t) * ep.map_s + ep.map_off + (mm % ep.map_t) : mm;
          const int outrow_l = mm < M ? mapped : -1;
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rl = it * 4 + rsub;
            tabrow[it] = __shfl_sync(0xffffffffu, tabrow_l, rl);
            seqrow[it] = __shfl_sync(0xffffffffu, seqrow_l, rl);
            outrow[it] = __shfl_sync(0xffffffffu, outrow_l, rl);
          }
        }
        const int ncopy = kSpec ? 2 : (ep.dup > 1 ? ep.dup : 1);
        // bias + the row-table quads of a chunk (HBM latency when the table is per row, as in the V2 input
        // projection): eight independent loads issued together, ahead of the TMEM load.  (Fetching them a whole
        // chunk ahead measured slower: 62 vs 54 us on the V2 shape -- the epilogue is issue-bound, not latency-bound.)
        auto prefetch = [&](int c, float4 (&pre)[8]) {
          const int n = n_base + c + c4 * 4;
          const bool col_ok = n < N;                         // N % 4 == 0: a quad is inside or outside
          float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (has_bias && col_ok) bias4 = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            pre[it] = bias4;
            if (has_tab && col_ok && outrow[it] >= 0)
              f4_add(pre[it], __ldg(reinterpret_cast<const float4*>(ep.rowtab + (size_t)tabrow[it] * ep.ld_tab + n)));
          }
        };
        // The specialised forms with two chunks per warp (BN = 128) request the table quads of BOTH chunks before they
        // wait for the accumulator: the V2 table is per row and comes from HBM, and with the lean code the epilogue is no
        // longer issue-bound (the generic form measured slower with a chunk of look-ahead).
        constexpr bool kAhead = kSpec && kColsPerWarp == 64;
        float4 pre[8], pre_b[8];
        if constexpr (kAhead) {
          prefetch(0, pre);
          prefetch(32, pre_b);
        }
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
#pragma unroll(kAhead ? 2 : 1)
        for (int c = 0; c < kColsPerWarp; c += 32) {
          const int n0 = n_base + c;
          if (n0 >= N) break;                                // warp-uniform
          uint32_t v[32];
          ptx::tmem_ld_32x32_nowait(taddr0 + c, v);
          if constexpr (kAhead) {
            if (c == 32) {
#pragma unroll
              for (int it = 0; it < 8; ++it) pre[it] = pre_b[it];
            }
          } else {
            prefetch(c, pre);                                // in flight during the TMEM load and the staging
          }
          const int n = n0 + c4 * 4;
          const bool col_ok = n < N;
          ptx::tmem_ld_wait_on(v);
          epi_stage(stg, lane, v);
          __syncwarp();
          float4 w[8];
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rl = it * 4 + rsub;
            w[it] = ptx::lds_v4(stg + rl * 128 + ((c4 ^ (rl & 7)) << 4));
            f4_add(w[it], pre[it]);
          }
#pragma unroll 2
          for (int cp = 0; cp < ncopy; ++cp) {
            float4 o[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {                 // all loads of the copy before any store
              o[it] = w[it];
              if (outrow[it] < 0 || !col_ok) continue;
              if (has_seq)
                f4_add(o[it], __ldg(reinterpret_cast<const float4*>(
                                  ep.seqbias + (size_t)(cp * ep.seq_count + seqrow[it]) * ep.ld_seq + n)));
              if (has_res)
                f4_add(o[it], *reinterpret_cast<const float4*>(
                                  ep.residual + ((size_t)outrow[it] + (size_t)cp * ep.dup_row_stride) * ep.ld_res + n));
            }
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              if (outrow[it] < 0 || !col_ok) continue;
              const size_t r = (size_t)outrow[it] + (size_t)cp * ep.dup_row_stride;
              if (has_f32) *reinterpret_cast<float4*>(ep.out_f32 + r * ep.ld_out + n) = o[it];
              if (has_h16) {
                uint2 pk;
                pk.x = pack_h16x2(o[it].x, o[it].y, ep.h16_fmt);
                pk.y = pack_h16x2(o[it].z, o[it].w, ep.h16_fmt);
                *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(ep.out_h16) + r * ep.ld_out_h16 + n) = pk;
              }
            }
          }
          __syncwarp();
        }
      } else if constexpr (EPI == EPI_OUT_T) {
        // bias + transposed fp32 output out_t[seq][n][s - t_skip] (OutputProcess + output[1:] + permute): in the
        // TMEM layout lane = row = consecutive frames of one sequence, so each column's store is one contiguous run.
        const int mm = m_warp + lane;
        const int seq = mm / ep.t_s, s_in = mm - seq * ep.t_s;
        const bool row_ok = mm < M && s_in >= ep.t_skip;
        float* orow = ep.out_t + (size_t)seq * N * ep.t_t + (s_in - ep.t_skip);
        // the first chunk's bias (32 warp-uniform loads) is requested before the accumulator wait, the next chunk's while
        // this chunk's stores drain: fetched inside the store loop they were 32 dependent L1 round trips per chunk
        float bv[32];
        auto load_bias = [&](int n0) {
#pragma unroll
          for (int i = 0; i < 32; ++i) bv[i] = (ep.bias && n0 + i < N) ? __ldg(ep.bias + n0 + i) : 0.f;
        };
        if (n_base < N) load_bias(n_base);
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
#pragma unroll 1
        for (int c = 0; c < kColsPerWarp; c += 32) {
          const int n0 = n_base + c;
          if (n0 >= N) break;                                // warp-uniform
          uint32_t v[32];
          ptx::tmem_ld_32x32_nowait(taddr0 + c, v);
          ptx::tmem_ld_wait_on(v);
          float o[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(v[i]) + bv[i];
          if (c + 32 < kColsPerWarp && n0 + 32 < N) load_bias(n0 + 32);
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (n0 + i < N) orow[(size_t)(n0 + i) * ep.t_t] = o[i];
          }
        }
      } else if constexpr (EPI == EPI_BIAS_H16 || EPI == EPI_BIAS_GELU_H16) {
        // 16-bit outputs leave through TMA stores: per warp and 64 columns, TMEM -> registers (row = lane) ->
        // + bias [-> GELU] -> 16-bit pairs -> this warp's 32 x 128 B staging block in the
        // 128B-swizzle layout -> one cp.async.bulk.tensor store of a [32 rows x 64 columns] box.  The stores are
        // asynchronous full-line writes issued by one lane; rows >= M and columns >= N are clipped by the TMA.
        // The bias of a 64-column block is fetched as 16 warp-uniform 16-byte loads (one sector each, L1 hits after
        // the first tile of a column block) into registers BEFORE the wait on the accumulator / the TMEM load.
        auto load_bias = [&](int n0, float4 (&bq)[16]) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            bq[j] = (n0 + 4 * j < N) ? __ldg(reinterpret_cast<const float4*>(ep.bias + n0 + 4 * j))
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        float4 bq[16];
        if (n_base < N) load_bias(n_base, bq);
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
#pragma unroll 1
        for (int r = 0; r < kColsPerWarp / 64; ++r) {
          const int n0 = n_base + r * 64;
          if (n0 >= N) break;                                // warp-uniform
          uint32_t v0[32], v1[32];
          ptx::tmem_ld_32x32_nowait(taddr0 + r * 64, v0);
          ptx::tmem_ld_32x32_nowait(taddr0 + r * 64 + 32, v1);
          if (r > 0) load_bias(n0, bq);
          if (lane == 0) ptx::bulk_wait_read_all();          // the previous store has finished reading the staging
          __syncwarp();
          ptx::tmem_ld_wait_on(v0);
          ptx::tmem_ld_wait_on(v1);
#pragma unroll
          for (int j = 0; j < 8; ++j) {                      // 16-byte chunk j = columns 8j .. 8j+7 of the 64
            const uint32_t* src = (j < 4) ? &v0[8 * j] : &v1[8 * (j - 4)];
            const float4 b0 = bq[2 * j], b1 = bq[2 * j + 1];
            float w[8];
            w[0] = __uint_as_float(src[0]) + b0.x; w[1] = __uint_as_float(src[1]) + b0.y;
            w[2] = __uint_as_float(src[2]) + b0.z; w[3] = __uint_as_float(src[3]) + b0.w;
            w[4] = __uint_as_float(src[4]) + b1.x; w[5] = __uint_as_float(src[5]) + b1.y;
            w[6] = __uint_as_float(src[6]) + b1.z; w[7] = __uint_as_float(src[7]) + b1.w;
            if constexpr (EPI == EPI_BIAS_GELU_H16) {
#pragma unroll
              for (int t = 0; t < 8; ++t) w[t] = gelu_fast(w[t]);
            }
            ptx::sts_v4u(stg + lane * 128 + ((j ^ (lane & 7)) << 4), pack_h16x2(w[0], w[1], ep.h16_fmt),
                         pack_h16x2(w[2], w[3], ep.h16_fmt), pack_h16x2(w[4], w[5], ep.h16_fmt),
                         pack_h16x2(w[6], w[7], ep.h16_fmt));
          }
          ptx::fence_proxy_async();                          // generic-proxy smem writes -> visible to the TMA
          __syncwarp();
          if (lane == 0 && !(dbg & 16)) {
            ptx::tma_store_2d(&tmap_out, stg, n0, m_warp);
            ptx::bulk_commit();
          }
        }
      } else {
        // Everything that does not depend on the accumulator is fetched BEFORE waiting for it: the bias of all
        // of this lane's columns (almost no L1 is left beside 225 KB of smem, so these are L2 round trips) and
        // the first residual block.  TMEM reads are software-pipelined one block ahead of the staging.
        constexpr int NCH = kColsPerWarp / 32;
        const int nch = min(NCH, (N - n_base + 31) / 32);    // specialised paths have N % 32 == 0
        float4 pre[8];
        static_assert(EPI == EPI_BIAS_RES_F32, "the last epilogue branch is the fp32 + residual one");
        float4 bias[NCH];
        if (nch > 0) {
#pragma unroll
          for (int c = 0; c < NCH; ++c)
            if (c < nch) bias[c] = __ldg(reinterpret_cast<const float4*>(ep.bias + n_base + c * 32 + (lane & 7) * 4));
          prefetch_residual(ep, lane, m_warp, n_base, M, pre);
        }
        ptx::mbar_wait(&tfull_bar[buf], use & 1);
        ptx::tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * BN + half * kColsPerWarp);
        uint32_t v[32];
        if (nch > 0) ptx::tmem_ld_32x32_nowait(taddr0, v);
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          if (c < nch) {                                     // warp-uniform
            const int n0 = n_base + c * 32;
            ptx::tmem_ld_wait_on(v);
            epi_stage(stg, lane, v);
            if (c + 1 < nch) ptx::tmem_ld_32x32_nowait(taddr0 + (c + 1) * 32, v);   // overlaps the stores below
            float4 cur[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) cur[it] = pre[it];
            if (c + 1 < nch) prefetch_residual(ep, lane, m_warp, n0 + 32, M, pre);
            epi_finish_f32(ep, stg, lane, m_warp, n0, M, bias[c], cur);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[buf]);
    }
  }

  if constexpr (EPI == EPI_BIAS_H16 || EPI == EPI_BIAS_GELU_H16) {
    if (warp < kEpiWarps && lane == 0) ptx::bulk_wait_all(); // this lane's TMA stores are complete
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
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

// 2D fp32 row-major [rows, cols] with row stride ld elements; box = [box_rows, 32 cols] (128 B), SW128.
int make_tmap_2d_f32(CUtensorMap* out, const void* base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return GD_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fp32) failed with CUresult %d (rows=%d cols=%d ld=%d)", (int)r, rows, cols, ld);
    return GD_ERR_CUDA;
  }
  return GD_OK;
}

// 2D 16-bit row-major [rows, cols]; box = [box_rows, 32 cols] (64-byte rows), 64B swizzle: 16-byte chunk c of box
// row r sits at chunk c ^ ((r >> 1) & 3).
int make_tmap_2d_h16_sw64(CUtensorMap* out, const void* base, int rows, int cols, int ld, int box_rows, int fmt) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return GD_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, fmt == GD_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (16-bit, 64B swizzle) failed with CUresult %d (rows=%d cols=%d ld=%d)", (int)r,
              rows, cols, ld);
    return GD_ERR_CUDA;
  }
  return GD_OK;
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
  if (ep->out_t) {
    const bool plain = !ep->rowtab && !ep->seqbias && !ep->residual && ep->act == GD_ACT_NONE && ep->dup <= 1 &&
                       ep->map_t <= 0 && !ep->out_f32 && !ep->out_h16;
    return plain ? EPI_OUT_T : EPI_GENERIC;
  }
  const bool quad_ok = N % 4 == 0 && ep->act == GD_ACT_NONE && (!ep->bias || al16(ep->bias)) &&
                       (!ep->rowtab || (al16(ep->rowtab) && ep->ld_tab % 4 == 0)) &&
                       (!ep->seqbias || (al16(ep->seqbias) && ep->ld_seq % 4 == 0)) &&
                       (!ep->residual || (al16(ep->residual) && ep->ld_res % 4 == 0)) &&
                       (!ep->out_f32 || (al16(ep->out_f32) && ep->ld_out % 4 == 0)) &&
                       (!ep->out_h16 || (al16(ep->out_h16) && ep->ld_out_h16 % 4 == 0));
  int fallback = quad_ok ? EPI_QUAD : EPI_GENERIC;
  if (quad_ok && ep->dup == 2 && ep->rowtab && !ep->residual && ep->out_f32) {       // the two input projections
    if (ep->bias && !ep->seqbias && ep->map_t > 0 && ep->out_h16) fallback = EPI_QUAD_IN1;
    if (!ep->bias && ep->seqbias && ep->map_t <= 0 && !ep->out_h16) fallback = EPI_QUAD_IN2;
  }
  if (N % 32 != 0 || !ep->bias || !al16(ep->bias) || ep->rowtab || ep->seqbias || ep->map_t > 0 || ep->dup > 1)
    return fallback;
  if (ep->out_h16 && !ep->out_f32 && !ep->residual && al16(ep->out_h16) && ep->ld_out_h16 % 8 == 0) {
    if (ep->act == GD_ACT_NONE) return EPI_BIAS_H16;
    if (ep->act == GD_ACT_GELU_ERF) return EPI_BIAS_GELU_H16;
    return fallback;
  }
  if (ep->out_f32 && !ep->out_h16 && ep->residual && ep->act == GD_ACT_NONE && al16(ep->out_f32) &&
      al16(ep->residual) && ep->ld_out % 4 == 0 && ep->ld_res % 4 == 0)
    return EPI_BIAS_RES_F32;
  return fallback;
}

// GD_GEMM_DBG (diagnostics only, results are wrong): bit 0 = skip the epilogue work, bit 1 = skip the MMAs,
// bit 2 = skip the TMA loads.  Used to time the three pipelines of the kernel in isolation.
static int gemm_dbg() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GD_GEMM_DBG");
    v = e ? atoi(e) : 0;
  }
  return v;
}

template <int BN, int EPI>
static int launch_gemm(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int fmt,
                       const gdEpilogue* ep, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
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
  CUtensorMap to = tw;                                       // only read by the 16-bit-output epilogues
  if (EPI == EPI_BIAS_H16 || EPI == EPI_BIAS_GELU_H16) {
    rc = make_tmap_2d(&to, ep->out_h16, M, N, ep->ld_out_h16, 32, ep->h16_fmt);
    if (rc) return rc;
  }
  const int tiles = ceil_div(M, kBM) * ceil_div(N, BN);
  const int workers = tiles < g_num_sms ? tiles : g_num_sms;
  const uint32_t idesc = ptx::umma_idesc_h16_f32(kBM, BN, fmt == GD_FP16);
  GD_CUDA_OK(launch_pdl(tiles < g_num_sms ? PDL_GEMM_SMALL : PDL_GEMM, gemm_h16_tcgen05_kernel<BN, EPI>, dim3(workers), dim3(kGemmThreads), Cfg::kSmemBytes, stream,
                        ta, tw, to, M, N, K, idesc, *ep, gemm_dbg()));
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
  // (unless the last 256-wide column tile would be mostly padding, e.g. N = 263)
  const int pad256 = ceil_div(N, 256) * 256 - N, pad128 = ceil_div(N, 128) * 128 - N;
  static int bn_env = -1;                                    // GD_GEMM_BN=128 forces the narrow tile (experiments)
  if (bn_env < 0) {
    const char* e = getenv("GD_GEMM_BN");
    bn_env = e ? atoi(e) : 0;
  }
  const bool use256 = (N >= 256) && (tiles256 >= 2L * g_num_sms) && (pad256 <= pad128) && bn_env != 128;
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
    GD_GEMM_CASE(EPI_QUAD)
    GD_GEMM_CASE(EPI_QUAD_IN1)
    GD_GEMM_CASE(EPI_QUAD_IN2)
    GD_GEMM_CASE(EPI_OUT_T)
    default: break;
  }
  return use256 ? launch_gemm<256, EPI_GENERIC>(A, lda, W, ldw, M, N, K, fmt, ep, st)
                : launch_gemm<128, EPI_GENERIC>(A, lda, W, ldw, M, N, K, fmt, ep, st);
#undef GD_GEMM_CASE
}
