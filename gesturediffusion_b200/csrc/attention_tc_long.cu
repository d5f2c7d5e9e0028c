// attention_tc_long.cu -- tcgen05/TMEM softmax self-attention for head_dim 128 and LONG sequences (S > 256: the
// 300 / 400 / 600-frame rows of BASELINE configs[4]; the encoder accepts any S <= 5000, mdm.py:216,277), with the
// keys walked in blocks of 128 and an online softmax (running row maximum / sum, the output accumulator rescaled in
// TMEM when the maximum moves).  attention_tc.cu keeps a whole row of scores in TMEM and therefore stops at S = 256.
//
// Work item = (sequence, head, PAIR of 128-query tiles): the two tiles of a pair live in the two TMEM "slots"
// (256 columns each: S / P in [0,128), O in [128,256)) with one softmax warpgroup each, and share every K / V block,
// which is fetched once per pair into a two-deep ring.  A sequence with an odd number of query tiles leaves slot 1
// of its last pair idle.
//   warp 0      TMA producer: the pair's Q tiles (128 x 128 each), then K_j, V_j (128 keys x 128) per block
//   warp 1      UMMA issuer, per block j:  S_s = Q_s K_j^T (SS, N = 128) for both slots, then, as each slot's
//               probabilities arrive, O_s (+)= P_s V_j (TS: A = P from TMEM, B = V MN-major)
//   warps 2-5 / 6-9   softmax warpgroup of slot 0 / 1, one thread per query row, per block:
//               m' = max(m, rowmax(S));  p = exp2((s - m') scale log2e);  l = l alpha + sum p with
//               alpha = exp2((m - m') scale log2e);  P -> TMEM as 16-bit pairs in place;  O *= alpha in TMEM (j > 0,
//               skipped when no row of the warp moved its maximum);  after the last block O / l -> 16 bit -> staging
//               -> row-contiguous global stores.
// Within a slot the chain S -> softmax -> PV is serial (P aliases S); the two slots interleave.  All tcgen05.mma of
// the issuing thread complete in order, so "S_s(j) complete" (the barrier the softmax waits on) implies
// "O_s += P_s V_(j-1) complete": the rescale of O needs no barrier of its own.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kAlThreads = 320;
constexpr int kAlQBytes = 128 * 128 * 2;          // one Q tile: 2 slabs of [128 rows x 128 B]
constexpr int kAlBlk = 128;                       // keys per block
constexpr int kAlKvBytes = kAlBlk * 128 * 2;      // one K (or V) block: 2 slabs of [128 rows x 128 B]
constexpr int kAlSlab = kAlBlk * 128;             // bytes of one 64-column slab
constexpr int kAlStgBytes = 8 * 32 * 128;         // 32 rows x 128 B per softmax warp
constexpr int kAlSmemBytes = 2 * kAlQBytes + 4 * kAlKvBytes + kAlStgBytes + 256 + 1024;

struct AlParams {
  void* out;
  int S, n_qt, n_pairs, n_blk, H, d, n_items, fmt;   // n_items = n_seq * H * n_pairs (pairs of query tiles)
  float scale_log2e;
  uint32_t idesc_s, idesc_o;
};

__device__ __forceinline__ float al_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAlThreads, 1)
attention_tc_long_kernel(const __grid_constant__ CUtensorMap tmap, const AlParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* q_smem = smem;                                   // 2 x 32 KB (slot 0 / 1)
  uint8_t* kv_smem = smem + 2 * kAlQBytes;                  // ring: buf b = [K_b 32 KB | V_b 32 KB]
  uint8_t* stg_smem = kv_smem + 4 * kAlKvBytes;             // 32 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg_smem + kAlStgBytes);
  uint64_t* q_full = bars;          uint64_t* q_empty = bars + 2;
  uint64_t* k_full = bars + 4;      uint64_t* k_empty = bars + 6;
  uint64_t* v_full = bars + 8;      uint64_t* v_empty = bars + 10;
  uint64_t* s_full = bars + 12;     uint64_t* p_ready = bars + 14;
  uint64_t* o_full = bars + 16;     uint64_t* slot_free = bars + 18;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  const int S = p.S, n_blk = p.n_blk;
  const int G = (int)gridDim.x, cta = (int)blockIdx.x;

  if (warp == 0 && lane == 0) ptx::prefetch_tensormap(&tmap);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&q_full[i], 1);   ptx::mbar_init(&q_empty[i], 1);
      ptx::mbar_init(&k_full[i], 1);   ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1);   ptx::mbar_init(&v_empty[i], 1);
      ptx::mbar_init(&s_full[i], 1);   ptx::mbar_init(&p_ready[i], 4);
      ptx::mbar_init(&o_full[i], 1);   ptx::mbar_init(&slot_free[i], 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::grid_dep_wait();                                      // prologue above overlaps the previous kernel's tail
  ptx::grid_dep_launch();

  // item -> (sequence, head, pair of query tiles); slot 1 of a sequence's last pair is idle when n_qt is odd
  auto seq_of = [&](int item) { return item / (p.H * p.n_pairs); };
  auto head_of = [&](int item) { return (item / p.n_pairs) % p.H; };
  auto pair_of = [&](int item) { return item % p.n_pairs; };
  auto has1_of = [&](int item) { return 2 * pair_of(item) + 1 < p.n_qt; };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int us[2] = {0, 0};
      int g = 0;                                             // K/V blocks fetched so far (ring position)
      for (int item = cta; item < p.n_items; item += G) {
        const int row0 = seq_of(item) * S, colq = head_of(item) * 128;
        const int n_slots = has1_of(item) ? 2 : 1;
        for (int s = 0; s < n_slots; ++s) {
          ptx::mbar_wait(&q_empty[s], (us[s] & 1) ^ 1);
          uint8_t* qs = q_smem + s * kAlQBytes;
          const int qrow = row0 + (2 * pair_of(item) + s) * 128;
          ptx::mbar_arrive_expect_tx(&q_full[s], kAlQBytes);
          ptx::tma_load_2d(qs, &tmap, &q_full[s], colq, qrow);
          ptx::tma_load_2d(qs + 16384, &tmap, &q_full[s], colq + 64, qrow);
          ++us[s];
        }
        for (int j = 0; j < n_blk; ++j, ++g) {
          const int b = g & 1, use = g >> 1;
          uint8_t* ks = kv_smem + b * 2 * kAlKvBytes;
          uint8_t* vs = ks + kAlKvBytes;
          const int krow = row0 + j * kAlBlk;
          ptx::mbar_wait(&k_empty[b], (use & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&k_full[b], kAlKvBytes);
          ptx::tma_load_2d(ks, &tmap, &k_full[b], p.d + colq, krow);
          ptx::tma_load_2d(ks + kAlSlab, &tmap, &k_full[b], p.d + colq + 64, krow);
          ptx::mbar_wait(&v_empty[b], (use & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&v_full[b], kAlKvBytes);
          ptx::tma_load_2d(vs, &tmap, &v_full[b], 2 * p.d + colq, krow);
          ptx::tma_load_2d(vs + kAlSlab, &tmap, &v_full[b], 2 * p.d + colq + 64, krow);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer
    // the whole warp runs the loop (uniform control flow, barrier waits by all lanes); one elected lane issues
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t q_u = __shfl_sync(0xffffffffu, ptx::smem_u32(q_smem), 0);
    const uint32_t kv_u = __shfl_sync(0xffffffffu, ptx::smem_u32(kv_smem), 0);
    int us[2] = {0, 0};
    int g0 = 0;                                              // ring position of the item's block 0
    for (int item = cta; item < p.n_items; item += G, g0 += n_blk) {
      const int n_slots = has1_of(item) ? 2 : 1;
      // S_s(j) = Q_s K_j^T into the slot's columns [0, 128)
      auto issue_s = [&](int s, int j) {
        const int b = (g0 + j) & 1, use = (g0 + j) >> 1;
        if (s == 0) ptx::mbar_wait(&k_full[b], use & 1);
        ptx::tc_fence_after();
        const uint32_t qa = q_u + (uint32_t)(s * kAlQBytes);
        const uint32_t ka = kv_u + (uint32_t)(b * 2 * kAlKvBytes);
        const uint32_t td = tmem_u + (uint32_t)(s * 256);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint64_t da = ptx::umma_desc_k_sw128(qa + (k >> 2) * 16384 + (k & 3) * 32);
            const uint64_t db = ptx::umma_desc_k_sw128(ka + (k >> 2) * kAlSlab + (k & 3) * 32);
            ptx::umma_h16(td, da, db, p.idesc_s, k != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&s_full[s]);
          if (s == n_slots - 1) ptx::umma_commit(&k_empty[b]);         // both slots have read K_j
          if (j == n_blk - 1) ptx::umma_commit(&q_empty[s]);           // the unit's last use of its Q tile
        }
        __syncwarp();
      };
      for (int s = 0; s < n_slots; ++s) {
        ptx::mbar_wait(&slot_free[s], (us[s] & 1) ^ 1);      // the previous unit's output has left the slot
        ptx::mbar_wait(&q_full[s], us[s] & 1);
        issue_s(s, 0);
      }
      for (int j = 0; j < n_blk; ++j) {
        const int b = (g0 + j) & 1, use = (g0 + j) >> 1;
        const int nvalid = min(kAlBlk, S - j * kAlBlk);
        const int ksteps = (nvalid + 15) >> 4;
        for (int s = 0; s < n_slots; ++s) {
          const int c = us[s] * n_blk + j;
          ptx::mbar_wait(&p_ready[s], c & 1);                // P_s(j) written, O_s rescaled
          if (s == 0) ptx::mbar_wait(&v_full[b], use & 1);
          ptx::tc_fence_after();
          const uint32_t va = kv_u + (uint32_t)(b * 2 * kAlKvBytes + kAlKvBytes);
          const uint32_t tp = tmem_u + (uint32_t)(s * 256);
          const uint32_t to = tp + 128;
          if (ptx::elect_one()) {
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t db = ptx::umma_desc_mn_sw128(va + k * 2048, (uint32_t)kAlSlab);
              ptx::umma_h16_ts(to, tp + (uint32_t)(k * 8), db, p.idesc_o, (j | k) != 0 ? 1u : 0u);
            }
            if (j == n_blk - 1) ptx::umma_commit(&o_full[s]);
            if (s == n_slots - 1) ptx::umma_commit(&v_empty[b]);       // both slots have read V_j
          }
          __syncwarp();
          if (j + 1 < n_blk) issue_s(s, j + 1);              // next block's scores while the other slot's softmax runs
        }
      }
      for (int s = 0; s < n_slots; ++s) ++us[s];
    }
  } else {
    // ------------------------------------------------------------------ softmax / output warpgroups
    const int slot = (warp - 2) >> 2;
    const int quarter = warp & 3;                           // TMEM lane quarter of this warp
    const uint32_t lane_bits = (uint32_t)(quarter * 32) << 16;
    const uint32_t ts = tmem_base + lane_bits + (uint32_t)(slot * 256);      // S / P columns of this slot
    const uint32_t to = ts + 128;                                             // O columns
    const uint32_t stg = ptx::smem_u32(stg_smem) + (uint32_t)(warp - 2) * 4096;
    const float sl2 = p.scale_log2e;
    uint16_t* out = reinterpret_cast<uint16_t*>(p.out);
    int us = 0;
    for (int item = cta; item < p.n_items; item += G) {
      if (slot == 1 && !has1_of(item)) continue;
      const int seq = seq_of(item), head = head_of(item), qt = 2 * pair_of(item) + slot;
      float m = -INFINITY, l = 0.f;
      for (int j = 0; j < n_blk; ++j) {
        const int c = us * n_blk + j;
        const int nvalid = min(kAlBlk, S - j * kAlBlk);     // keys of this block inside the sequence
        const int cend = min(kAlBlk, (nvalid + 31) & ~31);  // score columns the PV MMA may read (16 * ksteps <= cend)
        ptx::mbar_wait(&s_full[slot], c & 1);
        ptx::tc_fence_after();
        // ---- pass 1: block row maximum
        float bm = -INFINITY;
        for (int c0 = 0; c0 < cend; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld_32x32_nowait(ts + c0, r);
          ptx::tmem_ld_wait_on(r);
          if (c0 + 32 > nvalid) {
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (c0 + t >= nvalid) r[t] = 0xff800000u;     // -inf: keys past the sequence end
          }
#pragma unroll
          for (int t = 0; t < 32; ++t) bm = fmaxf(bm, __uint_as_float(r[t]));
        }
        const float m_new = fmaxf(m, bm);
        const float alpha = al_ex2((m - m_new) * sl2);      // j == 0: exp2(-inf) = 0
        const float mb = -m_new * sl2;
        // ---- pass 2: probabilities -> TMEM (16-bit pairs, in place), block row sum
        float sum = 0.f;
        for (int c0 = 0; c0 < cend; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld_32x32_nowait(ts + c0, r);
          ptx::tmem_ld_wait_on(r);
          if (c0 + 32 > nvalid) {
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (c0 + t >= nvalid) r[t] = 0xff800000u;
          }
          uint32_t pk[16];
#pragma unroll
          for (int t = 0; t < 16; ++t) {
            const float a = al_ex2(fmaf(__uint_as_float(r[2 * t]), sl2, mb));
            const float b = al_ex2(fmaf(__uint_as_float(r[2 * t + 1]), sl2, mb));
            sum += a + b;
            pk[t] = pack_h16x2(a, b, p.fmt);
          }
          ptx::tmem_st_32x16(ts + (uint32_t)(c0 >> 1), pk);
        }
        l = fmaf(l, alpha, sum);
        m = m_new;
        // ---- the accumulator follows the maximum: O *= alpha (the PV MMA of block j - 1 has completed, see the header)
        if (j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {
#pragma unroll 1
          for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t r[32];
            ptx::tmem_ld_32x32_nowait(to + c0, r);
            ptx::tmem_ld_wait_on(r);
            float v[32];
#pragma unroll
            for (int t = 0; t < 32; ++t) v[t] = __uint_as_float(r[t]) * alpha;
            ptx::tmem_st_32x32(to + c0, v);
          }
        }
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&p_ready[slot]);
      }
      const float inv = 1.0f / l;

      // ---- output: O / l -> 16 bit -> staging -> row-contiguous stores
      ptx::mbar_wait(&o_full[slot], us & 1);
      ptx::tc_fence_after();
      const int q_base = qt * 128 + quarter * 32;           // first query row of this warp inside the sequence
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32_nowait(to + h * 64, r0);
        ptx::tmem_ld_32x32_nowait(to + h * 64 + 32, r1);
        ptx::tmem_ld_wait_on(r0);
        ptx::tmem_ld_wait_on(r1);
        if (h == 1) {                                       // accumulator fully read: release the slot
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&slot_free[slot]);
        }
#pragma unroll
        for (int jq = 0; jq < 4; ++jq) {
          uint32_t w[4];
#pragma unroll
          for (int t = 0; t < 4; ++t)
            w[t] = pack_h16x2(__uint_as_float(r0[8 * jq + 2 * t]) * inv, __uint_as_float(r0[8 * jq + 2 * t + 1]) * inv, p.fmt);
          ptx::sts_v4u(stg + lane * 128 + ((jq ^ (lane & 7)) << 4), w[0], w[1], w[2], w[3]);
#pragma unroll
          for (int t = 0; t < 4; ++t)
            w[t] = pack_h16x2(__uint_as_float(r1[8 * jq + 2 * t]) * inv, __uint_as_float(r1[8 * jq + 2 * t + 1]) * inv, p.fmt);
          ptx::sts_v4u(stg + lane * 128 + (((4 + jq) ^ (lane & 7)) << 4), w[0], w[1], w[2], w[3]);
        }
        __syncwarp();
        const int j8 = lane & 7, rsub = lane >> 3;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int rl = it * 4 + rsub, q = q_base + rl;
          const uint4 val = ptx::lds_v4u(stg + rl * 128 + ((j8 ^ (rl & 7)) << 4));
          if (q < S)
            *reinterpret_cast<uint4*>(out + ((size_t)(seq * S + q) * p.d + head * 128 + h * 64 + j8 * 8)) = val;
        }
        __syncwarp();
      }
      ++us;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------- host side
int launch_attention_tc_long(const void* qkv, void* out, int n_seq, int S, int n_heads, int fmt, cudaStream_t stream) {
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(attention_tc_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAlSmemBytes));
    attr_set = true;
  }
  const int d = n_heads * 128;
  const long rows = (long)n_seq * S;
  GD_REQUIRE(rows < (1L << 31), "gd_attention_h16: too many rows");
  CUtensorMap tm;
  int rc = make_tmap_2d(&tm, qkv, (int)rows, 3 * d, 3 * d, 128, fmt);       // Q tiles and K / V blocks: 128-row boxes
  if (rc) return rc;
  AlParams p;
  p.out = out;
  p.S = S;
  p.n_qt = (S + 127) / 128;
  p.n_pairs = (p.n_qt + 1) / 2;
  p.n_blk = (S + kAlBlk - 1) / kAlBlk;
  p.H = n_heads;
  p.d = d;
  const long items = (long)n_seq * n_heads * p.n_pairs;
  GD_REQUIRE(items < (1L << 31), "gd_attention_h16: too many work items");
  p.n_items = (int)items;
  p.fmt = fmt;
  p.scale_log2e = (1.0f / sqrtf(128.0f)) * 1.4426950408889634f;
  p.idesc_s = ptx::umma_idesc_h16_f32(128, kAlBlk, fmt == GD_FP16);
  p.idesc_o = ptx::umma_idesc_h16_f32_bmn(128, 128, fmt == GD_FP16);
  const int sms = num_sms();
  GD_REQUIRE(sms > 0, "gd_attention_h16: cannot query the SM count");
  const int grid = p.n_items < sms ? p.n_items : sms;
  GD_CUDA_OK(launch_pdl(PDL_ATTENTION, attention_tc_long_kernel, dim3(grid), dim3(kAlThreads), kAlSmemBytes, stream, tm, p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

}  // namespace gd
