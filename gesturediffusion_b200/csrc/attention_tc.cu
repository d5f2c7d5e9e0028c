// attention_tc.cu -- tcgen05/TMEM softmax self-attention for head_dim 128 and S <= 256 (the encoder
// layers of both denoisers: S = T+1 = 197 / 121; mdm.py:90-96,216), one persistent CTA per SM.
//
// Work item = (sequence, head); unit = one 128-query tile of an item (1 or 2 per item).  Two TMEM
// "slots" of 256 columns alternate between units, each with its own softmax warpgroup:
//   warp 0      TMA producer: Q tile (128 x 128), K and V (Sp x 128, Sp = S rounded up to 16) straight
//               from the packed in_proj output [rows, 3d] with 128-byte swizzle
//   warp 1      UMMA issuer:  S = Q K^T   (SS, K-major both, N = Sp, 8 k-steps)      -> slot cols [0, Sp)
//                             O = P V     (TS: A = P read from TMEM, B = V MN-major) -> slot cols [128, 256)
//   warps 2-5 / 6-9  softmax warpgroup of slot 0 / 1, one thread per query row:
//               pass 1: row max from TMEM; pass 2: p = exp2((s - max) * scale * log2e), row sum, P packed
//               to 16 bit and written back to TMEM in place (cols [0, Sp/2)); after the PV MMA: O / sum ->
//               16 bit -> swizzled smem staging -> row-contiguous 128-byte global stores.
// Scores never touch shared or global memory; K/V are read once per item.  What bounds it (GD_ATT_DBG, DESIGN.md
// 4.7): the per-item chain S -> softmax -> PV -> output with both TMEM slots in lockstep.  Tried and measured slower
// on the same box: serving the slots as their barriers complete, half a cycle apart (30.6 vs 28.4 us), and
// software-pipelined TMEM loads + four-way reductions in the softmax passes (31.7 vs 28.5 us).  Numerics match the previous
// mma.sync kernel (fp32 scores / statistics, 16-bit P, fp32 accumulation).
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kAtThreads = 320;
constexpr int kAtQBytes = 128 * 128 * 2;          // one Q tile: 2 slabs of [128 rows x 128 B]
constexpr int kAtKvRegion = 128 * 1024;           // K,V (n_qt = 2) or 2 x (K,V) (n_qt = 1)
constexpr int kAtStgBytes = 8 * 32 * 128;         // 32 rows x 128 B per softmax warp
constexpr int kAtSmemBytes = 2 * kAtQBytes + kAtKvRegion + kAtStgBytes + 256 + 1024;

struct AtParams {
  void* out;
  int S, Sp, n_qt, H, d, n_items, fmt, rev;
  int dbg;           // GD_ATT_DBG timing diagnostics (results wrong): 1 no Q/K/V loads, 2 no softmax arithmetic, 4 no stores
  float scale_log2e;
  uint32_t idesc_s, idesc_o;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAtThreads, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                    const AtParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* q_smem = smem;                                   // 2 x 32 KB
  uint8_t* kv_smem = smem + 2 * kAtQBytes;                  // 128 KB
  uint8_t* stg_smem = kv_smem + kAtKvRegion;                // 32 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg_smem + kAtStgBytes);
  uint64_t* q_full = bars;          uint64_t* q_empty = bars + 2;
  uint64_t* k_full = bars + 4;      uint64_t* k_empty = bars + 6;
  uint64_t* v_full = bars + 8;      uint64_t* v_empty = bars + 10;
  uint64_t* s_full = bars + 12;     uint64_t* p_ready = bars + 14;
  uint64_t* o_full = bars + 16;     uint64_t* slot_free = bars + 18;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  // warp index through a shuffle: warp-uniform for the compiler (uniform role branches, MMA descriptors in the uniform
  // datapath instead of an ELECT / R2UR waterfall around every UTCHMMA)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int S = p.S, Sp = p.Sp, n_qt = p.n_qt;
  const int kv_bytes = Sp * 256;                            // one K (or V): 2 slabs of [Sp rows x 128 B]
  const int slab = Sp * 128;
  // Items are dealt round-robin; when the last, partial round leaves at most half of the CTAs busy (512 items on 148
  // CTAs: 68), its items are split into their two query tiles ("solo" units with their own K/V load) so that round
  // costs half an item: 3.5 instead of 4 item times.  Items are walked from the last sequence to the first (p.rev,
  // GD_ATT_REV=0 turns it off): the in_proj GEMM before this kernel wrote the last rows last, so they are the ones still
  // in L2 -- 0.6 % of the cfg2a step.
  const int G = (int)gridDim.x, cta = (int)blockIdx.x;
  const int full = p.n_items / G, rem = p.n_items - full * G;
  const bool split_tail = n_qt == 2 && rem > 0 && 2 * rem <= G;
  const int n_whole = full + ((!split_tail && cta < rem) ? 1 : 0);      // items this CTA processes whole
  const int U_whole = n_whole * n_qt;
  const int U = U_whole + ((split_tail && cta < 2 * rem) ? 1 : 0);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_q);
    ptx::prefetch_tensormap(&tmap_kv);
  }
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

  // unit u -> (item, query tile, kv buffer); both kv-buffer use counts equal u >> 1
  // (a solo unit has an even index, so slot / use / buffer bookkeeping is the same as for a first tile)
  auto item_fwd = [&](int u) { return u < U_whole ? cta + (n_qt == 2 ? (u >> 1) : u) * G : full * G + (cta >> 1); };
  auto item_of = [&](int u) { return p.rev ? p.n_items - 1 - item_fwd(u) : item_fwd(u); };
  auto qt_of = [&](int u) { return u < U_whole ? (n_qt == 2 ? (u & 1) : 0) : (cta & 1); };
  auto kvb_of = [&](int u) { return n_qt == 2 ? 0 : (u & 1); };
  auto first_of = [&](int u) { return u >= U_whole || qt_of(u) == 0; };            // first / last unit using the
  auto last_of = [&](int u) { return u >= U_whole || qt_of(u) == n_qt - 1; };      // item's K and V

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int u = 0; u < U; ++u) {
        const int slot = u & 1, use = u >> 1;
        const int item = item_of(u), qt = qt_of(u), b = kvb_of(u);
        const int seq = item / p.H, head = item % p.H;
        const int row0 = seq * S, colq = head * 128;
        ptx::mbar_wait(&q_empty[slot], (use & 1) ^ 1);
        uint8_t* qs = q_smem + slot * kAtQBytes;
        if (p.dbg & 1) {
          ptx::mbar_arrive(&q_full[slot]);
        } else {
          ptx::mbar_arrive_expect_tx(&q_full[slot], kAtQBytes);
          ptx::tma_load_2d(qs, &tmap_q, &q_full[slot], colq, row0 + qt * 128);
          ptx::tma_load_2d(qs + 16384, &tmap_q, &q_full[slot], colq + 64, row0 + qt * 128);
        }
        if (last_of(u)) {                                   // after the item's last Q tile: its K and V
          uint8_t* ks = kv_smem + b * 2 * kv_bytes;
          uint8_t* vs = ks + kv_bytes;
          ptx::mbar_wait(&k_empty[b], (use & 1) ^ 1);
          if (p.dbg & 1) {
            ptx::mbar_arrive(&k_full[b]);
          } else {
            ptx::mbar_arrive_expect_tx(&k_full[b], kv_bytes);
            ptx::tma_load_2d(ks, &tmap_kv, &k_full[b], p.d + colq, row0);
            ptx::tma_load_2d(ks + slab, &tmap_kv, &k_full[b], p.d + colq + 64, row0);
          }
          ptx::mbar_wait(&v_empty[b], (use & 1) ^ 1);
          if (p.dbg & 1) {
            ptx::mbar_arrive(&v_full[b]);
          } else {
            ptx::mbar_arrive_expect_tx(&v_full[b], kv_bytes);
            ptx::tma_load_2d(vs, &tmap_kv, &v_full[b], 2 * p.d + colq, row0);
            ptx::tma_load_2d(vs + slab, &tmap_kv, &v_full[b], 2 * p.d + colq + 64, row0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer
    // the whole warp runs the loop (uniform control flow, barrier waits by all lanes); one elected lane issues
    {
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t q_u = __shfl_sync(0xffffffffu, ptx::smem_u32(q_smem), 0);
      const uint32_t kv_u = __shfl_sync(0xffffffffu, ptx::smem_u32(kv_smem), 0);
      for (int u0 = 0; u0 < U; u0 += 2) {
        for (int e = 0; e < 2; ++e) {                       // S = Q K^T for the pair
          const int u = u0 + e;
          if (u >= U) break;
          const int slot = u & 1, use = u >> 1, b = kvb_of(u);
          ptx::mbar_wait(&slot_free[slot], (use & 1) ^ 1);
          ptx::mbar_wait(&q_full[slot], use & 1);
          if (first_of(u)) ptx::mbar_wait(&k_full[b], use & 1);
          ptx::tc_fence_after();
          const uint32_t qa = q_u + (uint32_t)(slot * kAtQBytes);
          const uint32_t ka = kv_u + (uint32_t)(b * 2 * kv_bytes);
          const uint32_t td = tmem_u + (uint32_t)(slot * 256);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const uint64_t da = ptx::umma_desc_k_sw128(qa + (k >> 2) * 16384 + (k & 3) * 32);
              const uint64_t db = ptx::umma_desc_k_sw128(ka + (k >> 2) * slab + (k & 3) * 32);
              ptx::umma_h16(td, da, db, p.idesc_s, k != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&s_full[slot]);
            ptx::umma_commit(&q_empty[slot]);
            if (last_of(u)) ptx::umma_commit(&k_empty[b]);
          }
          __syncwarp();
        }
        for (int e = 0; e < 2; ++e) {                       // O = P V for the pair
          const int u = u0 + e;
          if (u >= U) break;
          const int slot = u & 1, use = u >> 1, b = kvb_of(u);
          ptx::mbar_wait(&p_ready[slot], use & 1);
          if (first_of(u)) ptx::mbar_wait(&v_full[b], use & 1);
          ptx::tc_fence_after();
          const uint32_t va = kv_u + (uint32_t)(b * 2 * kv_bytes + kv_bytes);
          const uint32_t tp = tmem_u + (uint32_t)(slot * 256);
          const uint32_t to = tp + 128;
          const int ksteps = Sp >> 4;
          if (ptx::elect_one()) {
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t db = ptx::umma_desc_mn_sw128(va + k * 2048, (uint32_t)slab);
              ptx::umma_h16_ts(to, tp + (uint32_t)(k * 8), db, p.idesc_o, k != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&o_full[slot]);
            if (last_of(u)) ptx::umma_commit(&v_empty[b]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax / output warpgroups
    const int slot = (warp - 2) >> 2;
    const int quarter = warp & 3;                           // TMEM lane quarter of this warp
    const uint32_t lane_bits = (uint32_t)(quarter * 32) << 16;
    const uint32_t ts = tmem_base + lane_bits + (uint32_t)(slot * 256);
    const uint32_t stg = ptx::smem_u32(stg_smem) + (uint32_t)(warp - 2) * 4096;
    const float sl2 = p.scale_log2e;
    uint16_t* out = reinterpret_cast<uint16_t*>(p.out);
    for (int u = slot; u < U; u += 2) {
      const int use = u >> 1, item = item_of(u), qt = qt_of(u);
      const int seq = item / p.H, head = item % p.H;
      ptx::mbar_wait(&s_full[slot], use & 1);
      ptx::tc_fence_after();
      // ---- pass 1: row maximum
      float mx = -INFINITY;
      for (int c = 0; c < ((p.dbg & 2) ? 0 : Sp); c += 32) {
        uint32_t r[32];
        const bool wide = c + 32 <= Sp;
        if (wide) {
          ptx::tmem_ld_32x32_nowait(ts + c, r);
        } else {
          uint32_t r16[16];
          ptx::tmem_ld_32x16_nowait(ts + c, r16);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) { r[j] = r16[j]; r[16 + j] = 0xff800000u; }
        }
        ptx::tmem_ld_wait();
        if (c + 32 > S) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c + j >= S) r[j] = 0xff800000u;             // -inf: keys past the sequence end
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(r[j]));
      }
      // ---- pass 2: probabilities -> TMEM (16-bit pairs, in place), row sum
      const float mb = -mx * sl2;
      float sum = 0.f;
      for (int c = 0; c < ((p.dbg & 2) ? 0 : Sp); c += 32) {
        uint32_t r[32];
        const bool wide = c + 32 <= Sp;
        if (wide) {
          ptx::tmem_ld_32x32_nowait(ts + c, r);
        } else {
          uint32_t r16[16];
          ptx::tmem_ld_32x16_nowait(ts + c, r16);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) { r[j] = r16[j]; r[16 + j] = 0xff800000u; }
        }
        ptx::tmem_ld_wait();
        if (c + 32 > S) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c + j >= S) r[j] = 0xff800000u;
        }
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float a = ex2_approx(fmaf(__uint_as_float(r[2 * j]), sl2, mb));
          const float b = ex2_approx(fmaf(__uint_as_float(r[2 * j + 1]), sl2, mb));
          sum += a + b;
          pk[j] = pack_h16x2(a, b, p.fmt);
        }
        if (wide) {
          ptx::tmem_st_32x16(ts + (uint32_t)(c >> 1), pk);
        } else {
          uint32_t pk8[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) pk8[j] = pk[j];
          ptx::tmem_st_32x8(ts + (uint32_t)(c >> 1), pk8);
        }
      }
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&p_ready[slot]);
      if (p.dbg & 2) sum = 1.f;
      const float inv = 1.0f / sum;

      // ---- output: O / sum -> 16 bit -> staging -> row-contiguous stores
      ptx::mbar_wait(&o_full[slot], use & 1);
      ptx::tc_fence_after();
      const int q_base = qt * 128 + quarter * 32;           // first query row of this warp inside the sequence
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32_nowait(ts + 128 + h * 64, r0);
        ptx::tmem_ld_32x32_nowait(ts + 128 + h * 64 + 32, r1);
        ptx::tmem_ld_wait();
        if (h == 1) {                                       // accumulator fully read: release the slot
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&slot_free[slot]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t w[4];
#pragma unroll
          for (int t = 0; t < 4; ++t)
            w[t] = pack_h16x2(__uint_as_float(r0[8 * j + 2 * t]) * inv, __uint_as_float(r0[8 * j + 2 * t + 1]) * inv, p.fmt);
          ptx::sts_v4u(stg + lane * 128 + ((j ^ (lane & 7)) << 4), w[0], w[1], w[2], w[3]);
#pragma unroll
          for (int t = 0; t < 4; ++t)
            w[t] = pack_h16x2(__uint_as_float(r1[8 * j + 2 * t]) * inv, __uint_as_float(r1[8 * j + 2 * t + 1]) * inv, p.fmt);
          ptx::sts_v4u(stg + lane * 128 + (((4 + j) ^ (lane & 7)) << 4), w[0], w[1], w[2], w[3]);
        }
        __syncwarp();
        const int j8 = lane & 7, rsub = lane >> 3;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int rl = it * 4 + rsub, q = q_base + rl;
          const uint4 val = ptx::lds_v4u(stg + rl * 128 + ((j8 ^ (rl & 7)) << 4));
          if (q < S && !(p.dbg & 4))
            *reinterpret_cast<uint4*>(out + ((size_t)(seq * S + q) * p.d + head * 128 + h * 64 + j8 * 8)) = val;
        }
        __syncwarp();
      }
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
// Returns GD_ERR_UNSUPPORTED (without setting an error) when the shape is outside this kernel's range so
// that gd_attention_h16 can fall back to its mma.sync flash kernel (other head dims, S > 256).
// S <= 256: this file's kernel (a whole row of scores in TMEM); longer sequences: attention_tc_long.cu (keys in
// blocks of 128, online softmax) -- up to the encoder's positional table (5000 rows, mdm.py:277) and beyond.
bool attention_tc_supported(int S, int n_heads, int head_dim) {
  return head_dim == 128 && S >= 1 && n_heads >= 1;
}

int launch_attention_tc_long(const void* qkv, void* out, int n_seq, int S, int n_heads, int fmt, cudaStream_t stream);

int launch_attention_tc(const void* qkv, void* out, int n_seq, int S, int n_heads, int fmt, cudaStream_t stream) {
  {
    static int long_min = -1;                                // GD_ATT_LONG_MIN (experiments): smallest S on the block-wise kernel
    if (long_min < 0) {
      const char* e = getenv("GD_ATT_LONG_MIN");
      long_min = e ? atoi(e) : 257;
      if (long_min > 257 || long_min < 1) long_min = 257;
    }
    if (S >= long_min) return launch_attention_tc_long(qkv, out, n_seq, S, n_heads, fmt, stream);
  }
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtSmemBytes));
    attr_set = true;
  }
  const int d = n_heads * 128;
  const int Sp = (S + 15) & ~15;
  const long rows = (long)n_seq * S;
  GD_REQUIRE(rows < (1L << 31), "gd_attention_h16: too many rows");
  CUtensorMap tq, tkv;
  int rc = make_tmap_2d(&tq, qkv, (int)rows, 3 * d, 3 * d, 128, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tkv, qkv, (int)rows, 3 * d, 3 * d, Sp, fmt);
  if (rc) return rc;
  AtParams p;
  p.out = out;
  p.S = S;
  p.Sp = Sp;
  p.n_qt = (S + 127) / 128;
  p.H = n_heads;
  p.d = d;
  p.n_items = n_seq * n_heads;
  p.fmt = fmt;
  {
    static int rev = -1;
    if (rev < 0) {
      const char* e = getenv("GD_ATT_REV");
      rev = e ? atoi(e) : 1;
    }
    p.rev = rev;
    static int dbg = -1;
    if (dbg < 0) {
      const char* e = getenv("GD_ATT_DBG");
      dbg = e ? atoi(e) : 0;
    }
    p.dbg = dbg;
  }
  p.scale_log2e = (1.0f / sqrtf(128.0f)) * 1.4426950408889634f;
  p.idesc_s = ptx::umma_idesc_h16_f32(128, Sp, fmt == GD_FP16);
  p.idesc_o = ptx::umma_idesc_h16_f32_bmn(128, 128, fmt == GD_FP16);
  const int sms = num_sms();
  GD_REQUIRE(sms > 0, "gd_attention_h16: cannot query the SM count");
  // few items (small batches): one CTA per query tile instead of per item when that fits (the kernel's tail split)
  int grid = p.n_items < sms ? p.n_items : sms;
  if (p.n_qt == 2 && 2 * p.n_items <= sms) grid = 2 * p.n_items;
  GD_CUDA_OK(launch_pdl(PDL_ATTENTION, attention_tc_kernel, dim3(grid), dim3(kAtThreads), kAtSmemBytes, stream, tq, tkv, p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

}  // namespace gd
