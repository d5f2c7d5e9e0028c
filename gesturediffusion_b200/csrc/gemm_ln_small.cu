// gemm_ln_small.cu -- GEMM + bias + residual + LayerNorm for FEW rows (batch 1-16: a few row tiles):
//   X = LayerNorm(residual + A @ W^T + bias) * gamma + beta        (mdm.py:90-96, post-norm sub-layer tail)
// gemm_ln.cu gives a 128-row tile to a 2-CTA cluster; with four row tiles that is 8 of 148 SMs, each streaming 8-16
// operand stages of 48 KB behind one another, so below ~5 k rows the engine ran a plain GEMM (N spread over the SMs) and
// gd_layernorm as two launches.  Here a row tile belongs to a cluster of EIGHT CTAs: CTA r owns output columns
// [64 r, 64 r + 64) (UMMA 128 x 64 x 16), its K loop is 8 stages of 24 KB that are all in flight at once (K = 512), the
// epilogue thread holds its row's 64 values in registers (residual prefetched during the main loop), the eight partial
// (mean, M2) of a row travel to all eight CTAs with st.async (data + transaction bytes on the destination's mbarrier) and
// are combined by Chan's formula in rank order, and the outputs leave as plain 16-byte stores.  One tile per CTA: no
// persistence, no phases.  N = 512 only.
#include <stdlib.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace gd {

constexpr int kSmBM = 128, kSmBK = 64, kSmBN = 64, kSmCluster = 8, kSmThreads = 192, kSmStages = 8;
constexpr int kSmABytes = kSmBM * kSmBK * 2, kSmBBytes = kSmBN * kSmBK * 2;
constexpr int kSmStageBytes = kSmABytes + kSmBBytes;
constexpr int kSmStatOff = kSmStages * kSmStageBytes;            // [part][row] (mean, M2)
constexpr int kSmBarOff = kSmStatOff + kSmCluster * kSmBM * 8;
constexpr int kSmSmemBytes = kSmBarOff + 256;

struct SmParams {
  const float* bias; const float* gamma; const float* beta;
  const float* residual; int ld_res;
  float* out_f32; int ld_out;
  void* out_h16; int ld_out_h16; int h16_fmt;
  float eps;
  int M, K;
  uint32_t idesc;
};

__global__ void __cluster_dims__(kSmCluster, 1, 1) __launch_bounds__(kSmThreads, 1)
gemm_ln_small_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                             const SmParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmBarOff);
  uint64_t* full_bar = bars;                                 // [stages]
  uint64_t* empty_bar = full_bar + kSmStages;                // [stages]
  uint64_t* tfull_bar = empty_bar + kSmStages;
  uint64_t* stat_bar = tfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + 1);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  const uint32_t rank = ptx::cluster_ctarank();
  const int m0 = ((int)blockIdx.x / kSmCluster) * kSmBM;
  const int n0 = (int)rank * kSmBN;
  const int k_blocks = (p.K + kSmBK - 1) / kSmBK;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_w);
    for (int s = 0; s < kSmStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    ptx::mbar_init(tfull_bar, 1);
    ptx::mbar_init(stat_bar, 1);                             // one expect_tx arrival; the partial statistics are the bytes
    ptx::fence_barrier_init();
    ptx::mbar_arrive_expect_tx(stat_bar, kSmCluster * kSmBM * 8);
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, 64);
  ptx::tc_fence_before();
  __syncthreads();
  // Every CTA's stat_bar must be armed before a peer's st.async lands: all threads ARRIVE on the cluster barrier here and
  // WAIT where they need it -- the epilogue warps just before their first remote store, i.e. after the main loop, so the
  // cluster-wide rendezvous costs nothing on the way to the first TMA load.
  ptx::cluster_arrive();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::grid_dep_wait();                                      // prologue above overlaps the previous kernel's tail
  ptx::grid_dep_launch();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < k_blocks; ++kb) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * kSmStageBytes;
        ptx::mbar_arrive_expect_tx(&full_bar[stage], kSmStageBytes);
        ptx::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kSmBK, m0);
        ptx::tma_load_2d(sa + kSmABytes, &tmap_w, &full_bar[stage], kb * kSmBK, n0);
        if (++stage == kSmStages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
    ptx::cluster_wait();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ UMMA issuer (whole warp, one elected lane)
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t smem_u = __shfl_sync(0xffffffffu, ptx::smem_u32(smem), 0);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    for (int kb = 0; kb < k_blocks; ++kb) {
      ptx::mbar_wait(&full_bar[stage], phase);
      ptx::tc_fence_after();
      const uint32_t sa = smem_u + (uint32_t)(stage * kSmStageBytes);
      const uint32_t sb = sa + kSmABytes;
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < kSmBK / 16; ++k)
          ptx::umma_h16(tmem_u, ptx::umma_desc_k_sw128(sa + k * 32), ptx::umma_desc_k_sw128(sb + k * 32), p.idesc,
                        (kb | k) != 0 ? 1u : 0u);
        ptx::umma_commit(&empty_bar[stage]);
      }
      __syncwarp();
      if (++stage == kSmStages) { stage = 0; phase ^= 1; }
    }
    if (ptx::elect_one()) ptx::umma_commit(tfull_bar);
    __syncwarp();
    ptx::cluster_wait();
  } else {
    // ------------------------------------------------------------------ epilogue: warps 2..5, one row per thread
    const int quarter = warp & 3;                            // the TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const int m = m0 + row;
    const bool live = m < p.M;
    float y[kSmBN];
    {
      // bias + residual, fetched while the main loop runs
      const float* rp = p.residual + (size_t)(live ? m : 0) * p.ld_res + n0;
#pragma unroll
      for (int q = 0; q < kSmBN / 4; ++q) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + 4 * q));
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) r = *reinterpret_cast<const float4*>(rp + 4 * q);
        y[4 * q] = r.x + b.x;
        y[4 * q + 1] = r.y + b.y;
        y[4 * q + 2] = r.z + b.z;
        y[4 * q + 3] = r.w + b.w;
      }
    }
    ptx::mbar_wait(tfull_bar, 0);
    ptx::tc_fence_after();
    {
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
      uint32_t va[32], vb[32];
      ptx::tmem_ld_32x32_nowait(taddr, va);
      ptx::tmem_ld_32x32_nowait(taddr + 32, vb);
      ptx::tmem_ld_wait_on(va);
      ptx::tmem_ld_wait_on(vb);
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        y[e] += __uint_as_float(va[e]);
        y[32 + e] += __uint_as_float(vb[e]);
      }
    }
    ptx::tc_fence_before();
    // partial statistics of the 64 columns (two-pass: no E[x^2] - mean^2 cancellation)
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int e = 0; e < kSmBN; e += 4) {
      s0 += y[e];
      s1 += y[e + 1];
      s2 += y[e + 2];
      s3 += y[e + 3];
    }
    const float pmean = ((s0 + s1) + (s2 + s3)) * (1.0f / (float)kSmBN);
    float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
    for (int e = 0; e < kSmBN; e += 4) {
      const float d0 = y[e] - pmean, d1 = y[e + 1] - pmean, d2 = y[e + 2] - pmean, d3 = y[e + 3] - pmean;
      q0 = fmaf(d0, d0, q0);
      q1 = fmaf(d1, d1, q1);
      q2 = fmaf(d2, d2, q2);
      q3 = fmaf(d3, d3, q3);
    }
    const float pm2 = (q0 + q1) + (q2 + q3);
    // to every CTA of the cluster (this one included): slot [rank][row]
    const uint32_t stats_s = ptx::smem_u32(smem + kSmStatOff);
    const uint32_t sbar_s = ptx::smem_u32(stat_bar);
    const uint32_t off = (rank * kSmBM + (uint32_t)row) * 8u;
    ptx::cluster_wait();                                     // the peers' barriers are armed (arrive: kernel prologue)
#pragma unroll
    for (uint32_t r = 0; r < (uint32_t)kSmCluster; ++r)
      ptx::st_async_v2(ptx::mapa(stats_s, r) + off, pmean, pm2, ptx::mapa(sbar_s, r));
    ptx::mbar_wait(stat_bar, 0);
    float pm[kSmCluster], pq[kSmCluster];
    float mean = 0.f;
#pragma unroll
    for (int r = 0; r < kSmCluster; ++r) {
      const uint32_t a = stats_s + (uint32_t)(r * kSmBM + row) * 8u;
      asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(pm[r]), "=f"(pq[r]) : "r"(a) : "memory");
      mean += pm[r];
    }
    mean *= 1.0f / (float)kSmCluster;
    float m2 = 0.f;
#pragma unroll
    for (int r = 0; r < kSmCluster; ++r) m2 += pq[r] + (float)kSmBN * (pm[r] - mean) * (pm[r] - mean);
    const float rstd = rsqrtf(m2 * (1.0f / (float)(kSmCluster * kSmBN)) + p.eps);
    const float shift = -mean * rstd;
#pragma unroll
    for (int q = 0; q < kSmBN / 4; ++q) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + n0 + 4 * q));
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.beta + n0 + 4 * q));
      y[4 * q] = fmaf(fmaf(y[4 * q], rstd, shift), g.x, b.x);
      y[4 * q + 1] = fmaf(fmaf(y[4 * q + 1], rstd, shift), g.y, b.y);
      y[4 * q + 2] = fmaf(fmaf(y[4 * q + 2], rstd, shift), g.z, b.z);
      y[4 * q + 3] = fmaf(fmaf(y[4 * q + 3], rstd, shift), g.w, b.w);
    }
    if (live) {
      if (p.out_f32) {
        float* op = p.out_f32 + (size_t)m * p.ld_out + n0;
#pragma unroll
        for (int q = 0; q < kSmBN / 4; ++q)
          *reinterpret_cast<float4*>(op + 4 * q) = make_float4(y[4 * q], y[4 * q + 1], y[4 * q + 2], y[4 * q + 3]);
      }
      if (p.out_h16) {
        uint16_t* hp = reinterpret_cast<uint16_t*>(p.out_h16) + (size_t)m * p.ld_out_h16 + n0;
#pragma unroll
        for (int q = 0; q < kSmBN / 8; ++q) {
          uint4 w;
          w.x = pack_h16x2(y[8 * q], y[8 * q + 1], p.h16_fmt);
          w.y = pack_h16x2(y[8 * q + 2], y[8 * q + 3], p.h16_fmt);
          w.z = pack_h16x2(y[8 * q + 4], y[8 * q + 5], p.h16_fmt);
          w.w = pack_h16x2(y[8 * q + 6], y[8 * q + 7], p.h16_fmt);
          *reinterpret_cast<uint4*>(hp + 8 * q) = w;
        }
      }
    }
  }

  // No cluster barrier at the end: the only remote accesses are the st.async pushes, and a CTA has waited for all eight
  // partials to land in its own shared memory before it gets here; its own pushes carry their data with them.
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 64);
  }
}

// gd_gemm_h16_ln (gemm_ln.cu) sends N = 512 problems of few rows here; arguments are validated there.
int launch_gemm_ln_small(const void* A, int lda, const void* W, int ldw, int M, int K, int fmt, const float* bias,
                         const float* residual, int ld_res, const float* gamma, const float* beta, float eps,
                         float* out_f32, int ld_out, void* out_h16, int ld_out_h16, int h16_fmt, cudaStream_t stream) {
  SmParams p;
  p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.residual = residual; p.ld_res = ld_res;
  p.out_f32 = out_f32; p.ld_out = ld_out;
  p.out_h16 = out_h16; p.ld_out_h16 = ld_out_h16; p.h16_fmt = h16_fmt;
  p.eps = eps; p.M = M; p.K = K;
  p.idesc = ptx::umma_idesc_h16_f32(kSmBM, kSmBN, fmt == GD_FP16);
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device()];
  if (!attr_set) {
    GD_CUDA_OK(cudaFuncSetAttribute(gemm_ln_small_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmSmemBytes));
    attr_set = true;
  }
  CUtensorMap ta, tw;
  int rc = make_tmap_2d(&ta, A, M, K, lda, kSmBM, fmt);
  if (rc) return rc;
  rc = make_tmap_2d(&tw, W, kSmCluster * kSmBN, K, ldw, kSmBN, fmt);
  if (rc) return rc;
  const int m_tiles = ceil_div(M, kSmBM);
  GD_CUDA_OK(launch_pdl(PDL_GEMM_LN, gemm_ln_small_tcgen05_kernel, dim3(kSmCluster * m_tiles), dim3(kSmThreads),
                        kSmSmemBytes, stream, ta, tw, p));
  GD_LAUNCH_CHECK();
  return GD_OK;
}

// Rows up to which the few-row kernel is used: as many row tiles as 8-CTA clusters can be resident at once (a second
// wave doubles the kernel's time: 0.30 -> 0.41 ms per denoise step between 1 464 and 1 952 rows with a fixed 2 304),
// capped at 2 304; GD_LN_SMALL_MAX_ROWS overrides (0 = never).
int gemm_ln_small_max_rows() {
  static int rows_dev[kMaxDevices];
  static bool known_dev[kMaxDevices] = {};
  const int dev = current_device();
  if (!known_dev[dev]) {
    int rows = 2304;
    const char* e = getenv("GD_LN_SMALL_MAX_ROWS");
    if (e) {
      rows = atoi(e);
    } else {
      (void)cudaFuncSetAttribute(gemm_ln_small_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmSmemBytes);
      cudaLaunchConfig_t qc = {};
      qc.gridDim = dim3(kSmCluster * 32);
      qc.blockDim = dim3(kSmThreads);
      qc.dynamicSmemBytes = kSmSmemBytes;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, gemm_ln_small_tcgen05_kernel, &qc) == cudaSuccess && n > 0) {
        if (n * kSmBM < rows) rows = n * kSmBM;
      } else {
        (void)cudaGetLastError();
      }
    }
    rows_dev[dev] = rows;
    known_dev[dev] = true;
  }
  return rows_dev[dev];
}

}  // namespace gd

extern "C" int gd_gemm_ln_small_max_rows(void) { return gd::gemm_ln_small_max_rows(); }
