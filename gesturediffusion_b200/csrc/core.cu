// core.cu -- library state shared by all entry points: error string, launch counter,
// epilogue validation, ABI version.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace gd {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
cudaAccessPolicyWindow g_l2_window = {};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int validate_epilogue(const gdEpilogue* ep, int M, int N) {
  GD_REQUIRE(ep->out_f32 || ep->out_h16 || ep->out_t, "epilogue: no output pointer");
  if (ep->rowtab) GD_REQUIRE(ep->tab_period > 0 && ep->ld_tab >= N, "epilogue: bad rowtab period/ld");
  if (ep->seqbias) GD_REQUIRE(ep->seq_rows > 0 && ep->ld_seq >= N, "epilogue: bad seqbias rows/ld");
  if (ep->residual) GD_REQUIRE(ep->ld_res >= N, "epilogue: ld_res < N");
  if (ep->out_f32) GD_REQUIRE(ep->ld_out >= N, "epilogue: ld_out < N");
  if (ep->out_h16)
    GD_REQUIRE(ep->ld_out_h16 >= N && (ep->h16_fmt == GD_BF16 || ep->h16_fmt == GD_FP16),
               "epilogue: ld_out_h16 < N or bad h16_fmt");
  if (ep->out_t) GD_REQUIRE(ep->t_s > 0 && ep->t_t > 0 && ep->t_skip >= 0, "epilogue: bad transposed-store dims");
  if (ep->map_t > 0) GD_REQUIRE(ep->map_s >= ep->map_t + ep->map_off, "epilogue: rowmap overlaps");
  GD_REQUIRE(ep->act >= GD_ACT_NONE && ep->act <= GD_ACT_SILU, "epilogue: unknown activation %d", ep->act);
  (void)M;
  return GD_OK;
}

}  // namespace gd

extern "C" int gd_abi_version(void) { return GD_ABI_VERSION; }
extern "C" const char* gd_last_error(void) { return gd::g_err; }
extern "C" int64_t gd_launch_count(void) { return (int64_t)gd::g_launches.load(); }

/* L2 residency of the residual stream: [base, base + bytes) is attached as a persisting access-policy window to every
 * tcgen05 kernel launched afterwards (bytes = 0 turns it off).  Sets the device's persisting-L2 carve-out to
 * min(bytes, device maximum) and scales the hit ratio when the window is larger than the carve-out.  Returns the
 * carve-out in bytes through *granted (may be NULL).  Host-side state only; call outside stream capture. */
extern "C" int gd_set_l2_window(void* base, int64_t bytes, int64_t* granted) {
  using namespace gd;
  if (granted) *granted = 0;
  if (base == nullptr || bytes <= 0) {
    g_l2_window = cudaAccessPolicyWindow{};
    return GD_OK;
  }
  int dev = 0, max_persist = 0, max_window = 0;
  GD_CUDA_OK(cudaGetDevice(&dev));
  GD_CUDA_OK(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
  GD_CUDA_OK(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev));
  if (max_persist <= 0 || max_window <= 0) {
    g_l2_window = cudaAccessPolicyWindow{};
    return GD_OK;                                            // not supported: silently off (a hint, not semantics)
  }
  const size_t carve = (size_t)bytes < (size_t)max_persist ? (size_t)bytes : (size_t)max_persist;
  GD_CUDA_OK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
  const size_t win = (size_t)bytes < (size_t)max_window ? (size_t)bytes : (size_t)max_window;
  cudaAccessPolicyWindow w = {};
  w.base_ptr = base;
  w.num_bytes = win;
  w.hitRatio = win <= carve ? 1.0f : (float)((double)carve / (double)win);
  w.hitProp = cudaAccessPropertyPersisting;
  w.missProp = cudaAccessPropertyStreaming;
  g_l2_window = w;
  if (granted) *granted = (int64_t)carve;
  return GD_OK;
}
