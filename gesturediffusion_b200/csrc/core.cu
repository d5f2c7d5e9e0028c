// core.cu -- library state shared by all entry points: error string, launch counter,
// epilogue validation, ABI version.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace gd {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

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

