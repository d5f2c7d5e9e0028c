/* gd_b200.h -- C ABI of the B200-native GestureDiffusion/MDM sampling hot path.
 *
 * The reference (AI-Unicamp/GestureDiffusion) has no native code and no FFI: its
 * "plugin API" is the Python surface of diffusion/respace.py, diffusion/gaussian_diffusion.py,
 * model/mdm.py, model/mdm_old.py and model/cfg_sampler.py.  This header is the boundary
 * that a Python (ctypes) binding of that surface calls; every entry point names the
 * reference code it replaces (file:line, relative to the reference root).
 *
 * Conventions
 *  - plain pointers and sizes only; all pointers are DEVICE pointers unless noted;
 *  - `stream` is a cudaStream_t passed as void*; nothing synchronises, allocates or reads
 *    back inside a call, so every call is CUDA-graph capturable;
 *  - return 0 on success, negative gdStatus otherwise; gd_last_error() gives the message
 *    (thread-local, host pointer);
 *  - row-major everywhere; "act" buffers are fp16 (default) or bf16 in the 16-bit modes and fp32 in fp32 mode;
 *  - weights use the nn.Linear layout W[N][K] (K contiguous) exactly as in the reference
 *    state_dict, so y = x @ W^T + b.
 */
#ifndef GD_B200_H
#define GD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GD_ABI_VERSION 1

typedef enum gdStatus {
  GD_OK = 0,
  GD_ERR_INVALID = -1,   /* bad argument (shape, alignment, null pointer) */
  GD_ERR_CUDA = -2,      /* a CUDA runtime/driver call failed */
  GD_ERR_UNSUPPORTED = -3
} gdStatus;

typedef enum gdAct { GD_ACT_NONE = 0, GD_ACT_GELU_ERF = 1, GD_ACT_SILU = 2 } gdAct;

/* Element formats of "act" buffers.  The two 16-bit formats run the same tcgen05 kind::f16 path at
 * the same rate and bytes; GD_FP16 has 3 more mantissa bits (measured per-step parity 5e-4 vs 4e-3
 * for GD_BF16), GD_BF16 has fp32's exponent range.  fp16 stores saturate to +-65504. */
typedef enum gdFmt { GD_F32 = 0, GD_BF16 = 1, GD_FP16 = 2 } gdFmt;

/* Fused GEMM epilogue.  For accumulator element (m, n) of  A[M,K] @ W[N,K]^T :
 *   v  = acc + bias[n] + rowtab[((m % tab_period) + tab_offset) * ld_tab + n]
 *   for copy c in [0, dup):                      (dup = 0 or 1 means a single copy)
 *     w = v + seqbias[(c * seq_count + m / seq_rows) * ld_seq + n]
 *     r = rowmap(m) + c * dup_row_stride         rowmap(m) = (m / map_t) * map_s + map_off + m % map_t
 *                                                 (identity when map_t == 0)
 *     w += residual[r * ld_res + n];  w = act(w)
 *     out_f32[r * ld_out + n] = w;  out_h16[r * ld_out_h16 + n] = (h16_fmt)(w)
 *   transposed store (out_t != NULL, replaces the row-major stores):
 *     seq = m / t_s, s = m % t_s;  if (s >= t_skip) out_t[((seq * N + n) * t_t) + s - t_skip] = w
 * Every pointer may be NULL to disable that term.  This covers, with one GEMM each:
 *   InputProcess + additive PositionalEncoding (mdm_old.py:102-112), project_to_lat with its
 *   per-sample coarse term (mdm.py:151-169), in_proj/out_proj/linear1/linear2 with bias, erf-GELU
 *   and the residual add of nn.TransformerEncoderLayer (mdm.py:90-96), and OutputProcess with
 *   its [T,B,J]->[B,J,1,T] permute and the output[1:] slice (mdm.py:219-222, 372-380). */
typedef struct gdEpilogue {
  const float* bias;
  const float* rowtab;   int32_t tab_period, tab_offset, ld_tab;
  const float* seqbias;  int32_t seq_rows, seq_count, ld_seq;
  const float* residual; int32_t ld_res;
  int32_t act;
  int32_t map_t, map_s, map_off;
  int32_t dup, dup_row_stride;
  float* out_f32;        int32_t ld_out;
  void*  out_h16;        int32_t ld_out_h16; int32_t h16_fmt;   /* GD_BF16 or GD_FP16 */
  float* out_t;          int32_t t_s, t_t, t_skip;
} gdEpilogue;

int gd_abi_version(void);
const char* gd_last_error(void);                 /* host string, thread-local */

/* Dense projections (replace every aten::addmm on the path: mdm.py:350-356, 151-169, 90-96,
 * 372-380; mdm_old.py:102-118).  h16: A and W both bf16 or both fp16 (`fmt`), tcgen05.mma (UMMA
 * 128xBNx16 kind::f16, fp32 accumulate in TMEM), operands staged by TMA with 128B swizzle; requires
 * sm_100a, lda/ldw multiples of 8 elements and 16-byte aligned bases.  f32: SIMT fp32
 * reference-precision path ("fp32 mode" + one-off setup). */
int gd_gemm_h16(const void* A, int32_t lda, const void* W, int32_t ldw,
                int32_t M, int32_t N, int32_t K, int32_t fmt, const gdEpilogue* ep, void* stream);
int gd_gemm_f32(const float* A, int32_t lda, const float* W, int32_t ldw,
                int32_t M, int32_t N, int32_t K, const gdEpilogue* ep, void* stream);

/* Post-norm sub-layer tail in one kernel (out_proj + residual + norm1, linear2 + residual + norm2 of
 * nn.TransformerEncoderLayer, mdm.py:90-96; norm_first=False, eps 1e-5):
 *   X = LayerNorm(residual + A[M,K] @ W[N,K]^T + bias) * gamma + beta
 * out_f32 [M, ld_out] and / or out_h16 [M, ld_out_h16] (h16_fmt).  N must be 256 or 512 (= 2 x the per-CTA
 * accumulator width: a row is normalised across the two CTAs of a thread-block cluster through distributed
 * shared memory); other widths return GD_ERR_UNSUPPORTED and callers use gd_gemm_h16 + gd_layernorm.
 * residual may alias out_f32 (each element is read before it is written, by the same warp).  ld_res / ld_out are
 * multiples of 4 elements, ld_out_h16 of 8 (the outputs leave through TMA stores). */
/* Largest M for which gd_gemm_h16_ln (N = 512) runs its few-row form -- one 8-CTA cluster per 128-row tile, all tiles
 * resident at once: 128 x the number of such clusters the current device holds (cudaOccupancyMaxActiveClusters), at most
 * 2304; environment GD_LN_SMALL_MAX_ROWS overrides (0 = never).  Callers that choose between this call and
 * gd_gemm_h16 + gd_layernorm for small batches (the engine) ask here.  No reference counterpart (dispatch helper). */
int gd_gemm_ln_small_max_rows(void);
int gd_gemm_h16_ln(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                   int32_t fmt, const float* bias, const float* residual, int32_t ld_res,
                   const float* gamma, const float* beta, float eps, float* out_f32, int32_t ld_out,
                   void* out_h16, int32_t ld_out_h16, int32_t h16_fmt, void* stream);

/* Same sub-layer tail with the fp32 residual stream carried through HBM as a 16-bit pair (the kernel's memory phases
 * run at HBM speed): x = hi + lo with hi = round16(x) -- the operand the next GEMM reads anyway -- and
 * lo = round16(x - hi), exact to ~2^-22 relative for fp16 / 2^-16 for bf16.  The residual comes either as fp32
 * (res_f32, first sub-layer of the stack; res_hi = res_lo = NULL) or as a pair (res_f32 = NULL); the result leaves as
 * out_hi and, unless NULL (last sub-layer), out_lo.  res_hi / res_lo may alias out_hi / out_lo.  All 16-bit arrays
 * use h16_fmt and leading dimensions that are multiples of 8.  Same N restriction as gd_gemm_h16_ln. */
int gd_gemm_h16_ln_split(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                         int32_t fmt, const float* bias, const float* res_f32, int32_t ld_res,
                         const void* res_hi, const void* res_lo, int32_t ld_res16,
                         const float* gamma, const float* beta, float eps,
                         void* out_hi, void* out_lo, int32_t ld_out16, int32_t h16_fmt, void* stream);

/* The whole feed-forward sub-layer of nn.TransformerEncoderLayer (mdm.py:90-96; torch's _ff_block + norm2, post-norm,
 * erf-GELU) in one kernel, on the split residual stream of gd_gemm_h16_ln_split:
 *   x' = LayerNorm((x_hi + x_lo) + gelu(x_hi @ W1^T + b1) @ W2^T + b2) * gamma + beta
 * x_hi / x_lo / out_hi / out_lo are [M, N] 16-bit (fmt) with leading dimensions ld_x / ld_out, W1 is [F, N] (linear1),
 * W2 is [N, F] (linear2); out_lo may be NULL (last sub-layer); the outputs may alias the inputs.  The hidden
 * activation gelu(linear1) never goes to HBM.  N = 512 and F a multiple of 256 (>= 512); GD_ERR_UNSUPPORTED
 * otherwise (callers use gd_gemm_h16 + gd_gemm_h16_ln_split). */
int gd_ffn_h16_ln_split(const void* x_hi, const void* x_lo, int32_t ld_x, const void* W1, int32_t ldw1,
                        const float* b1, const void* W2, int32_t ldw2, const float* b2, int32_t M, int32_t N,
                        int32_t F, int32_t fmt, const float* gamma, const float* beta, float eps,
                        void* out_hi, void* out_lo, int32_t ld_out, void* stream);


/* Debug aid (tools/trace_ffn.py): device buffer of >= 3 * 4096 uint64 that CTA 0 of every later
 * gd_ffn_h16_ln_split launch fills with (tag << 48 | SM clock) records; NULL (default) turns it off. */
int gd_ffn_trace(void* buf);

/* Full softmax self-attention of nn.MultiheadAttention inside nn.TransformerEncoderLayer
 * (mdm.py:90-96,216; no mask, scale 1/sqrt(head_dim)).  qkv is the packed in_proj output
 * [n_seq*S, 3*d] (q | k | v, heads contiguous inside each); out is [n_seq*S, d]. */
int gd_attention_h16(const void* qkv, void* out, int32_t n_seq, int32_t S, int32_t n_heads,
                     int32_t head_dim, int32_t fmt, void* stream);
int gd_attention_f32(const float* qkv, float* out, int32_t n_seq, int32_t S, int32_t n_heads,
                     int32_t head_dim, void* stream);

/* LayerNorm (norm1/norm2 of nn.TransformerEncoderLayer, eps 1e-5): rows x d fp32 in,
 * fp32 and/or 16-bit (`h16_fmt`) out (either may be NULL). */
int gd_layernorm(const float* in, const float* gamma, const float* beta, float* out_f32,
                 void* out_h16, int32_t h16_fmt, int32_t rows, int32_t d, float eps, void* stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The fp32-mode post-norm
 * sub-layer tail  X = LayerNorm(residual + A[M,K] @ W[N,K]^T + bias) * gamma + beta  = gd_gemm_f32 (residual
 * epilogue) -> [M, N] fp32 scratch -> gd_layernorm, with the scratch in a file-scope `static __device__` array inside
 * the library instead of a caller workspace; the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limit: M * N <= gd_noalloc_max_elems() = 1 Mi fp32 elements (4 MiB, held by every GPU that loads the library until
 * exit); larger problems return GD_ERR_INVALID with a message.  N as gd_layernorm (multiple of 128 in [128, 1024]).
 * All calls on a device SHARE the array, which keeps the previous call's contents (nothing needs to start at zero:
 * the GEMM overwrites the whole [M, N] region before it is read); calls that overlap on the GPU -- different streams,
 * concurrent graph replays -- race on it.  The device must be current when the call is made. */
int gd_noalloc_max_elems(void);
int gd_linear_res_ln_f32_noalloc(const float* A, int32_t lda, const float* W, int32_t ldw, int32_t M, int32_t N,
                                 int32_t K, const float* bias, const float* residual, int32_t ld_res,
                                 const float* gamma, const float* beta, float eps, float* out_f32, void* out_h16,
                                 int32_t h16_fmt, void* stream);

/* RoPE + causal windowed LocalAttention + second RoPE (mdm.py:176-213; local_attention.py:43-62,
 * 92-172): z [n_seq, T, d] fp32 -> encoder-input rows 1..T of each sequence (token 0 is written by
 * gd_cond_tokens): x_f32 / x_h16 [n_seq, T+1, d].  q = k = v, `heads` heads of e = d/heads, window
 * `window`, one window of look-back; rope_cos/rope_sin are fp32 [T+1, e] tables of
 * cos/sin(cat(pos x inv_freq, pos x inv_freq)) built from the checkpoint's rel_pos.inv_freq: rows 0..T-1
 * rotate the attention input, rows 1..T rotate its output (the cond token sits at position 0).
 * T must be a multiple of `window` (the reference fails otherwise, local_attention.py:100-112). */
int gd_local_attention(const float* z, const float* rope_cos, const float* rope_sin, float* x_f32,
                       void* x_h16, int32_t h16_fmt, int32_t n_seq, int32_t T, int32_t d, int32_t heads,
                       int32_t window, void* stream);

/* Per-step conditioning token (mdm_old.py:94-99,109-112; mdm.py:130,154-160,197):
 *   tok[c, b, :] = temb_table[t[b], :] + seed_emb[c, b, :] (+ pe0[:])      c in [0, n_branch)
 * written to row (c*B + b) * S of x_f32 / x_h16 (row stride d), and optionally
 *   coarse_out[c*B + b, :] = coarse_tab[t[b], :] + coarse_seed[c, b, :]
 * (the W_c . coa term of project_to_lat, by linearity).  t is int64 [B] of ORIGINAL-schedule
 * timesteps (respace.py:124-129 already applied). */
int gd_cond_tokens(const int64_t* t, const float* temb_table, const float* seed_emb, const float* pe0,
                   const float* coarse_tab, const float* coarse_seed, float* coarse_out,
                   float* x_f32, void* x_h16, int32_t h16_fmt, int32_t B, int32_t n_branch, int32_t S,
                   int32_t d, void* stream);

/* The fused guidance + posterior + noise update (cfg_sampler.py:28; gaussian_diffusion.py:253-275,
 * 307-311, 354-376, 532-547).  All motion tensors are [B, J, T] fp32 (== [B,J,1,T]).
 *   x0   = uncond ? uncond + scale[b]*(cond - uncond) : cond          (model_mean_type START_X)
 *          (EPSILON: x0 = c_recip*x_t - c_recipm1*eps, on the guided eps)
 *   x0   = inpaint_mask ? inpaint_mask ? inpainted : x0 ...            (optional blend)
 *   x0   = clip ? clamp(x0,-1,1) : x0
 *   mean = coef1*x0 + coef2*x_t ;  x_prev = mean + (step != 0) * exp(0.5*logvar) * noise
 * Coefficients are read on the device from coef[6][n_steps] fp32 (coef1, coef2, logvar,
 * sqrt_recip_acp, sqrt_recipm1_acp, reserved) at index *step_ptr -- the float64 tables of
 * gaussian_diffusion.py:160-197 cast once to fp32 exactly like _extract_into_tensor (:1605).
 * Noise: `noise` != NULL reads injected noise (const_noise -> sample 0's noise for all b), else
 * Philox4x32-10 keyed (seed, sample_base + b (or 0 if const_noise), *step_ptr, element) is
 * generated in-kernel.  Optional outputs: pred_xstart, mean, and the next step's GEMM operand
 * a_next[(b*T + t) * ld_a + j] (format a_fmt: GD_F32 / GD_BF16 / GD_FP16), i.e. InputProcess's permute (mdm.py:350-352) fused
 * into this kernel's store.  If next_t != NULL the kernel also advances the loop state on the
 * device: step_ptr[0] -= 1 and next_t[b] = timestep_map[step-1] (respace.py:124-129). */
typedef struct gdPosteriorArgs {
  const float* cond; const float* uncond; const float* scale;
  const float* x_t; const float* noise;
  const uint8_t* inpaint_mask; const float* inpainted;
  const float* coef; int32_t n_steps; int32_t* step_ptr;
  const int64_t* timestep_map; int64_t* next_t;
  float* x_prev; float* pred_xstart; float* mean;
  void* a_next; int32_t ld_a; int32_t a_fmt;
  int32_t B, J, T;
  int32_t eps_mode, clip, const_noise;
  uint64_t seed; int64_t sample_base;
  const uint64_t* seed_dev;   /* optional: Philox seed read from device memory (one captured graph, many calls) */
  int32_t ddim;               /* 1: DDIM update (gaussian_diffusion.py:732-782): coef rows 0/1/2 hold
                                 sqrt(alpha_bar_prev), sqrt(1 - alpha_bar_prev - sigma^2), sigma(eta) per step */
} gdPosteriorArgs;
int gd_posterior_step(const gdPosteriorArgs* args, void* stream);

/* InputProcess's permute for the first step / standalone forward (mdm.py:350-352, mdm_old.py:102-103):
 * src [B, C, T] fp32 -> dst[(b*T + t) * ld + col_off + c] (dst_fmt: GD_F32 / GD_BF16 / GD_FP16). */
int gd_pack_rows(const float* src, void* dst, int32_t B, int32_t C, int32_t T, int32_t ld,
                 int32_t col_off, int32_t dst_fmt, void* stream);

/* Post-processing of one generated chunk (sample/generate.py:132-146 with data_loaders/gesture/data/dataset.py:
 * 118-119): inverse normalisation x * std + mean per feature, then the 6-interleaved GENEA layout
 * [rot_x, rot_y, rot_z, pos_x, pos_y, pos_z] per joint is split into
 *   rot[b][joint][c][t_off + t] and pos[b][joint][c][t_off + t]   ([B, J/6, 3, T_total] fp32, either may be NULL),
 * i.e. the reference's permuted [B, n_joints, 3, chunk_len] tensors written straight into the take at the chunk's
 * frame offset.  x is [B, J, T] fp32 (= [B, J, 1, T]); mean / std are [J] or NULL (0 / 1). */
int gd_unnormalize_split(const float* x, const float* mean, const float* stdv, float* pos, float* rot, int32_t B,
                         int32_t J, int32_t T, int32_t t_off, int32_t T_total, void* stream);

/* q_sample (gaussian_diffusion.py:233-251) for skip_timesteps/init_image:
 * out = a[step]*x0 + b[step]*noise with a/b fp32 device tables. */
int gd_q_sample(const float* x0, const float* noise, const float* sqrt_acp, const float* sqrt_1macp,
                int32_t step, float* out, int64_t n, void* stream);

/* Standalone Philox4x32-10 + Box-Muller N(0,1) fill with the same keying as gd_posterior_step
 * (element index = (b*J + j)*T + t). */
int gd_philox_normal(float* out, int32_t B, int32_t J, int32_t T, uint64_t seed, int64_t sample_base,
                     int32_t step, void* stream);

/* Small device helpers used by the host loop (no torch ops on the hot path). */
int gd_set_loop_state(int32_t* step_ptr, int64_t* t, const int64_t* timestep_map, int32_t step,
                      int32_t B, void* stream);

/* Count of kernel launches issued through this library by the calling process (for bench.py's
 * gpu_launches claim). */
int64_t gd_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GD_B200_H */
