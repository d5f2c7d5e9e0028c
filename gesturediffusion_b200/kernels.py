"""Tensor-level wrappers over the C ABI (one function per entry point of include/gd_b200.h).

These only validate device/dtype/contiguity and unwrap pointers; all arithmetic happens in the
sm_100a kernels.  Every wrapper raises on CPU tensors -- there is no fallback path.
"""
import ctypes as C

import torch

from . import _cabi
from ._cabi import GD_ACT_GELU_ERF, GD_ACT_NONE, GD_ACT_SILU, GdEpilogue, GdPosteriorArgs, fmt_of  # noqa: F401

H16 = (torch.bfloat16, torch.float16)

# Optional per-launch CUDA-event timing (bench.py's roofline leg / profiles): when PROFILE is a list,
# every wrapper records (name, meta, start_event, end_event) on the launching stream.
PROFILE = None


class _Timed:
    def __init__(self, name, **meta):
        self.name, self.meta = name, meta

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if PROFILE is not None:
            self.e1.record()
            PROFILE.append((self.name, self.meta, self.e0, self.e1))
        return False


def profile_summary(records):
    """Aggregate PROFILE records -> {(name, meta-tuple): (count, total_ms)} (synchronises)."""
    torch.cuda.synchronize()
    out = {}
    for name, meta, e0, e1 in records:
        key = (name, tuple(sorted(meta.items())))
        c, t = out.get(key, (0, 0.0))
        out[key] = (c + 1, t + e0.elapsed_time(e1))
    return out


def _dev(t, dtype=None, name="tensor"):
    if t is None:
        return None
    if dtype == "h16":
        if t.dtype not in H16:
            raise _cabi.GdError("%s must be bfloat16 or float16, got %s" % (name, t.dtype))
        dtype = None
    if not t.is_cuda:
        raise _cabi.GdError("%s must be a CUDA tensor (gesturediffusion_b200 has no CPU path)" % name)
    if dtype is not None and t.dtype != dtype:
        raise _cabi.GdError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not (t.is_contiguous() or (t.dim() == 2 and t.stride(1) == 1)):
        raise _cabi.GdError("%s must be contiguous (or a row-strided 2-D view)" % name)
    return t.data_ptr()


def epilogue(bias=None, rowtab=None, tab_period=0, tab_offset=0, seqbias=None, seq_rows=0, seq_count=0,
             residual=None, act=GD_ACT_NONE, rowmap=None, dup=1, dup_row_stride=0, out_f32=None,
             out_h16=None, out_t=None, t_s=0, t_t=0, t_skip=0):
    """Build a GdEpilogue from tensors.  2-D tensors give their row stride as the leading dim."""
    ep = GdEpilogue()
    ep.bias = _dev(bias, torch.float32, "bias")
    ep.rowtab = _dev(rowtab, torch.float32, "rowtab")
    ep.tab_period, ep.tab_offset = tab_period, tab_offset
    ep.ld_tab = rowtab.stride(0) if rowtab is not None else 0
    ep.seqbias = _dev(seqbias, torch.float32, "seqbias")
    ep.seq_rows, ep.seq_count = seq_rows, seq_count
    ep.ld_seq = seqbias.stride(0) if seqbias is not None else 0
    ep.residual = _dev(residual, torch.float32, "residual")
    ep.ld_res = residual.stride(0) if residual is not None else 0
    ep.act = act
    if rowmap is not None:
        ep.map_t, ep.map_s, ep.map_off = rowmap
    ep.dup, ep.dup_row_stride = dup, dup_row_stride
    ep.out_f32 = _dev(out_f32, torch.float32, "out_f32")
    ep.ld_out = out_f32.stride(0) if out_f32 is not None else 0
    ep.out_h16 = _dev(out_h16, "h16", "out_h16")
    ep.ld_out_h16 = out_h16.stride(0) if out_h16 is not None else 0
    ep.h16_fmt = fmt_of(out_h16)
    ep.out_t = _dev(out_t, torch.float32, "out_t")
    ep.t_s, ep.t_t, ep.t_skip = t_s, t_t, t_skip
    return ep


def gemm(a, w, ep, M=None, N=None, K=None):
    """C = A[M,K] @ W[N,K]^T with fused epilogue.  dtype of `a`/`w` picks the kernel:
    bf16 / fp16 (both operands alike) -> tcgen05/TMEM/TMA, fp32 -> SIMT."""
    lib = _cabi.load()
    M = a.shape[0] if M is None else M
    N = w.shape[0] if N is None else N
    K = a.shape[1] if K is None else K
    if a.dtype in H16:
      with _Timed("gemm_h16", M=M, N=N, K=K):
        rc = lib.gd_gemm_h16(_dev(a, a.dtype, "A"), a.stride(0), _dev(w, a.dtype, "W"), w.stride(0),
                             M, N, K, fmt_of(a), C.byref(ep), _cabi.stream_handle())
        _cabi.check(rc, "gd_gemm_h16")
    else:
      with _Timed("gemm_f32", M=M, N=N, K=K):
        rc = lib.gd_gemm_f32(_dev(a, torch.float32, "A"), a.stride(0), _dev(w, torch.float32, "W"), w.stride(0),
                             M, N, K, C.byref(ep), _cabi.stream_handle())
        _cabi.check(rc, "gd_gemm_f32")


def gemm_ln_supported(a, N):
    """gd_gemm_h16_ln covers 16-bit operands and N in {256, 512}."""
    return a.dtype in H16 and N in (256, 512)


def gemm_ln(a, w, bias, residual, gamma, beta, out_f32=None, out_h16=None, eps=1e-5, K=None):
    """LayerNorm(residual + a @ w^T + bias) * gamma + beta in one kernel (tcgen05 GEMM, 2-CTA cluster epilogue)."""
    lib = _cabi.load()
    M, N = a.shape[0], w.shape[0]
    K = a.shape[1] if K is None else K
    with _Timed("gemm_ln_h16", M=M, N=N, K=K):
        rc = lib.gd_gemm_h16_ln(_dev(a, a.dtype, "A"), a.stride(0), _dev(w, a.dtype, "W"), w.stride(0), M, N, K,
                                fmt_of(a), _dev(bias, torch.float32, "bias"),
                                _dev(residual, torch.float32, "residual"), residual.stride(0),
                                _dev(gamma, torch.float32, "gamma"), _dev(beta, torch.float32, "beta"), eps,
                                _dev(out_f32, torch.float32, "out_f32"),
                                out_f32.stride(0) if out_f32 is not None else 0,
                                _dev(out_h16, "h16", "out_h16"), out_h16.stride(0) if out_h16 is not None else 0,
                                fmt_of(out_h16), _cabi.stream_handle())
    _cabi.check(rc, "gd_gemm_h16_ln")


def gemm_ln_split(a, w, bias, gamma, beta, out_hi, out_lo=None, res_f32=None, res_hi=None, res_lo=None, eps=1e-5,
                  K=None):
    """gemm_ln with the residual stream carried as a 16-bit (hi, lo) pair: residual = res_f32 or res_hi + res_lo,
    result written as out_hi = round16(x) and out_lo = round16(x - out_hi) (out_lo may be None)."""
    lib = _cabi.load()
    M, N = a.shape[0], w.shape[0]
    K = a.shape[1] if K is None else K
    r16 = res_hi if res_hi is not None else out_hi
    # rw = residual + output bytes per element through HBM (bench.py's algorithmic-bytes figure)
    with _Timed("gemm_ln_h16", M=M, N=N, K=K, rw=4 + (4 if out_lo is not None else 2)):
        rc = lib.gd_gemm_h16_ln_split(_dev(a, a.dtype, "A"), a.stride(0), _dev(w, a.dtype, "W"), w.stride(0), M, N, K,
                                      fmt_of(a), _dev(bias, torch.float32, "bias"),
                                      _dev(res_f32, torch.float32, "res_f32"),
                                      res_f32.stride(0) if res_f32 is not None else 0,
                                      _dev(res_hi, out_hi.dtype, "res_hi"), _dev(res_lo, out_hi.dtype, "res_lo"),
                                      r16.stride(0),
                                      _dev(gamma, torch.float32, "gamma"), _dev(beta, torch.float32, "beta"), eps,
                                      _dev(out_hi, "h16", "out_hi"), _dev(out_lo, out_hi.dtype, "out_lo"),
                                      out_hi.stride(0), fmt_of(out_hi), _cabi.stream_handle())
    _cabi.check(rc, "gd_gemm_h16_ln_split")


def ffn_ln_supported(x, N, F):
    """gd_ffn_h16_ln_split covers 16-bit operands, N = 512 and F a multiple of 256 (>= 512)."""
    return x.dtype in H16 and N == 512 and F >= 512 and F % 256 == 0


def ffn_ln_split(x_hi, x_lo, w1, b1, w2, b2, gamma, beta, out_hi, out_lo=None, eps=1e-5):
    """LayerNorm((x_hi + x_lo) + gelu(x_hi @ w1^T + b1) @ w2^T + b2) * gamma + beta in one kernel (the hidden
    activation stays on chip); result as out_hi = round16(x'), out_lo = round16(x' - out_hi) (out_lo may be None)."""
    lib = _cabi.load()
    M, N, F = x_hi.shape[0], w2.shape[0], w1.shape[0]
    with _Timed("ffn_ln_h16", M=M, N=N, F=F, rw=4 + (4 if out_lo is not None else 2)):
        rc = lib.gd_ffn_h16_ln_split(_dev(x_hi, x_hi.dtype, "x_hi"), _dev(x_lo, x_hi.dtype, "x_lo"), x_hi.stride(0),
                                     _dev(w1, x_hi.dtype, "W1"), w1.stride(0), _dev(b1, torch.float32, "b1"),
                                     _dev(w2, x_hi.dtype, "W2"), w2.stride(0), _dev(b2, torch.float32, "b2"),
                                     M, N, F, fmt_of(x_hi), _dev(gamma, torch.float32, "gamma"),
                                     _dev(beta, torch.float32, "beta"), eps, _dev(out_hi, x_hi.dtype, "out_hi"),
                                     _dev(out_lo, x_hi.dtype, "out_lo"), out_hi.stride(0), _cabi.stream_handle())
    _cabi.check(rc, "gd_ffn_h16_ln_split")


def attention(qkv, out, n_seq, S, n_heads, head_dim):
    lib = _cabi.load()
    with _Timed("attention", n_seq=n_seq, S=S, H=n_heads, dh=head_dim):
      if qkv.dtype in H16:
        rc = lib.gd_attention_h16(_dev(qkv, qkv.dtype, "qkv"), _dev(out, qkv.dtype, "out"), n_seq, S,
                                  n_heads, head_dim, fmt_of(qkv), _cabi.stream_handle())
        _cabi.check(rc, "gd_attention_h16")
      else:
        rc = lib.gd_attention_f32(_dev(qkv, torch.float32, "qkv"), _dev(out, torch.float32, "out"), n_seq, S,
                                  n_heads, head_dim, _cabi.stream_handle())
        _cabi.check(rc, "gd_attention_f32")


def layernorm(x, gamma, beta, out_f32=None, out_h16=None, eps=1e-5, rows=None):
    lib = _cabi.load()
    rows = x.shape[0] if rows is None else rows
    with _Timed("layernorm", rows=rows, d=x.shape[-1]):
      rc = lib.gd_layernorm(_dev(x, torch.float32, "in"), _dev(gamma, torch.float32, "gamma"),
                          _dev(beta, torch.float32, "beta"), _dev(out_f32, torch.float32, "out_f32"),
                          _dev(out_h16, "h16", "out_h16"), fmt_of(out_h16), rows, x.shape[-1], eps,
                          _cabi.stream_handle())
    _cabi.check(rc, "gd_layernorm")


def linear_res_ln_f32_noalloc(a, w, bias, residual, gamma, beta, out_f32=None, out_h16=None, eps=1e-5):
    """LayerNorm(residual + a @ w^T + bias) * gamma + beta on the fp32 path with the [M, N] intermediate in the
    library's built-in static scratch (gd_linear_res_ln_f32_noalloc: no workspace argument, no allocation; M * N is
    limited to gd_noalloc_max_elems(); calls on one device share the scratch and must not overlap on the GPU)."""
    lib = _cabi.load()
    M, N, K = a.shape[0], w.shape[0], a.shape[1]
    rc = lib.gd_linear_res_ln_f32_noalloc(_dev(a, torch.float32, "A"), a.stride(0), _dev(w, torch.float32, "W"),
                                          w.stride(0), M, N, K, _dev(bias, torch.float32, "bias"),
                                          _dev(residual, torch.float32, "residual"), residual.stride(0),
                                          _dev(gamma, torch.float32, "gamma"), _dev(beta, torch.float32, "beta"), eps,
                                          _dev(out_f32, torch.float32, "out_f32"), _dev(out_h16, "h16", "out_h16"),
                                          fmt_of(out_h16), _cabi.stream_handle())
    _cabi.check(rc, "gd_linear_res_ln_f32_noalloc")


def local_attention(z, rope_cos, rope_sin, x_f32, x_h16, n_seq, T, d, heads, window=10):
    lib = _cabi.load()
    with _Timed("local_attention", n_seq=n_seq, T=T, d=d):
      rc = lib.gd_local_attention(_dev(z, torch.float32, "z"), _dev(rope_cos, torch.float32, "rope_cos"),
                                _dev(rope_sin, torch.float32, "rope_sin"), _dev(x_f32, torch.float32, "x_f32"),
                                _dev(x_h16, "h16", "x_h16"), fmt_of(x_h16), n_seq, T, d, heads, window,
                                _cabi.stream_handle())
    _cabi.check(rc, "gd_local_attention")


def cond_tokens(t, temb_table, seed_emb, pe0, x_f32, x_h16, B, n_branch, S, d, coarse_tab=None,
                coarse_seed=None, coarse_out=None):
    lib = _cabi.load()
    rc = lib.gd_cond_tokens(_dev(t, torch.int64, "t"), _dev(temb_table, torch.float32, "temb_table"),
                            _dev(seed_emb, torch.float32, "seed_emb"), _dev(pe0, torch.float32, "pe0"),
                            _dev(coarse_tab, torch.float32, "coarse_tab"),
                            _dev(coarse_seed, torch.float32, "coarse_seed"),
                            _dev(coarse_out, torch.float32, "coarse_out"), _dev(x_f32, torch.float32, "x_f32"),
                            _dev(x_h16, "h16", "x_h16"), fmt_of(x_h16), B, n_branch, S, d, _cabi.stream_handle())
    _cabi.check(rc, "gd_cond_tokens")


def on_device_of(t):
    """Context: make `t`'s GPU the current CUDA device for the launches inside (the C ABI uses torch's current stream)."""
    from .sampler import DeviceScope
    return DeviceScope(t.device)


def posterior_step(cond, x_t, coef, step_ptr, B, J, T, uncond=None, scale=None, noise=None, x_prev=None,
                   pred_xstart=None, mean=None, a_next=None, inpaint_mask=None, inpainted=None,
                   timestep_map=None, next_t=None, eps_mode=False, clip=False, const_noise=False, seed=0,
                   sample_base=0, seed_dev=None, ddim=False):
    lib = _cabi.load()
    a = GdPosteriorArgs()
    a.cond, a.uncond, a.scale = _dev(cond, torch.float32, "cond"), _dev(uncond, torch.float32, "uncond"), \
        _dev(scale, torch.float32, "scale")
    a.x_t, a.noise = _dev(x_t, torch.float32, "x_t"), _dev(noise, torch.float32, "noise")
    a.inpaint_mask, a.inpainted = _dev(inpaint_mask, torch.uint8, "inpaint_mask"), \
        _dev(inpainted, torch.float32, "inpainted")
    a.coef, a.n_steps, a.step_ptr = _dev(coef, torch.float32, "coef"), coef.shape[1], \
        _dev(step_ptr, torch.int32, "step_ptr")
    a.timestep_map, a.next_t = _dev(timestep_map, torch.int64, "timestep_map"), _dev(next_t, torch.int64, "next_t")
    a.x_prev, a.pred_xstart, a.mean = _dev(x_prev, torch.float32, "x_prev"), \
        _dev(pred_xstart, torch.float32, "pred_xstart"), _dev(mean, torch.float32, "mean")
    if a_next is not None:
        a.a_next = _dev(a_next, None, "a_next")
        a.ld_a = a_next.stride(0)
        a.a_fmt = fmt_of(a_next)
    a.B, a.J, a.T = B, J, T
    a.eps_mode, a.clip, a.const_noise = int(eps_mode), int(clip), int(const_noise)
    a.seed, a.sample_base = seed, sample_base
    a.seed_dev = _dev(seed_dev, torch.int64, "seed_dev")
    a.ddim = int(ddim)
    with _Timed("posterior_step", B=B, J=J, T=T, cfg=uncond is not None, philox=noise is None), on_device_of(x_t):
      rc = lib.gd_posterior_step(C.byref(a), _cabi.stream_handle())
    _cabi.check(rc, "gd_posterior_step")


def pack_rows(src, dst, B, Cc, T, col_off=0):
    lib = _cabi.load()
    rc = lib.gd_pack_rows(_dev(src, torch.float32, "src"), _dev(dst, None, "dst"), B, Cc, T, dst.stride(0), col_off,
                          fmt_of(dst), _cabi.stream_handle())
    _cabi.check(rc, "gd_pack_rows")


def unnormalize_split(x, mean, std, pos, rot, t_off):
    """x [B, J, T] (or [B, J, 1, T]) fp32 -> pos / rot [B, J/6, 3, T_total] at frame offset t_off (generate.py:132-146)."""
    lib = _cabi.load()
    B, J, T = x.shape[0], x.shape[1], x.shape[-1]
    ref = pos if pos is not None else rot
    with on_device_of(x):
      rc = lib.gd_unnormalize_split(_dev(x, torch.float32, "x"), _dev(mean, torch.float32, "mean") if mean is not None else None,
                                  _dev(std, torch.float32, "std") if std is not None else None,
                                  _dev(pos, torch.float32, "pos") if pos is not None else None,
                                  _dev(rot, torch.float32, "rot") if rot is not None else None,
                                  B, J, T, t_off, ref.shape[-1], _cabi.stream_handle())
    _cabi.check(rc, "gd_unnormalize_split")


def q_sample(x0, noise, sqrt_acp, sqrt_1macp, step, out):
    lib = _cabi.load()
    with on_device_of(x0):
        rc = lib.gd_q_sample(_dev(x0, torch.float32, "x0"), _dev(noise, torch.float32, "noise"),
                             _dev(sqrt_acp, torch.float32, "sqrt_acp"), _dev(sqrt_1macp, torch.float32, "sqrt_1macp"),
                             step, _dev(out, torch.float32, "out"), x0.numel(), _cabi.stream_handle())
    _cabi.check(rc, "gd_q_sample")


def philox_normal(out, B, J, T, seed, sample_base, step):
    lib = _cabi.load()
    rc = lib.gd_philox_normal(_dev(out, torch.float32, "out"), B, J, T, seed, sample_base, step,
                              _cabi.stream_handle())
    _cabi.check(rc, "gd_philox_normal")


def set_loop_state(step_ptr, t, timestep_map, step, B):
    lib = _cabi.load()
    rc = lib.gd_set_loop_state(_dev(step_ptr, torch.int32, "step_ptr"), _dev(t, torch.int64, "t"),
                               _dev(timestep_map, torch.int64, "timestep_map"), step, B, _cabi.stream_handle())
    _cabi.check(rc, "gd_set_loop_state")
