"""ctypes binding of include/gd_b200.h (the C-ABI shared library libgd_b200.so).

No CPU fallback: if the library is missing or a call fails this raises.  Tensors are passed as
raw device pointers (``tensor.data_ptr()``) and the current torch CUDA stream handle.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GD_B200_LIB: another build of the same library (A/B timing of two kernel versions on one box)
LIB_PATH = os.environ.get("GD_B200_LIB") or os.path.join(_HERE, "_lib", "libgd_b200.so")

GD_ACT_NONE, GD_ACT_GELU_ERF, GD_ACT_SILU = 0, 1, 2
GD_F32, GD_BF16, GD_FP16 = 0, 1, 2

EXPORTED_SYMBOLS = [
    "gd_abi_version", "gd_last_error", "gd_gemm_h16", "gd_gemm_f32", "gd_gemm_h16_ln", "gd_gemm_h16_ln_split", "gd_ffn_h16_ln_split", "gd_ffn_trace", "gd_attention_h16",
    "gd_attention_f32", "gd_layernorm", "gd_local_attention", "gd_cond_tokens", "gd_posterior_step",
    "gd_pack_rows", "gd_unnormalize_split", "gd_q_sample", "gd_philox_normal", "gd_set_loop_state", "gd_launch_count",
    "gd_noalloc_max_elems", "gd_linear_res_ln_f32_noalloc", "gd_gemm_ln_small_max_rows",
]

_p, _i32, _i64, _u64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float


class GdEpilogue(C.Structure):
    _fields_ = [
        ("bias", _p),
        ("rowtab", _p), ("tab_period", _i32), ("tab_offset", _i32), ("ld_tab", _i32),
        ("seqbias", _p), ("seq_rows", _i32), ("seq_count", _i32), ("ld_seq", _i32),
        ("residual", _p), ("ld_res", _i32),
        ("act", _i32),
        ("map_t", _i32), ("map_s", _i32), ("map_off", _i32),
        ("dup", _i32), ("dup_row_stride", _i32),
        ("out_f32", _p), ("ld_out", _i32),
        ("out_h16", _p), ("ld_out_h16", _i32), ("h16_fmt", _i32),
        ("out_t", _p), ("t_s", _i32), ("t_t", _i32), ("t_skip", _i32),
    ]


class GdPosteriorArgs(C.Structure):
    _fields_ = [
        ("cond", _p), ("uncond", _p), ("scale", _p),
        ("x_t", _p), ("noise", _p),
        ("inpaint_mask", _p), ("inpainted", _p),
        ("coef", _p), ("n_steps", _i32), ("step_ptr", _p),
        ("timestep_map", _p), ("next_t", _p),
        ("x_prev", _p), ("pred_xstart", _p), ("mean", _p),
        ("a_next", _p), ("ld_a", _i32), ("a_fmt", _i32),
        ("B", _i32), ("J", _i32), ("T", _i32),
        ("eps_mode", _i32), ("clip", _i32), ("const_noise", _i32),
        ("seed", _u64), ("sample_base", _i64),
        ("seed_dev", _p), ("ddim", _i32),
    ]


class GdError(RuntimeError):
    pass


_lib = None


def load():
    """Load the shared library (building is done by gesturediffusion_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GdError("libgd_b200.so not found at %s -- run `python -m gesturediffusion_b200.build` "
                      "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.gd_abi_version.restype = _i32
    lib.gd_last_error.restype = C.c_char_p
    lib.gd_launch_count.restype = _i64
    E, PA = C.POINTER(GdEpilogue), C.POINTER(GdPosteriorArgs)
    sigs = {
        "gd_gemm_h16": [_p, _i32, _p, _i32, _i32, _i32, _i32, _i32, E, _p],
        "gd_gemm_f32": [_p, _i32, _p, _i32, _i32, _i32, _i32, E, _p],
        "gd_gemm_h16_ln": [_p, _i32, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _i32, _p, _p, _f32, _p, _i32, _p, _i32,
                           _i32, _p],
        "gd_gemm_h16_ln_split": [_p, _i32, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _i32, _p, _p, _i32, _p, _p, _f32,
                                 _p, _p, _i32, _i32, _p],
        "gd_ffn_h16_ln_split": [_p, _p, _i32, _p, _i32, _p, _p, _i32, _p, _i32, _i32, _i32, _i32, _p, _p, _f32, _p, _p,
                                _i32, _p],
        "gd_ffn_trace": [_p],
        "gd_attention_h16": [_p, _p, _i32, _i32, _i32, _i32, _i32, _p],
        "gd_attention_f32": [_p, _p, _i32, _i32, _i32, _i32, _p],
        "gd_layernorm": [_p, _p, _p, _p, _p, _i32, _i32, _i32, _f32, _p],
        "gd_noalloc_max_elems": [],
        "gd_gemm_ln_small_max_rows": [],
        "gd_linear_res_ln_f32_noalloc": [_p, _i32, _p, _i32, _i32, _i32, _i32, _p, _p, _i32, _p, _p, _f32, _p, _p, _i32,
                                         _p],
        "gd_local_attention": [_p, _p, _p, _p, _p, _i32, _i32, _i32, _i32, _i32, _i32, _p],
        "gd_cond_tokens": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _i32, _i32, _i32, _i32, _p],
        "gd_posterior_step": [PA, _p],
        "gd_pack_rows": [_p, _p, _i32, _i32, _i32, _i32, _i32, _i32, _p],
        "gd_unnormalize_split": [_p, _p, _p, _p, _p, _i32, _i32, _i32, _i32, _i32, _p],
        "gd_q_sample": [_p, _p, _p, _p, _i32, _p, _i64, _p],
        "gd_philox_normal": [_p, _i32, _i32, _i32, _u64, _i64, _i32, _p],
        "gd_set_loop_state": [_p, _p, _p, _i32, _i32, _p],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = _i32
    if lib.gd_abi_version() != 1:
        raise GdError("libgd_b200.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().gd_last_error().decode("utf-8", "replace")
        raise GdError("%s failed (%d): %s" % (what, rc, msg))


def fmt_of(t):
    """gdFmt code of a torch tensor's dtype."""
    import torch
    if t is None:
        return GD_BF16
    code = {torch.float32: GD_F32, torch.bfloat16: GD_BF16, torch.float16: GD_FP16}.get(t.dtype)
    if code is None:
        raise GdError("unsupported dtype %s" % t.dtype)
    return code


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_handle():
    import torch
    return torch.cuda.current_stream().cuda_stream


def launch_count():
    return int(load().gd_launch_count())
