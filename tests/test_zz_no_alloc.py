"""GPU: gd_linear_res_ln_f32_noalloc -- the allocation-free variant of the fp32 sub-layer tail (built-in static device
scratch, include/gd_b200.h) -- through the C ABI against the existing entry points gd_gemm_f32 + gd_layernorm with a
caller-provided scratch, on two calls in a row (the second reuses the scratch the first left behind), plus the size
limit.  Collected last on purpose (zz)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(M, N, K, seed, dev):
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g).to(dev)
    return dict(a=mk(M, K), w=mk(N, K) / K ** 0.5, bias=mk(N), res=mk(M, N), gamma=1 + 0.1 * mk(N), beta=0.1 * mk(N))


def _existing_path(c):
    from gesturediffusion_b200 import kernels as K
    M, N = c["res"].shape
    y = torch.empty(M, N, device=c["a"].device)                   # caller-provided scratch (what the engine's Y32 is)
    K.gemm(c["a"], c["w"], K.epilogue(bias=c["bias"], residual=c["res"], out_f32=y))
    out32 = torch.empty_like(y)
    out16 = torch.empty(M, N, device=y.device, dtype=torch.float16)
    K.layernorm(y, c["gamma"], c["beta"], out_f32=out32, out_h16=out16)
    return out32, out16


def test_noalloc_matches_existing_entry_points_twice_in_a_row(cuda_device):
    from gesturediffusion_b200 import _cabi
    from gesturediffusion_b200 import kernels as K
    lib = _cabi.load()
    assert lib.gd_noalloc_max_elems() == 1 << 20
    # two different problems back to back: the second (smaller rows, other N) finds the first one's data in the scratch
    for M, N, Kd, seed in ((1500, 512, 512, 1), (197, 256, 1024, 2)):
        c = _case(M, N, Kd, seed, cuda_device)
        want32, want16 = _existing_path(c)
        got32 = torch.empty_like(want32)
        got16 = torch.empty_like(want16)
        mem0 = torch.cuda.memory_allocated()
        K.linear_res_ln_f32_noalloc(c["a"], c["w"], c["bias"], c["res"], c["gamma"], c["beta"], out_f32=got32,
                                    out_h16=got16)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == mem0
        # same kernels on the same values: bit-identical (the tolerance of the fp32 tests, 1e-5, holds trivially)
        assert torch.equal(got32, want32) and torch.equal(got16, want16)
        ref = torch.nn.functional.layer_norm(c["res"].double() + c["a"].double() @ c["w"].double().T + c["bias"].double(),
                                             (N,), c["gamma"].double(), c["beta"].double(), 1e-5)
        err = float((got32.double() - ref).norm() / ref.norm())
        assert err < 1e-5, err


def test_noalloc_rejects_sizes_beyond_the_scratch(cuda_device):
    from gesturediffusion_b200 import _cabi
    from gesturediffusion_b200 import kernels as K
    M, N, Kd = 2049, 512, 128                                      # 2049 * 512 > 1 Mi elements
    c = _case(M, N, Kd, 3, cuda_device)
    out = torch.empty(M, N, device=cuda_device)
    with pytest.raises(_cabi.GdError, match="exceeds the built-in scratch"):
        K.linear_res_ln_f32_noalloc(c["a"], c["w"], c["bias"], c["res"], c["gamma"], c["beta"], out_f32=out)
    # the largest accepted problem: exactly the limit
    M = (1 << 20) // N
    c = _case(M, N, Kd, 4, cuda_device)
    want32, _ = _existing_path(c)
    got = torch.empty_like(want32)
    K.linear_res_ln_f32_noalloc(c["a"], c["w"], c["bias"], c["res"], c["gamma"], c["beta"], out_f32=got)
    torch.cuda.synchronize()
    assert torch.equal(got, want32)
