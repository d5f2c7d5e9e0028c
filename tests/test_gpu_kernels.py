"""GPU: every CUDA kernel, called through the C ABI, against a plain torch restatement of the same op
on the same seeded inputs (fp64 on the GPU for the reference value).  Tolerances: fp32 kernels 1e-5
relative L2 (north star 'fp32 mode'), 16-bit (bf16 / fp16) tensor-core kernels against the reference evaluated on
the SAME bf16-rounded operands with wide accumulation (operand rounding itself is a model-level effect
measured in test_gpu_model.py)."""
import math

import pytest
import torch

from helpers import record_parity, rel_l2

pytestmark = pytest.mark.gpu


def _K():
    from gesturediffusion_b200 import kernels
    return kernels


TDT = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}
H16_ROUND = {"bf16": 3e-3, "fp16": 4e-4}        # relative L2 of one 16-bit output rounding (worst case)


def _gelu(x):
    return 0.5 * x * (1 + torch.erf(x / math.sqrt(2)))


# ------------------------------------------------------------------ GEMMs
GEMM_SHAPES = [(128, 128, 64), (256, 512, 512), (1000, 1536, 512), (197, 263, 512), (392, 512, 289),
               (3000, 1024, 512), (40000, 512, 1024), (60, 24, 24), (130, 384, 128)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("dtype", ["bf16", "fp16", "fp32"])
def test_gemm_bias(cuda_device, M, N, K, dtype):
    Kn = _K()
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    ld = (K + 7) // 8 * 8
    a = torch.zeros(M, ld, device="cuda")
    w = torch.zeros(N, ld, device="cuda")
    a[:, :K] = torch.randn(M, K, device="cuda", generator=g)
    w[:, :K] = torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)
    bias = torch.randn(N, device="cuda", generator=g)
    a, w = a.to(TDT[dtype]), w.to(TDT[dtype])
    ref = a[:, :K].double() @ w[:, :K].double().T + bias.double()
    out = torch.full((M, N), float("nan"), device="cuda")
    Kn.gemm(a, w, Kn.epilogue(bias=bias, out_f32=out), K=K)
    torch.cuda.synchronize()
    assert rel_l2(out, ref) < (2e-6 if dtype == "fp32" else 1e-5)     # fp32 accumulation in both


@pytest.mark.parametrize("M,N,K", [(128, 512, 4980), (1, 512, 2630), (64, 512, 2630), (70, 130, 1037), (256, 512, 1024)])
def test_gemm_f32_split_k(cuda_device, M, N, K):
    """Few tiles, long K (the seed-pose embedding, mdm.py:186-187): a cluster of CTAs shares a tile, partial tiles are added
    in rank order through distributed shared memory.  fp32 against float64, bit-repeatable."""
    Kn = _K()
    g = torch.Generator(device="cuda").manual_seed(M + 3 * K)
    a = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    ref = a.double() @ w.double().T + bias.double() + res.double()
    outs = []
    for _ in range(3):
        out = torch.full((M, N), float("nan"), device="cuda")
        Kn.gemm(a, w, Kn.epilogue(bias=bias, residual=res, out_f32=out), K=K)
        outs.append(out)
    torch.cuda.synchronize()
    err = rel_l2(outs[0], ref)
    record_parity("kernel gemm_f32 split-K M=%d N=%d K=%d: rel_l2 vs float64 %.3e" % (M, N, K, err))
    assert err < 2e-6
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("M,N,K", [(700, 1024, 512), (129, 512, 64), (25216, 1024, 512)])
@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_gemm_specialised_epilogues(cuda_device, M, N, K, dtype):
    """The three compile-time epilogues of the tcgen05 GEMM (bias->16 bit, bias+erf-GELU->16 bit in the tanh-of-fitted-
    polynomial form of common.cuh::gelu_fast, bias+residual->fp32) against float64.  The GELU path is held to its stated
    accuracy: per element one 16-bit ulp of the result + the fit's 2.5e-5 + the MUFU.TANH bound 0.5 |x| 2^-11, and in
    relative L2 about one extra output rounding."""
    Kn = _K()
    tdt = TDT[dtype]
    g = torch.Generator(device="cuda").manual_seed(M + K)
    a = (torch.randn(M, K, device="cuda", generator=g) * 1.5).to(tdt)
    w = (torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)).to(tdt)
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    base = a.double() @ w.double().T + bias.double()
    oh = torch.empty(M, N, device="cuda", dtype=tdt)
    Kn.gemm(a, w, Kn.epilogue(bias=bias, out_h16=oh))
    assert rel_l2(oh.float(), base) < H16_ROUND[dtype]
    Kn.gemm(a, w, Kn.epilogue(bias=bias, act=Kn.GD_ACT_GELU_ERF, out_h16=oh))
    ref = _gelu(base)
    e_gelu = rel_l2(oh.float(), ref)
    record_parity("gelu epilogue %s M=%d N=%d K=%d: rel-L2 vs float64 erf-GELU %.2e" % (dtype, M, N, K, e_gelu))
    assert e_gelu < 1.5 * H16_ROUND[dtype]
    exact = ref.to(torch.float32).to(tdt)
    ulp = exact.float().abs() * (2.0 ** -10 if dtype == "fp16" else 2.0 ** -7)
    diff = (oh.float() - exact.float()).abs()
    bound = 1.01 * ulp + 3e-5 + 0.5 * base.abs().float() * 2.0 ** -11
    assert bool((diff <= bound).all()), float((diff - bound).max())
    o32 = torch.empty(M, N, device="cuda")
    Kn.gemm(a, w, Kn.epilogue(bias=bias, residual=res, out_f32=o32))
    assert rel_l2(o32, base + res.double()) < 1e-5


@pytest.mark.parametrize("dtype", ["bf16", "fp16", "fp32"])
def test_gemm_epilogue_residual_gelu_h16out(cuda_device, dtype):
    Kn = _K()
    M, N, K = 700, 1024, 512
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    a, w = a.to(TDT[dtype]), w.to(TDT[dtype])
    base = a.double() @ w.double().T + bias.double()
    out32 = torch.empty(M, N, device="cuda")
    hdt = torch.float16 if dtype == "fp16" else torch.bfloat16
    outb = torch.empty(M, N, device="cuda", dtype=hdt)
    Kn.gemm(a, w, Kn.epilogue(bias=bias, residual=res, out_f32=out32))
    assert rel_l2(out32, base + res.double()) < 1e-5
    Kn.gemm(a, w, Kn.epilogue(bias=bias, act=Kn.GD_ACT_GELU_ERF, out_f32=out32, out_h16=outb))
    assert rel_l2(out32, _gelu(base)) < 1e-5
    assert rel_l2(outb.float(), _gelu(base)) < H16_ROUND["fp16" if dtype == "fp16" else "bf16"]   # output rounding only
    Kn.gemm(a, w, Kn.epilogue(bias=bias, act=Kn.GD_ACT_SILU, out_f32=out32))
    assert rel_l2(out32, base * torch.sigmoid(base)) < 1e-5


@pytest.mark.parametrize("dtype", ["bf16", "fp16", "fp32"])
def test_gemm_epilogue_rowmap_dup_tables_transposed(cuda_device, dtype):
    """The input-projection epilogue (row remap to rows 1..T of each sequence, duplicated per CFG branch,
    + positional table + per-sequence bias) and the output-projection epilogue (transposed [.,J,T] store
    skipping token 0)."""
    Kn = _K()
    B, T, d, J, nb = 3, 20, 128, 24, 2
    S = T + 1
    g = torch.Generator(device="cuda").manual_seed(11)
    a = torch.randn(B * T, 32, device="cuda", generator=g)
    w = torch.randn(d, 32, device="cuda", generator=g) / 5
    bias = torch.randn(d, device="cuda", generator=g)
    tab = torch.randn(50, d, device="cuda", generator=g)
    seqb = torch.randn(nb * B, d, device="cuda", generator=g)
    a, w = a.to(TDT[dtype]), w.to(TDT[dtype])
    X = torch.zeros(nb * B * S, d, device="cuda")
    Kn.gemm(a, w, Kn.epilogue(bias=bias, rowtab=tab, tab_period=T, tab_offset=1, seqbias=seqb, seq_rows=T,
                              seq_count=B, rowmap=(T, S, 1), dup=nb, dup_row_stride=B * S, out_f32=X))
    base = (a.double() @ w.double().T + bias.double()).view(B, T, d) + tab[1:T + 1].double()
    ref = torch.zeros(nb, B, S, d, dtype=torch.float64, device="cuda")
    for c in range(nb):
        ref[c, :, 1:] = base + seqb[c * B:(c + 1) * B].double().unsqueeze(1)
    assert rel_l2(X, ref.view(-1, d)) < 1e-5
    assert float(X.view(nb, B, S, d)[:, :, 0].abs().max()) == 0.0      # token-0 rows untouched

    xin = torch.randn(nb * B * S, d, device="cuda", generator=g)
    wo = torch.randn(J, d, device="cuda", generator=g) / 10
    bo = torch.randn(J, device="cuda", generator=g)
    xin, wo = xin.to(TDT[dtype]), wo.to(TDT[dtype])
    out = torch.full((nb * B, J, T), float("nan"), device="cuda")
    Kn.gemm(xin, wo, Kn.epilogue(bias=bo, out_t=out, t_s=S, t_t=T, t_skip=1))
    ref = (xin.double() @ wo.double().T + bo.double()).view(nb * B, S, J)[:, 1:].transpose(1, 2)
    assert rel_l2(out, ref) < 1e-5


# ------------------------------------------------------------------ attention / LN
@pytest.mark.parametrize("n_seq,S,H,dh", [(3, 21, 4, 32), (2, 121, 4, 128), (2, 197, 4, 128), (1, 64, 2, 64),
                                         (2, 301, 4, 128),
                                         # tcgen05 kernel (dh 128, S <= 256): edge lengths, one and two query tiles,
                                         # and enough (sequence, head) items for several units per CTA
                                         (1, 1, 1, 128), (3, 16, 2, 128), (2, 128, 4, 128), (2, 129, 4, 128),
                                         (1, 256, 4, 128), (5, 61, 3, 128), (160, 121, 4, 128), (100, 197, 4, 128),
                                         # two query tiles: split last round (512 items on 148 CTAs), all-solo small batch
                                         # (68 items -> 136 CTAs), exact rounds (148 items), odd item count
                                         (128, 197, 4, 128), (17, 197, 4, 128), (37, 200, 4, 128), (51, 197, 3, 128)])
@pytest.mark.parametrize("dtype", ["bf16", "fp16", "fp32"])
def test_attention(cuda_device, n_seq, S, H, dh, dtype):
    Kn = _K()
    d = H * dh
    g = torch.Generator(device="cuda").manual_seed(S)
    qkv = torch.randn(n_seq * S, 3 * d, device="cuda", generator=g)
    tdt = TDT[dtype]
    qkv = qkv.to(tdt)
    out = torch.zeros(n_seq * S, d, device="cuda", dtype=tdt)
    Kn.attention(qkv, out, n_seq, S, H, dh)
    q, k, v = qkv.double().view(n_seq, S, 3, H, dh).permute(2, 0, 3, 1, 4)
    ref = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(dh), dim=-1) @ v
    ref = ref.permute(0, 2, 1, 3).reshape(n_seq * S, d)
    # 16-bit: P and O are rounded to the 16-bit format
    assert rel_l2(out.float(), ref) < {"fp32": 2e-6, "bf16": 6e-3, "fp16": 8e-4}[dtype]


# S > 256 at head_dim 128: the tcgen05 kernel with the keys in blocks of 128 and an online softmax
# (attention_tc_long.cu).  257 = one key in the third block; 301 / 401 / 601 = the 300 / 400 / 600-frame sweep rows
# (odd / even / odd number of query tiles: idle second slot in the last pair); 1000 = 8 blocks; 160 sequences x 4 heads
# x 3 pairs = several items per CTA (ring and phase bookkeeping across items); large scores exercise the rescale.
@pytest.mark.parametrize("n_seq,S,H,scale", [(3, 257, 4, 1.0), (5, 301, 4, 1.0), (2, 401, 4, 1.0), (3, 601, 4, 1.0),
                                             (2, 1000, 2, 1.0), (160, 601, 4, 1.0), (7, 384, 1, 1.0), (4, 601, 4, 4.0)])
@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_attention_long_sequences(cuda_device, n_seq, S, H, scale, dtype):
    Kn = _K()
    dh = 128
    d = H * dh
    g = torch.Generator(device="cuda").manual_seed(S + n_seq)
    qkv = torch.randn(n_seq * S, 3 * d, device="cuda", generator=g)
    qkv[:, :2 * d] *= scale                                  # larger |q.k|: sharper rows, the running maximum moves
    tdt = TDT[dtype]
    qkv = qkv.to(tdt)
    out = torch.full((n_seq * S, d), 7.0, device="cuda", dtype=tdt)
    Kn.attention(qkv, out, n_seq, S, H, dh)
    out2 = torch.full_like(out, -3.0)
    Kn.attention(qkv, out2, n_seq, S, H, dh)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)                            # every element written, repeatable
    errs = []
    for s0 in range(0, n_seq, 16):                           # reference in slices (float64 scores of 160 x 601^2 x 4)
        s1 = min(n_seq, s0 + 16)
        blk = qkv[s0 * S:s1 * S].double().view(s1 - s0, S, 3, H, dh).permute(2, 0, 3, 1, 4)
        ref = torch.softmax(blk[0] @ blk[1].transpose(-1, -2) / math.sqrt(dh), dim=-1) @ blk[2]
        ref = ref.permute(0, 2, 1, 3).reshape((s1 - s0) * S, d)
        errs.append(rel_l2(out[s0 * S:s1 * S].float(), ref))
    record_parity("attention long n_seq=%d S=%d H=%d x%.0f %s: rel-L2 %.2e" % (n_seq, S, H, scale, dtype, max(errs)))
    assert max(errs) < {"bf16": 6e-3, "fp16": 8e-4}[dtype]


@pytest.mark.parametrize("rows,d", [(1, 128), (1000, 512), (777, 1024)])
def test_layernorm(cuda_device, rows, d):
    Kn = _K()
    g = torch.Generator(device="cuda").manual_seed(rows)
    x = torch.randn(rows, d, device="cuda", generator=g) * 3 + 1
    gamma = torch.randn(d, device="cuda", generator=g)
    beta = torch.randn(d, device="cuda", generator=g)
    o32 = torch.empty_like(x)
    ref = torch.nn.functional.layer_norm(x.double(), (d,), gamma.double(), beta.double(), 1e-5)
    for hdt in (torch.bfloat16, torch.float16):
        ob = torch.empty(rows, d, device="cuda", dtype=hdt)
        Kn.layernorm(x, gamma, beta, out_f32=o32, out_h16=ob)
        assert rel_l2(o32, ref) < 1e-6
        assert torch.equal(ob, o32.to(hdt))


@pytest.mark.parametrize("M,N,K", [(1, 512, 512), (127, 512, 512), (129, 256, 64), (1000, 512, 1024),
                                   (128 * 150 + 77, 512, 512), (4100, 256, 512),
                                   # few rows, N = 512: the 8-CTA-cluster kernel (gemm_ln_small.cu), incl. a K tail,
                                   # K > 8 stages, its last row count and the first one past it
                                   (394, 512, 512), (394, 512, 1024), (300, 512, 296), (2304, 512, 512), (2305, 512, 512),
                                   # balanced row tiles (110 / 105 / 122 rows instead of 128: fourth-quarter tail boxes)
                                   (24325, 512, 512), (30976, 512, 1024), (17923, 512, 512), (17923, 256, 512)])
@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_gemm_ln_fused(cuda_device, M, N, K, dtype):
    """gd_gemm_h16_ln == LayerNorm(residual + A W^T + bias) (float64 reference on the rounded operands), in place on
    the residual stream; covers single / partial tiles, several tiles per cluster and both cluster widths."""
    Kn = _K()
    tdt = TDT[dtype]
    g = torch.Generator(device="cuda").manual_seed(M + N)
    a = torch.randn(M, K, device="cuda", generator=g).to(tdt)
    w = (torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)).to(tdt)
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g) * 2 + 0.5
    gamma = 1 + 0.1 * torch.randn(N, device="cuda", generator=g)
    beta = 0.1 * torch.randn(N, device="cuda", generator=g)
    y = res.double() + a.double() @ w.double().t() + bias.double()
    ref = torch.nn.functional.layer_norm(y, (N,), gamma.double(), beta.double(), 1e-5)
    x32 = res.clone()
    xh = torch.zeros(M, N, device="cuda", dtype=tdt)
    Kn.gemm_ln(a, w, bias, x32, gamma, beta, out_f32=x32, out_h16=xh)
    assert rel_l2(x32, ref) < 5e-6
    assert torch.equal(xh, x32.to(tdt))
    # separate outputs, residual untouched
    o32 = torch.empty_like(res)
    Kn.gemm_ln(a, w, bias, res, gamma, beta, out_f32=o32)
    assert torch.equal(o32, x32)


def test_gemm_ln_repeatable_under_load(cuda_device):
    """The cluster exchange (st.async + transaction bytes), the residual ring and the aliased staging blocks carry
    no race: 24 back-to-back launches at the bench shape (3 tiles per cluster, ragged last tile absent / present),
    for both operand-stage variants, give the same bits every time, only-16-bit and only-fp32 outputs included."""
    Kn = _K()
    for M, Kd in ((25216, 512), (25216 - 37, 1024)):
        g = torch.Generator(device="cuda").manual_seed(M)
        a = torch.randn(M, Kd, device="cuda", generator=g).half()
        w = (torch.randn(512, Kd, device="cuda", generator=g) / math.sqrt(Kd)).half()
        bias, gamma, beta = (torch.randn(512, device="cuda", generator=g) for _ in range(3))
        res = torch.randn(M, 512, device="cuda", generator=g)
        first32 = first16 = None
        for it in range(24):
            o32 = torch.full((M, 512), float("nan"), device="cuda")
            o16 = torch.zeros(M, 512, device="cuda", dtype=torch.float16)
            if it % 3 == 0:
                Kn.gemm_ln(a, w, bias, res, gamma, beta, out_f32=o32, out_h16=o16)
            elif it % 3 == 1:
                Kn.gemm_ln(a, w, bias, res, gamma, beta, out_f32=o32)
                o16 = o32.half()
            else:
                Kn.gemm_ln(a, w, bias, res, gamma, beta, out_h16=o16)
                o32 = None
            if first32 is None:
                first32, first16 = o32, o16
                y = res.double() + a.double() @ w.double().t() + bias.double()
                ref = torch.nn.functional.layer_norm(y, (512,), gamma.double(), beta.double(), 1e-5)
                assert rel_l2(o32, ref) < 5e-6
            else:
                assert o32 is None or torch.equal(o32, first32)
                assert torch.equal(o16, first16)


@pytest.mark.parametrize("M,N,K", [(1, 512, 512), (127, 512, 512), (129, 256, 64), (1000, 512, 1024),
                                   (128 * 150 + 77, 512, 512), (25216, 512, 1024),
                                   (24325, 512, 512), (30976, 512, 512), (17923, 256, 512)])   # balanced row tiles
@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_gemm_ln_split_residual(cuda_device, M, N, K, dtype):
    """gd_gemm_h16_ln_split: the residual stream as a 16-bit (hi, lo) pair.  fp32 residual in -> pair out equals the
    fp32 kernel's result split exactly (hi = round16(x), lo = round16(x - hi)); pair in (in place) -> pair out
    equals the float64 LayerNorm of the reconstructed residual; hi-only output for the last sub-layer."""
    Kn = _K()
    tdt = TDT[dtype]
    g = torch.Generator(device="cuda").manual_seed(M + N + 1)
    a = torch.randn(M, K, device="cuda", generator=g).to(tdt)
    w = (torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)).to(tdt)
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g) * 2 + 0.5
    gamma = 1 + 0.1 * torch.randn(N, device="cuda", generator=g)
    beta = 0.1 * torch.randn(N, device="cuda", generator=g)
    x32 = torch.empty_like(res)
    Kn.gemm_ln(a, w, bias, res, gamma, beta, out_f32=x32)
    hi = torch.zeros(M, N, device="cuda", dtype=tdt)
    lo = torch.full((M, N), 7.0, device="cuda", dtype=tdt)
    Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi, out_lo=lo, res_f32=res)
    if N == 512 and M <= 2304:
        # few rows: gd_gemm_h16_ln runs its 8-CTA-cluster kernel (other order of the fp32 sums), the split form the 2-CTA
        # one -- equal up to an fp32 rounding step, i.e. a different 16-bit value in rare elements
        assert (hi != x32.to(tdt)).float().mean() < 2e-3
    else:
        assert torch.equal(hi, x32.to(tdt))
        assert torch.equal(lo, (x32 - hi.float()).to(tdt))
    recon = hi.float() + lo.float()
    assert rel_l2(recon, x32) < (3e-7 if dtype == "fp16" else 2e-5)
    # pair in, in place, pair out: the next sub-layer
    y = recon.double() + a.double() @ w.double().t() + bias.double()
    ref = torch.nn.functional.layer_norm(y, (N,), gamma.double(), beta.double(), 1e-5)
    hi2, lo2 = hi.clone(), lo.clone()
    Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi2, out_lo=lo2, res_hi=hi2, res_lo=lo2)
    assert rel_l2(hi2.float() + lo2.float(), ref) < (5e-6 if dtype == "fp16" else 3e-5)
    x32b = torch.empty_like(res)
    Kn.gemm_ln(a, w, bias, recon, gamma, beta, out_f32=x32b)      # fp32 kernel on the same reconstructed residual
    # same values up to the order of the fp32 sum (at N = 512 the pair-in kernel adds the residual on the tensor pipe):
    # a 16-bit result may differ from the fp32 kernel's by one rounding step in rare elements
    assert (hi2 != x32b.to(tdt)).float().mean() < 2e-3
    assert rel_l2(hi2.float(), x32b.to(tdt).float()) < (3e-5 if dtype == "fp16" else 3e-4)
    # hi only (last sub-layer), separate output, inputs untouched
    hi3 = torch.zeros_like(hi)
    Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi3, res_hi=hi, res_lo=lo)
    assert torch.equal(hi3, hi2)
    assert torch.equal(hi.float() + lo.float(), recon)


def test_gemm_ln_split_repeatable_and_argument_checks(cuda_device):
    from gesturediffusion_b200 import _cabi
    Kn = _K()
    M, Kd = 25216 - 37, 1024
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randn(M, Kd, device="cuda", generator=g).half()
    w = (torch.randn(512, Kd, device="cuda", generator=g) / math.sqrt(Kd)).half()
    bias, gamma, beta = (torch.randn(512, device="cuda", generator=g) for _ in range(3))
    res = torch.randn(M, 512, device="cuda", generator=g)
    rh = res.half()
    rl = (res - rh.float()).half()
    first = None
    for it in range(16):
        hi = torch.zeros(M, 512, device="cuda", dtype=torch.float16)
        lo = torch.zeros_like(hi)
        Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi, out_lo=lo, res_hi=rh, res_lo=rl)
        if first is None:
            first = (hi, lo)
        else:
            assert torch.equal(hi, first[0]) and torch.equal(lo, first[1])
    with pytest.raises(_cabi.GdError):          # both residual forms at once
        Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi, out_lo=lo, res_f32=res, res_hi=rh, res_lo=rl)
    with pytest.raises(_cabi.GdError):          # half a pair
        Kn.gemm_ln_split(a, w, bias, gamma, beta, out_hi=hi, out_lo=lo, res_hi=rh)


@pytest.mark.parametrize("M,F", [(1, 512), (127, 1024), (300, 1024), (128 * 150 + 77, 1024), (25216, 1024),
                                 (128 * 75, 2048)])
@pytest.mark.parametrize("dtype", ["bf16", "fp16"])
def test_ffn_ln_split(cuda_device, M, F, dtype):
    """gd_ffn_h16_ln_split (linear1 + GELU + linear2 + residual + LayerNorm in one kernel, H kept on chip) against
    the float64 sub-layer on the same 16-bit inputs and against the two-kernel path it replaces."""
    Kn = _K()
    tdt = TDT[dtype]
    N = 512
    g = torch.Generator(device="cuda").manual_seed(M + F)
    x = torch.randn(M, N, device="cuda", generator=g)
    hi = x.to(tdt)
    lo = (x - hi.float()).to(tdt)
    w1 = (torch.randn(F, N, device="cuda", generator=g) / math.sqrt(N)).to(tdt)
    w2 = (torch.randn(N, F, device="cuda", generator=g) / math.sqrt(F)).to(tdt)
    b1 = 0.5 * torch.randn(F, device="cuda", generator=g)
    b2 = torch.randn(N, device="cuda", generator=g)
    gamma = 1 + 0.1 * torch.randn(N, device="cuda", generator=g)
    beta = 0.1 * torch.randn(N, device="cuda", generator=g)
    # the path it replaces
    h = torch.empty(M, F, device="cuda", dtype=tdt)
    Kn.gemm(hi, w1, Kn.epilogue(bias=b1, act=Kn.GD_ACT_GELU_ERF, out_h16=h))
    hi_ref, lo_ref = torch.zeros_like(hi), torch.zeros_like(lo)
    Kn.gemm_ln_split(h, w2, b2, gamma, beta, out_hi=hi_ref, out_lo=lo_ref, res_hi=hi, res_lo=lo)
    # float64 sub-layer with the hidden activation rounded to 16 bit as both paths do
    xr = hi.double() + lo.double()
    hd = torch.nn.functional.gelu(hi.double() @ w1.double().t() + b1.double()).to(tdt).double()
    ref = torch.nn.functional.layer_norm(xr + hd @ w2.double().t() + b2.double(), (N,), gamma.double(),
                                         beta.double(), 1e-5)
    out_hi = torch.full_like(hi, 3.0)
    out_lo = torch.full_like(lo, 3.0)
    Kn.ffn_ln_split(hi, lo, w1, b1, w2, b2, gamma, beta, out_hi=out_hi, out_lo=out_lo)
    got = out_hi.float() + out_lo.float()
    # vs float64: the (hi, lo) output rounding + the GELU form's own error (gelu_fast: fit 2.5e-5 + MUFU.TANH 2^-11, about
    # 2e-4 relative on H, 5e-5 here); vs the two-kernel path (same GELU, same roundings): accumulation order only
    tol = 1e-4 if dtype == "fp16" else 3e-4
    tol_same = 3e-5 if dtype == "fp16" else 3e-4
    e64, e2k = rel_l2(got, ref), rel_l2(got, hi_ref.float() + lo_ref.float())
    record_parity("ffn_ln_split %s M=%d F=%d: vs float64 %.2e, vs two-kernel path %.2e" % (dtype, M, F, e64, e2k))
    assert e64 < tol
    assert rel_l2(hi_ref.float() + lo_ref.float(), ref) < tol
    assert e2k < tol_same
    # in place, hi only (last sub-layer)
    hi2, lo2 = hi.clone(), lo.clone()
    Kn.ffn_ln_split(hi2, lo2, w1, b1, w2, b2, gamma, beta, out_hi=hi2)
    assert torch.equal(hi2, out_hi)
    assert torch.equal(lo2, lo)
    # in place, pair out; repeatable
    hi3, lo3 = hi.clone(), lo.clone()
    Kn.ffn_ln_split(hi3, lo3, w1, b1, w2, b2, gamma, beta, out_hi=hi3, out_lo=lo3)
    assert torch.equal(hi3, out_hi) and torch.equal(lo3, out_lo)


def test_ffn_ln_split_repeatable_under_load_and_argument_checks(cuda_device):
    from gesturediffusion_b200 import _cabi
    Kn = _K()
    M, N, F = 25216 - 37, 512, 1024
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(M, N, device="cuda", generator=g)
    hi = x.half()
    lo = (x - hi.float()).half()
    w1 = (torch.randn(F, N, device="cuda", generator=g) / math.sqrt(N)).half()
    w2 = (torch.randn(N, F, device="cuda", generator=g) / math.sqrt(F)).half()
    b1, b2, gamma, beta = (torch.randn(n, device="cuda", generator=g) for n in (F, N, N, N))
    first = None
    for it in range(16):
        oh, ol = torch.zeros_like(hi), torch.zeros_like(lo)
        Kn.ffn_ln_split(hi, lo, w1, b1, w2, b2, gamma, beta, out_hi=oh, out_lo=ol)
        if first is None:
            first = (oh, ol)
        else:
            assert torch.equal(oh, first[0]) and torch.equal(ol, first[1])
    with pytest.raises(_cabi.GdError, match="unsupported"):
        Kn.ffn_ln_split(hi[:, :256], lo[:, :256], w1[:, :256], b1, w2[:256], b2[:256], gamma[:256], beta[:256],
                        out_hi=torch.zeros(M, 256, device="cuda", dtype=torch.float16))
    with pytest.raises(_cabi.GdError, match="unsupported"):
        Kn.ffn_ln_split(hi, lo, w1[:384], b1[:384], w2[:, :384], b2, gamma, beta, out_hi=torch.zeros_like(hi))


def test_gemm_ln_rejects_other_widths(cuda_device):
    from gesturediffusion_b200 import _cabi
    Kn = _K()
    a = torch.zeros(8, 64, device="cuda", dtype=torch.float16)
    w = torch.zeros(384, 64, device="cuda", dtype=torch.float16)
    v = torch.zeros(384, device="cuda")
    r = torch.zeros(8, 384, device="cuda")
    with pytest.raises(_cabi.GdError, match="not in"):
        Kn.gemm_ln(a, w, v, r, v, v, out_f32=r)


# ------------------------------------------------------------------ local attention + RoPE
@pytest.mark.parametrize("n_seq,T,d,heads", [(3, 20, 128, 8), (2, 120, 512, 8), (1, 600, 512, 8), (2, 10, 128, 8)])
def test_local_attention_rope(cuda_device, n_seq, T, d, heads):
    from oracle.model_oracle import local_window_attention, rope_angles, rope_apply
    Kn = _K()
    e = d // heads
    g = torch.Generator().manual_seed(T)
    z = torch.randn(n_seq, T, d, generator=g)
    ang = rope_angles(T + 1, e, torch.float64)
    zh = z.double().view(n_seq, T, heads, e).transpose(1, 2).reshape(n_seq * heads, T, e)
    oh = local_window_attention(rope_apply(zh, ang[:T]), 10)
    oh = rope_apply(oh, ang[1:T + 1])
    ref = oh.view(n_seq, heads, T, e).transpose(1, 2).reshape(n_seq, T, d)
    x32 = torch.zeros(n_seq, T + 1, d, device="cuda")
    xb = torch.zeros(n_seq, T + 1, d, device="cuda", dtype=torch.float16)
    angf = rope_angles(T + 1, e, torch.float32)
    Kn.local_attention(z.cuda(), angf.cos().cuda().contiguous(), angf.sin().cuda().contiguous(), x32, xb,
                       n_seq, T, d, heads, 10)
    assert rel_l2(x32[:, 1:], ref) < 1e-5
    assert float(x32[:, 0].abs().max()) == 0.0
    assert torch.equal(xb, x32.half())


def test_local_attention_rejects_bad_length(cuda_device):
    from gesturediffusion_b200 import _cabi
    Kn = _K()
    z = torch.zeros(1, 25, 128, device="cuda")
    tab = torch.zeros(26, 16, device="cuda")
    with pytest.raises(_cabi.GdError, match="divisible by window size"):
        Kn.local_attention(z, tab, tab, torch.zeros(1, 26, 128, device="cuda"), None, 1, 25, 128, 8, 10)


# ------------------------------------------------------------------ fused posterior step
def _coef(n):
    from gesturediffusion_b200 import create_gaussian_diffusion
    d = create_gaussian_diffusion(timestep_respacing=[n])
    return d, d.device_tables(torch.device("cuda"))


@pytest.mark.parametrize("B,J,T", [(3, 24, 20), (2, 263, 196), (2, 498, 120), (1, 7, 13)])
@pytest.mark.parametrize("step", [0, 3])
def test_posterior_step_injected_noise(cuda_device, B, J, T, step):
    Kn = _K()
    d, tabs = _coef(5)
    g = torch.Generator(device="cuda").manual_seed(B * 100 + T)
    c, u, x, nz = (torch.randn(B, J, T, device="cuda", generator=g) for _ in range(4))
    scale = torch.rand(B, device="cuda", generator=g) * 3
    tabs["step"].fill_(step)
    xp, x0, mean = (torch.empty(B, J, T, device="cuda") for _ in range(3))
    ld = (J + 26 + 7) // 8 * 8
    a_next = torch.zeros(B * T, ld, device="cuda", dtype=torch.float16 if step else torch.bfloat16)
    Kn.posterior_step(c, x, tabs["coef"], tabs["step"], B, J, T, uncond=u, scale=scale, noise=nz, x_prev=xp,
                      pred_xstart=x0, mean=mean, a_next=a_next)
    r0 = u + scale.view(-1, 1, 1) * (c - u)
    c1 = torch.tensor(d.posterior_mean_coef1[step]).float().cuda()
    c2 = torch.tensor(d.posterior_mean_coef2[step]).float().cuda()
    lv = torch.tensor(d.posterior_log_variance_clipped[step]).float().cuda()
    rmean = c1 * r0 + c2 * x
    rxp = rmean + (0.0 if step == 0 else 1.0) * torch.exp(0.5 * lv) * nz
    assert rel_l2(x0, r0) < 1e-6 and rel_l2(mean, rmean) < 1e-6 and rel_l2(xp, rxp) < 1e-6
    assert torch.equal(a_next[:, :J].view(B, T, J), xp.transpose(1, 2).to(a_next.dtype))
    assert float(a_next[:, J:].float().abs().max()) == 0.0
    assert int(tabs["step"].item()) == step                      # no loop-state update requested


def test_posterior_step_inpaint_clip_eps_const_noise_and_loop_state(cuda_device):
    Kn = _K()
    d, tabs = _coef(6)
    B, J, T = 4, 24, 20
    g = torch.Generator(device="cuda").manual_seed(1)
    c, x, nz, motion = (torch.randn(B, J, T, device="cuda", generator=g) for _ in range(4))
    mask = (torch.rand(B, J, T, device="cuda", generator=g) < 0.4)
    step = 4
    tabs["step"].fill_(step)
    tnext = torch.zeros(B, dtype=torch.int64, device="cuda")
    xp, x0 = torch.empty_like(x), torch.empty_like(x)
    Kn.posterior_step(c, x, tabs["coef"], tabs["step"], B, J, T, noise=nz, x_prev=xp, pred_xstart=x0,
                      inpaint_mask=mask.to(torch.uint8), inpainted=motion, clip=True, const_noise=True,
                      timestep_map=tabs["timestep_map"], next_t=tnext)
    r0 = torch.where(mask, motion, c).clamp(-1, 1)
    f = lambda a: torch.tensor(a[step]).float().cuda()
    rxp = f(d.posterior_mean_coef1) * r0 + f(d.posterior_mean_coef2) * x + \
        torch.exp(0.5 * f(d.posterior_log_variance_clipped)) * nz[:1]
    assert rel_l2(x0, r0) < 1e-6 and rel_l2(xp, rxp) < 1e-6
    assert int(tabs["step"].item()) == step - 1
    assert tnext.tolist() == [d.timestep_map[step - 1]] * B
    # epsilon parameterisation
    tabs["step"].fill_(2)
    Kn.posterior_step(c, x, tabs["coef"], tabs["step"], B, J, T, pred_xstart=x0, eps_mode=True)
    r = torch.tensor(d.sqrt_recip_alphas_cumprod[2]).float().cuda() * x - \
        torch.tensor(d.sqrt_recipm1_alphas_cumprod[2]).float().cuda() * c
    assert rel_l2(x0, r) < 1e-6


def test_philox_noise_statistics_and_shard_invariance(cuda_device):
    """In-kernel Philox: N(0,1) moments, determinism, and the per-sample stream does not depend on how
    the batch is sharded (sample_base) -- the multi-GPU invariance of SURVEY 8.5."""
    Kn = _K()
    B, J, T = 8, 263, 196
    full = torch.empty(B, J, T, device="cuda")
    Kn.philox_normal(full, B, J, T, 1234, 0, 17)
    assert abs(float(full.mean())) < 5e-3 and abs(float(full.std()) - 1) < 5e-3
    assert abs(float((full ** 4).mean()) - 3.0) < 0.05
    part = torch.empty(3, J, T, device="cuda")
    Kn.philox_normal(part, 3, J, T, 1234, 5, 17)
    assert torch.equal(part, full[5:8])
    other = torch.empty(B, J, T, device="cuda")
    Kn.philox_normal(other, B, J, T, 1234, 0, 18)
    assert float((other - full).abs().max()) > 1.0
    # the fused kernel draws the same values: x_prev - mean == sigma * z
    d, tabs = _coef(50)
    tabs["step"].fill_(17)
    c, x = torch.zeros(B, J, T, device="cuda"), torch.zeros(B, J, T, device="cuda")
    xp = torch.empty_like(x)
    Kn.posterior_step(c, x, tabs["coef"], tabs["step"], B, J, T, x_prev=xp, seed=1234, sample_base=0)
    sigma = math.exp(0.5 * d.posterior_log_variance_clipped[17])
    assert rel_l2(xp, full * sigma) < 1e-6


def test_pack_rows_and_q_sample(cuda_device):
    Kn = _K()
    B, C, T = 3, 26, 50
    src = torch.randn(B, C, T, device="cuda")
    for ddt in (torch.bfloat16, torch.float16, torch.float32):
        dst = torch.zeros(B * T, 64, device="cuda", dtype=ddt)
        Kn.pack_rows(src, dst, B, C, T, col_off=30)
        assert torch.equal(dst[:, 30:56].view(B, T, C), src.transpose(1, 2).to(ddt))
        assert float(dst[:, :30].float().abs().max()) == 0.0
    d, tabs = _coef(10)
    x0, nz = torch.randn(1000, device="cuda"), torch.randn(1000, device="cuda")
    out = torch.empty(1000, device="cuda")
    Kn.q_sample(x0, nz, tabs["sqrt_acp"], tabs["sqrt_1macp"], 7, out)
    ref = float(d.sqrt_alphas_cumprod[7]) * x0.double() + float(d.sqrt_one_minus_alphas_cumprod[7]) * nz.double()
    assert rel_l2(out, ref) < 1e-6
