"""GPU: the drop-in API end to end (MDM / MDM_Old / ClassifierFreeSampleModel / SpacedDiffusion), through
the C ABI, against (a) golden outputs of the unmodified reference and (b) the CPU oracle on fresh inputs.

Tolerances: the north star's 1e-5 per-step relative L2 in fp32 mode and 2e-3 in the 16-bit mode, 1e-2 for the final
motion of a chain.  The 16-bit mode held to 2e-3 is fp16 (the package default, what bench.py times); bf16 operands (the
same kernels) are bounded by bf16's own rounding at 3-4.5e-3 per forward (tools/sim_16bit_rounding.py) and are a stated
deviation (bench.py `precision_note`), tested here only against that looser bound.  fp32 mode holds 1e-5
for the guided combination and for whole chains (50 and 1000 steps) too.  Every measured error is appended to gpurun_out/parity.txt (helpers.record_parity; kept as profiles/r02_parity.txt).
Noise is injected identically on both sides."""
import pytest
import torch

from helpers import (CHAIN_CASES, MODEL_CASES, build_product_model, load_chain_case, load_model_case, record_parity,
                     rel_l2, to_dev)

pytestmark = pytest.mark.gpu

# "fp16" is the default 16-bit tensor-core mode and is held to the north star's 16-bit bound (2e-3 per-step,
# 1e-2 chain) -- also for the guided x0 prediction.  "bf16" (same kernels, bf16 operands) is bounded by its
# operand rounding: 3-4.5e-3 per forward measured AND reproduced by a CPU simulation of the roundings
# (tools/sim_16bit_rounding.py); it keeps the chain bound.  "fp32" is the 1e-5 mode.
TOL = {"fp32": 1e-5, "fp16": 2e-3, "bf16": 6e-3}
GUIDED_TOL = {"fp32": 1e-5, "fp16": 2e-3, "bf16": 1e-2}     # u + 2.5 (c - u) amplifies the c-u rounding (bf16 only)
CHAIN_TOL = {"fp32": 1e-5, "fp16": 1e-2, "bf16": 1e-2}     # fp32 measured <= 1.9e-6 over 50 / 1000 steps
PRECISIONS = ["fp32", "fp16", "bf16"]


def _diffusion(respacing):
    from gesturediffusion_b200 import create_gaussian_diffusion
    return create_gaussian_diffusion(timestep_respacing=respacing)


@pytest.mark.parametrize("case", MODEL_CASES)
@pytest.mark.parametrize("precision", PRECISIONS)
def test_forward_matches_reference_golden(cuda_device, case, precision):
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    cfg, sd, x, y, gold = load_model_case(case)
    m = build_product_model(cfg, sd, cuda_device, precision)
    xd, yd, t = x.to(cuda_device), to_dev(y, cuda_device), gold["t"].to(cuda_device)
    cond = m(xd, t, yd)
    assert cond.shape == x.shape and cond.dtype == torch.float32 and cond.is_cuda
    yu = dict(yd)
    yu["uncond"] = True
    unc = m(xd, t, yu)
    guided = ClassifierFreeSampleModel(m)(xd, t, yd)
    errs = (rel_l2(cond, gold["cond"]), rel_l2(unc, gold["uncond"]), rel_l2(guided, gold["guided"]))
    record_parity("forward %s %s: cond %.2e uncond %.2e guided %.2e" % ((case, precision) + errs))
    assert errs[0] < TOL[precision] and errs[1] < TOL[precision]
    assert errs[2] < GUIDED_TOL[precision]


@pytest.mark.parametrize("arch", ["v1", "v2"])
@pytest.mark.parametrize("precision", PRECISIONS)
def test_forward_matches_oracle_fresh_inputs(cuda_device, arch, precision):
    """Fresh seeds / other shapes than the fixtures, checked against the CPU oracle (float64)."""
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, model_config
    from oracle.model_oracle import make_model_fn
    cfg = model_config(arch=arch, njoints=37, latent_dim=256, num_layers=3, num_heads=4, ff_size=512)
    sd = make_state_dict(cfg, seed=7)
    for B, T in ((1, 10), (5, 40), (2, 130)):
        x, y = make_inputs(cfg, B, T, seed=B * 31 + T)
        t = torch.randint(0, 1000, (B,), generator=torch.Generator().manual_seed(T))
        ref = make_model_fn(sd, cfg, torch.float64)(x.double(), t, y)
        m = build_product_model(cfg, sd, cuda_device, precision)
        out = m(x.to(cuda_device), t.to(cuda_device), to_dev(y, cuda_device))
        err = rel_l2(out, ref)
        record_parity("fresh %s %s B=%d T=%d: %.2e" % (arch, precision, B, T, err))
        assert err < TOL[precision]


@pytest.mark.parametrize("case", ["v1_humanml", "v2_gesture"])
@pytest.mark.parametrize("precision", ["fp16", "bf16"])
def test_forward_with_fused_ffn_kernel(cuda_device, case, precision):
    """The one-kernel feed-forward sub-layer (gd_ffn_h16_ln_split, the default; GD_FUSE_FFN=0 selects linear1 + GELU and
    linear2 + residual + LayerNorm as two kernels) against the reference golden and against the two-kernel path."""
    cfg, sd, x, y, gold = load_model_case(case)
    m = build_product_model(cfg, sd, cuda_device, precision)
    xd, yd, t = x.to(cuda_device), to_dev(y, cuda_device), gold["t"].to(cuda_device)
    eng = m.engine()
    assert eng.fuse_ffn and eng.fuse_ffn_min_rows > 0 and eng.fuse_ln_min_rows > 0   # defaults: fused from ~5 k token rows on
    rows = x.shape[0] * (x.shape[-1] + 1)
    front = 3 + (1 if cfg["arch"] == "v2" else 0)
    saved = (eng.fuse_ffn_min_rows, eng.fuse_ln_min_rows, eng.fuse_ln_small_max_rows)
    # the golden cases are small: at d = 512 the few-row GEMM + LayerNorm kernel (8-CTA cluster per row tile) ...
    assert eng.fuse_ln_small_max_rows >= rows
    assert eng.kernels_per_step(rows) == front + (5 if eng.d == 512 else 7) * cfg["num_layers"]
    small = m(xd, t, yd).clone()
    try:
        eng.fuse_ln_small_max_rows = 0                      # ... otherwise plain GEMM + LayerNorm kernels
        assert eng.kernels_per_step(rows) == front + 7 * cfg["num_layers"]
        base = m(xd, t, yd).clone()
        e_small = rel_l2(small, gold["cond"])
        record_parity("few-row ln %s %s: %.2e (vs unfused path %.2e)" % (case, precision, e_small, rel_l2(small, base)))
        assert e_small < TOL[precision]
        assert rel_l2(small, base) < TOL[precision]
        eng.fuse_ln_min_rows = 0                            # cluster GEMM+LayerNorm kernel, two-kernel feed-forward path
        eng.fuse_ffn_min_rows = 1 << 30
        assert eng.kernels_per_step(rows) == 3 + (1 if cfg["arch"] == "v2" else 0) + 5 * cfg["num_layers"]
        two = m(xd, t, yd).clone()
        eng.fuse_ffn_min_rows = 0                           # + the one-kernel feed-forward sub-layer
        assert eng.kernels_per_step(rows) == 3 + (1 if cfg["arch"] == "v2" else 0) + 4 * cfg["num_layers"]
        fused = m(xd, t, yd).clone()
    finally:
        eng.fuse_ffn_min_rows, eng.fuse_ln_min_rows, eng.fuse_ln_small_max_rows = saved
    e_two = rel_l2(two, gold["cond"])
    record_parity("fused ln %s %s: %.2e (vs unfused path %.2e)" % (case, precision, e_two, rel_l2(two, base)))
    assert e_two < TOL[precision]
    base = two
    err = rel_l2(fused, gold["cond"])
    record_parity("fused ffn %s %s: %.2e (vs two-kernel path %.2e)" % (case, precision, err, rel_l2(fused, base)))
    assert err < TOL[precision]
    assert rel_l2(fused, base) < TOL[precision]


def test_v2_rejects_length_not_multiple_of_window(cuda_device):
    cfg, sd, x, y, gold = load_model_case("v2_small")
    m = build_product_model(cfg, sd, cuda_device, "fp32")
    x25 = torch.zeros(1, cfg["njoints"], 1, 25, device=cuda_device)
    y25 = {"seed": torch.zeros(1, cfg["njoints"], 1, 10, device=cuda_device),
           "mfcc": torch.zeros(1, 26, 1, 25, device=cuda_device)}
    with pytest.raises(ValueError, match="divisible by window size"):
        m(x25, torch.zeros(1, dtype=torch.long, device=cuda_device), y25)


@pytest.mark.parametrize("case", CHAIN_CASES)
@pytest.mark.parametrize("precision", PRECISIONS)
def test_chain_matches_reference_golden(cuda_device, case, precision):
    """p_sample_loop_progressive with injected noise vs the reference's trace (pred_xstart and sample at
    recorded steps, and the final motion)."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    cfg, sd, x, y, noises, meta, kwargs, gold = load_chain_case(case)
    m = build_product_model(cfg, sd, cuda_device, precision)
    model = ClassifierFreeSampleModel(m) if meta["cfg"] else m
    dif = _diffusion(meta["respacing"])
    if kwargs["init_image"] is not None:
        kwargs["init_image"] = kwargs["init_image"].to(cuda_device)
    outs = list(dif.p_sample_loop_progressive(model, tuple(x.shape), noise=x.to(cuda_device), clip_denoised=False,
                                              model_kwargs={"y": to_dev(y, cuda_device)}, step_noise=noises,
                                              **kwargs))
    assert len(outs) == meta["n_exec"]
    first_pred = rel_l2(outs[0]["pred_xstart"], gold["trace_pred"][0])
    final = rel_l2(outs[-1]["sample"], gold["final"])
    record_parity("chain %s %s: first pred_xstart %.2e final %.2e" % (case, precision, first_pred, final))
    assert first_pred < GUIDED_TOL[precision] if meta["cfg"] else first_pred < TOL[precision]
    assert final < CHAIN_TOL[precision]
    for j, i in enumerate(gold["trace_idx"]):
        assert rel_l2(outs[i]["sample"], gold["trace_sample"][j]) < CHAIN_TOL[precision]


def test_loop_api_variants_agree(cuda_device):
    """p_sample_loop == last yield of the progressive loop == eager (no CUDA graph) loop; dump_steps; and the
    generic per-step path used for arbitrary callables gives the same numbers as the captured graph."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    cfg, sd, x, y, noises, meta, kwargs, gold = load_chain_case("chain_v2_small_cfg5")
    m = build_product_model(cfg, sd, cuda_device, "fp32")
    model = ClassifierFreeSampleModel(m)
    dif = _diffusion([5])
    args = dict(noise=x.to(cuda_device), clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)},
                step_noise=noises)
    a = dif.p_sample_loop(model, tuple(x.shape), **args)
    prog = list(dif.p_sample_loop_progressive(model, tuple(x.shape), **args))
    assert torch.equal(a, prog[-1]["sample"])
    dumps = dif.p_sample_loop(model, tuple(x.shape), dump_steps=[1, 3], **args)
    assert len(dumps) == 2 and torch.equal(dumps[0], prog[1]["sample"]) and torch.equal(dumps[1], prog[3]["sample"])
    dif.use_cuda_graph = False
    b = dif.p_sample_loop(model, tuple(x.shape), **args)
    assert torch.equal(a, b)
    dif.use_cuda_graph = True
    # generic path: wrap the model in a plain callable so the loop cannot take the engine fast path
    class Plain(torch.nn.Module):
        def __init__(self, inner):
            super().__init__()
            self.inner = inner

        def forward(self, x, t, **kw):
            return self.inner(x, t, **kw)
    c = dif.p_sample_loop(Plain(model), tuple(x.shape), **args)
    assert rel_l2(c, a) < 1e-6
    assert rel_l2(a, gold["final"]) < CHAIN_TOL["fp32"]


def test_p_mean_variance_contract(cuda_device):
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    from oracle.diffusion_oracle import OracleDiffusion, oracle_cfg
    from oracle.model_oracle import make_model_fn
    cfg, sd, x, y, gold = load_model_case("v1_small")
    m = build_product_model(cfg, sd, cuda_device, "fp32")
    dif = _diffusion([50])
    t = torch.full((x.shape[0],), 31, dtype=torch.long)
    out = dif.p_mean_variance(ClassifierFreeSampleModel(m), x.to(cuda_device), t.to(cuda_device),
                              clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)})
    ref = OracleDiffusion(respacing=[50]).p_mean_variance(oracle_cfg(make_model_fn(sd, cfg)), x, t, y)
    assert set(out) == {"mean", "variance", "log_variance", "pred_xstart"}
    for k in out:
        assert out[k].shape == x.shape
        assert rel_l2(out[k], ref[k]) < 1e-5, k


def test_torch_rng_stream_is_the_references(cuda_device):
    """rng='torch': the loop consumes torch's CUDA generator exactly like the reference (one randn for
    x_T, one randn_like per step), so re-drawing that stream and injecting it reproduces the run bit for bit."""
    cfg, sd, x, y, gold = load_model_case("v1_small")
    m = build_product_model(cfg, sd, cuda_device, "fp32")
    dif = _diffusion([4])
    shape = tuple(x.shape)
    kw = dict(clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)})
    torch.manual_seed(77)
    a = dif.p_sample_loop(m, shape, **kw)
    torch.manual_seed(77)
    x_T = torch.randn(*shape, device=cuda_device)
    noises = [torch.randn_like(x_T) for _ in range(4)]
    b = dif.p_sample_loop(m, shape, noise=x_T, step_noise=noises, **kw)
    assert torch.equal(a, b)


def test_philox_sharding_invariance_single_gpu(cuda_device):
    """Two 'ranks' emulated on one GPU: shard-then-concatenate equals the unsharded run on the same
    per-sample Philox streams (multi-GPU row)."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    from gesturediffusion_b200 import dist as gdist
    from gesturediffusion_b200.synthetic import make_inputs
    cfg, sd, _, _, _ = load_model_case("v2_small")
    x, y = make_inputs(cfg, 5, 20, seed=3)
    m = build_product_model(cfg, sd, cuda_device, "fp16")
    model = ClassifierFreeSampleModel(m)
    dif = _diffusion([6])
    dif.rng, dif.seed = "philox", 2024
    shape = tuple(x.shape)
    kw = {"y": to_dev(y, cuda_device)}
    dif.sample_base = 0
    full = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs=kw)
    parts = []
    for r in range(2):
        sh, kwr, _, base = gdist.shard_inputs(shape, kw, r, 2)
        dif.sample_base = base
        parts.append(dif.p_sample_loop(model, sh, clip_denoised=False, model_kwargs=kwr))
    got = torch.cat(parts, dim=0)
    assert got.shape == full.shape
    assert rel_l2(got, full) < 1e-6


@pytest.mark.parametrize("arch,J,T,B", [("v1", 263, 196, 64), ("v2", 498, 120, 64)])
def test_full_size_properties(cuda_device, arch, J, T, B):
    """BASELINE-size workloads (cfg2a at its full batch, cfg3 at half its batch), 50-step CFG sampling with in-kernel
    noise, through properties that do not need the CPU reference: the 16-bit tensor-core path agrees with the fp32
    SIMT path of the same package within the chain tolerance, sharding the batch changes no bit (both halves still have the
    4096 token rows from which the one-kernel feed-forward path is used, so the shards run the same kernels), and a
    repeated call replays the same graph to the same bits."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    from gesturediffusion_b200 import dist as gdist
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, model_config
    cfg = model_config(arch=arch, njoints=J)
    sd = make_state_dict(cfg, seed=0)
    x, y = make_inputs(cfg, B, T, seed=11)
    shape, kw = tuple(x.shape), {"y": to_dev(y, cuda_device)}
    outs = {}
    for precision in ("fp16", "fp32"):
        model = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, precision))
        dif = _diffusion([50])
        dif.rng, dif.seed, dif.sample_base = "philox", 7, 0
        outs[precision] = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs=kw)
        assert bool(torch.isfinite(outs[precision]).all())
        if precision == "fp16":
            again = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs=kw)
            assert torch.equal(again, outs["fp16"])
            parts = []
            for r in range(2):
                sh, kwr, _, base = gdist.shard_inputs(shape, kw, r, 2)
                dif.sample_base = base
                parts.append(dif.p_sample_loop(model, sh, clip_denoised=False, model_kwargs=kwr))
            assert torch.equal(torch.cat(parts, dim=0), outs["fp16"])
    err = rel_l2(outs["fp16"], outs["fp32"])
    record_parity("full size %s: fp16 vs fp32 mode after 50 steps %.2e" % (arch, err))
    assert err < CHAIN_TOL["fp16"]
