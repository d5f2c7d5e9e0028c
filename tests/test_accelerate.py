"""`gesturediffusion_b200.accelerate()` -- the three-line switch INTEGRATION.md section 1 leads with -- on live objects of
the UNMODIFIED reference (model/mdm.py, model/mdm_old.py, model/cfg_sampler.py, diffusion/respace.py): the conversion
carries weights and float64 schedule tables across bit-identically (CPU part), and the converted objects run a 5-step
guided chain to the reference's golden result (GPU part).  The reference is /root/reference in the build container and
the byte-identical archive oracle/_ref/ (oracle/stage_ref.py) on the GPU box."""
import numpy as np
import pytest
import torch

from helpers import build_product_model, load_chain_case, load_model_case, record_parity, rel_l2, to_dev
from oracle import ref_shims

needs_ref = pytest.mark.skipif(not ref_shims.reference_available(), reason="reference neither live nor staged")

TABLES = ["betas", "alphas_cumprod", "alphas_cumprod_prev", "alphas_cumprod_next", "sqrt_alphas_cumprod",
          "sqrt_one_minus_alphas_cumprod", "log_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod",
          "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
          "posterior_mean_coef1", "posterior_mean_coef2"]


def _reference_objects(case, respacing, use_cfg=True):
    cfg, sd, x, y, gold = load_model_case(case)
    m = ref_shims.build_reference_model(cfg["arch"], cfg["njoints"], latent_dim=cfg["latent_dim"],
                                        num_layers=cfg["num_layers"], num_heads=cfg["num_heads"],
                                        ff_size=cfg["ff_size"], seed_poses=cfg["seed_poses"], use_text=cfg["use_text"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not missing and not unexpected
    model = ref_shims.load_reference().ClassifierFreeSampleModel(m) if use_cfg else m
    return cfg, sd, model, ref_shims.build_reference_diffusion(respacing)


@needs_ref
@pytest.mark.parametrize("case", ["v1_small", "v2_small", "v2_small_text"])
@pytest.mark.parametrize("respacing", [[5], "ddim7", [1000]])
def test_accelerate_conversion_is_bit_identical(case, respacing):
    """CPU part: no kernel runs -- only the conversion (class mapping, state_dict, schedule tables, index map).  The
    converted module stays on the CPU; its engine is only built on first use (forward), which this test never does."""
    import gesturediffusion_b200 as gdb
    from gesturediffusion_b200 import convert
    cfg, sd, ref_model, ref_dif = _reference_objects(case, respacing)
    model, dif = convert.accelerate(ref_model, ref_dif, device="cpu", precision="fp16")
    assert type(model).__name__ == "ClassifierFreeSampleModel"
    assert type(model).__module__.startswith("gesturediffusion_b200")
    inner, ref_inner = model.model, ref_model.model
    assert type(inner).__name__ == type(ref_inner).__name__
    ref_sd = {k: v for k, v in ref_inner.state_dict().items() if not k.startswith("clip_model.")}
    new_sd = inner.state_dict()
    assert list(new_sd.keys()) == list(ref_sd.keys())
    for k in ref_sd:
        assert torch.equal(new_sd[k], ref_sd[k]), k
    for attr in ("njoints", "nfeats", "data_rep"):                   # what cfg_sampler.py:17-21 re-exports
        assert getattr(model, attr) == getattr(ref_model, attr)
    assert inner.cond_mask_prob == ref_inner.cond_mask_prob
    assert dif.timestep_map == list(ref_dif.timestep_map)
    assert dif.num_timesteps == ref_dif.num_timesteps and dif.original_num_steps == ref_dif.original_num_steps
    assert dif.model_mean_type.name == ref_dif.model_mean_type.name
    assert dif.model_var_type.name == ref_dif.model_var_type.name
    for t in TABLES:
        a, b = getattr(dif, t), getattr(ref_dif, t)
        assert a.dtype == np.float64 and np.array_equal(a, b), t
    assert gdb.accelerate is convert.accelerate


@needs_ref
def test_accelerate_rejects_unknown_model():
    from gesturediffusion_b200 import convert
    with pytest.raises(TypeError):
        convert.accelerate(torch.nn.Linear(2, 2), None, device="cpu")


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("chain", ["chain_v1_small_cfg5", "chain_v2_small_cfg5"])
def test_accelerated_objects_reproduce_reference_chain(cuda_device, chain):
    """GPU part: reference objects -> accelerate() -> p_sample_loop with the call signature of sample/generate.py:119-130
    against the chain the reference itself produced (tests/golden)."""
    import gesturediffusion_b200 as gdb
    cfg, sd, x, y, noises, meta, kwargs, gold = load_chain_case(chain)
    _, _, ref_model, ref_dif = _reference_objects(meta["case"], meta["respacing"])
    model, dif = gdb.accelerate(ref_model, ref_dif, device=cuda_device, precision="fp16")
    out = dif.p_sample_loop(model, tuple(x.shape), clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)},
                            skip_timesteps=0, init_image=None, progress=False, dump_steps=None,
                            noise=x.to(cuda_device), const_noise=False, step_noise=noises)
    err = rel_l2(out, gold["final"])
    record_parity("accelerate %s fp16: final %.2e" % (chain, err))
    assert err < 1e-2
    # and the package's own classes built from the same state_dict give the same bits (one engine path)
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    own = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, "fp16"))
    out2 = create_gaussian_diffusion(timestep_respacing=meta["respacing"]).p_sample_loop(
        own, tuple(x.shape), clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)}, noise=x.to(cuda_device),
        step_noise=noises)
    assert torch.equal(out, out2)
