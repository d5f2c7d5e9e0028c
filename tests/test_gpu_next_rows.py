"""GPU: DDIM sampler and chunked-take driver of the package against the reference's own
outputs (tests/golden), through the C ABI; tolerances as in test_gpu_model.py."""
import pytest
import torch

from helpers import (DDIM_CASES, TAKE_CASES, build_product_model, load_ddim_case, load_take_case, record_parity, rel_l2, to_dev)

pytestmark = pytest.mark.gpu

PRECISIONS = ["fp32", "fp16", "bf16"]
CHAIN_TOL = {"fp32": 1e-5, "fp16": 1e-2, "bf16": 1e-2}


def _diffusion(resp):
    from gesturediffusion_b200 import create_gaussian_diffusion
    return create_gaussian_diffusion(timestep_respacing=resp)


@pytest.mark.parametrize("case", DDIM_CASES)
@pytest.mark.parametrize("precision", PRECISIONS)
def test_ddim_matches_reference_golden(cuda_device, case, precision):
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    cfg, sd, x, y, noises, meta, gold = load_ddim_case(case)
    m = build_product_model(cfg, sd, cuda_device, precision)
    model = ClassifierFreeSampleModel(m) if meta["cfg"] else m
    dif = _diffusion(meta["respacing"])
    args = dict(noise=x.to(cuda_device), clip_denoised=meta["clip"], model_kwargs={"y": to_dev(y, cuda_device)},
                eta=meta["eta"], step_noise=noises)
    outs = list(dif.ddim_sample_loop_progressive(model, tuple(x.shape), **args))
    assert len(outs) == meta["n_exec"]
    errs = [rel_l2(o["sample"], gold["trace_sample"][i]) for i, o in enumerate(outs)]
    record_parity("ddim %s %s: per-step sample err max %.2e, final %.2e" % (case, precision, max(errs), errs[-1]))
    assert max(errs) < CHAIN_TOL[precision]
    assert rel_l2(outs[-1]["pred_xstart"], gold["trace_pred"][-1]) < CHAIN_TOL[precision]
    final = dif.ddim_sample_loop(model, tuple(x.shape), **args)
    assert torch.equal(final, outs[-1]["sample"])
    dif.use_cuda_graph = False                      # eager launches give the same bits as the replayed graph
    assert torch.equal(dif.ddim_sample_loop(model, tuple(x.shape), **args), final)


def test_ddim_generic_callable_path_and_refusals(cuda_device):
    """Arbitrary callables go through ddim_sample step by step and agree with the captured graph; the reference's
    refusals are kept (gaussian_diffusion.py:900-903)."""
    cfg, sd, x, y, noises, meta, gold = load_ddim_case("ddim_v2_small_uncond8_eta1")
    m = build_product_model(cfg, sd, cuda_device, "fp32")
    dif = _diffusion(meta["respacing"])
    args = dict(noise=x.to(cuda_device), clip_denoised=True, model_kwargs={"y": to_dev(y, cuda_device)}, eta=1.0,
                step_noise=noises)
    fast = dif.ddim_sample_loop(m, tuple(x.shape), **args)
    slow = dif.ddim_sample_loop(lambda *a, **k: m(*a, **k), tuple(x.shape), device=cuda_device, **args)
    assert rel_l2(slow, fast) < 1e-6
    with pytest.raises(NotImplementedError):
        dif.ddim_sample_loop(m, tuple(x.shape), dump_steps=[1], **args)
    with pytest.raises(NotImplementedError):
        dif.ddim_sample_loop(m, tuple(x.shape), const_noise=True, **args)
    # one step through the public ddim_sample with torch's generator where the reference draws it
    t = torch.full((x.shape[0],), 3, dtype=torch.long, device=cuda_device)
    torch.manual_seed(5)
    a = dif.ddim_sample(m, x.to(cuda_device), t, model_kwargs={"y": to_dev(y, cuda_device)}, eta=0.3)
    torch.manual_seed(5)
    b = dif.ddim_sample(m, x.to(cuda_device), t, model_kwargs={"y": to_dev(y, cuda_device)}, eta=0.3)
    assert torch.equal(a["sample"], b["sample"]) and a["sample"].shape == x.shape


@pytest.mark.parametrize("case", TAKE_CASES)
@pytest.mark.parametrize("precision", PRECISIONS)
def test_take_matches_reference_golden(cuda_device, case, precision):
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    from gesturediffusion_b200.generate import generate_take
    cfg, sd, B, T, ys, x_T, mean, std, step_noise, meta, gold = load_take_case(case)
    model = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, precision))
    dif = _diffusion(meta["respacing"])
    out = generate_take(model, dif, ys, B, T, seed_poses=cfg["seed_poses"], guidance_param=meta["guidance"],
                        mean=mean, std=std, init_noise=[v.to(cuda_device) for v in x_T], step_noise=step_noise)
    for k in ("sample", "motion", "motion_rot"):
        assert out[k].is_cuda and tuple(out[k].shape) == tuple(gold[k].shape)
        err = rel_l2(out[k], gold[k])
        record_parity("take %s %s %s: %.2e" % (case, precision, k, err))
        assert err < CHAIN_TOL[precision]


def test_take_philox_chunks_differ_and_reuse_one_graph(cuda_device):
    """Throughput mode: in-kernel noise, a fresh seed per chunk, and still a single captured step graph."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel
    from gesturediffusion_b200.generate import generate_take
    cfg, sd, B, T, ys, x_T, mean, std, step_noise, meta, gold = load_take_case("take_v2_small_cfg4x3")
    model = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, "fp16"))
    dif = _diffusion(meta["respacing"])
    dif.rng, dif.seed = "philox", 11
    same = [ys[0]] * 3
    out = generate_take(model, dif, same, B, T, mean=mean, std=std)
    again = generate_take(model, dif, same, B, T, mean=mean, std=std)
    assert torch.equal(out["motion"], again["motion"]) and dif.seed == 11
    assert not torch.equal(out["sample"][..., :T], out["sample"][..., T:2 * T])
    eng = model.model.engine()
    graphs = [st["graphs"] for ws in eng._ws.values() for k, st in ws.items()
              if isinstance(k, tuple) and k and k[0] == "loop"]
    assert sum(len(g) for g in graphs) == 1


def test_unnormalize_split_kernel(cuda_device):
    from gesturediffusion_b200 import kernels as K
    from oracle.generate_oracle import oracle_split_positions_rotations
    B, J, T, Tt = 3, 498, 120, 360
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, J, 1, T, generator=g)
    mean, std = torch.randn(J, generator=g), torch.rand(J, generator=g) + 0.5
    pos = torch.zeros(B, J // 6, 3, Tt, device=cuda_device)
    rot = torch.zeros_like(pos)
    K.unnormalize_split(x.to(cuda_device).reshape(B, J, T), mean.to(cuda_device), std.to(cuda_device), pos, rot, T)
    rp, rr = oracle_split_positions_rotations(x, mean, std)
    assert torch.equal(pos[..., T:2 * T].cpu(), rp.contiguous()) and torch.equal(rot[..., T:2 * T].cpu(), rr.contiguous())
    assert float(pos[..., :T].abs().max()) == 0.0 and float(pos[..., 2 * T:].abs().max()) == 0.0
    from gesturediffusion_b200 import _cabi
    with pytest.raises(_cabi.GdError, match="multiple of 6"):
        K.unnormalize_split(torch.zeros(1, 7, 4, device=cuda_device), None, None, torch.zeros(1, 1, 3, 4, device=cuda_device), None, 0)


def test_checkpoint_to_take_end_to_end(cuda_device, tmp_path):
    """args.json + checkpoint -> load_checkpoint -> gg_collate'd chunks -> generate_take (sample/generate.py:62-189
    through the package), checked chunk by chunk against separate p_sample_loop calls + the oracle post-processing."""
    import json
    import os
    import numpy as np
    from gesturediffusion_b200.data_loaders.tensors import gg_collate
    from gesturediffusion_b200.generate import generate_take
    from gesturediffusion_b200.synthetic import make_state_dict, make_step_noise, model_config
    from gesturediffusion_b200.utils import model_util
    from oracle.generate_oracle import oracle_split_positions_rotations
    cfg = model_config(arch="v2", njoints=498, latent_dim=128, num_layers=2, num_heads=4, ff_size=1024)
    path = os.path.join(str(tmp_path), "model.pt")
    torch.save(make_state_dict(cfg, seed=3), path)
    with open(os.path.join(str(tmp_path), "args.json"), "w") as fh:
        json.dump(dict(dataset="genea2023", layers=2, latent_dim=128, cond_mask_prob=0.1, use_text=False,
                       use_audio=False, mfcc_input=True, use_wav_enc=False, seed_poses=10, noise_schedule="cosine",
                       sigma_small=True, lambda_vel=0.0, lambda_rcxyz=0.0, lambda_fc=0.0), fh)
    model, dif, args = model_util.load_checkpoint(path, device=cuda_device, precision="fp32", timestep_respacing=[4])
    B, T, chunks = 2, 120, 2
    r = np.random.RandomState(1)
    conds = [gg_collate([(r.randn(T, 498), "w", T, r.randn(800), r.randn(T, 26), r.randn(10, 498))
                         for _ in range(B)])[1]["y"] for _ in range(chunks)]
    g = torch.Generator().manual_seed(2)
    mean, std = torch.randn(498, generator=g), torch.rand(498, generator=g) + 0.5
    x_T = [torch.randn(B, 498, 1, T, generator=g) for _ in range(chunks)]
    noises = [make_step_noise((B, 498, 1, T), 4, seed=50 + k) for k in range(chunks)]
    out = generate_take(model, dif, conds, B, T, guidance_param=args.guidance_param, mean=mean, std=std,
                        init_noise=[v.to(cuda_device) for v in x_T], step_noise=noises)
    assert tuple(out["motion"].shape) == (B, 83, 3, chunks * T) and bool(torch.isfinite(out["motion"]).all())
    prev = None
    for k in range(chunks):
        y = to_dev({kk: v for kk, v in conds[k].items()}, cuda_device)
        if k:
            y["seed"] = prev[..., -10:]
        y["scale"] = torch.ones(B, device=cuda_device) * args.guidance_param
        prev = dif.p_sample_loop(model, (B, 498, 1, T), clip_denoised=False, model_kwargs={"y": y},
                                 noise=x_T[k].to(cuda_device), step_noise=noises[k])
        assert torch.equal(prev, out["sample"][..., k * T:(k + 1) * T])
        pos, rot = oracle_split_positions_rotations(prev, mean, std)
        assert torch.equal(out["motion"][..., k * T:(k + 1) * T].cpu(), pos.contiguous())
        assert torch.equal(out["motion_rot"][..., k * T:(k + 1) * T].cpu(), rot.contiguous())
