"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference
through oracle/ref_shims.py).  Run in the build container only:

    python tests/golden/make_golden.py

Weights and inputs are regenerated from seeds by gesturediffusion_b200.synthetic (deterministic on
any machine with this torch build), so the fixtures hold only configs, seeds and reference OUTPUTS.
The GPU box has no /root/reference: tests there compare against these files.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, make_step_noise, model_config  # noqa: E402
from oracle import ref_shims  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

# name -> (model_config kwargs, batch, frames)
MODEL_CASES = {
    "v1_small": (dict(arch="v1", njoints=24, latent_dim=128, num_layers=2, num_heads=4, ff_size=256), 3, 20),
    "v2_small": (dict(arch="v2", njoints=24, latent_dim=128, num_layers=2, num_heads=4, ff_size=256), 3, 20),
    "v2_small_text": (dict(arch="v2", njoints=24, latent_dim=128, num_layers=2, num_heads=4, ff_size=256,
                           use_text=True), 2, 30),
    "v1_humanml": (dict(arch="v1", njoints=263), 1, 196),          # BASELINE config 1 shape
    "v1_kit": (dict(arch="v1", njoints=251), 2, 196),              # BASELINE config 4 shape
    "v2_gesture": (dict(arch="v2", njoints=498), 2, 120),          # BASELINE config 3 shape
    "v2_text": (dict(arch="v2", njoints=263, use_text=True), 2, 200),   # BASELINE config 2b shape (ref breaks at B=1: mdm.py:122 squeeze)
}
# name -> (model case, respacing, cfg?, extra)
CHAIN_CASES = {
    "chain_v1_small_cfg5": ("v1_small", [5], True, {}),
    "chain_v2_small_cfg5": ("v2_small", [5], True, {}),
    "chain_v1_small_uncond4_skip": ("v1_small", [6], False, dict(skip_timesteps=2, init_image=True)),
    "chain_v2_small_const_noise": ("v2_small", [4], True, dict(const_noise=True)),
    "chain_v1_small_inpaint": ("v1_small", [4], True, dict(inpaint=True)),
    "chain_v1_humanml_uncond50": ("v1_humanml", [50], False, {}),  # BASELINE config 1 in full
    "chain_v2_gesture_cfg50": ("v2_gesture", [50], True, {}),      # BASELINE config 3 at B=2
    # the reference's DEFAULT schedule: all 1000 steps (utils/model_util.py:42,49-50), CFG -- BASELINE config 2's schedule
    "chain_v1_small_cfg1000": ("v1_small", [1000], True, {}),
    "chain_v2_small_cfg1000": ("v2_small", [1000], True, {}),
}


def build(case):
    kw, B, T = MODEL_CASES[case]
    cfg = model_config(**kw)
    sd = make_state_dict(cfg, seed=0)
    m = ref_shims.build_reference_model(cfg["arch"], cfg["njoints"], latent_dim=cfg["latent_dim"],
                                        num_layers=cfg["num_layers"], num_heads=cfg["num_heads"],
                                        ff_size=cfg["ff_size"], seed_poses=cfg["seed_poses"],
                                        use_text=cfg["use_text"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not missing and not unexpected, (missing, unexpected)
    x, y = make_inputs(cfg, B, T)
    return cfg, sd, m, x, y


def schedule_fixture():
    ns = ref_shims.load_reference()
    out = {}
    specs = {"s1000_50": (1000, [50]), "s1000_ddim50": (1000, "ddim50"), "s300_10_15_20": (300, [10, 15, 20]),
             "s1000_1000": (1000, [1000]), "s1000_str": (1000, "25,25"), "s1000_7": (1000, [7]),
             "s1000_1": (1000, [1])}
    for name, (n, spec) in specs.items():
        out["space_" + name] = np.array(sorted(ns.space_timesteps(n, spec)), dtype=np.int64)
    tables = ["betas", "alphas_cumprod", "alphas_cumprod_prev", "alphas_cumprod_next", "sqrt_alphas_cumprod",
              "sqrt_one_minus_alphas_cumprod", "log_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod",
              "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
              "posterior_mean_coef1", "posterior_mean_coef2"]
    for sched in ("cosine", "linear"):
        for resp in ([1000], [50], "ddim50"):
            dif = ref_shims.build_reference_diffusion(resp, schedule=sched)
            tag = "%s_%s" % (sched, resp if isinstance(resp, str) else resp[0])
            out["tmap_" + tag] = np.array(dif.timestep_map, dtype=np.int64)
            for t in tables:
                out["%s_%s" % (t, tag)] = getattr(dif, t)
    np.savez(os.path.join(OUT, "schedule.npz"), **out)
    print("schedule.npz", len(out), "arrays")


def model_fixtures():
    for case in MODEL_CASES:
        cfg, sd, m, x, y = build(case)
        B = x.shape[0]
        t = torch.tensor([999, 500, 17, 0][:B]) if B > 1 else torch.tensor([421])
        with torch.no_grad():
            cond = m(x, t, y)
            yu = dict(y)
            yu["uncond"] = True
            unc = m(x, t, yu)
            guided = ref_shims.load_reference().ClassifierFreeSampleModel(m)(x, t, y)
        np.savez(os.path.join(OUT, "model_%s.npz" % case), t=t.numpy(), cond=cond.numpy(), uncond=unc.numpy(),
                 guided=guided.numpy(), meta=json.dumps(dict(cfg=cfg, batch=B, frames=x.shape[-1])))
        print("model_%s.npz" % case, tuple(cond.shape))


def chain_fixtures(only=None):
    ns = ref_shims.load_reference()
    for name, (case, resp, use_cfg, extra) in CHAIN_CASES.items():
        if only and name not in only:
            continue
        cfg, sd, m, x, y = build(case)
        dif = ref_shims.build_reference_diffusion(resp)
        model = ns.ClassifierFreeSampleModel(m) if use_cfg else m
        skip = extra.get("skip_timesteps", 0)
        n_exec = dif.num_timesteps - skip
        noises = make_step_noise(tuple(x.shape), n_exec)
        kwargs = dict(clip_denoised=False, model_kwargs={"y": dict(y)}, skip_timesteps=skip,
                      const_noise=extra.get("const_noise", False))
        g = torch.Generator().manual_seed(99)
        if extra.get("init_image"):
            kwargs["init_image"] = torch.randn(*x.shape, generator=g)
        if extra.get("inpaint"):
            mask = torch.rand(*x.shape, generator=g) < 0.3
            kwargs["model_kwargs"]["y"]["inpainting_mask"] = mask
            kwargs["model_kwargs"]["y"]["inpainted_motion"] = torch.randn(*x.shape, generator=g)
        trace = []
        with ref_shims.injected_noise(noises):
            for out in dif.p_sample_loop_progressive(model, tuple(x.shape), noise=x, **kwargs):
                trace.append((out["sample"].clone(), out["pred_xstart"].clone()))
        keep = range(len(trace)) if len(trace) <= 8 else [0, 1, len(trace) // 2, len(trace) - 1]
        arrays = dict(final=trace[-1][0].numpy(), trace_idx=np.array(list(keep)),
                      trace_sample=np.stack([trace[i][0].numpy() for i in keep]),
                      trace_pred=np.stack([trace[i][1].numpy() for i in keep]),
                      meta=json.dumps(dict(case=case, respacing=resp, cfg=use_cfg, extra=extra, n_exec=n_exec)))
        np.savez(os.path.join(OUT, "%s.npz" % name), **arrays)
        print("%s.npz" % name, "steps", len(trace))


# name -> (model case, respacing, cfg?, eta, clip_denoised)
DDIM_CASES = {
    "ddim_v1_small_cfg6_eta0": ("v1_small", "ddim6", True, 0.0, True),
    "ddim_v2_small_cfg5_eta05": ("v2_small", "ddim5", True, 0.5, False),
    "ddim_v2_small_uncond8_eta1": ("v2_small", [8], False, 1.0, True),
}
# name -> (model case, respacing, chunks, guidance)
TAKE_CASES = {"take_v2_small_cfg4x3": ("v2_small", [4], 3, 2.5), "take_v1_small_cfg3x2": ("v1_small", [3], 2, 2.5)}


def ddim_fixtures():
    """ddim_sample_loop_progressive of the reference (gaussian_diffusion.py:933-993) with injected noise."""
    ns = ref_shims.load_reference()
    for name, (case, resp, use_cfg, eta, clip) in DDIM_CASES.items():
        cfg, sd, m, x, y = build(case)
        dif = ref_shims.build_reference_diffusion(resp)
        model = ns.ClassifierFreeSampleModel(m) if use_cfg else m
        noises = make_step_noise(tuple(x.shape), dif.num_timesteps)
        trace = []
        with ref_shims.injected_noise(noises):
            for out in dif.ddim_sample_loop_progressive(model, tuple(x.shape), noise=x, clip_denoised=clip,
                                                        model_kwargs={"y": dict(y)}, eta=eta):
                trace.append((out["sample"].clone(), out["pred_xstart"].clone()))
        np.savez(os.path.join(OUT, "%s.npz" % name), final=trace[-1][0].numpy(),
                 trace_sample=np.stack([t[0].numpy() for t in trace]),
                 trace_pred=np.stack([t[1].numpy() for t in trace]),
                 meta=json.dumps(dict(case=case, respacing=resp, cfg=use_cfg, eta=eta, clip=clip,
                                      n_exec=dif.num_timesteps)))
        print("%s.npz" % name, "steps", len(trace))


def take_inputs(cfg, B, T, chunks):
    """Per-chunk conditioning, chunk-specific noise and dataset statistics of a synthetic take."""
    ys = [make_inputs(cfg, B, T, seed=500 + k)[1] for k in range(chunks)]
    x_T = [make_inputs(cfg, B, T, seed=600 + k)[0] for k in range(chunks)]
    g = torch.Generator().manual_seed(77)
    mean = torch.randn(cfg["njoints"], generator=g)
    std = torch.rand(cfg["njoints"], generator=g) + 0.5
    return ys, x_T, mean, std


def take_fixtures():
    """The chunk loop of sample/generate.py:91-189 run on the reference model / diffusion objects (the script's
    own data loading, argument parsing and rendering are not part of the path)."""
    from oracle.generate_oracle import oracle_generate_take
    ns = ref_shims.load_reference()
    for name, (case, resp, chunks, guidance) in TAKE_CASES.items():
        cfg, sd, m, x, _ = build(case)
        B, T = x.shape[0], x.shape[-1]
        dif = ref_shims.build_reference_diffusion(resp)
        model = ns.ClassifierFreeSampleModel(m)
        ys, x_T, mean, std = take_inputs(cfg, B, T, chunks)

        def sample_fn(y, k):
            noises = make_step_noise(tuple(x.shape), dif.num_timesteps, seed=4321 + k)
            with ref_shims.injected_noise(noises):
                return dif.p_sample_loop(model, (B, m.njoints, m.nfeats, T), clip_denoised=False,
                                         model_kwargs={"y": y}, skip_timesteps=0, init_image=None, progress=False,
                                         dump_steps=None, noise=x_T[k], const_noise=False)
        out = oracle_generate_take(sample_fn, ys, B, cfg["seed_poses"], guidance, mean, std)
        np.savez(os.path.join(OUT, "%s.npz" % name), sample=out["sample"].numpy(), motion=out["motion"],
                 motion_rot=out["motion_rot"],
                 meta=json.dumps(dict(case=case, respacing=resp, chunks=chunks, guidance=guidance,
                                      n_exec=dif.num_timesteps)))
        print("%s.npz" % name, out["motion"].shape)


if __name__ == "__main__":
    torch.manual_seed(0)
    if len(sys.argv) > 1:                       # python make_golden.py chain_x chain_y: only these chain fixtures
        chain_fixtures(only=set(sys.argv[1:]))
        sys.exit(0)
    schedule_fixture()
    model_fixtures()
    chain_fixtures()
    ddim_fixtures()
    take_fixtures()
