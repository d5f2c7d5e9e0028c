"""Shared test helpers: golden-fixture access and seeded reconstruction of weights / inputs."""
import json
import os

import numpy as np
import torch

from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, make_step_noise, model_config

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

MODEL_CASES = ["v1_small", "v2_small", "v2_small_text", "v1_humanml", "v1_kit", "v2_gesture", "v2_text"]
SMALL_MODEL_CASES = ["v1_small", "v2_small", "v2_small_text"]
CHAIN_CASES = ["chain_v1_small_cfg5", "chain_v2_small_cfg5", "chain_v1_small_uncond4_skip",
               "chain_v2_small_const_noise", "chain_v1_small_inpaint", "chain_v1_humanml_uncond50",
               "chain_v2_gesture_cfg50", "chain_v1_small_cfg1000", "chain_v2_small_cfg1000"]


def record_parity(line):
    """Print a measured parity error and append it to the artefact the round keeps (profiles/r02_parity.txt is a copy
    of gpurun_out/parity.txt from the GPU box; GD_PARITY_LOG overrides the path, empty string turns it off)."""
    print(line)
    path = os.environ.get("GD_PARITY_LOG")
    if path is None:
        out = os.path.join(os.path.dirname(os.path.dirname(GOLDEN)), "gpurun_out")
        path = os.path.join(out, "parity.txt") if os.path.isdir(out) else ""
    if path:
        with open(path, "a") as fh:
            fh.write(line + "\n")


def rel_l2(a, b):
    a, b = a.double().cpu().reshape(-1), b.double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm())


def load_model_case(name):
    g = np.load(os.path.join(GOLDEN, "model_%s.npz" % name))
    meta = json.loads(str(g["meta"]))
    cfg = model_config(**{k: meta["cfg"][k] for k in ("arch", "njoints", "latent_dim", "num_layers", "num_heads",
                                                      "ff_size", "seed_poses", "use_text")})
    sd = make_state_dict(cfg, seed=0)
    x, y = make_inputs(cfg, meta["batch"], meta["frames"])
    gold = {k: torch.from_numpy(g[k]) for k in ("t", "cond", "uncond", "guided")}
    return cfg, sd, x, y, gold


def load_chain_case(name):
    g = np.load(os.path.join(GOLDEN, "%s.npz" % name))
    meta = json.loads(str(g["meta"]))
    cfg, sd, x, y, _ = load_model_case(meta["case"])
    extra = meta["extra"]
    noises = make_step_noise(tuple(x.shape), meta["n_exec"])
    gen = torch.Generator().manual_seed(99)
    kwargs = dict(skip_timesteps=extra.get("skip_timesteps", 0), const_noise=extra.get("const_noise", False),
                  init_image=None)
    y = dict(y)
    if extra.get("init_image"):
        kwargs["init_image"] = torch.randn(*x.shape, generator=gen)
    if extra.get("inpaint"):
        y["inpainting_mask"] = torch.rand(*x.shape, generator=gen) < 0.3
        y["inpainted_motion"] = torch.randn(*x.shape, generator=gen)
    gold = dict(final=torch.from_numpy(g["final"]), trace_idx=g["trace_idx"].tolist(),
                trace_sample=torch.from_numpy(g["trace_sample"]), trace_pred=torch.from_numpy(g["trace_pred"]))
    return cfg, sd, x, y, noises, meta, kwargs, gold


def build_product_model(cfg, sd, device, precision):
    """The package's drop-in model class for `cfg`, loaded with `sd`, on `device`."""
    from gesturediffusion_b200.synthetic import build_model
    return build_model(cfg, sd, device, precision)


def to_dev(y, device):
    return {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in y.items()}


DDIM_CASES = ["ddim_v1_small_cfg6_eta0", "ddim_v2_small_cfg5_eta05", "ddim_v2_small_uncond8_eta1"]
TAKE_CASES = ["take_v2_small_cfg4x3", "take_v1_small_cfg3x2"]


def load_ddim_case(name):
    g = np.load(os.path.join(GOLDEN, "%s.npz" % name))
    meta = json.loads(str(g["meta"]))
    cfg, sd, x, y, _ = load_model_case(meta["case"])
    noises = make_step_noise(tuple(x.shape), meta["n_exec"])
    gold = dict(final=torch.from_numpy(g["final"]), trace_sample=torch.from_numpy(g["trace_sample"]),
                trace_pred=torch.from_numpy(g["trace_pred"]))
    return cfg, sd, x, y, noises, meta, gold


def take_inputs(cfg, B, T, chunks):
    """Same construction as tests/golden/make_golden.py::take_inputs."""
    ys = [make_inputs(cfg, B, T, seed=500 + k)[1] for k in range(chunks)]
    x_T = [make_inputs(cfg, B, T, seed=600 + k)[0] for k in range(chunks)]
    g = torch.Generator().manual_seed(77)
    mean = torch.randn(cfg["njoints"], generator=g)
    std = torch.rand(cfg["njoints"], generator=g) + 0.5
    return ys, x_T, mean, std


def load_take_case(name):
    g = np.load(os.path.join(GOLDEN, "%s.npz" % name))
    meta = json.loads(str(g["meta"]))
    cfg, sd, x, _, _ = load_model_case(meta["case"])
    B, T = x.shape[0], x.shape[-1]
    ys, x_T, mean, std = take_inputs(cfg, B, T, meta["chunks"])
    step_noise = [make_step_noise(tuple(x.shape), meta["n_exec"], seed=4321 + k) for k in range(meta["chunks"])]
    gold = {k: torch.from_numpy(g[k]) for k in ("sample", "motion", "motion_rot")}
    return cfg, sd, B, T, ys, x_T, mean, std, step_noise, meta, gold
