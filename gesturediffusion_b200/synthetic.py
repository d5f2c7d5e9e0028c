"""Seeded synthetic weights and inputs in the reference's formats.

There is no network for checkpoints or datasets, so benchmarks and parity tests use
random-init weights of the reference architecture and synthetic inputs of the dataset
shapes.  ``make_state_dict`` emits exactly the ``state_dict`` key set of
the reference modules (model/mdm.py:11-103, model/mdm_old.py:7-67) so that the very same
dict can be loaded into the reference (``load_state_dict``), into the oracle, and into the
B200 engine; generation is deterministic in (config, seed) on any machine with this torch.
"""
import math

import torch

MFCC_DIM = 26          # model/mdm.py:57, model/mdm_old.py:35
PE_MAX_LEN = 5000      # model/mdm.py:278


def sinusoid_pe(d_model, max_len=PE_MAX_LEN):
    """The 'pe' buffer of PositionalEncoding (model/mdm.py:282-289), shape [max_len,1,d]."""
    pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
    pe = torch.zeros(max_len, d_model)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.unsqueeze(1)


def model_config(arch="v2", njoints=498, latent_dim=512, num_layers=8, num_heads=4, ff_size=1024,
                 seed_poses=10, use_text=False, text_dim=64, clip_dim=512, cl_head=8):
    return dict(arch=arch, njoints=njoints, latent_dim=latent_dim, num_layers=num_layers,
                num_heads=num_heads, ff_size=ff_size, seed_poses=seed_poses, use_text=use_text,
                text_dim=text_dim, clip_dim=clip_dim, cl_head=cl_head)


def make_state_dict(cfg, seed=0):
    g = torch.Generator().manual_seed(seed)
    d, J, F = cfg["latent_dim"], cfg["njoints"], cfg["ff_size"]
    sd = {}

    def linear(name, n_out, n_in):
        bound = 1.0 / math.sqrt(n_in)
        sd[name + ".weight"] = (torch.rand(n_out, n_in, generator=g) * 2 - 1) * bound
        sd[name + ".bias"] = (torch.rand(n_out, generator=g) * 2 - 1) * bound

    pe = sinusoid_pe(d)
    sd["sequence_pos_encoder.pe"] = pe
    sd["embed_timestep.sequence_pos_encoder.pe"] = pe.clone()
    linear("embed_timestep.time_embed.0", d, d)
    linear("embed_timestep.time_embed.2", d, d)
    if cfg["arch"] == "v2":
        seed_out = d - cfg["text_dim"] if cfg["use_text"] else d
        if cfg["use_text"]:
            linear("embed_text", cfg["text_dim"], cfg["clip_dim"])
        linear("seed_pose_encoder.seed_embed", seed_out, J * cfg["seed_poses"])
        linear("input_process.poseEmbedding", d, J)
        linear("project_to_lat", d, 2 * d + MFCC_DIM)
        e = d // cfg["cl_head"]
        sd["rel_pos.inv_freq"] = 1.0 / (10000 ** (torch.arange(0, e, 2).float() / e))
    else:
        linear("seed_pose_encoder.seed_embed", d, J * cfg["seed_poses"])
        linear("input_process.poseEmbedding", d, J + MFCC_DIM)
    for l in range(cfg["num_layers"]):
        p = "seqTransEncoder.layers.%d." % l
        bound = math.sqrt(6.0 / (d + 3 * d))          # xavier_uniform on [3d, d]
        sd[p + "self_attn.in_proj_weight"] = (torch.rand(3 * d, d, generator=g) * 2 - 1) * bound
        sd[p + "self_attn.in_proj_bias"] = 0.02 * torch.randn(3 * d, generator=g)
        linear(p + "self_attn.out_proj", d, d)
        linear(p + "linear1", F, d)
        linear(p + "linear2", d, F)
        for nm in ("norm1", "norm2"):
            sd[p + nm + ".weight"] = 1.0 + 0.1 * torch.randn(d, generator=g)
            sd[p + nm + ".bias"] = 0.1 * torch.randn(d, generator=g)
    linear("output_process.poseFinal", J, d)
    return sd


def make_inputs(cfg, batch, frames, seed=1234, guidance=2.5, device="cpu"):
    """x_T, y dict in the layouts of data_loaders/tensors.py:56-66 (synthetic, z-normal)."""
    g = torch.Generator().manual_seed(seed)
    J = cfg["njoints"]
    x = torch.randn(batch, J, 1, frames, generator=g)
    y = {
        "seed": torch.randn(batch, J, 1, cfg["seed_poses"], generator=g),
        "mfcc": torch.randn(batch, MFCC_DIM, 1, frames, generator=g),
        "scale": torch.full((batch,), float(guidance)),
    }
    if cfg.get("use_text", False):
        y["text"] = torch.randn(batch, cfg["clip_dim"], generator=g)
    if device != "cpu":
        x = x.to(device)
        y = {k: v.to(device) for k, v in y.items()}
    return x, y


def make_step_noise(shape, n_steps, seed=4321):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(*shape, generator=g) for _ in range(n_steps)]


def build_model(cfg, state_dict, device, precision="fp16"):
    """The package's drop-in denoiser class for `cfg`, loaded with `state_dict`, on `device`."""
    from .model.mdm import MDM
    from .model.mdm_old import MDM_Old
    common = dict(njoints=cfg["njoints"], nfeats=1, pose_rep="rot6d", data_rep="genea_vec",
                  latent_dim=cfg["latent_dim"], ff_size=cfg["ff_size"], num_layers=cfg["num_layers"],
                  num_heads=cfg["num_heads"], dropout=0.1, activation="gelu", dataset="genea2023",
                  cond_mask_prob=0.1, seed_poses=cfg["seed_poses"], mfcc_input=True, use_wav_enc=False,
                  use_text=cfg["use_text"])
    if cfg["arch"] == "v2":
        m = MDM(**common)
    else:
        m = MDM_Old(translation=True, glob=True, glob_rot=True, **common)
    missing, unexpected = m.load_state_dict(state_dict, strict=False)
    assert not missing and not unexpected, (missing, unexpected)
    m.precision = precision
    m.eval()
    return m.to(device)
