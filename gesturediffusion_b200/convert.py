"""Switching an existing reference pipeline over: build the B200 objects from reference instances.

`accelerate(ref_model, ref_diffusion)` reads the configuration and weights off objects created by the
reference's own factory (utils/model_util.py:12-71: `create_model_and_diffusion`, `load_model_wo_clip`)
and returns engine-backed equivalents with the same call signatures, so sample/generate.py:74-130 keeps
working unchanged.  `create_gaussian_diffusion` mirrors utils/model_util.py:37-71 with an explicit
`timestep_respacing`.
"""
import torch

from .diffusion import gaussian_diffusion as gd
from .diffusion.respace import SpacedDiffusion, space_timesteps
from .model.cfg_sampler import ClassifierFreeSampleModel
from .model.mdm import MDM
from .model.mdm_old import MDM_Old


def create_gaussian_diffusion(noise_schedule="cosine", steps=1000, timestep_respacing="", sigma_small=True,
                              predict_xstart=True, rescale_timesteps=False, **loss_kwargs):
    betas = gd.get_named_beta_schedule(noise_schedule, steps, 1.)
    if not timestep_respacing:
        timestep_respacing = [steps]
    return SpacedDiffusion(
        use_timesteps=space_timesteps(steps, timestep_respacing), betas=betas,
        model_mean_type=gd.ModelMeanType.START_X if predict_xstart else gd.ModelMeanType.EPSILON,
        model_var_type=gd.ModelVarType.FIXED_SMALL if sigma_small else gd.ModelVarType.FIXED_LARGE,
        loss_type=gd.LossType.MSE, rescale_timesteps=rescale_timesteps, **loss_kwargs)


def _convert_model(ref, device, precision):
    inner = getattr(ref, "model", None) if type(ref).__name__ == "ClassifierFreeSampleModel" else None
    if inner is not None:
        return ClassifierFreeSampleModel(_convert_model(inner, device, precision))
    name = type(ref).__name__
    if name not in ("MDM", "MDM_Old"):
        raise TypeError("cannot accelerate a %s" % name)
    common = dict(njoints=ref.njoints, nfeats=ref.nfeats, pose_rep=ref.pose_rep, data_rep=ref.data_rep,
                  latent_dim=ref.latent_dim, ff_size=ref.ff_size, num_layers=ref.num_layers,
                  num_heads=ref.num_heads, dropout=0.1, activation=ref.activation, dataset=ref.dataset,
                  cond_mask_prob=ref.cond_mask_prob, seed_poses=ref.seed_poses)
    if name == "MDM":
        new = MDM(text_dim=ref.text_dim, clip_dim=ref.clip_dim, use_text=ref.use_text,
                  mfcc_input=ref.mfcc_input, use_wav_enc=ref.use_wav_enc, **common)
    elif name == "MDM_Old":
        new = MDM_Old(translation=ref.translation, glob=ref.glob, glob_rot=ref.glob_rot, **common)
    else:
        raise TypeError("cannot accelerate a %s" % name)
    sd = {k: v for k, v in ref.state_dict().items() if not k.startswith("clip_model.")}
    missing, unexpected = new.load_state_dict(sd, strict=False)
    assert len(unexpected) == 0 and len(missing) == 0, (missing, unexpected)
    new.precision = precision
    new.eval()
    return new.to(device)


def _convert_diffusion(ref):
    kw = dict(betas=None, model_mean_type=gd.ModelMeanType[ref.model_mean_type.name],
              model_var_type=gd.ModelVarType[ref.model_var_type.name], loss_type=gd.LossType[ref.loss_type.name],
              rescale_timesteps=ref.rescale_timesteps)
    if hasattr(ref, "timestep_map"):
        # The reference object already holds the re-derived betas: initialise the tables from them
        # directly (bit-identical) and carry the index map across.
        new = SpacedDiffusion.__new__(SpacedDiffusion)
        gd.GaussianDiffusion.__init__(new, **{**kw, "betas": ref.betas})
        new.timestep_map = list(ref.timestep_map)
        new.use_timesteps = set(ref.use_timesteps)
        new.original_num_steps = ref.original_num_steps
        return new
    return gd.GaussianDiffusion(**{**kw, "betas": ref.betas})


def accelerate(ref_model, ref_diffusion=None, device="cuda", precision="fp16"):
    """Returns (model, diffusion) engine-backed equivalents of reference instances."""
    model = _convert_model(ref_model, torch.device(device), precision)
    if ref_diffusion is None:
        return model
    return model, _convert_diffusion(ref_diffusion)
