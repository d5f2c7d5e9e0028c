"""Checkpoint / `args.json` loader path of the reference (utils/model_util.py:6-71, utils/parser_util.py:7-33)
for the package's engine-backed objects, so a trained checkpoint directory runs through the B200 path with the
calls sample/generate.py makes:

    args = load_args_from_model(model_path, guidance_param=2.5)          # parser_util.parse_and_load_from_model
    model, diffusion = create_model_and_diffusion(args, data=None)       # model_util.py:12-15
    load_model_wo_clip(model, torch.load(model_path, map_location="cpu"))  # model_util.py:6-9
    model = ClassifierFreeSampleModel(model) if args.guidance_param != 1 else model
    model.to(device); model.eval()

Names, argument meaning and failure behaviour follow the reference; argparse itself (command line) stays out.
"""
import argparse
import json
import os

from ..convert import create_gaussian_diffusion as _create_diffusion
from ..model.mdm import MDM

# the defaults of the 'dataset', 'model' and 'diffusion' argument groups (parser_util.py:62-108): the keys that
# parse_and_load_from_model overwrites from args.json
GROUP_DEFAULTS = {
    "dataset": dict(dataset="humanml", data_dir="", num_frames=120),
    "model": dict(arch="trans_enc", emb_trans_dec=False, layers=8, latent_dim=256, cond_mask_prob=.1,
                  lambda_rcxyz=0.0, lambda_vel=0.0, lambda_fc=0.0, unconstrained=False, use_text=False,
                  use_audio=False, mfcc_input=False, use_wav_enc=False, seed_poses=10),
    "diffusion": dict(noise_schedule="cosine", diffusion_steps=1000, sigma_small=True),
}


def load_model_wo_clip(model, state_dict):
    """utils/model_util.py:6-9."""
    missing_keys, unexpected_keys = model.load_state_dict(state_dict, strict=False)
    assert len(unexpected_keys) == 0
    assert all([k.startswith('clip_model.') for k in missing_keys])


def get_model_args(args, data=None):
    """utils/model_util.py:18-34 (only the GENEA datasets define the layout there)."""
    if args.dataset not in ['genea2022', 'genea2023']:
        raise ValueError("dataset %r: the reference only defines njoints / data_rep for genea2022 and genea2023 "
                         "(utils/model_util.py:23-26)" % (args.dataset,))
    return {'modeltype': '', 'njoints': 498, 'nfeats': 1,
            'translation': True, 'pose_rep': 'rot6d', 'glob': True, 'glob_rot': True,
            'latent_dim': args.latent_dim, 'ff_size': 1024, 'num_layers': args.layers, 'num_heads': 4,
            'dropout': 0.1, 'activation': "gelu", 'data_rep': 'genea_vec', 'cond_mask_prob': args.cond_mask_prob,
            'clip_version': 'ViT-B/32', 'dataset': args.dataset,
            'use_text': args.use_text, 'mfcc_input': args.mfcc_input,
            'use_wav_enc': args.use_wav_enc, 'seed_poses': args.seed_poses, 'use_audio': args.use_audio}


def create_gaussian_diffusion(args, timestep_respacing=''):
    """utils/model_util.py:37-71; `timestep_respacing` is the knob the reference hard-codes to '' (:42)."""
    return _create_diffusion(noise_schedule=args.noise_schedule, steps=1000, timestep_respacing=timestep_respacing,
                             sigma_small=args.sigma_small, predict_xstart=True, rescale_timesteps=False,
                             lambda_vel=args.lambda_vel, lambda_rcxyz=args.lambda_rcxyz, lambda_fc=args.lambda_fc)


def create_model_and_diffusion(args, data=None, timestep_respacing=''):
    """utils/model_util.py:12-15."""
    kw = get_model_args(args, data)
    for k in ('modeltype', 'translation', 'glob', 'glob_rot', 'clip_version', 'use_audio'):
        kw.pop(k)                              # accepted-and-ignored by the reference's MDM (mdm.py:15-19 **kargs)
    model = MDM(**kw)
    return model, create_gaussian_diffusion(args, timestep_respacing)


def load_args_from_model(model_path, **overrides):
    """parser_util.parse_and_load_from_model (:7-33) without argparse: defaults of the dataset / model / diffusion
    groups, overwritten by `args.json` next to the checkpoint; other options (guidance_param, batch_size, ...) come
    from `overrides`.  cond_mask_prob == 0 forces guidance_param = 1 (:31-32)."""
    args_path = os.path.join(os.path.dirname(model_path), 'args.json')
    assert os.path.exists(args_path), 'Arguments json file was not found!'
    with open(args_path, 'r') as fr:
        model_args = json.load(fr)
    values = dict(model_path=model_path, guidance_param=2.5, seed=10, batch_size=64)
    values.update(overrides)
    for group in GROUP_DEFAULTS.values():
        for a, default in group.items():
            if a in model_args:
                values[a] = model_args[a]
            else:
                print('Warning: was not able to load [{}], using default value [{}] instead.'.format(a, default))
                values[a] = default
    args = argparse.Namespace(**values)
    if args.cond_mask_prob == 0:
        args.guidance_param = 1
    return args


def load_checkpoint(model_path, device="cuda", precision="fp16", timestep_respacing='', **overrides):
    """The object construction of sample/generate.py:62-75 in one call: (model, diffusion, args), the model wrapped
    for classifier-free guidance when args.guidance_param != 1, on `device`, in eval mode."""
    import torch
    from ..model.cfg_sampler import ClassifierFreeSampleModel
    args = load_args_from_model(model_path, **overrides)
    model, diffusion = create_model_and_diffusion(args, None, timestep_respacing)
    state_dict = torch.load(model_path, map_location='cpu')
    load_model_wo_clip(model, state_dict)
    model.precision = precision
    if args.guidance_param != 1:
        model = ClassifierFreeSampleModel(model)
    model.to(device)
    model.eval()
    return model, diffusion, args
