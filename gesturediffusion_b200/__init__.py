"""gesturediffusion_b200 -- B200-native (sm_100a) denoising sampling loop of GestureDiffusion/MDM,
a drop-in behind the reference's Python API for that path:

    reference                                      this package
    diffusion.respace.SpacedDiffusion          ->  gesturediffusion_b200.diffusion.respace.SpacedDiffusion
    diffusion.respace.space_timesteps          ->  gesturediffusion_b200.diffusion.respace.space_timesteps
    diffusion.gaussian_diffusion (gd.*)        ->  gesturediffusion_b200.diffusion.gaussian_diffusion
    model.mdm.MDM / model.mdm_old.MDM_Old      ->  gesturediffusion_b200.model.mdm.MDM / .mdm_old.MDM_Old
    model.cfg_sampler.ClassifierFreeSampleModel->  gesturediffusion_b200.model.cfg_sampler.ClassifierFreeSampleModel
    utils.model_util (loader path)             ->  gesturediffusion_b200.utils.model_util
    data_loaders.tensors.gg_collate            ->  gesturediffusion_b200.data_loaders.tensors.gg_collate
    the chunk loop of sample/generate.py       ->  gesturediffusion_b200.generate.generate_take

Host code is Python/PyTorch (device memory, streams, torch.distributed); all arithmetic of the path
runs in hand-written CUDA kernels behind the C ABI in include/gd_b200.h.  No CPU fallback.
"""
from .diffusion import gaussian_diffusion  # noqa: F401
from .diffusion.respace import SpacedDiffusion, space_timesteps  # noqa: F401
from .model.cfg_sampler import ClassifierFreeSampleModel  # noqa: F401
from .model.mdm import MDM  # noqa: F401
from .model.mdm_old import MDM_Old  # noqa: F401
from .convert import accelerate, create_gaussian_diffusion  # noqa: F401

__all__ = ["gaussian_diffusion", "SpacedDiffusion", "space_timesteps", "ClassifierFreeSampleModel", "MDM",
           "MDM_Old", "accelerate", "create_gaussian_diffusion"]
