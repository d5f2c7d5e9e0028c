"""Shared machinery of the engine-backed denoisers (MDM, MDM_Old).

The nn.Module objects only *hold* parameters under the reference's state_dict key names (so
`load_state_dict` of a reference checkpoint works, model_util.py:6-9); their torch `forward`s are
never used.  `forward(x, timesteps, y)` packs the current parameters into a DenoiserEngine for the
parameters' CUDA device and runs the sm_100a kernels.  There is no CPU path: calling forward on a
CPU model raises.
"""
import math

import numpy as np
import torch
import torch.nn as nn

from .. import kernels as K
from ..engine import DenoiserEngine


class PositionalEncoding(nn.Module):
    """Sinusoid table `pe` [max_len, 1, d] (reference model/mdm.py:277-294); only the buffer is used."""

    def __init__(self, d_model, dropout=0.1, max_len=5000):
        super().__init__()
        self.dropout = nn.Dropout(p=dropout)
        pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
        div = torch.exp(torch.arange(0, d_model, 2).float() * (-np.log(10000.0) / d_model))
        pe = torch.zeros(max_len, d_model)
        pe[:, 0::2] = torch.sin(pos * div)
        pe[:, 1::2] = torch.cos(pos * div)
        self.register_buffer('pe', pe.unsqueeze(0).transpose(0, 1))


class TimestepEmbedder(nn.Module):
    def __init__(self, latent_dim, sequence_pos_encoder):
        super().__init__()
        self.latent_dim = latent_dim
        self.sequence_pos_encoder = sequence_pos_encoder
        self.time_embed = nn.Sequential(nn.Linear(latent_dim, latent_dim), nn.SiLU(),
                                        nn.Linear(latent_dim, latent_dim))


class InputProcess(nn.Module):
    def __init__(self, data_rep, input_feats, latent_dim):
        super().__init__()
        self.data_rep, self.input_feats, self.latent_dim = data_rep, input_feats, latent_dim
        self.poseEmbedding = nn.Linear(input_feats, latent_dim)
        if data_rep == 'rot_vel':
            self.velEmbedding = nn.Linear(input_feats, latent_dim)


class OutputProcess(nn.Module):
    def __init__(self, data_rep, input_feats, latent_dim, njoints, nfeats):
        super().__init__()
        self.data_rep, self.input_feats, self.latent_dim = data_rep, input_feats, latent_dim
        self.njoints, self.nfeats = njoints, nfeats
        self.poseFinal = nn.Linear(latent_dim, input_feats)
        if data_rep == 'rot_vel':
            self.velFinal = nn.Linear(latent_dim, input_feats)


class SeedPoseEncoder(nn.Module):
    def __init__(self, njoints, seed_poses, latent_dim):
        super().__init__()
        self.njoints, self.seed_poses, self.latent_dim = njoints, seed_poses, latent_dim
        self.seed_embed = nn.Linear(njoints * seed_poses, latent_dim)


class SinusoidalEmbeddings(nn.Module):
    """Holds `inv_freq` (reference model/local_attention.py:43-48)."""

    def __init__(self, dim):
        super().__init__()
        self.register_buffer('inv_freq', 1. / (10000 ** (torch.arange(0, dim, 2).float() / dim)))


class Rotation2xyz:
    """Post-processing hook kept for interface parity (reference model/rotation2xyz.py:10-23): the
    gesture pipeline only ever uses the identity branch; SMPL forward kinematics is out of scope."""

    def __init__(self, device='cpu', dataset='amass'):
        self.device, self.dataset = device, dataset
        self.smpl_model = nn.Module()

    def __call__(self, x, mask=None, pose_rep="xyz", **kwargs):
        if pose_rep == "xyz":
            return x
        raise NotImplementedError


def make_encoder(latent_dim, num_heads, ff_size, dropout, activation, num_layers):
    """Parameter container with nn.TransformerEncoder's key names and default init."""
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        layer = nn.TransformerEncoderLayer(d_model=latent_dim, nhead=num_heads, dim_feedforward=ff_size,
                                           dropout=dropout, activation=activation)
        return nn.TransformerEncoder(layer, num_layers=num_layers)


class EngineBackedDenoiser(nn.Module):
    """forward()/engine plumbing common to both denoisers."""

    arch = None
    precision = "fp16"          # "fp16" / "bf16" (tcgen05 tensor cores) or "fp32" (SIMT parity mode)

    def _engine_cfg(self):
        return dict(arch=self.arch, njoints=self.input_feats, latent_dim=self.latent_dim,
                    num_layers=self.num_layers, num_heads=self.num_heads, ff_size=self.ff_size,
                    seed_poses=self.seed_poses, use_text=getattr(self, "use_text", False),
                    text_dim=getattr(self, "text_dim", 64), clip_dim=getattr(self, "clip_dim", 512),
                    cl_head=getattr(self, "cl_head", 8))

    def _param_signature(self):
        dev = next(self.parameters()).device
        ver = 0
        for p in self.parameters():
            ver += p._version
        return (str(dev), self.precision, ver)

    def engine(self):
        """The DenoiserEngine for the current parameters (rebuilt when they move or change)."""
        sig = self._param_signature()
        if getattr(self, "_engine_sig", None) != sig:
            dev = next(self.parameters()).device
            if dev.type != "cuda":
                raise K._cabi.GdError("gesturediffusion_b200 models run on CUDA only: call .to('cuda') first "
                                      "(there is no CPU fallback)")
            sd = {k: v for k, v in self.state_dict().items() if not k.startswith("clip_model.")}
            from ..sampler import DeviceScope
            with DeviceScope(dev):                          # packing launches set-up GEMMs on the model's GPU
                self.__dict__["_engine"] = DenoiserEngine(sd, self._engine_cfg(), dev, self.precision)
            self.__dict__["_engine_sig"] = sig
        return self.__dict__["_engine"]

    def set_precision(self, precision):
        self.precision = precision
        return self

    def parameters_wo_clip(self):
        return [p for name, p in self.named_parameters() if not name.startswith('clip_model.')]

    def mask_cond(self, cond, force_mask=False):
        if force_mask:
            return torch.zeros_like(cond)
        return cond

    def forward(self, x, timesteps, y=None):
        """x [B, njoints, nfeats, frames] (x_t), timesteps [B] int (original schedule), y dict with
        'seed' [B,J,1,seed_poses], 'mfcc' [B,26,1,T], optional 'uncond', 'text'.  Returns [B,J,1,T]."""
        eng = self.engine()
        B, J, F_, T = x.shape
        if J * F_ != self.input_feats:
            raise ValueError("x must be [B, %d, 1, T]" % self.input_feats)
        uncond = bool(y.get('uncond', False))
        from ..sampler import DeviceScope
        with torch.no_grad(), DeviceScope(eng.device):
            ws = eng.workspace(B, T, 1)
            eng.prepare_cond(ws, y, [uncond])
            eng.load_x(ws, x.to(eng.device, torch.float32).contiguous())
            ws["t"].copy_(timesteps.to(eng.device).long())
            eng.forward_step(ws)
            return ws["OUT"].view(B, J, 1, T).clone()
