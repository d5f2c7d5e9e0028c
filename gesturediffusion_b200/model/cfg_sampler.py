"""Drop-in for the reference's model/cfg_sampler.py `ClassifierFreeSampleModel` (:8-28).

forward(x, t, y) = uncond + y['scale'] * (cond - uncond).  Instead of two model calls plus a deepcopy of
every conditioning tensor per step, both branches are evaluated as ONE batch of 2B sequences (only the
conditioning token / coarse bias differs between them) and combined by the fused posterior kernel.
"""
import torch
import torch.nn as nn

from .. import kernels as K


class ClassifierFreeSampleModel(nn.Module):

    def __init__(self, model):
        super().__init__()
        self.model = model
        assert self.model.cond_mask_prob > 0, \
            'Cannot run a guided diffusion on a model that has not been trained with no conditions'
        self.rot2xyz = self.model.rot2xyz
        self.njoints = self.model.njoints
        self.nfeats = self.model.nfeats
        self.data_rep = self.model.data_rep

    def forward(self, x, timesteps, y=None):
        eng = self.model.engine()
        B, J, F_, T = x.shape
        from ..sampler import DeviceScope
        with torch.no_grad(), DeviceScope(eng.device):
            ws = eng.workspace(B, T, 2)
            eng.prepare_cond(ws, y, [bool(y.get('uncond', False)), True])
            xin = x.to(eng.device, torch.float32).contiguous()
            eng.load_x(ws, xin)
            ws["t"].copy_(timesteps.to(eng.device).long())
            eng.forward_step(ws)
            out = torch.empty(B, J * F_, 1, T, device=eng.device)
            scale = y['scale'].to(eng.device, torch.float32).reshape(-1).contiguous()
            ident = _identity_tables(eng.device)
            K.posterior_step(ws["OUT"][:B], xin, ident["coef"], ident["step"], B, J * F_, T,
                             uncond=ws["OUT"][B:], scale=scale, pred_xstart=out)
            return out


_ident = {}


def _identity_tables(device):
    """coef1 = 1, coef2 = 0, step 0 (no noise): turns the posterior kernel into the pure CFG combine."""
    key = str(device)
    if key not in _ident:
        coef = torch.zeros(6, 1, device=device)
        coef[0, 0] = 1.0
        _ident[key] = dict(coef=coef, step=torch.zeros(1, dtype=torch.int32, device=device))
    return _ident[key]
