"""Import the UNMODIFIED reference from /root/reference (build container only).

Test infrastructure: used by tests/golden/make_golden.py to generate the golden
fixtures and by tests that are skipped when /root/reference is absent (the GPU
box).  Three in-process shims, no edits to the reference:
  1. ``np.float``/``np.int`` aliases (data_loaders/humanml/common/quaternion.py:13 is
     pulled in by diffusion/gaussian_diffusion.py:18);
  2. a stub ``clip`` module (model/mdm.py:5) -- OpenAI CLIP is not installed;
  3. a stub ``model.smpl`` (model/rotation2xyz.py:5) -- smplx/body models absent, yet
     MDM.__init__ builds Rotation2xyz -> SMPL() (model/mdm.py:103).
"""
import contextlib
import io
import os
import sys
import types

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_path.zip")


def _pick_root():
    """/root/reference in the build container; on the GPU box (no /root/reference) the archive of byte-identical
    copies of the same modules that oracle/stage_ref.py wrote to oracle/_ref/ (git-ignored, travels with the snapshot)."""
    env = os.environ.get("GD_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/diffusion"):
        return "/root/reference"
    return _STAGED


REF_ROOT = _pick_root()


def reference_available():
    if REF_ROOT.endswith(".zip"):
        return os.path.isfile(REF_ROOT)            # zipimport: the archive itself is the sys.path entry
    return os.path.isdir(os.path.join(REF_ROOT, "diffusion"))


def reference_kind():
    """'live' (the reference tree itself), 'staged' (oracle/_ref archive) or None."""
    if not reference_available():
        return None
    return "staged" if os.path.realpath(REF_ROOT) == os.path.realpath(_STAGED) else "live"


_loaded = {}


def load_reference():
    """Returns a namespace with the reference's hot-path classes."""
    if _loaded:
        return _loaded["ns"]
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    import numpy as np
    import torch

    if not hasattr(np, "float"):
        np.float = float
    if not hasattr(np, "int"):
        np.int = int

    clip = types.ModuleType("clip")
    clip.load = lambda *a, **k: (torch.nn.Identity(), None)
    clip.tokenize = lambda *a, **k: None
    clip.model = types.SimpleNamespace(convert_weights=lambda m: None)
    sys.modules.setdefault("clip", clip)

    smpl = types.ModuleType("model.smpl")

    class SMPL(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()

    smpl.SMPL = SMPL
    smpl.JOINTSTYPE_ROOT = {}
    sys.modules["model.smpl"] = smpl

    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    with contextlib.redirect_stdout(io.StringIO()):
        from diffusion import gaussian_diffusion as gd
        from diffusion.respace import SpacedDiffusion, space_timesteps
        from model.cfg_sampler import ClassifierFreeSampleModel
        from model.mdm import MDM
        from model.mdm_old import MDM_Old
        from model import local_attention
    ns = types.SimpleNamespace(gd=gd, SpacedDiffusion=SpacedDiffusion, space_timesteps=space_timesteps,
                               ClassifierFreeSampleModel=ClassifierFreeSampleModel, MDM=MDM,
                               MDM_Old=MDM_Old, local_attention=local_attention)
    _loaded["ns"] = ns
    return ns


def build_reference_model(arch, njoints, latent_dim=512, num_layers=8, num_heads=4, ff_size=1024,
                          seed_poses=10, use_text=False, cond_mask_prob=0.1):
    """Instantiate MDM (V2) / MDM_Old (V1) exactly as utils/model_util.py:18-34 would."""
    ns = load_reference()
    common = dict(njoints=njoints, nfeats=1, pose_rep="rot6d", data_rep="genea_vec",
                  latent_dim=latent_dim, ff_size=ff_size, num_layers=num_layers, num_heads=num_heads,
                  dropout=0.1, activation="gelu", dataset="genea2023", cond_mask_prob=cond_mask_prob,
                  seed_poses=seed_poses, mfcc_input=True, use_wav_enc=False, use_text=use_text,
                  clip_version="ViT-B/32")
    with contextlib.redirect_stdout(io.StringIO()):
        if arch == "v2":
            m = ns.MDM(**common)
            if use_text:
                m.encode_text = lambda raw: raw          # synthetic CLIP-like embedding passthrough
        else:
            m = ns.MDM_Old(translation=True, glob=True, glob_rot=True, **common)
    m.eval()        # returns None in the reference (mdm.py:273-275) -- statement form only
    return m


def build_reference_diffusion(respacing=None, schedule="cosine", steps=1000, sigma_small=True,
                              predict_xstart=True):
    """utils/model_util.py:37-71 with an explicit respacing."""
    ns = load_reference()
    gd = ns.gd
    betas = gd.get_named_beta_schedule(schedule, steps, 1.0)
    if not respacing:
        respacing = [steps]
    return ns.SpacedDiffusion(
        use_timesteps=ns.space_timesteps(steps, respacing), betas=betas,
        model_mean_type=gd.ModelMeanType.START_X if predict_xstart else gd.ModelMeanType.EPSILON,
        model_var_type=gd.ModelVarType.FIXED_SMALL if sigma_small else gd.ModelVarType.FIXED_LARGE,
        loss_type=gd.LossType.MSE, rescale_timesteps=False)


@contextlib.contextmanager
def injected_noise(noises):
    """Make the reference's th.randn_like (gaussian_diffusion.py:532) return our pre-drawn
    per-step noise tensors, in order."""
    ns = load_reference()
    th = ns.gd.th
    it = iter(noises)
    orig = th.randn_like
    th.randn_like = lambda x, *a, **k: next(it).to(x)
    try:
        yield
    finally:
        th.randn_like = orig
