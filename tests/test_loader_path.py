"""the checkpoint / args.json loader path and the GENEA batch layout, against the live reference
where it is importable (this container) and against their documented layouts elsewhere."""
import argparse
import json
import os

import numpy as np
import pytest
import torch

from gesturediffusion_b200.data_loaders.tensors import gg_collate
from gesturediffusion_b200.synthetic import make_state_dict, model_config
from gesturediffusion_b200.utils import model_util
from oracle import ref_shims

ARGS_JSON = dict(dataset="genea2023", data_dir="", num_frames=120, arch="trans_enc", emb_trans_dec=False, layers=2,
                 latent_dim=128, cond_mask_prob=0.1, lambda_rcxyz=0.0, lambda_vel=0.0, lambda_fc=0.0,
                 unconstrained=False, use_text=False, use_audio=False, mfcc_input=True, use_wav_enc=False,
                 seed_poses=10, noise_schedule="cosine", diffusion_steps=1000, sigma_small=True, save_dir="x", lr=1e-4)


def _checkpoint(tmp_path, **changes):
    cfg = model_config(arch="v2", njoints=498, latent_dim=128, num_layers=2, num_heads=4, ff_size=1024)
    sd = make_state_dict(cfg, seed=3)
    path = os.path.join(tmp_path, "model000001.pt")
    torch.save(sd, path)
    with open(os.path.join(tmp_path, "args.json"), "w") as fh:
        json.dump({**ARGS_JSON, **changes}, fh)
    return path, cfg, sd


def _items(n, T=120, J=498, seed=0):
    r = np.random.RandomState(seed)
    return [(r.randn(T, J), "word %d" % i, T, r.randn(1600), r.randn(T, 26), r.randn(10, J)) for i in range(n)]


def test_load_checkpoint_builds_the_reference_objects(tmp_path):
    path, cfg, sd = _checkpoint(str(tmp_path))
    model, diffusion, args = model_util.load_checkpoint(path, device="cpu", precision="fp32", guidance_param=2.5)
    assert type(model).__name__ == "ClassifierFreeSampleModel" and type(model.model).__name__ == "MDM"
    m = model.model
    assert (m.njoints, m.nfeats, m.latent_dim, m.num_layers, m.num_heads, m.ff_size) == (498, 1, 128, 2, 4, 1024)
    assert m.data_rep == "genea_vec" and m.cond_mask_prob == 0.1 and not m.training
    assert diffusion.num_timesteps == 1000 and args.dataset == "genea2023" and args.guidance_param == 2.5
    got = m.state_dict()
    assert set(got) == set(sd) and all(torch.equal(got[k], sd[k]) for k in sd)
    # respacing knob + cond_mask_prob == 0 disables guidance (parser_util.py:31-32)
    path0, _, _ = _checkpoint(str(tmp_path), cond_mask_prob=0)
    model0, dif50, args0 = model_util.load_checkpoint(path0, device="cpu", timestep_respacing=[50])
    assert args0.guidance_param == 1 and type(model0).__name__ == "MDM" and dif50.num_timesteps == 50


def test_loader_failures_match_the_reference(tmp_path):
    path, cfg, sd = _checkpoint(str(tmp_path))
    os.remove(os.path.join(str(tmp_path), "args.json"))
    with pytest.raises(AssertionError, match="Arguments json file was not found"):
        model_util.load_args_from_model(path)
    args = argparse.Namespace(**{**ARGS_JSON, "dataset": "humanml"})
    with pytest.raises(ValueError, match="genea2022 and genea2023"):
        model_util.get_model_args(args)
    model, _ = model_util.create_model_and_diffusion(argparse.Namespace(**ARGS_JSON))
    with pytest.raises(AssertionError):                                   # unexpected key (model_util.py:8)
        model_util.load_model_wo_clip(model, {**sd, "bogus.weight": torch.zeros(1)})
    model_util.load_model_wo_clip(model, {k: v for k, v in sd.items()})   # complete state dict: fine


def test_gg_collate_layout():
    motion, cond = gg_collate(_items(3))
    y = cond["y"]
    assert tuple(motion.shape) == (3, 498, 1, 120) and motion.dtype == torch.float32
    assert tuple(y["mfcc"].shape) == (3, 26, 1, 120) and tuple(y["seed"].shape) == (3, 498, 1, 10)
    assert tuple(y["mask"].shape) == (3, 1, 1, 120) and bool(y["mask"].all()) and y["lengths"].tolist() == [120] * 3
    assert y["text"] == ["word 0", "word 1", "word 2"] and tuple(y["audio"].shape) == (3, 1600)


@pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference not present (GPU box)")
def test_against_live_reference(tmp_path):
    ref_shims.load_reference()
    import importlib
    ref_mu = importlib.import_module("utils.model_util")
    ref_t = importlib.import_module("data_loaders.tensors")
    args = argparse.Namespace(**ARGS_JSON)
    assert model_util.get_model_args(args) == ref_mu.get_model_args(args, None)
    path, cfg, sd = _checkpoint(str(tmp_path))
    ours, dif = model_util.create_model_and_diffusion(args)
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        theirs, ref_dif = ref_mu.create_model_and_diffusion(args, None)
    ref_mu.load_model_wo_clip(theirs, sd)
    model_util.load_model_wo_clip(ours, sd)
    a, b = ours.state_dict(), {k: v for k, v in theirs.state_dict().items() if not k.startswith("clip_model.")}
    assert set(a) == set(b) and all(torch.equal(a[k], b[k]) for k in a)
    assert dif.timestep_map == ref_dif.timestep_map and np.array_equal(dif.betas, ref_dif.betas)
    assert dif.model_var_type.name == ref_dif.model_var_type.name
    m1, c1 = gg_collate(_items(4, seed=5))
    m2, c2 = ref_t.gg_collate(_items(4, seed=5))
    assert torch.equal(m1, m2) and set(c1["y"]) == set(c2["y"])
    for k in c1["y"]:
        assert (torch.equal(c1["y"][k], c2["y"][k]) if torch.is_tensor(c1["y"][k]) else c1["y"][k] == c2["y"][k])
