"""CPU: host-side logic of the drop-in -- model parameter containers, error behaviour without a GPU,
batch sharding and the start/end collectives (gloo, world_size 2)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from helpers import build_product_model, load_model_case
from gesturediffusion_b200 import ClassifierFreeSampleModel, _cabi, create_gaussian_diffusion
from gesturediffusion_b200 import dist as gdist
from oracle import ref_shims


@pytest.mark.parametrize("case", ["v1_small", "v2_small", "v2_small_text"])
def test_state_dict_roundtrip(case):
    cfg, sd, x, y, gold = load_model_case(case)
    m = build_product_model(cfg, sd, "cpu", "fp32")
    out = m.state_dict()
    assert sorted(out.keys()) == sorted(sd.keys())
    for k in sd:
        assert torch.equal(out[k], sd[k]), k


@pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference not present (GPU box)")
@pytest.mark.parametrize("arch", ["v1", "v2"])
def test_same_seed_same_init_as_reference(arch):
    """Same constructor order => identical random init and key ORDER as the reference classes."""
    from gesturediffusion_b200 import MDM, MDM_Old
    torch.manual_seed(3)
    ref = ref_shims.build_reference_model(arch, 30, latent_dim=128, num_layers=2, ff_size=256)
    torch.manual_seed(3)
    kw = dict(njoints=30, nfeats=1, pose_rep="rot6d", data_rep="genea_vec", latent_dim=128, ff_size=256,
              num_layers=2, num_heads=4, dropout=0.1, activation="gelu", dataset="genea2023", cond_mask_prob=0.1,
              seed_poses=10, mfcc_input=True, use_wav_enc=False, use_text=False)
    ours = MDM(**kw) if arch == "v2" else MDM_Old(translation=True, glob=True, glob_rot=True, **kw)
    a, b = ref.state_dict(), ours.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_cpu_model_refuses_to_run():
    cfg, sd, x, y, gold = load_model_case("v1_small")
    m = build_product_model(cfg, sd, "cpu", "fp16")
    with pytest.raises(_cabi.GdError):
        m(x, gold["t"], y)
    d = create_gaussian_diffusion(timestep_respacing=[4])
    with pytest.raises(_cabi.GdError):
        d.p_sample_loop(ClassifierFreeSampleModel(m), tuple(x.shape), model_kwargs={"y": y}, clip_denoised=False)


def test_cfg_wrapper_contract():
    cfg, sd, x, y, gold = load_model_case("v1_small")
    m = build_product_model(cfg, sd, "cpu", "fp16")
    w = ClassifierFreeSampleModel(m)
    assert w.njoints == m.njoints and w.nfeats == m.nfeats and w.data_rep == m.data_rep and w.rot2xyz is m.rot2xyz
    assert len(list(w.parameters())) == len(list(m.parameters()))
    m.cond_mask_prob = 0.0
    with pytest.raises(AssertionError):
        ClassifierFreeSampleModel(m)
    assert m.eval() is m                      # callers use .eval() as a statement (generate.py:77)


def test_shard_ranges_cover_batch():
    for B in (1, 7, 8, 1024, 1027):
        for G in (1, 2, 3, 4, 8):
            spans = [gdist.shard_range(B, r, G) for r in range(G)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(G - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_shard_inputs_slices_batch_tensors_only():
    B = 5
    y = {"seed": torch.arange(B * 3.0).view(B, 3), "scale": torch.ones(B), "uncond": False,
         "text": ["a", "b", "c", "d", "e"], "table": torch.zeros(7, 2)}
    shape, kw, nz, base = gdist.shard_inputs((B, 4, 1, 10), {"y": y}, 1, 2, noise=torch.zeros(B, 4, 1, 10))
    assert shape == (2, 4, 1, 10) and base == 3 and nz.shape[0] == 2
    assert torch.equal(kw["y"]["seed"], y["seed"][3:5]) and kw["y"]["text"] == ["d", "e"]
    assert kw["y"]["table"].shape == (7, 2) and kw["y"]["uncond"] is False


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    r, w = gdist.init_process_group("gloo")
    assert (r, w) == (rank, world)
    # start: weights + seed come from rank 0
    arena = torch.full((1000,), float(rank + 1))
    gdist.broadcast_tensors([arena], src=0)
    assert bool((arena == 1.0).all())
    assert gdist.broadcast_seed(1234 + rank, "cpu") == 1234
    # end: uneven shards (B = 5 over 2 ranks) are gathered in global order on every rank
    B = 5
    full = torch.arange(B * 2 * 1 * 3.0).view(B, 2, 1, 3)
    lo, hi = gdist.shard_range(B, rank, world)
    out = gdist.gather_motions(full[lo:hi].clone(), B)
    assert torch.equal(out, full)
    torch.save(out, os.path.join(tmp, "r%d.pt" % rank))
    torch.distributed.destroy_process_group()


def test_gloo_world2_broadcast_and_gather(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(a, b)
