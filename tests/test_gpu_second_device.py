"""GPU: a model on a GPU that is NOT the caller's current CUDA device (the C ABI launches on torch's current
stream and the library's one-off set-up is per device).  Needs two GPUs; skipped on a one-GPU box."""
import pytest
import torch

from helpers import build_product_model, load_chain_case, record_parity, rel_l2, to_dev

pytestmark = pytest.mark.gpu


def test_model_on_second_gpu_while_first_is_current(cuda_device):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    cfg, sd, x, y, noises, meta, kwargs, gold = load_chain_case("chain_v2_small_cfg5")
    dev1 = torch.device("cuda:1")
    torch.cuda.set_device(0)
    outs = {}
    for dev in (cuda_device, dev1):                         # same process, both GPUs, current device stays cuda:0
        model = ClassifierFreeSampleModel(build_product_model(cfg, sd, dev, "fp16"))
        dif = create_gaussian_diffusion(timestep_respacing=meta["respacing"])
        prog = []
        for step_out in dif.p_sample_loop_progressive(model, tuple(x.shape), noise=x.to(dev), clip_denoised=False,
                                                      model_kwargs={"y": to_dev(y, dev)}, step_noise=noises):
            assert torch.cuda.current_device() == 0          # the consumer keeps its own current device between steps
            prog.append(step_out["sample"])
        assert prog[-1].device == dev
        outs[str(dev)] = prog[-1].cpu()
        # eager single evaluation and p_mean_variance on that device too
        t = torch.full((x.shape[0],), 3, dtype=torch.long, device=dev)
        pm = dif.p_mean_variance(model, x.to(dev), t, clip_denoised=False, model_kwargs={"y": to_dev(y, dev)})
        assert pm["mean"].device == dev and bool(torch.isfinite(pm["mean"]).all())
        assert torch.cuda.current_device() == 0
    assert torch.equal(outs["cuda:0"], outs["cuda:1"])       # same kernels, same inputs: same bits on both GPUs
    err = rel_l2(outs["cuda:1"], gold["final"])
    record_parity("second device chain_v2_small_cfg5 fp16: final %.2e" % err)
    assert err < 1e-2
