"""GPU: parity of the PRODUCTION schedule against the CPU oracle at BASELINE batch sizes, and the shapes the golden
fixtures do not reach.

* one denoise step of cfg2a (V1 263 x 196, batch 64, CFG) and cfg3 (V2 498 x 120, batch 128, CFG): the one-kernel
  feed-forward path is engaged by SIZE (>= 4096 token rows), 197 / 242 row tiles over 74 clusters, exactly what bench.py
  times -- against oracle/ (float64) on the same seeded inputs; 2e-3 per-step relative L2 in fp16, 1e-5 in fp32
  (model/cfg_sampler.py:23-28, diffusion/gaussian_diffusion.py:277-388);
* guided FastLoop chains whose B*J*T is not a multiple of 4 (the uncond half of OUT is then only 4-byte aligned:
  the float4 path of the posterior kernel must not be taken).
"""
import pytest
import torch

from helpers import build_product_model, record_parity, rel_l2, to_dev

pytestmark = pytest.mark.gpu


def _oracle_step(cfg, sd, x, y, t, respacing, dtype):
    from oracle.diffusion_oracle import OracleDiffusion, oracle_cfg
    from oracle.model_oracle import make_model_fn
    fn = oracle_cfg(make_model_fn(sd, cfg, dtype))
    xd = x.to(dtype)
    return OracleDiffusion(respacing=respacing).p_mean_variance(fn, xd, t, y)


@pytest.mark.parametrize("name,arch,J,T,B", [("cfg2a", "v1", 263, 196, 64), ("cfg3", "v2", 498, 120, 128)])
def test_one_denoise_step_at_bench_batch_vs_oracle(cuda_device, name, arch, J, T, B):
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, model_config
    cfg = model_config(arch=arch, njoints=J)
    sd = make_state_dict(cfg, seed=0)
    x, y = make_inputs(cfg, B, T, seed=21)
    t = torch.full((B,), 37, dtype=torch.long)
    torch.set_num_threads(max(torch.get_num_threads(), 8))
    ref = _oracle_step(cfg, sd, x, y, t, [50], torch.float64)
    for precision, tol in (("fp16", 2e-3), ("fp32", 1e-5)):
        net = build_product_model(cfg, sd, cuda_device, precision)
        eng = net.engine()
        rows = 2 * B * (T + 1)
        if precision == "fp16":
            # the production kernel selection, not an override: fused feed-forward kernel from 4096 rows on
            assert eng.fuse_ffn and 0 < eng.fuse_ffn_min_rows <= rows
            assert eng.kernels_per_step(rows) == 2 + (1 if arch == "v2" else 0) + 4 * cfg["num_layers"] + 1
        dif = create_gaussian_diffusion(timestep_respacing=[50])
        out = dif.p_mean_variance(ClassifierFreeSampleModel(net), x.to(cuda_device), t.to(cuda_device),
                                  clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)})
        e_x0, e_mean = rel_l2(out["pred_xstart"], ref["pred_xstart"]), rel_l2(out["mean"], ref["mean"])
        record_parity("bench-batch step %s B=%d %s: pred_xstart %.2e mean %.2e (tol %.0e)"
                      % (name, B, precision, e_x0, e_mean, tol))
        assert e_x0 < tol and e_mean < tol
        del net, eng
        torch.cuda.empty_cache()


@pytest.mark.parametrize("arch,J,T,B", [("v1", 263, 10, 1), ("v1", 251, 30, 1), ("v1", 37, 13, 3), ("v2", 37, 10, 1)])
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_guided_loop_with_unaligned_uncond_half(cuda_device, arch, J, T, B, precision):
    """B*J*T % 4 != 0: `uncond = OUT[B:]` is not 16-byte aligned.  Guided 4-step chain vs the CPU oracle,
    through the captured graph and through the eager per-step path."""
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, make_step_noise, model_config
    from oracle.diffusion_oracle import OracleDiffusion, oracle_cfg
    from oracle.model_oracle import make_model_fn
    assert (B * J * T) % 4 != 0
    cfg = model_config(arch=arch, njoints=J, latent_dim=128, num_layers=2, num_heads=4, ff_size=256)
    sd = make_state_dict(cfg, seed=5)
    x, y = make_inputs(cfg, B, T, seed=9)
    noises = make_step_noise(tuple(x.shape), 4)
    ref = OracleDiffusion(respacing=[4]).p_sample_loop(oracle_cfg(make_model_fn(sd, cfg)), x, y, noises)
    model = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, precision))
    outs = []
    for graph in (True, False):
        dif = create_gaussian_diffusion(timestep_respacing=[4])
        dif.use_cuda_graph = graph
        outs.append(dif.p_sample_loop(model, tuple(x.shape), noise=x.to(cuda_device), clip_denoised=False,
                                      model_kwargs={"y": to_dev(y, cuda_device)}, step_noise=noises))
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1])
    err = rel_l2(outs[0], ref)
    record_parity("unaligned uncond %s J=%d T=%d B=%d %s: final %.2e" % (arch, J, T, B, precision, err))
    assert err < (1e-2 if precision == "fp16" else 1e-5)
    # the non-uniform-t eager path slices model_out per sample (gaussian_diffusion.py _posterior): same fault class
    t = torch.arange(B, dtype=torch.long) + 1
    dif = create_gaussian_diffusion(timestep_respacing=[4])
    out = dif.p_mean_variance(model, x.to(cuda_device), t.to(cuda_device), clip_denoised=False,
                              model_kwargs={"y": to_dev(y, cuda_device)})
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out["mean"]).all())


def test_step_graph_cache_is_keyed_on_schedule_content(cuda_device):
    """Two diffusion objects with different schedules on one model (the graph used to be keyed on id(coef)):
    each must replay its own graph; re-creating an equal schedule reuses the captured one; the cache stays bounded."""
    import gc
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion, sampler
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, model_config
    cfg = model_config(arch="v1", njoints=24, latent_dim=128, num_layers=2, num_heads=4, ff_size=256)
    sd = make_state_dict(cfg, seed=0)
    x, y = make_inputs(cfg, 3, 20)
    model = ClassifierFreeSampleModel(build_product_model(cfg, sd, cuda_device, "fp32"))
    kw = dict(noise=x.to(cuda_device), clip_denoised=False, model_kwargs={"y": to_dev(y, cuda_device)})

    def run(resp, graph=True):
        dif = create_gaussian_diffusion(timestep_respacing=resp)
        dif.rng, dif.seed, dif.use_cuda_graph = "philox", 11, graph
        out = dif.p_sample_loop(model, tuple(x.shape), **kw)
        del dif
        gc.collect()
        return out
    want = {n: run([n], graph=False) for n in (3, 5, 7)}
    for _ in range(3):                                   # interleave schedules; objects die between calls
        for n in (5, 3, 7):
            assert torch.equal(run([n]), want[n]), n
    for n in range(8, 8 + sampler.MAX_GRAPHS_PER_STATE + 3):
        run([n])
    eng = model.model.engine()
    for ws in eng._ws.values():
        for k, st in ws.items():
            if isinstance(k, tuple) and k and k[0] == "loop":
                assert len(st["graphs"]) <= sampler.MAX_GRAPHS_PER_STATE
    assert torch.equal(run([5]), want[5])
