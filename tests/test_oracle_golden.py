"""CPU: pins the oracle to the reference.  (a) oracle vs the committed golden fixtures generated from the
unmodified reference; (b) when /root/reference is present, oracle vs the live reference."""
import numpy as np
import pytest
import torch

from helpers import CHAIN_CASES, GOLDEN, MODEL_CASES, load_chain_case, load_model_case, rel_l2
from oracle import ref_shims
from oracle.diffusion_oracle import (OracleDiffusion, oracle_betas, oracle_cfg, oracle_respace,
                                     oracle_space_timesteps, oracle_tables)
from oracle.model_oracle import make_model_fn

FP32_TOL = 5e-6      # two independent fp32 evaluation orders of the same network


@pytest.mark.parametrize("case", MODEL_CASES)
def test_model_oracle_matches_reference_golden(case):
    cfg, sd, x, y, gold = load_model_case(case)
    fn = make_model_fn(sd, cfg)
    assert rel_l2(fn(x, gold["t"], y), gold["cond"]) < FP32_TOL
    assert rel_l2(fn(x, gold["t"], y, uncond=True), gold["uncond"]) < FP32_TOL
    assert rel_l2(oracle_cfg(fn)(x, gold["t"], y), gold["guided"]) < 2e-5   # scale 2.5 amplifies c-u rounding


@pytest.mark.parametrize("case", CHAIN_CASES)
def test_chain_oracle_matches_reference_golden(case):
    cfg, sd, x, y, noises, meta, kwargs, gold = load_chain_case(case)
    fn = make_model_fn(sd, cfg)
    model = oracle_cfg(fn) if meta["cfg"] else fn
    od = OracleDiffusion(respacing=meta["respacing"])
    trace = []
    final = od.p_sample_loop(model, x, y, noises, trace=trace, **kwargs)
    assert rel_l2(final, gold["final"]) < 1e-4
    for j, i in enumerate(gold["trace_idx"]):
        assert rel_l2(trace[i]["pred_xstart"], gold["trace_pred"][j]) < 1e-4
        assert rel_l2(trace[i]["sample"], gold["trace_sample"][j]) < 1e-4


def test_schedule_oracle_bit_exact():
    g = np.load(GOLDEN + "/schedule.npz")
    specs = {"s1000_50": (1000, [50]), "s1000_ddim50": (1000, "ddim50"), "s300_10_15_20": (300, [10, 15, 20]),
             "s1000_1000": (1000, [1000]), "s1000_str": (1000, "25,25"), "s1000_7": (1000, [7]),
             "s1000_1": (1000, [1])}
    for name, (n, spec) in specs.items():
        assert sorted(oracle_space_timesteps(n, spec)) == g["space_" + name].tolist()
    for sched in ("cosine", "linear"):
        for resp, tag in (([1000], "1000"), ([50], "50"), ("ddim50", "ddim50")):
            betas, tmap = oracle_respace(oracle_betas(sched, 1000), oracle_space_timesteps(1000, resp))
            assert tmap == g["tmap_%s_%s" % (sched, tag)].tolist()
            for k, v in oracle_tables(betas).items():
                assert np.array_equal(v, g["%s_%s_%s" % (k, sched, tag)]), (k, sched, tag)


@pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference not present (GPU box)")
def test_oracle_matches_live_reference():
    cfg, sd, x, y, gold = load_model_case("v2_small")
    m = ref_shims.build_reference_model("v2", cfg["njoints"], latent_dim=cfg["latent_dim"],
                                        num_layers=cfg["num_layers"], ff_size=cfg["ff_size"])
    m.load_state_dict(sd, strict=False)
    with torch.no_grad():
        ref = m(x, gold["t"], y)
    assert rel_l2(make_model_fn(sd, cfg)(x, gold["t"], y), ref) < FP32_TOL
    assert rel_l2(ref, gold["cond"]) == 0.0          # fixtures are reproducible from the committed script


def test_staged_reference_archive_is_the_reference():
    """oracle/_ref/reference_path.zip (written by oracle/stage_ref.py, git-ignored, travels to the GPU box) holds
    byte-identical copies of the reference modules the sampling path imports: every archived file has its manifest
    digest and -- where /root/reference exists -- equals the source file."""
    import os
    from oracle import stage_ref
    if not os.path.exists(os.path.join(stage_ref.DEST, stage_ref.ARCHIVE)):
        pytest.skip("oracle/_ref not staged (run __graft_entry__.build() where /root/reference exists)")
    src = stage_ref.SRC if os.path.isdir(os.path.join(stage_ref.SRC, "diffusion")) else None
    assert stage_ref.verify(src=src)
    import json
    man = json.load(open(os.path.join(stage_ref.DEST, "MANIFEST.json")))["files"]
    for needed in ("diffusion/gaussian_diffusion.py", "diffusion/respace.py", "model/mdm.py", "model/mdm_old.py",
                   "model/cfg_sampler.py", "model/local_attention.py"):
        assert needed in man, needed
