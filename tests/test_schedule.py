"""CPU: the product's schedule / respacing host logic is bit-exact with the reference (golden)."""
import numpy as np
import pytest

from helpers import GOLDEN
from gesturediffusion_b200 import create_gaussian_diffusion, space_timesteps
from gesturediffusion_b200.diffusion import gaussian_diffusion as gd

TABLES = ["betas", "alphas_cumprod", "alphas_cumprod_prev", "alphas_cumprod_next", "sqrt_alphas_cumprod",
          "sqrt_one_minus_alphas_cumprod", "log_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod",
          "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
          "posterior_mean_coef1", "posterior_mean_coef2"]


def test_space_timesteps_bit_exact():
    g = np.load(GOLDEN + "/schedule.npz")
    specs = {"s1000_50": (1000, [50]), "s1000_ddim50": (1000, "ddim50"), "s300_10_15_20": (300, [10, 15, 20]),
             "s1000_1000": (1000, [1000]), "s1000_str": (1000, "25,25"), "s1000_7": (1000, [7]),
             "s1000_1": (1000, [1])}
    for name, (n, spec) in specs.items():
        assert sorted(space_timesteps(n, spec)) == g["space_" + name].tolist(), name
    # 50 of 1000: running float sum + banker's rounding, first few known values
    assert sorted(space_timesteps(1000, [50]))[:5] == [0, 20, 41, 61, 82]


def test_space_timesteps_errors():
    with pytest.raises(ValueError):
        space_timesteps(10, [20])
    with pytest.raises(ValueError):
        space_timesteps(1000, "ddim999")        # no integer stride gives exactly 999 steps


@pytest.mark.parametrize("sched", ["cosine", "linear"])
@pytest.mark.parametrize("resp,tag", [("", "1000"), ([50], "50"), ("ddim50", "ddim50")])
def test_tables_and_timestep_map_bit_exact(sched, resp, tag):
    g = np.load(GOLDEN + "/schedule.npz")
    d = create_gaussian_diffusion(noise_schedule=sched, timestep_respacing=resp)
    assert d.timestep_map == g["tmap_%s_%s" % (sched, tag)].tolist()
    assert d.original_num_steps == 1000 and d.num_timesteps == len(d.timestep_map)
    for t in TABLES:
        assert getattr(d, t).dtype == np.float64
        assert np.array_equal(getattr(d, t), g["%s_%s_%s" % (t, sched, tag)]), t


def test_factory_defaults_match_reference_factory():
    d = create_gaussian_diffusion()
    assert d.model_mean_type == gd.ModelMeanType.START_X
    assert d.model_var_type == gd.ModelVarType.FIXED_SMALL
    assert d.rescale_timesteps is False and d.num_timesteps == 1000


def test_out_of_scope_paths_raise():
    d = create_gaussian_diffusion(timestep_respacing=[5])
    with pytest.raises(NotImplementedError):
        d.training_losses(None, None, None)
    with pytest.raises(NotImplementedError):
        d.plms_sample_loop(None, (1, 2, 1, 10))
    with pytest.raises(NotImplementedError):                       # the reference's own refusals (:900-903)
        d.ddim_sample_loop(None, (1, 2, 1, 10), dump_steps=[0])
    with pytest.raises(NotImplementedError):
        d.ddim_sample_loop(None, (1, 2, 1, 10), const_noise=True)
    with pytest.raises(NotImplementedError):
        d._check_supported(denoised_fn=lambda x: x)
