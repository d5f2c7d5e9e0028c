"""Drop-in for the reference's diffusion/respace.py: ``space_timesteps`` (:8-61),
``SpacedDiffusion`` (:73-121) and ``_WrappedModel`` (:124-129).

The retained-timestep sets and ``timestep_map`` are bit-exact with the reference: section strides are
accumulated as a running float and rounded with Python's ``round`` (half-to-even), never multiplied
out.  The map is cached on the device once instead of being rebuilt from a Python list every call.
"""
import numpy as np
import torch as th

from .gaussian_diffusion import GaussianDiffusion


def space_timesteps(num_timesteps, section_counts):
    """Timesteps of the original process to keep.  `section_counts`: list of per-section counts,
    a comma-separated string of them, or "ddimN" for the fixed integer stride giving N steps."""
    if isinstance(section_counts, str):
        if section_counts.startswith("ddim"):
            desired_count = int(section_counts[len("ddim"):])
            for stride in range(1, num_timesteps):
                if len(range(0, num_timesteps, stride)) == desired_count:
                    return set(range(0, num_timesteps, stride))
            raise ValueError(f"cannot create exactly {num_timesteps} steps with an integer stride")
        section_counts = [int(x) for x in section_counts.split(",")]
    size_per, extra = divmod(num_timesteps, len(section_counts))
    kept = []
    start = 0
    for idx, count in enumerate(section_counts):
        size = size_per + (1 if idx < extra else 0)
        if size < count:
            raise ValueError(f"cannot divide section of {size} steps into {count}")
        frac_stride = 1 if count <= 1 else (size - 1) / (count - 1)
        cursor = 0.0
        for _ in range(count):
            kept.append(start + round(cursor))
            cursor += frac_stride
        start += size
    return set(kept)


class SpacedDiffusion(GaussianDiffusion):
    """A diffusion process that keeps only `use_timesteps` of a base process; betas are re-derived
    from the base alphas_cumprod so the marginals at the kept steps are unchanged."""

    def __init__(self, use_timesteps, **kwargs):
        self.use_timesteps = set(use_timesteps)
        self.timestep_map = []
        self.original_num_steps = len(kwargs["betas"])

        base = GaussianDiffusion(**kwargs)
        last = 1.0
        new_betas = []
        for i, acp in enumerate(base.alphas_cumprod):
            if i in self.use_timesteps:
                new_betas.append(1 - acp / last)
                last = acp
                self.timestep_map.append(i)
        kwargs["betas"] = np.array(new_betas)
        super().__init__(**kwargs)

    def p_mean_variance(self, model, *args, **kwargs):
        return super().p_mean_variance(self._wrap_model(model), *args, **kwargs)

    def _wrap_for_call(self, model):
        return self._wrap_model(model)

    def _wrap_model(self, model):
        if isinstance(model, _WrappedModel):
            return model
        return _WrappedModel(model, self.timestep_map, self.rescale_timesteps, self.original_num_steps)

    def _scale_timesteps(self, t):
        return t        # scaling is done by the wrapped model

    def model_timesteps(self, t):
        """Original-schedule timesteps the denoiser sees for respaced indices t."""
        return self._wrap_model(None).map_timesteps(t)


class _WrappedModel:
    def __init__(self, model, timestep_map, rescale_timesteps, original_num_steps):
        self.model = model
        self.timestep_map = timestep_map
        self.rescale_timesteps = rescale_timesteps
        self.original_num_steps = original_num_steps
        self._maps = {}

    def map_timesteps(self, ts):
        key = (str(ts.device), ts.dtype)
        if key not in self._maps:
            self._maps[key] = th.tensor(self.timestep_map, device=ts.device, dtype=ts.dtype)
        new_ts = self._maps[key][ts]
        if self.rescale_timesteps:
            new_ts = new_ts.float() * (1000.0 / self.original_num_steps)
        return new_ts

    def __call__(self, x, ts, **kwargs):
        return self.model(x, self.map_timesteps(ts), **kwargs)
