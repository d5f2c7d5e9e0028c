"""FastLoop: the denoising loop of `p_sample_loop_progressive` (gaussian_diffusion.py:663-730) for
engine-backed models, with everything per-step kept on the device.

Per step the reference issues ~3000 eager ATen ops, two model calls, a deepcopy of the conditioning
dict and >= 8 implicit host->device copies.  Here one step is
    cond_tokens -> input GEMM [-> local attention] -> 8 x (QKV GEMM, attention, out-proj GEMM + residual,
    LayerNorm, FFN1 GEMM + GELU, FFN2 GEMM + residual, LayerNorm) -> output GEMM (transposed store)
    -> fused CFG + posterior + noise kernel (which also writes the next step's GEMM operand and
       advances the device-side step counter / model timesteps)
captured ONCE into a CUDA graph and replayed `num_timesteps` times; the step index lives in device
memory, so the host only replays.
"""
import torch

from . import kernels as K
from .diffusion.gaussian_diffusion import ModelMeanType


# launch accounting for bench.py's `gpu_launches` claim: eager launches are counted by the library
# (gd_launch_count); kernels replayed from a captured graph are counted here.
STATS = {"graph_replays": 0, "graph_kernels": 0}
MAX_GRAPHS_PER_STATE = 8


class DeviceScope:
    """Makes the engine's GPU the current CUDA device while kernels are launched (the C ABI takes torch's CURRENT stream,
    and the library's one-off set-up is per current device), and restores the caller's device around every `yield` and at
    the end -- a model on cuda:1 works while cuda:0 stays the caller's current device.  A no-op when they agree."""

    def __init__(self, device):
        self.idx = device.index if device.index is not None else torch.cuda.current_device()
        self.prev = None

    def enter(self):
        cur = torch.cuda.current_device()
        if cur != self.idx:
            self.prev = cur
            torch.cuda.set_device(self.idx)

    def exit(self):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)
            self.prev = None

    def __enter__(self):
        self.enter()
        return self

    def __exit__(self, *exc):
        self.exit()
        return False


def is_engine_model(model):
    from .model.base import EngineBackedDenoiser
    from .model.cfg_sampler import ClassifierFreeSampleModel
    if isinstance(model, ClassifierFreeSampleModel):
        return isinstance(model.model, EngineBackedDenoiser)
    return isinstance(model, EngineBackedDenoiser)


class FastLoop:
    def __init__(self, diffusion, model, device):
        from .model.cfg_sampler import ClassifierFreeSampleModel
        self.diffusion = diffusion
        self.cfg = isinstance(model, ClassifierFreeSampleModel)
        self.denoiser = model.model if self.cfg else model
        self.engine = self.denoiser.engine()
        self.device = self.engine.device

    # -------------------------------------------------------------- per-call state
    def _state(self, ws, shape, want_pred, has_noise_buf, inpaint):
        key = ("loop", want_pred, has_noise_buf, inpaint)
        st = ws.get(key)
        if st is None:
            B, J, T = ws["B"], self.engine.J, ws["T"]
            dev = self.device
            st = dict(x=torch.zeros(B, J, T, device=dev), scale=torch.ones(B, device=dev),
                      step=torch.zeros(1, dtype=torch.int32, device=dev),
                      seed=torch.zeros(1, dtype=torch.int64, device=dev), graphs={})
            if want_pred:
                st["pred"] = torch.zeros(B, J, T, device=dev)
            if has_noise_buf:
                st["noise"] = torch.zeros(B, J, T, device=dev)
            if inpaint:
                st["mask"] = torch.zeros(B, J, T, dtype=torch.uint8, device=dev)
                st["motion"] = torch.zeros(B, J, T, device=dev)
            ws[key] = st
        return st

    def _step(self, ws, st, tabs, flags, coef):
        """One denoising step: model evaluation + fused update (in place on st['x'])."""
        eng, d = self.engine, self.diffusion
        B, J, T = ws["B"], eng.J, ws["T"]
        eng.forward_step(ws)
        out = ws["OUT"]
        K.posterior_step(out[:B], st["x"], coef, st["step"], B, J, T,
                         uncond=out[B:] if self.cfg else None, scale=st["scale"] if self.cfg else None,
                         noise=st.get("noise"), x_prev=st["x"], pred_xstart=st.get("pred"),
                         a_next=ws["a_in"], inpaint_mask=st.get("mask"), inpainted=st.get("motion"),
                         timestep_map=tabs["timestep_map"], next_t=ws["t"],
                         eps_mode=d.model_mean_type == ModelMeanType.EPSILON, clip=flags["clip"],
                         const_noise=flags["const_noise"], seed=d.seed, sample_base=d.sample_base,
                         seed_dev=st["seed"], ddim=flags["ddim"])

    # -------------------------------------------------------------- the loop
    def run(self, shape, noise, model_kwargs, clip_denoised, skip_timesteps, init_image, const_noise, step_noise,
            progress, want_pred, yield_steps, ddim_eta=None):
        d, eng = self.diffusion, self.engine
        dev = self.device
        B, J, F_, T = shape
        if J * F_ != eng.J:
            raise ValueError("shape %s does not match the model's %d input features" % (shape, eng.J))
        y = model_kwargs.get("y", {}) if model_kwargs else {}
        nb = 2 if self.cfg else 1
        # cfg_sampler.py:23-28 passes y (with whatever y['uncond'] says) to the first call and forces uncond on the second
        branches = [bool(y.get("uncond", False)), True] if self.cfg else [bool(y.get("uncond", False))]
        rng = "inject" if step_noise is not None else d.rng
        if rng not in ("torch", "philox", "inject"):
            raise ValueError("diffusion.rng must be 'torch' or 'philox'")
        inpaint = d._inpaint_of(model_kwargs)
        n_exec = d.num_timesteps - skip_timesteps
        if n_exec <= 0:
            raise ValueError("skip_timesteps leaves no steps to run")
        if step_noise is not None and len(step_noise) < n_exec:
            raise ValueError("step_noise needs %d entries" % n_exec)

        scope = DeviceScope(dev)
        scope.enter()
        try:
          with torch.no_grad():
            ws = eng.workspace(B, T, nb)
            tabs = d.device_tables(dev)
            st = self._state(ws, shape, want_pred, rng != "philox", inpaint is not None)
            eng.prepare_cond(ws, y, branches)
            if self.cfg:
                st["scale"].copy_(y["scale"].to(dev, torch.float32).reshape(-1))
            if inpaint is not None:
                st["mask"].copy_(inpaint[0].to(dev).reshape(B, eng.J, T).to(torch.uint8))
                st["motion"].copy_(inpaint[1].to(dev, torch.float32).reshape(B, eng.J, T))

            # initial image (reference :690-703)
            if noise is not None:
                img = noise.to(dev, torch.float32).reshape(B, eng.J, T).contiguous()
            elif rng == "philox":
                img = torch.empty(B, eng.J, T, device=dev)
                K.philox_normal(img, B, eng.J, T, int(d.seed) & 0x7FFFFFFFFFFFFFFF, d.sample_base, d.num_timesteps)
            else:
                img = torch.randn(*shape, device=dev).reshape(B, eng.J, T)
            if skip_timesteps and init_image is None:
                init_image = torch.zeros(*shape, device=dev)
            first = n_exec - 1
            if init_image is not None:
                q = torch.empty_like(img)
                K.q_sample(init_image.to(dev, torch.float32).reshape(B, eng.J, T).contiguous(), img,
                           tabs["sqrt_acp"], tabs["sqrt_1macp"], first, q)
                img = q

            # DDIM (ddim_sample_loop_progressive, :933-993) replays the same step with another coefficient table
            coef = tabs["coef"] if ddim_eta is None else d.ddim_coef(dev, ddim_eta)
            flags = dict(clip=bool(clip_denoised), const_noise=bool(const_noise), ddim=ddim_eta is not None)
            # the Philox seed is read from device memory, so one captured graph serves every seed (chunked takes)
            st["seed"].fill_(int(d.seed) & 0x7FFFFFFFFFFFFFFF)
            # A captured graph bakes in the pointers of coef / timestep_map and n_steps: key it on the schedule CONTENT
            # (not on object identity, which CPython reuses) and keep those tensors alive next to the graph.
            gkey = (flags["clip"], flags["const_noise"], d.model_mean_type, d.sample_base, flags["ddim"], tabs["sig"],
                    None if ddim_eta is None else float(ddim_eta))
            graph = None
            if d.use_cuda_graph:
                entry = st["graphs"].pop(gkey, None)
                if entry is not None:
                    graph, coef, tmap_keep = entry          # replay against the tensors the graph was captured with
                    tabs = dict(tabs, timestep_map=tmap_keep)
                    st["graphs"][gkey] = entry              # re-insert: most recently used last
                if graph is None:
                    # warm-up run outside capture (sets kernel attributes, allocates nothing afterwards)
                    st["x"].copy_(img)
                    eng.load_x(ws, st["x"])
                    K.set_loop_state(st["step"], ws["t"], tabs["timestep_map"], first, B)
                    self._step(ws, st, tabs, flags, coef)
                    torch.cuda.current_stream().synchronize()
                    graph = torch.cuda.CUDAGraph()
                    # capture on a stream of THIS engine's GPU (torch's default capture stream is created once per process,
                    # on whatever device was current then, and entering it would switch the current device)
                    if getattr(eng, "_capture_stream", None) is None:
                        eng._capture_stream = torch.cuda.Stream(device=dev)
                    with torch.cuda.graph(graph, stream=eng._capture_stream):
                        self._step(ws, st, tabs, flags, coef)
                    st["graphs"][gkey] = (graph, coef, tabs["timestep_map"])
                    while len(st["graphs"]) > MAX_GRAPHS_PER_STATE:      # bounded: evict the least recently used
                        st["graphs"].pop(next(iter(st["graphs"])))

            st["x"].copy_(img)
            eng.load_x(ws, st["x"])
            K.set_loop_state(st["step"], ws["t"], tabs["timestep_map"], first, B)

            order = range(n_exec)
            if progress:
                from tqdm.auto import tqdm
                order = tqdm(order)
            for k in order:
                if rng == "inject":
                    st["noise"].copy_(step_noise[k].to(dev, torch.float32).reshape(B, eng.J, T))
                elif rng == "torch":
                    st["noise"].normal_()
                if graph is not None:
                    graph.replay()
                    STATS["graph_replays"] += 1
                    STATS["graph_kernels"] += self.launches_per_step(ws["R"])
                else:
                    self._step(ws, st, tabs, flags, coef)
                last = k == n_exec - 1
                if yield_steps == "all" or last or (yield_steps is not None and k in yield_steps):
                    out = {"sample": st["x"].view(B, J, F_, T).clone()}
                    if want_pred:
                        out["pred_xstart"] = st["pred"].view(B, J, F_, T).clone()
                else:
                    out = None
                scope.exit()                                 # the consumer runs on ITS current device
                yield out
                scope.enter()
        finally:
            scope.exit()

    def launches_per_step(self, rows):
        return self.engine.kernels_per_step(rows) + 1
