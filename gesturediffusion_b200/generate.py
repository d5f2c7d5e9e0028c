"""Chunked autoregressive generation of a take: the sampling loop of sample/generate.py:91-184 behind the
package's drop-in objects.

The reference generates a take as `chunks_per_take` sequential `p_sample_loop` calls; chunk k+1 is conditioned
on the last `seed_poses` frames of chunk k (generate.py:104-107), every chunk is de-normalised with the dataset
statistics (`inv_transform`, data_loaders/gesture/data/dataset.py:118-119), split into joint positions and
rotations (generate.py:138-146) and the chunks are concatenated along the frame axis at the end (:185-189).
Between chunks it moves the sample to the host and back.

Here the whole take stays on the device: the seed hand-off is a device slice, the de-normalisation + split is
one small kernel (`gd_unnormalize_split`) that writes each chunk straight into the take's position / rotation
tensors at its frame offset, and one captured step graph serves all chunks (the Philox seed lives in device
memory).  The only device->host copy is the caller's, at the end.
"""
import torch

from . import kernels as K


def generate_take(model, diffusion, chunk_conditions, num_samples, num_frames, seed_poses=10, guidance_param=2.5,
                  mean=None, std=None, init_noise=None, step_noise=None, progress=False):
    """Generate `len(chunk_conditions)` chunks of `num_frames` frames for `num_samples` takes.

    model / diffusion   the package's denoiser (optionally wrapped in ClassifierFreeSampleModel when
                        guidance_param != 1, as generate.py:72-73 does) and SpacedDiffusion, on a CUDA device.
    chunk_conditions    one `y` dict per chunk, what `gg_collate` returns under model_kwargs['y'] for that chunk
                        (generate.py:99-101): 'seed' [B, J, 1, seed_poses] (used for chunk 0 only), 'mfcc'
                        [B, 26, 1, num_frames], optional 'text'; host or device tensors.
    mean, std           dataset statistics [J] for inv_transform (None: identity).
    init_noise, step_noise   optional per-chunk lists of injected noise (parity runs): x_T [B, J, 1, T] and the
                        per-step noise tensors of each chunk.
    Returns dict(sample=[B, J, 1, chunks*T] network-space output, motion=[B, J/6, 3, chunks*T] positions,
    motion_rot=[B, J/6, 3, chunks*T] rotations), all on the device -- `motion` is what the reference stores as
    `all_motions` before `rot2xyz(pose_rep='xyz')`, which is the identity for this representation.
    """
    chunks = list(chunk_conditions)
    if not chunks:
        raise ValueError("generate_take needs at least one chunk of conditioning")
    dev = next((v.device for v in _tensors(model) if v.is_cuda), None)
    if dev is None:
        raise K._cabi.GdError("generate_take needs a model on a CUDA device (gesturediffusion_b200 has no CPU path)")
    J, F = model.njoints, model.nfeats
    if (J * F) % 6 != 0:
        raise ValueError("the GENEA feature layout is 6 values per joint; got %d features" % (J * F))
    B, T, n = num_samples, num_frames, len(chunks)
    to_dev = lambda v: v.to(dev, torch.float32).contiguous() if v is not None else None   # noqa: E731
    mean_d, std_d = to_dev(mean), to_dev(std)
    sample_all = torch.empty(B, J, F, n * T, device=dev)
    motion = torch.empty(B, J * F // 6, 3, n * T, device=dev)
    motion_rot = torch.empty_like(motion)
    base_seed = getattr(diffusion, "seed", 0)
    sample_out = None
    try:
        for k, cond in enumerate(chunks):
            y = {key: (val.to(dev) if torch.is_tensor(val) else val) for key, val in cond.items()}
            if k > 0:                                                   # generate.py:104-107
                y["seed"] = sample_out[..., -seed_poses:].contiguous()
            if guidance_param != 1:                                     # generate.py:114-115
                y["scale"] = torch.ones(B, device=dev) * guidance_param
            diffusion.seed = base_seed + k                              # fresh noise per chunk (Philox mode)
            sample_out = diffusion.p_sample_loop(
                model, (B, J, F, T), clip_denoised=False, model_kwargs={"y": y}, skip_timesteps=0, init_image=None,
                progress=progress, dump_steps=None, noise=None if init_noise is None else init_noise[k],
                const_noise=False, **({} if step_noise is None else {"step_noise": step_noise[k]}))
            sample_all[..., k * T:(k + 1) * T] = sample_out
            K.unnormalize_split(sample_out.reshape(B, J * F, T), mean_d, std_d, motion, motion_rot, k * T)
    finally:
        diffusion.seed = base_seed
    return {"sample": sample_all, "motion": motion, "motion_rot": motion_rot}


def _tensors(model):
    inner = getattr(model, "model", model)
    yield from inner.parameters()
    yield from inner.buffers()
