"""Data-parallel sampling across the GPUs of one box.

Sampling is embarrassingly parallel over the batch: every op on the path is per-sample and the
timestep is batch-uniform.  One process per GPU; rank r takes a contiguous slice of the samples of
x_T and of every tensor in `y` whose leading dimension is the batch.  Collectives (NCCL over NVLink on
GPUs, gloo on CPU for the host-logic tests) are used ONLY
  * at start: broadcast of the packed weight arenas (one per dtype) and of the int64 noise seed
    (what the reference's dead `dist_util.sync_params`, utils/dist_util.py:61-67, intended);
  * at end: all-gather of the fp32 [B/G, J, 1, T] motions.
There is no per-step collective.  Noise is Philox keyed by (seed, GLOBAL sample index, step), so the
result does not depend on the number of ranks; `const_noise` uses global sample 0's stream on every
rank without communication.
"""
import os

import torch
import torch.distributed as dist


def init_process_group(backend=None):
    """Initialise torch.distributed from the torchrun environment (RANK/WORLD_SIZE/MASTER_*)."""
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world == 1:
        return 0, 1
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    kw = {}
    if backend == "nccl":
        kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world


def shard_range(batch, rank, world):
    """Contiguous [lo, hi) of the global batch owned by `rank`; the first batch % world ranks get one
    extra sample."""
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_inputs(shape, model_kwargs, rank, world, noise=None):
    """Local (shape, model_kwargs, noise, sample_base) for this rank."""
    B = shape[0]
    lo, hi = shard_range(B, rank, world)
    y = (model_kwargs or {}).get("y", {})
    y_loc = {}
    for k, v in y.items():
        if torch.is_tensor(v) and v.dim() >= 1 and v.shape[0] == B:
            y_loc[k] = v[lo:hi]
        elif isinstance(v, (list, tuple)) and len(v) == B:
            y_loc[k] = v[lo:hi]
        else:
            y_loc[k] = v
    kw = dict(model_kwargs or {})
    kw["y"] = y_loc
    return (hi - lo,) + tuple(shape[1:]), kw, (None if noise is None else noise[lo:hi]), lo


def broadcast_seed(seed, device, src=0):
    t = torch.tensor([int(seed)], dtype=torch.int64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(t, src=src)
    return int(t.item())


def broadcast_tensors(tensors, src=0):
    """In-place broadcast of a list of tensors from rank `src`."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        for t in tensors:
            dist.broadcast(t, src=src)


def broadcast_engine(engine, src=0):
    """Replicate rank `src`'s packed weights: one broadcast per arena (bf16, fp32), then re-derive the
    step-invariant tables locally from the received weights."""
    broadcast_tensors(list(engine.arenas.values()), src=src)
    engine._derive_tables()


def gather_motions(local, batch, group=None):
    """All-gather the per-rank [b_r, J, 1, T] motions into [batch, J, 1, T] (identical on all ranks).
    Shards may differ by one sample, so they are padded to the largest shard for the collective."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    sizes = [shard_range(batch, r, world)[1] - shard_range(batch, r, world)[0] for r in range(world)]
    mx = max(sizes)
    pad = local
    if local.shape[0] < mx:
        pad = torch.cat((local, local.new_zeros((mx - local.shape[0],) + tuple(local.shape[1:]))), dim=0)
    out = local.new_empty((world * mx,) + tuple(local.shape[1:]))
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    if min(sizes) == mx:
        return out                                   # equal shards: the collective's output already is the batch
    parts = [out[r * mx: r * mx + sizes[r]] for r in range(world)]
    return torch.cat(parts, dim=0)


def sharded_p_sample_loop(diffusion, model, shape, model_kwargs=None, noise=None, seed=0, gather=True, **kwargs):
    """p_sample_loop over the global batch `shape[0]`, sharded over the ranks of the default process
    group.  Uses the in-kernel Philox noise (rank-count invariant).  Returns the full [B,J,1,T] batch on
    every rank when `gather`, else this rank's shard."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    device = next(model.parameters()).device
    loc_shape, loc_kwargs, loc_noise, base = shard_inputs(shape, model_kwargs, rank, world, noise)
    saved = (diffusion.rng, diffusion.seed, diffusion.sample_base)
    try:
        diffusion.rng = "philox"
        diffusion.seed = broadcast_seed(seed, device)
        diffusion.sample_base = base
        out = diffusion.p_sample_loop(model, loc_shape, noise=loc_noise, model_kwargs=loc_kwargs, **kwargs)
    finally:
        diffusion.rng, diffusion.seed, diffusion.sample_base = saved     # the caller's diffusion object is unchanged
    return gather_motions(out, shape[0]) if gather else out
