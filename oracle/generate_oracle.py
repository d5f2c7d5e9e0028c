"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the chunked generation loop of the reference,
sample/generate.py:91-189, and of its post-processing (data_loaders/gesture/data/dataset.py:118-119,
generate.py:132-146).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package.

`sample_fn(y, k)` is whatever produces chunk k's network-space sample [B, J, 1, T] from the conditioning dict
(the reference's diffusion.p_sample_loop with injected noise, or the oracle sampler): the glue around it is
what is restated here.
"""
import numpy as np
import torch


def oracle_inv_transform(data, mean, std):
    """dataset.py:118-119."""
    return data * std + mean


def oracle_split_positions_rotations(sample_out, mean, std):
    """generate.py:132-146: inv_transform on [B, 1, T, J], GENEA 6-interleave split, views and permutes.
    Returns (positions, rotations), each [B, J/6, 3, T]."""
    J = sample_out.shape[1]
    n_joints = J // 6
    sample = oracle_inv_transform(sample_out.cpu().permute(0, 2, 3, 1), mean, std).float()       # [B, 1, T, J]
    idx_positions = np.asarray([[i * 6 + 3, i * 6 + 4, i * 6 + 5] for i in range(n_joints)]).flatten()
    idx_rotations = np.asarray([[i * 6, i * 6 + 1, i * 6 + 2] for i in range(n_joints)]).flatten()
    sample, sample_rot = sample[..., idx_positions], sample[..., idx_rotations]
    sample = sample.view(sample.shape[:-1] + (-1, 3))
    sample = sample.view(-1, *sample.shape[2:]).permute(0, 2, 3, 1)
    sample_rot = sample_rot.view(sample_rot.shape[:-1] + (-1, 3))
    sample_rot = sample_rot.view(-1, *sample_rot.shape[2:]).permute(0, 2, 3, 1)
    return sample, sample_rot


def oracle_generate_take(sample_fn, chunk_conditions, num_samples, seed_poses, guidance_param, mean, std):
    """generate.py:91-130 (seed hand-off :104-107, guidance scale :114-115) and :185-189 (concatenation)."""
    all_motions, all_motions_rot, all_samples = [], [], []
    sample_out = None
    for chunk, cond in enumerate(chunk_conditions):
        y = dict(cond)
        if chunk > 0:
            y["seed"] = sample_out[..., -seed_poses:]
        if guidance_param != 1:
            y["scale"] = torch.ones(num_samples) * guidance_param
        sample_out = sample_fn(y, chunk)
        pos, rot = oracle_split_positions_rotations(sample_out, mean, std)
        all_samples.append(sample_out)
        all_motions.append(pos.numpy())
        all_motions_rot.append(rot.numpy())
    return dict(sample=torch.cat(all_samples, dim=3), motion=np.concatenate(all_motions, axis=3),
                motion_rot=np.concatenate(all_motions_rot, axis=3))
