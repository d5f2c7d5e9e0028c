"""Batch layout of the GENEA loader: data_loaders/tensors.py:3-66 (`lengths_to_mask`, `collate_tensors`, `collate`,
`gg_collate`) restated for the chunked driver, so items produced by the reference's dataset
(data_loaders/gesture/data/dataset.py:58-99: motion [T, J], text, length, audio [T_a], mfcc [T, 26],
seed [n_seed, J]) become the `(motion, {'y': ...})` pair the sampling path consumes:
motion [B, J, 1, T], y['mfcc'] [B, 26, 1, T], y['seed'] [B, J, 1, n_seed], y['mask'], y['lengths'], y['text'],
y['audio'].  Host-side only (the tensors are moved by the caller)."""
import torch


def lengths_to_mask(lengths, max_len):
    return torch.arange(max_len, device=lengths.device).expand(len(lengths), max_len) < lengths.unsqueeze(1)


def collate_tensors(batch):
    dims = batch[0].dim()
    size = (len(batch),) + tuple(max(b.size(i) for b in batch) for i in range(dims))
    canvas = batch[0].new_zeros(size=size)
    for i, b in enumerate(batch):
        canvas[i][tuple(slice(0, n) for n in b.shape)] += b
    return canvas


def collate(batch):
    items = [b for b in batch if b is not None]
    data = collate_tensors([b['inp'] for b in items])
    lengths = torch.as_tensor([b['lengths'] if 'lengths' in b else len(b['inp'][0][0]) for b in items])
    cond = {'y': {'mask': lengths_to_mask(lengths, data.shape[-1]).unsqueeze(1).unsqueeze(1), 'lengths': lengths}}
    if 'text' in items[0]:
        cond['y']['text'] = [b['text'] for b in items]
    if 'mfcc' in items[0]:
        cond['y']['mfcc'] = torch.cat([b['mfcc'] for b in items], dim=0)
    if 'audio' in items[0]:
        cond['y']['audio'] = torch.cat([b['audio'] for b in items], dim=0)
    if 'seed' in items[0]:
        cond['y']['seed'] = torch.cat([b['seed'].unsqueeze(0) for b in items], dim=0)
    return data, cond


def gg_collate(batch):
    """tensors.py:56-66: dataset tuples -> collate()."""
    adapted = [{
        'inp': torch.as_tensor(b[0].T).float().unsqueeze(1),                      # [T, J] -> [J, 1, T]
        'text': b[1],
        'lengths': b[2],
        'audio': torch.as_tensor(b[3]).unsqueeze(0),                              # [T_a] -> [1, T_a]
        'mfcc': torch.as_tensor(b[4].T).float().unsqueeze(1).unsqueeze(0),        # [T, 26] -> [1, 26, 1, T]
        'seed': torch.as_tensor(b[5].T).float().unsqueeze(1),                     # [n_seed, J] -> [J, 1, n_seed]
    } for b in batch]
    return collate(adapted)
