"""Synthetic point-cloud batches of the shapes BASELINE.json names.

Per cloud: points on a bumpy unit sphere, then the training augmentations' ranges — anisotropic
scale U(0.8,1.25)^3 and translation U(-0.1,0.1)^3 (reference: experiments/train_modelnet.py:37-40,
deltaconv/transforms/random_scale.py:29-33, random_translate_global.py:29-33).  Normals are the
sphere normal transformed like RandomScale transforms normals.  Always generated on the CPU
from a seeded generator, so every consumer (oracle, CUDA path, any rank) sees identical bytes.
"""
import torch

from .data import Data


def synthetic_batch(num_clouds, points, config=2, rank=0, normals=True, num_classes=40,
                    per_point_labels=False, categories=0, sizes=None):
    gen = torch.Generator().manual_seed(1234 + 100 * config + rank)
    sizes = [points] * num_clouds if sizes is None else list(sizes)
    pos, nrm, bat = [], [], []
    for b, P in enumerate(sizes):
        u = torch.randn(P, 3, generator=gen)
        u = u / u.norm(dim=1, keepdim=True).clamp(1e-12)
        radius = 1 + 0.15 * torch.sin(3 * torch.atan2(u[:, 1], u[:, 0])) * u[:, 2]
        scale = torch.empty(3).uniform_(0.8, 1.25, generator=gen)
        shift = torch.empty(3).uniform_(-0.1, 0.1, generator=gen)
        pos.append((radius[:, None] * u) * scale + shift)
        n = u / scale
        nrm.append(n / n.norm(dim=1, keepdim=True).clamp(1e-12))
        bat.append(torch.full((P,), b, dtype=torch.long))
    pos, nrm, bat = torch.cat(pos).contiguous(), torch.cat(nrm).contiguous(), torch.cat(bat)
    N = pos.size(0)
    fields = dict(pos=pos, batch=bat)
    if normals:
        fields['norm'] = nrm
    if per_point_labels:
        fields['y'] = torch.randint(0, num_classes, (N,), generator=gen)
    else:
        fields['y'] = torch.randint(0, num_classes, (len(sizes),), generator=gen)
    if categories:
        cat = torch.randint(0, categories, (len(sizes),), generator=gen)
        fields['category'] = torch.nn.functional.one_hot(cat, categories).float()
    return Data(**fields)
