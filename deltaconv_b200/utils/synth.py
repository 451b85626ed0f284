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


def synthetic_mesh(label, gen, subdivisions=3):
    """A closed triangle mesh whose shape depends on `label`: a subdivided octahedron pushed onto an ellipsoid.  Stands in for
    a dataset shape (`pos` [V,3], `face` [3,F], `y` [1]) wherever the reference's mesh datasets would be read."""
    v = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    f = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]
    for _ in range(subdivisions):
        mid, nf = {}, []

        def m(a, b):
            key = (min(a, b), max(a, b))
            if key not in mid:
                v.append(tuple((v[a][i] + v[b][i]) / 2 for i in range(3)))
                mid[key] = len(v) - 1
            return mid[key]
        for a, b, c in f:
            ab, bc, ca = m(a, b), m(b, c), m(c, a)
            nf += [(a, ab, ca), (ab, b, bc), (ca, bc, c), (ab, bc, ca)]
        f = nf
    pos = torch.tensor(v, dtype=torch.float32)
    axes = torch.tensor([1.0, 0.4 + 0.15 * (label % 4), 1.0 - 0.2 * (label % 4) + 0.05 * (label // 4)]) * (0.9 + 0.2 * torch.rand(3, generator=gen))
    pos = pos / pos.norm(dim=1, keepdim=True) * axes
    return Data(pos=pos, face=torch.tensor(f, dtype=torch.long).t().contiguous(), y=torch.tensor([label]))
