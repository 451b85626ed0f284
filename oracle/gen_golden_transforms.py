"""Golden vectors for the deterministic transforms: the UNMODIFIED reference classes (deltaconv/transforms/normalize_scale.py,
normalize_axes.py) imported from /root/reference in the build container, applied cloud by cloud, stored in
tests/golden/transforms.npz.  Test infrastructure: run once here, the fixture travels, /root/reference does not.

    python oracle/gen_golden_transforms.py
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference/deltaconv/transforms'


def load(name):
    spec = importlib.util.spec_from_file_location(f'ref_{name}', os.path.join(REF, f'{name}.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class Bag:
    pass


def _mesh():
    v = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    f = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]
    for _ in range(2):                                            # two rounds of 1-to-4 subdivision
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
    verts = torch.tensor(v, dtype=torch.float32)
    verts = verts / verts.norm(dim=1, keepdim=True) * torch.tensor([1.0, 0.7, 0.4]) + torch.tensor([0.2, -0.1, 0.3])
    return verts, torch.tensor(f, dtype=torch.long).t().contiguous()


def main():
    ns = load('normalize_scale').NormalizeScale
    na = load('normalize_axes').NormalizeAxes
    gen = torch.Generator().manual_seed(77)
    sizes = [40, 7, 133]
    clouds = [torch.randn(n, 3, generator=gen) * torch.tensor([0.3, 2.0, 1.0]).roll(i) + torch.tensor([1.0, -2.0, 0.5]) * i
              for i, n in enumerate(sizes)]
    out = dict(sizes=np.array(sizes), pos=torch.cat(clouds).numpy())
    for key, make in (('scale_l2', lambda: ns()), ('scale_inf', lambda: ns(norm_ord=float('inf'))), ('scale_fixed', lambda: ns(scaling_factor=3.0)),
                      ('axes', lambda: na())):
        parts = []
        for c in clouds:
            d = Bag()
            d.pos = c.clone()
            parts.append(make()(d).pos)
        out[key] = torch.cat(parts).numpy()
    # the two mesh transforms: a small closed mesh (subdivided octahedron pushed to an ellipsoid), seeded sampling
    sp = load('sample_points').SamplePoints
    nar = load('normalize_area').NormalizeArea
    verts, faces = _mesh()
    out['mesh_pos'], out['mesh_face'] = verts.numpy(), faces.numpy()
    labels = torch.arange(verts.size(0)) % 5
    for key, kw in (('sample_plain', {}), ('sample_full', dict(include_normals=True, include_labels=True, remove_faces=False))):
        d = Bag()
        d.pos, d.face, d.y = verts.clone(), faces.clone(), labels.clone()
        torch.manual_seed(4321)
        d = sp(200, **kw)(d)
        out[key + '_pos'] = d.pos.numpy()
        if kw:
            out[key + '_norm'], out[key + '_y'] = d.norm.numpy(), d.y.numpy()
    d = Bag()
    d.pos, d.face = verts.clone(), faces.clone()
    out['area_pos'] = nar()(d).pos.numpy()
    np.savez(os.path.join(ROOT, 'tests', 'golden', 'transforms.npz'), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    sys.exit(main())
