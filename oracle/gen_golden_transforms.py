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
    np.savez(os.path.join(ROOT, 'tests', 'golden', 'transforms.npz'), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    sys.exit(main())
