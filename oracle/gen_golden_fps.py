"""Writes tests/golden/geodesic_fps.npz: sample sequences of the REFERENCE's geodesic farthest-point sampler
(deltaconv/cpp/sampling.cpp compiled as oracle/_ref/libref_fps.so by `make -C oracle`) on small seeded clouds.
Run here, where /root/reference exists:  python oracle/gen_golden_fps.py
The reference picks its first sample at random; it is element 0 of every stored sequence."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import fps   # noqa: E402


def cloud(kind, n, rng):
    if kind == 'sphere':          # bumpy sphere, the bench's synthetic shape
        u = rng.standard_normal((n, 3)); u /= np.linalg.norm(u, axis=1, keepdims=True)
        r = 1 + 0.15 * np.sin(3 * np.arctan2(u[:, 1], u[:, 0])) * u[:, 2]
        x = u * r[:, None]
    elif kind == 'two_parts':     # two far-apart blobs: the kNN graph is disconnected, distances stay infinite
        x = rng.standard_normal((n, 3)) * 0.1
        x[n // 3:] += np.array([5.0, 0.0, 0.0])
    elif kind == 'plane':         # flat patch with a fold
        x = rng.uniform(-1, 1, (n, 3)); x[:, 2] = np.abs(x[:, 0]) * 0.5
    else:
        raise ValueError(kind)
    return x.astype(np.float32)


def main():
    assert fps.have_reference(), 'build oracle/_ref/libref_fps.so first (make -C oracle)'
    rng = np.random.default_rng(20240611)
    out = {}
    cases = [('sphere', 400, 128), ('sphere', 2048, 512), ('two_parts', 300, 64), ('plane', 700, 200), ('sphere', 40, 64)]
    for c, (kind, n, m) in enumerate(cases):
        x = cloud(kind, n, rng)
        out[f'pos_{c}'] = x
        out[f'samples_{c}'] = fps.reference_fps(x, m)
        out[f'knn_{c}'] = fps.reference_knn(x)[:, 1:].astype(np.int32) if n <= 700 else np.zeros((0, 10), np.int32)
    out['kinds'] = np.array([k for k, _, _ in cases])
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'geodesic_fps.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
