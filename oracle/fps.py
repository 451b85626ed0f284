"""TEST INFRASTRUCTURE — ctypes doors to the two CPU checkers of geodesic farthest-point sampling:

  * `restated_*`: oracle/geodesic_fps.c, the C restatement (always built by `make -C oracle`);
  * `reference_*`: oracle/_ref/libref_fps.so, the reference's own deltaconv/cpp/sampling.cpp compiled from where it lies
    (only built where /root/reference exists; the built file travels to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_P = ctypes.c_void_p


def _load(path):
    return ctypes.CDLL(path) if os.path.exists(path) else None


_restated = _load(os.path.join(_HERE, 'liboracle_fps.so'))
_reference = _load(os.path.join(_HERE, '_ref', 'libref_fps.so'))


def have_restated():
    return _restated is not None


def have_reference():
    return _reference is not None


def _xyz(points):
    x = np.ascontiguousarray(np.asarray(points, dtype=np.float64))
    assert x.ndim == 2 and x.shape[1] == 3
    return x


def restated_fps(points, n_samples, start, want_dist=False):
    """samples [n_samples] int32 (and the distance vector after the pass from the last sample)"""
    x = _xyz(points)
    out = np.empty(n_samples, np.int32)
    dist = np.empty(x.shape[0], np.float64) if want_dist else None
    rc = _restated.oracle_geodesic_fps(_P(x.ctypes.data), x.shape[0], int(n_samples), int(start), _P(out.ctypes.data),
                                       _P(dist.ctypes.data) if want_dist else None)
    if rc != 0:
        raise ValueError('oracle_geodesic_fps: invalid arguments')
    return (out, dist) if want_dist else out


def restated_knn(points):
    """([n,10] int32 neighbours, -1 = none; [n,10] float64 edge lengths)"""
    x = _xyz(points)
    nbr = np.empty((x.shape[0], 10), np.int32)
    length = np.empty((x.shape[0], 10), np.float64)
    _restated.oracle_fps_knn(_P(x.ctypes.data), x.shape[0], _P(nbr.ctypes.data), _P(length.ctypes.data))
    return nbr, length


def reference_fps(points, n_samples):
    """constructGeodesicFPS of the reference; its random first sample is element 0 of the result"""
    x = _xyz(points)
    out = np.empty(n_samples, np.int32)
    _reference.ref_geodesic_fps(_P(x.ctypes.data), x.shape[0], int(n_samples), _P(out.ctypes.data))
    return out


def reference_knn(points, k=10):
    """generateKNN of the reference: [n, k+1] int64, the point itself in column 0"""
    x = _xyz(points)
    out = np.empty((x.shape[0], k + 1), np.int64)
    _reference.ref_generate_knn(_P(x.ctypes.data), x.shape[0], int(k), _P(out.ctypes.data))
    return out


def reference_dijkstra(points, source, dist, k=10):
    """computeDijkstra of the reference from `source` into a copy of `dist` (min-combined)"""
    x = _xyz(points)
    d = np.array(dist, dtype=np.float64, copy=True)
    _reference.ref_dijkstra(_P(x.ctypes.data), x.shape[0], int(k), int(source), _P(d.ctypes.data))
    return d
