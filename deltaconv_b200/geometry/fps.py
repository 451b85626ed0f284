"""`geodesic_fps(points, n_samples)` with the reference's signature, checks and return value (deltaconv/geometry/fps.py:5-19),
computed on the GPU by dc_geodesic_fps instead of the pybind11 module `deltaconv_bindings` (deltaconv/cpp/core.cpp:17-27).
For many clouds at once use `deltaconv_b200.ops.geodesic_fps` (or the `GeodesicFPS` transform on a batch): one launch samples
them all, which is where the GPU pays off."""
import warnings

import numpy as np
import torch

from .. import ops

__all__ = ['geodesic_fps']


def geodesic_fps(points, n_samples, start=None):
    """Indices [n_samples] (int32 numpy, like the binding's squeezed Eigen vector) of a geodesic farthest-point sampling of
    `points` (V,3).  `start`: first sample; None draws one uniformly, as the reference does (sampling.cpp:33-40)."""
    if type(points) is not np.ndarray:
        raise ValueError("`points` should be a numpy array")
    if (len(points.shape) != 2) or (points.shape[1] != 3):
        raise ValueError("`points` should have shape (V,3), shape is " + str(points.shape))
    if n_samples > points.shape[0]:
        warnings.warn("Number of samples is larger than number of points.")
    if points.shape[0] < 11:
        # geometry-central's kNearestNeighbors, knn.cpp:53
        raise RuntimeError("k+1 is greater than number of points")
    pos = torch.from_numpy(np.ascontiguousarray(points, dtype=np.float64)).cuda()
    idx = ops.geodesic_fps(pos, int(n_samples), start=None if start is None else [int(start)])
    return idx[0].to(torch.int32).cpu().numpy()
