"""CPU: the reference's public geometry helpers (deltaconv_b200/geometry/helpers.py, plain torch) against the
outputs of the unmodified reference stored in tests/golden."""
import pytest
import torch

from helpers import load, t, rel_err

GEOM = ['geom_normals_k10', 'geom_ragged_k20', 'geom_estimated_k16', 'geom_unnormalized_shape']


@pytest.mark.parametrize('name', GEOM)
def test_helpers_match_reference(name):
    from deltaconv_b200.geometry import helpers as H
    g, _ = load(name)
    dt = torch.float64
    pos, batch, k = t(g.pos, dt), t(g.batch), int(g.k)
    ei = t(g.ref64_edge_index)
    normal, xb, yb = t(g.ref64_normal), t(g.ref64_xb), t(g.ref64_yb)
    coords = H.coords_projected(pos, normal, xb, yb, ei, k)
    assert (coords - t(g.ref64_coords)).abs().max() < 1e-12
    assert (H.coords_projected(pos, normal, xb, yb, ei) - coords).abs().max() == 0          # k derived from edge_index
    dist = (pos[ei[1]] - pos[ei[0]]).norm(dim=1)
    w = H.gaussian_weights(dist, k, batch, float(g.width))
    assert (w - t(g.ref64_weights)).abs().max() < 1e-12
    lam_shape = None if float(g.lam_shape) < 0 else float(g.lam_shape)
    if lam_shape is None:
        wls = surf = H.weighted_least_squares(coords, w, k, float(g.lam))
    else:
        wls, surf = H.weighted_least_squares(coords, w, k, float(g.lam), lam_shape)
    assert rel_err(wls, t(g.ref64_wls)) < 1e-9
    M = H.fit_vector_mapping(pos, normal, xb, yb, ei, surf, coords)
    assert (M - t(g.ref64_mapping)).abs().max() < 1e-9


def test_gaussian_weights_without_batch():
    from deltaconv_b200.geometry import helpers as H
    d = torch.rand(50 * 8)
    w = H.gaussian_weights(d, 8).view(50, 8)
    assert torch.allclose(w.sum(1), torch.ones(50), atol=1e-6)
