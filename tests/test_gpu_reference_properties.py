"""The reference's own property tests (test/geometry/test_grad_div_mls.py, test/nn/test_nonlin.py,
test/geometry/test_operators.py, test/geometry/test_fps.py), re-run against the CUDA path.  Same constructions and thresholds as the
reference unless a comment says otherwise.  Run with `-m gpu`."""
import pytest
import torch
import torch.linalg as LA

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def dc():
    import deltaconv_b200
    return deltaconv_b200


def quadric_surface(N, seed):
    """f(x,y) = sum c_i * [1, x, y, x^2, xy, y^2] over [-1,1]^2 with analytic frames (test_grad_div_mls.py:289-327)."""
    torch.manual_seed(seed)
    coords = torch.rand(N, 2) * 2 - 1
    x, y = coords.T
    B = torch.stack([torch.ones_like(x), x, y, x * x, x * y, y * y], 1)
    c = torch.rand(6)
    f = (B * c).sum(1, keepdim=True)
    pos = torch.cat([coords, f], 1)
    dfdx = torch.stack([torch.ones_like(x), torch.zeros_like(x), c[1] + 2 * c[3] * x + c[4] * y], 1)
    dfdy = torch.stack([torch.zeros_like(x), torch.ones_like(x), c[2] + c[4] * x + 2 * c[5] * y], 1)
    normal = torch.linalg.cross(dfdx, dfdy, dim=1)
    normal = normal / LA.norm(normal, dim=1, keepdim=True).clamp(1e-8)
    xb = dfdx / LA.norm(dfdx, dim=1, keepdim=True).clamp(1e-8)
    yb = torch.linalg.cross(normal, xb, dim=1)
    return pos, f, normal, xb, yb


def test_build_grad_div_properties():
    """test_build_grad_div (test_grad_div_mls.py:278-401), same regularizer (1e-8) and thresholds as the reference."""
    geo = dc().geometry
    N, k = 1000, 20
    pos, f, normal, xb, yb = (a.to(DEV) for a in quadric_surface(N, 42))
    ei = geo.knn_graph(pos, k, loop=True, flow='target_to_source')
    grad, div = geo.build_grad_div(pos, normal, xb, yb, ei, regularizer=1e-8, normalized=False)
    assert grad.size(0) == 2 * N and grad.size(1) == N and div.size(0) == N and div.size(1) == 2 * N
    assert (grad @ torch.rand(N, 1, device=DEV)).isnan().sum() == 0
    assert (div @ torch.rand(2 * N, 1, device=DEV)).isnan().sum() == 0
    ones = pos.new_ones(N, 1)
    assert torch.allclose(grad @ ones, pos.new_zeros(2 * N, 1), atol=1e-2)                    # 3a
    assert geo.laplacian(ones, grad, div).abs().mean() < 1e-2                                   # 3b
    assert LA.norm(geo.laplacian(torch.rand(N, 1, device=DEV), grad, div), ord=1) > 0           # 3c
    assert geo.curl(grad @ pos[:, 0:1], div).pow(2).mean() < 1e-2                               # 3d
    assert geo.curl(grad @ pos[:, 0:1], div).pow(2).median() < 1e-2
    assert (div @ geo.J(grad @ pos[:, 0:1])).pow(2).mean() < 1e-2                               # 3e
    gx, gy = (grad @ f).view(N, 2).T                                                            # 4
    assert torch.allclose(gx, xb[:, 2], atol=1e-2) and torch.allclose(gy, yb[:, 2], atol=1e-2)
    mc = geo.laplacian(pos, grad, div)
    assert torch.allclose(-geo.batch_dot(mc, normal), LA.norm(mc, dim=1, keepdims=True), atol=1e-2)
    # 5. gauge equivariance under a random in-plane rotation of the frames
    xb_r = geo.rotate_around(xb, normal, torch.rand(N, device=DEV) * 2 * torch.pi)
    yb_r = torch.linalg.cross(normal, xb_r, dim=1)
    G, Dv = geo.build_grad_div(pos, normal, xb, yb, ei, regularizer=1e-8)
    G_r, D_r = geo.build_grad_div(pos, normal, xb_r, yb_r, ei, regularizer=1e-8)
    x = torch.rand(N, 1, device=DEV)

    def ambient(op, field, fx, fy):
        gx, gy = (op @ field).view(-1, 2).T
        return gx[:, None] * fx + gy[:, None] * fy
    # 5a/5b: grad x as an ambient vector, and div grad x, in both gauges.  The regulariser is not rotation invariant,
    # so even the fp64 reference arithmetic shows a small, sample-dependent gauge difference (the reference asserts
    # 1e-3 for its own seed); require that the CUDA path adds nothing to it.
    from oracle import geometry as og
    idx = ei._dc_graph.idx.cpu().long()
    p64, n64, x64 = pos.cpu().double(), normal.cpu().double(), x.cpu().double()
    lap, amb = [], []
    for fx, fy in ((xb, yb), (xb_r, yb_r)):
        fx64, fy64 = fx.cpu().double(), fy.cpu().double()
        Go, Do = og.fit_operators(p64, n64, fx64, fy64, idx, None, 1.0, 1e-8)
        gv = og.apply_grad(idx, Go, x64)
        lap.append(og.apply_div(idx, Do, gv))
        g2 = gv.view(-1, 2)
        amb.append(g2[:, :1] * fx64 + g2[:, 1:] * fy64)
    inherent_a = (amb[0] - amb[1]).abs().max().item()
    ours_a = (ambient(G, x, xb, yb) - ambient(G_r, x, xb_r, yb_r)).abs().max().item()
    assert ours_a <= inherent_a + 1e-4 and ours_a < 2e-2, (ours_a, inherent_a)
    inherent = (lap[0] - lap[1]).abs().max().item()
    ours = (Dv @ (G @ x) - D_r @ (G_r @ x)).abs().max().item()
    assert ours <= inherent + 1e-4 and ours < 5e-3, (ours, inherent)
    assert ((Dv @ (G @ x)).cpu().double() - lap[0]).abs().max() < 1e-5


def test_coords_weights_and_mapping():
    """coords_projected recovers (x, y) on a paraboloid patch (test_grad_div_mls.py:58-84), gaussian weights are
    normalised and decrease with distance (:87-104), and the vector mapping carries neighbour frames into the
    centre frame (:158-275) — read out of the fused build kernel through `operator_parts`."""
    geo = dc().geometry
    N, k = 600, 16
    pos, f, normal, xb, yb = (a.to(DEV) for a in quadric_surface(N, 7))
    ei = geo.knn_graph(pos, k, loop=True, flow='target_to_source')
    parts = geo.operator_parts(pos, normal, xb, yb, ei, regularizer=1e-5)
    row, col = ei
    d = pos[col] - pos[row]
    dt = d - normal[row] * (d * normal[row]).sum(1, keepdim=True)
    want = torch.stack([(dt * xb[row]).sum(1), (dt * yb[row]).sum(1)], 1)
    assert torch.allclose(parts['coords'], want, atol=1e-6)
    w = parts['weights'].view(N, k)
    assert torch.allclose(w.sum(1), torch.ones(N, device=DEV), atol=1e-5)
    dist = d.norm(dim=1).view(N, k)
    order = dist.argsort(dim=1)
    assert (w.gather(1, order).diff(dim=1) <= 1e-7).all()                                      # monotone in distance
    assert (w[:, 0] >= w.max(1).values - 1e-7).all()                                            # self (distance 0) heaviest
    # mapping: identity on the centre slot, and close to a rotation for near neighbours on a smooth patch
    M = parts['mapping'].view(N, k, 2, 2)
    eye = torch.eye(2, device=DEV).expand(N, 2, 2)
    dev0 = (M[:, 0] - eye).abs().amax(dim=(1, 2))          # = fitted slope at the centre: ~0 in the interior, larger on the boundary
    assert dev0.median() < 1e-3 and dev0.max() < 0.2
    assert (torch.linalg.det(M[:, 1]) - 1).abs().median() < 5e-2


def test_vector_nonlin_behaviour():
    """test_vectornonlin (test/nn/test_nonlin.py:40-88): identity at zero bias; norms thresholded; directions kept."""
    nn = dc().nn
    N, C = 1024, 16
    ang = torch.rand(N, C, device=DEV) * 2 * torch.pi
    nrm = torch.rand(N, C, device=DEV) * 5
    v = torch.stack([nrm * torch.cos(ang), nrm * torch.sin(ang)], 1).view(-1, C)
    m = nn.VectorNonLin(C).to(DEV)
    assert torch.allclose(m(v), v, atol=1e-6)                                                   # zero bias => identity
    with torch.no_grad():
        m.bias.fill_(-2.5)
    out = m(v).view(N, 2, C)
    out_n = out.norm(dim=1)
    assert torch.allclose(out_n, (nrm - 2.5).clamp(min=0), atol=1e-5)
    keep = out_n > 1e-3
    cosang = (out * v.view(N, 2, C)).sum(1) / (out_n * nrm).clamp(1e-12)
    assert torch.allclose(cosang[keep], torch.ones_like(cosang[keep]), atol=1e-5)               # direction preserved
    bn = nn.VectorNonLin(C, batchnorm=nn.BatchNorm1d(C)).to(DEV).train()
    out = bn(v).view(N, 2, C).norm(dim=1)
    assert (out >= 0).all() and (out == 0).float().mean() > 0.2                                 # BN centres the norms, ReLU zeroes ~half


def test_batchnorm_wrapper_behaviour():
    """test_batchnorm (test/nn/test_nonlin.py:7-37): constant input -> 0, affine-invariance of the normalised output."""
    nn = dc().nn
    bn = nn.BatchNorm1d(8).to(DEV).train()
    assert repr(bn) == 'BatchNorm1d(8)'
    x = torch.full((100, 8), 3.0, device=DEV)
    assert torch.allclose(bn(x), torch.zeros_like(x), atol=1e-3)
    y = torch.randn(500, 8, device=DEV)
    assert torch.allclose(bn(y), bn(2.5 * y + 1.0), atol=1e-4)


def test_geodesic_fps_like_the_reference_test():
    """test/geometry/test_fps.py: every point can be sampled exactly once, the requested number comes back, wrong inputs
    raise ValueError — same calls, on the GPU sampler behind the reference's signature."""
    import numpy as np
    geodesic_fps = dc().geometry.geodesic_fps
    n, n_samples = 1024, 512
    rng = np.random.default_rng(0)
    pos = rng.standard_normal((n, 3))                     # float64, as np.random.randn gives the reference
    samples1 = geodesic_fps(pos, n)
    assert samples1.shape[0] == n
    assert np.unique(samples1).shape[0] == n
    samples2 = geodesic_fps(pos, n_samples)
    assert samples2.shape[0] == n_samples
    assert np.unique(samples2).shape[0] == n_samples
    with pytest.raises(ValueError):
        geodesic_fps(torch.rand(n, 3), n)
    with pytest.raises(ValueError):
        geodesic_fps(rng.standard_normal((n, 2, 3)), n)
