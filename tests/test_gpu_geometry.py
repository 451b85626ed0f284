"""GPU parity: kNN, frames, operator build and operator application of the CUDA library against
the oracle (CPU, fp64) and the reference-generated golden fixtures.  Run with `-m gpu`."""
import pytest
import torch

from helpers import load, t, row_scaled_err, rel_err, check_frames, OPERATOR_ROW_REL
from oracle import geometry as og

pytestmark = pytest.mark.gpu

GEOM = ['geom_normals_k10', 'geom_ragged_k20', 'geom_estimated_k16', 'geom_unnormalized_shape']
DEV = 'cuda'


def dc():
    import deltaconv_b200
    return deltaconv_b200


def synth(B, P, cfg=2, **kw):
    return dc().utils.synthetic_batch(B, P, config=cfg, **kw)


# ----------------------------------------------------------------------------------------------
# kNN: bit-exact indices (north_star: "kNN indices must be bit-exact")
# ----------------------------------------------------------------------------------------------

@pytest.mark.parametrize('sizes,k', [([128, 128], 10), ([50, 64, 80], 20), ([1024] * 8, 20), ([300, 1, 77, 2048], 1),
                                     ([700, 900], 40), ([150, 200], 70), ([5000, 300], 12), ([2048] * 2, 30),
                                     ([1024] * 4, 33), ([2048, 1000], 40), ([1024, 333], 64), ([4500, 64], 64), ([3000], 48)])
def test_knn_bit_exact(sizes, k):
    sizes = [max(s, k) for s in sizes]
    d = synth(len(sizes), 0, cfg=7, sizes=sizes)
    want = og.knn_ell(d.pos, k, d.batch)
    ei = dc().geometry.knn_graph(d.pos.to(DEV), k, d.batch.to(DEV), loop=True, flow='target_to_source')
    assert ei.dtype == torch.int64 and tuple(ei.shape) == (2, d.pos.size(0) * k)
    assert torch.equal(ei.cpu(), og.edge_index_from_ell(want))
    assert torch.equal(ei._dc_graph.idx.cpu().long(), want)


def test_knn_golden_graphs_and_flows():
    for name in GEOM:
        g, _ = load(name)
        pos, batch, k = t(g.pos).to(DEV), t(g.batch).to(DEV), int(g.k)
        ei = dc().geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
        assert torch.equal(ei.cpu(), t(g.ref32_edge_index))
        flipped = dc().geometry.knn_graph(pos, k, batch, loop=True, flow='source_to_target')
        assert torch.equal(flipped.cpu(), t(g.ref32_edge_index).flip(0))
        noloop = dc().geometry.knn_graph(pos, k - 1, batch, loop=False, flow='target_to_source')
        assert torch.equal(noloop.cpu()[1].view(-1, k - 1), t(g.ref32_edge_index)[1].view(-1, k)[:, 1:])


def test_knn_exact_ties_resolve_to_lower_index():
    # duplicated points and a regular lattice produce exact distance ties
    base = torch.stack(torch.meshgrid(*[torch.arange(6.)] * 3, indexing='ij'), -1).reshape(-1, 3)
    pos = torch.cat([base, base[:40]])
    want = og.knn_ell(pos, 16)
    got = dc().ops.knn(pos.to(DEV), 16).idx
    assert torch.equal(got.cpu().long(), want)


def test_knn_cloud_smaller_than_k_pads_with_sentinel():
    d = synth(2, 0, cfg=5, sizes=[5, 40])
    N = 45
    idx = dc().ops.knn(d.pos.to(DEV), 8, d.batch.to(DEV)).idx.cpu()
    assert (idx[:5, 5:] == N).all() and (idx[:5, :5] < 5).all() and (idx[5:] < N).all()
    assert torch.equal(idx[5:].long(), og.knn_ell(d.pos[5:], 8) + 5)


def test_knn_is_deterministic_and_full_size():
    d = synth(32, 1024)
    pos, batch = d.pos.to(DEV), d.batch.to(DEV)
    a = dc().ops.knn(pos, 20, batch).idx
    b = dc().ops.knn(pos, 20, batch).idx
    assert torch.equal(a, b)
    N = pos.size(0)
    assert torch.equal(a[:, 0].long().cpu(), torch.arange(N))                # self first
    own = (a.long() // 1024) == (torch.arange(N, device=DEV)[:, None] // 1024)
    assert own.all()                                                          # clouds never mix
    d2 = (pos[a.long()] - pos[:, None]).pow(2).sum(-1)
    assert (d2[:, 1:] >= d2[:, :-1] - 1e-6).all()                             # sortedness


def test_edge_index_roundtrip_and_foreign_edge_index():
    d = synth(2, 200, cfg=9)
    want = og.knn_ell(d.pos, 9, d.batch)
    ei = og.edge_index_from_ell(want).to(DEV)                                 # no _dc_graph attached
    g = dc().ops.graph_of(ei, 400)
    assert torch.equal(g.idx.cpu().long(), want)
    assert torch.equal(g.edge_index().cpu(), ei.cpu())


def test_transposed_structure():
    d = synth(3, 0, cfg=4, sizes=[100, 257, 64])
    g = dc().ops.knn(d.pos.to(DEV), 11, d.batch.to(DEV))
    in_ptr, in_edge = (x.cpu().long() for x in g.transpose())
    idx = g.idx.cpu().long().reshape(-1)
    N = g.N
    order = torch.argsort(idx * (N * 11) + torch.arange(idx.numel()), stable=True)
    assert torch.equal(in_edge[:idx.numel()], order)
    assert torch.equal(in_ptr, torch.cat([torch.zeros(1, dtype=torch.long), torch.bincount(idx, minlength=N).cumsum(0)]))


# ----------------------------------------------------------------------------------------------
# frames
# ----------------------------------------------------------------------------------------------

def test_tangent_basis_matches_reference():
    g, _ = load('geom_normals_k10')
    xb, yb = dc().geometry.build_tangent_basis(t(g.input_normal).to(DEV))
    assert (xb.cpu() - t(g.ref64_xb)).abs().max() < 1e-6 and (yb.cpu() - t(g.ref64_yb)).abs().max() < 1e-6
    n = torch.nn.functional.normalize(torch.randn(5000, 3), dim=1)
    n[:4] = torch.tensor([[1., 0, 0], [0, 1., 0], [0, 0, 1.], [-1., 0, 0]])
    xb, yb = dc().geometry.build_tangent_basis(n.to(DEV))
    check_frames(n, xb.cpu(), yb.cpu(), 1e-5)


def test_estimate_basis_gauge_invariant():
    g, _ = load('geom_estimated_k16')
    pos, batch = t(g.pos).to(DEV), t(g.batch).to(DEV)
    ei10 = dc().geometry.knn_graph(pos, 10, batch, loop=True, flow='target_to_source')
    assert torch.equal(ei10.cpu(), t(g.ref32_edge_index10))
    n, xb, yb = (a.cpu() for a in dc().geometry.estimate_basis(pos, ei10, orientation=pos))
    check_frames(n, xb, yb, 1e-5)
    assert ((n.double() * t(g.ref64_normal)).sum(1) - 1).abs().max() < 1e-5
    assert ((xb.double() * t(g.ref64_xb)).sum(1).abs() - 1).abs().max() < 1e-4
    # a rotated plane (reference test: test_grad_div_mls.py:27-55): |n . n_gt| = 1
    q = torch.linalg.qr(torch.randn(3, 3)).Q
    plane = torch.cat([torch.rand(800, 2) * 2 - 1, torch.zeros(800, 1)], 1) @ q.T
    ei = dc().geometry.knn_graph(plane.to(DEV), 10, loop=True, flow='target_to_source')
    n, xb, yb = (a.cpu() for a in dc().geometry.estimate_basis(plane.to(DEV), ei, k=10))
    check_frames(n, xb, yb, 1e-5)
    assert ((n @ q[:, 2]).abs() - 1).abs().max() < 1e-4


# ----------------------------------------------------------------------------------------------
# operator build: entries within 1e-5 (row-scaled) of the reference in fp64
# ----------------------------------------------------------------------------------------------

def _build_from_golden(g, debug=False):
    pos, batch, k = t(g.pos).to(DEV), t(g.batch).to(DEV), int(g.k)
    ei = dc().geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
    normal, xb, yb = (t(getattr(g, f'ref64_{n}')).float().to(DEV) for n in ('normal', 'xb', 'yb'))
    kw = dict(kernel_width=float(g.width), regularizer=float(g.lam), normalized=bool(int(g.normalized)),
              shape_regularizer=None if float(g.lam_shape) < 0 else float(g.lam_shape))
    if debug:
        return dc().geometry.operator_parts(pos, normal, xb, yb, ei, batch, **kw)
    return dc().geometry.build_grad_div(pos, normal, xb, yb, ei, batch, **kw)


@pytest.mark.parametrize('name', GEOM)
def test_operator_entries_vs_reference(name):
    g, _ = load(name)
    grad, div = _build_from_golden(g)
    N = g.pos.shape[0]
    assert grad.size(0) == 2 * N and grad.size(1) == N and div.size(0) == N and div.size(1) == 2 * N
    gr, gc, gv = (a.cpu() for a in grad.coo())
    dr, dcol, dv = (a.cpu() for a in div.coo())
    assert torch.equal(gr, t(g.ref64_grad_row)) and torch.equal(gc, t(g.ref64_grad_col))
    assert torch.equal(dr, t(g.ref64_div_row)) and torch.equal(dcol, t(g.ref64_div_col))
    assert row_scaled_err(gv, t(g.ref64_grad_val), gr, 2 * N) < OPERATOR_ROW_REL
    assert row_scaled_err(dv, t(g.ref64_div_val), dr, N) < OPERATOR_ROW_REL
    parts = _build_from_golden(g, debug=True)
    assert (parts['coords'].cpu() - t(g.ref64_coords)).abs().max() < 1e-6
    assert (parts['weights'].cpu() - t(g.ref64_weights)).abs().max() < 1e-6
    assert (parts['mapping'].cpu() - t(g.ref64_mapping)).abs().max() < 2e-5


@pytest.mark.parametrize('B,P,k,lam', [(8, 1024, 20, 1e-3), (4, 2048, 30, 1e-3), (2, 1024, 40, 1e-2), (3, 500, 70, 1e-3)])
def test_operator_entries_vs_oracle_fp64(B, P, k, lam):
    d = synth(B, P, cfg=1)
    idx = og.knn_ell(d.pos, k, d.batch)
    nd = d.norm.double()
    xb, yb = og.tangent_basis_from_normal(nd)
    G64, D64 = og.fit_operators(d.pos.double(), nd, xb, yb, idx, d.batch, 1.0, lam)
    pos, batch, normal = d.pos.to(DEV), d.batch.to(DEV), d.norm.to(DEV)
    ei = dc().geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
    xg, yg = dc().geometry.build_tangent_basis(normal)
    grad, div = dc().geometry.build_grad_div(pos, normal, xg, yg, ei, batch, regularizer=lam)
    assert torch.equal(grad.graph.idx.cpu().long(), idx)
    for ours, ref in ((grad.G, G64), (div.D, D64)):
        scale = ref.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-300)
        assert ((ours.cpu().double() - ref).abs() / scale).max() < OPERATOR_ROW_REL


def test_build_without_batch_and_sentinel_padding():
    g, _ = load('geom_normals_k10')
    N, k = g.pos.shape[0], int(g.k)
    pos = t(g.pos)[:128].to(DEV)                               # first cloud only, batch=None
    normal = t(g.input_normal)[:128].to(DEV)
    xb, yb = dc().geometry.build_tangent_basis(normal)
    idx = og.knn_ell(pos.cpu(), k)
    idx[::5, -2:] = 128                                        # sentinel slots
    ei = og.edge_index_from_ell(idx).to(DEV)
    grad, div = dc().geometry.build_grad_div(pos, normal, xb, yb, ei, regularizer=1e-3)
    Go, Do = og.fit_operators(pos.cpu().double(), normal.cpu().double(), xb.cpu().double(), yb.cpu().double(), idx,
                              None, 1.0, 1e-3)
    for ours, ref in ((grad.G, Go), (div.D, Do)):
        scale = ref.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-300)
        assert ((ours.cpu().double() - ref).abs() / scale).max() < OPERATOR_ROW_REL
    x = torch.randn(128, 7)
    assert rel_err((grad @ x.to(DEV)).cpu(), og.apply_grad(idx, Go, x.double())) < 1e-5


# ----------------------------------------------------------------------------------------------
# operator application, forward and backward
# ----------------------------------------------------------------------------------------------

def _ops_and_oracle(B=3, P=300, k=12, cfg=6):
    d = synth(B, P, cfg=cfg)
    idx = og.knn_ell(d.pos, k, d.batch)
    nd = d.norm.double()
    xb, yb = og.tangent_basis_from_normal(nd)
    G64, D64 = og.fit_operators(d.pos.double(), nd, xb, yb, idx, d.batch)
    pos, batch, normal = d.pos.to(DEV), d.batch.to(DEV), d.norm.to(DEV)
    ei = dc().geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
    xg, yg = dc().geometry.build_tangent_basis(normal)
    grad, div = dc().geometry.build_grad_div(pos, normal, xg, yg, ei, batch)
    # apply the fp64 oracle with the *CUDA-built* entries so only the SpMM is under test
    return grad, div, idx, grad.G.cpu().double(), div.D.cpu().double()


@pytest.mark.parametrize('C', [1, 3, 5, 16, 64, 100, 128, 256])
def test_spmm_forward(C):
    grad, div, idx, G, D = _ops_and_oracle()
    N = idx.size(0)
    x, v = torch.randn(N, C), torch.randn(2 * N, C)
    geo = dc().geometry
    assert rel_err((grad @ x.to(DEV)).cpu(), og.apply_grad(idx, G, x.double())) < 2e-6
    assert rel_err((div @ v.to(DEV)).cpu(), og.apply_div(idx, D, v.double())) < 2e-6
    assert rel_err(geo.curl(v.to(DEV), div).cpu(), og.curl(idx, D, v.double())) < 2e-6
    assert rel_err(geo.laplacian(x.to(DEV), grad, div).cpu(), og.laplacian(idx, G, D, x.double())) < 5e-6
    assert rel_err(geo.hodge_laplacian(v.to(DEV), grad, div).cpu(), og.hodge_laplacian(idx, G, D, v.double())) < 5e-6
    assert rel_err(geo.norm(v.to(DEV)).cpu(), og.vec_norm(v.double())) < 1e-6
    assert torch.equal(geo.J(v.to(DEV)).cpu(), og.rot90(v))
    assert torch.equal(geo.I_J(v.to(DEV)).cpu(), og.with_rot90(v))


def test_spmm_on_column_slices_of_a_wider_buffer():
    grad, div, idx, G, D = _ops_and_oracle()
    N = idx.size(0)
    wide = torch.randn(N, 48, device=DEV)
    x = wide[:, 16:32]
    assert rel_err((grad @ x).cpu(), og.apply_grad(idx, G, x.cpu().double())) < 2e-6


@pytest.mark.parametrize('C', [1, 6, 64])
def test_spmm_backward(C):
    grad, div, idx, G, D = _ops_and_oracle()
    N = idx.size(0)
    ops = dc().ops
    x0, v0 = torch.randn(N, C), torch.randn(2 * N, C)
    wv, wx = torch.randn(2 * N, C), torch.randn(N, C)
    # grad @ x
    x = x0.to(DEV).requires_grad_(True)
    ((grad @ x) * wv.to(DEV)).sum().backward()
    xo = x0.double().requires_grad_(True)
    (og.apply_grad(idx, G, xo) * wv.double()).sum().backward()
    assert rel_err(x.grad.cpu(), xo.grad) < 2e-6
    # div @ v and curl
    for fn_gpu, fn_cpu in ((lambda q: div @ q, lambda q: og.apply_div(idx, D, q)),
                           (lambda q: dc().geometry.curl(q, div), lambda q: og.curl(idx, D, q))):
        v = v0.to(DEV).requires_grad_(True)
        (fn_gpu(v) * wx.to(DEV)).sum().backward()
        vo = v0.double().requires_grad_(True)
        (fn_cpu(vo) * wx.double()).sum().backward()
        assert rel_err(v.grad.cpu(), vo.grad) < 2e-6
    # the fused concatenations DeltaConv uses
    x = x0.to(DEV).requires_grad_(True)
    v = v0.to(DEV).requires_grad_(True)
    cat, dcu = ops.scalar_cat(x, v, div)
    x_new = torch.randn(N, C + 3)
    xn = x_new.to(DEV).requires_grad_(True)
    vcat = ops.vector_cat(v, dcu, xn, grad)
    w1, w2 = torch.randn(N, 4 * C), torch.randn(2 * N, 3 * C + 3)
    ((cat * w1.to(DEV)).sum() + (vcat * w2.to(DEV)).sum()).backward()
    xo, vo, xno = x0.double().requires_grad_(True), v0.double().requires_grad_(True), x_new.double().requires_grad_(True)
    dv_, cv_ = og.apply_div(idx, D, vo), og.curl(idx, D, vo)
    cat_o = torch.cat([xo, dv_, cv_, og.vec_norm(vo)], 1)
    hodge = -(og.apply_grad(idx, G, dv_) + og.rot90(og.apply_grad(idx, G, cv_)))
    vcat_o = torch.cat([vo, hodge, og.apply_grad(idx, G, xno)], 1)
    ((cat_o * w1.double()).sum() + (vcat_o * w2.double()).sum()).backward()
    assert rel_err(cat.detach().cpu(), cat_o.detach()) < 2e-6 and rel_err(vcat.detach().cpu(), vcat_o.detach()) < 5e-6
    assert rel_err(x.grad.cpu(), xo.grad) < 2e-6
    assert rel_err(v.grad.cpu(), vo.grad) < 5e-6
    assert rel_err(xn.grad.cpu(), xno.grad) < 2e-6


# ----------------------------------------------------------------------------------------------
# full-size (BASELINE config 2: 32 x 1024, k=20) size-independent properties
# ----------------------------------------------------------------------------------------------

def test_full_size_properties_and_determinism():
    d = synth(32, 1024)
    pos, batch, normal = d.pos.to(DEV), d.batch.to(DEV), d.norm.to(DEV)
    geo = dc().geometry

    def build():
        ei = geo.knn_graph(pos, 20, batch, loop=True, flow='target_to_source')
        xb, yb = geo.build_tangent_basis(normal)
        return geo.build_grad_div(pos, normal, xb, yb, ei, batch)
    grad, div = build()
    grad2, div2 = build()
    assert torch.equal(grad.G, grad2.G) and torch.equal(div.D, div2.D)          # bitwise run-to-run
    N = pos.size(0)
    assert not grad.G.isnan().any() and not div.D.isnan().any()
    ones = torch.ones(N, 1, device=DEV)
    assert (grad @ ones).abs().max() < 1e-2                                     # grad of a constant (test_grad_div_mls.py:350)
    assert geo.laplacian(ones, grad, div).abs().mean() < 1e-2                   # :353
    # per-cloud infinity norm is 1 after normalisation (grad_div_mls.py:258-260)
    s = grad.G.abs().sum(1).norm(dim=1).view(32, 1024).amax(1)
    assert (s - 1).abs().max() < 1e-5
    # linearity and curl(grad f) ~ 0 (:359)
    a, b = torch.randn(N, 64, device=DEV), torch.randn(N, 64, device=DEV)
    lhs, rhs = grad @ (2 * a - 3 * b), 2 * (grad @ a) - 3 * (grad @ b)
    assert (lhs - rhs).abs().max() < 1e-4 * rhs.abs().max()
    assert geo.curl(grad @ pos[:, :1], div).pow(2).mean() < 1e-2
    # <grad x, w> == <x, grad^T w>: the transposed kernels are exact adjoints of the forward ones
    x = a.clone().requires_grad_(True)
    w = torch.randn(2 * N, 64, device=DEV)
    y = grad @ x
    (y * w).sum().backward()
    assert abs((y.detach() * w).sum().item() - (x.detach() * x.grad).sum().item()) < 1e-3 * (y.detach() * w).abs().sum().item()


@pytest.mark.parametrize('C', [4, 16, 64, 200])
def test_staged_spmm_is_bitwise_equal_to_the_gather_kernels(C):
    """Shared-memory staged SpMMs (cloud layout known) vs L2-gather SpMMs: same arithmetic in the same order."""
    from deltaconv_b200 import _lib
    lib = _lib.load()
    grad, div, idx, G, D = _ops_and_oracle(B=3, P=300, k=12, cfg=6)
    assert grad.graph.max_cloud == 300 and grad.graph.cloud_ptr is not None
    N = idx.size(0)
    ops = dc().ops
    x0, v0 = torch.randn(N, C, device=DEV), torch.randn(2 * N, C, device=DEV)
    results = []
    for staged in (1, 0):
        lib.dc_set_spmm_staged(staged)
        try:
            x, v = x0.clone().requires_grad_(True), v0.clone().requires_grad_(True)
            cat, dcu = ops.scalar_cat(x, v, div)
            xn = (x * 2).contiguous()
            vcat = ops.vector_cat(v, dcu, xn, grad)
            y = grad @ x
            (cat.sum() * 0.5 + (vcat * vcat).sum() + (y * y).sum()).backward()
            results.append([t_.detach().clone() for t_ in (cat, vcat, y, x.grad, v.grad)])
        finally:
            lib.dc_set_spmm_staged(0)
    for a, b in zip(*results):
        assert torch.equal(a, b)


@pytest.mark.parametrize('sizes,k', [([5000, 7000], 20), ([300, 9000, 17, 4200], 10), ([6000], 40), ([20000], 16), ([4500, 4500], 70), ([8000], 33), ([3000, 5000], 64)])
def test_grid_knn_is_bit_identical_to_brute_force(sizes, k):
    """dc_knn_grid (uniform grid + ring search, used for clouds of more than 1536 points) returns exactly the graph of the
    brute-force kernel: indices and squared distances, ties by lower index.  Clouds: uniform volume, a thin shell, a planar
    sheet with duplicated points, a line — the grid must stay exact where its cells are badly filled."""
    from deltaconv_b200 import ops
    torch.manual_seed(11)
    parts = []
    for c, n in enumerate(sizes):
        kind = c % 4
        if kind == 0:
            p = torch.rand(n, 3) * torch.tensor([2.0, 1.0, 0.5])
        elif kind == 1:
            p = torch.nn.functional.normalize(torch.randn(n, 3), dim=1) * (1 + 0.001 * torch.randn(n, 1)) + 5.0
        elif kind == 2:
            p = torch.cat([torch.rand(n, 2), torch.zeros(n, 1)], 1)
            p[n // 2:] = p[:n - n // 2].clone()                          # exact duplicates: ties decided by index
        else:
            p = torch.linspace(0, 1, n)[:, None] * torch.tensor([[1.0, 2.0, -1.0]])
        parts.append(p)
    pos = torch.cat(parts).to(DEV)
    batch = torch.repeat_interleave(torch.arange(len(sizes)), torch.tensor(sizes)).to(DEV)
    out = {}
    for method in ('brute', 'grid'):
        ops.KNN_METHOD = method
        try:
            g = ops.knn(pos, k, batch, num_clouds=len(sizes), want_edge_index=True)
        finally:
            ops.KNN_METHOD = 'auto'
        out[method] = (g.idx.clone(), g.edge_index().clone())
    assert torch.equal(out['brute'][0], out['grid'][0])
    assert torch.equal(out['brute'][1], out['grid'][1])


def test_deferred_normalisation_and_knn_distances_give_the_same_operators():
    """geometry.build_grad_div stores the entries before the per-cloud normalisation plus one factor per point (applied by
    the operator kernels) and takes the mean edge lengths from the kNN's squared distances.  Against the plain build
    (normalised entries written by the library, neighbour positions gathered again): same entries to fp32 rounding, same
    operator applications forward and backward."""
    from deltaconv_b200 import ops
    d = synth(3, 0, cfg=11, sizes=[500, 130, 1024])
    pos, nrm, batch = d.pos.to(DEV), d.norm.to(DEV), d.batch.to(DEV)
    graph = ops.knn(pos, 16, batch, num_clouds=3)
    assert graph.d2 is not None and torch.isfinite(graph.d2).all()
    nb = graph.idx.long()
    assert torch.allclose(graph.d2, (pos[nb] - pos[:, None]).pow(2).sum(-1), rtol=1e-5, atol=1e-7)
    xb, yb = ops.tangent_basis(nrm)
    plain_graph = ops.EllGraph(graph.idx, graph.cloud_ptr, graph.B, graph.max_cloud)          # no distances attached
    G1, D1 = ops.build_operators(pos, nrm, xb, yb, plain_graph, graph.cloud_ptr, 3)
    G2, D2, scale = ops.build_operators(pos, nrm, xb, yb, graph, graph.cloud_ptr, 3, deferred=True)
    assert scale is not None and (scale > 0).all()
    row = G1.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-30)
    assert ((G2 * scale[:, None, None] - G1).abs() / row).max() < 2e-6
    row = D1.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-30)
    assert ((D2 * scale[:, None, None] - D1).abs() / row).max() < 2e-6
    ops1 = (ops.EllOperator('grad', plain_graph, G1, D1), ops.EllOperator('div', plain_graph, G1, D1))
    ops2 = (ops.EllOperator('grad', graph, G2, D2, scale), ops.EllOperator('div', graph, G2, D2, scale))
    torch.manual_seed(2)
    x0, v0 = torch.randn(pos.size(0), 8, device=DEV), torch.randn(2 * pos.size(0), 8, device=DEV)
    outs = []
    for grad, div in (ops1, ops2):
        x, v = x0.clone().requires_grad_(True), v0.clone().requires_grad_(True)
        cat, dcu = ops.scalar_cat(x, v, div)
        vcat = ops.vector_cat(v, dcu, x, grad)
        lap = dc().geometry.laplacian(x, grad, div)
        (cat.sum() + (vcat * vcat).sum() + (lap * lap).sum()).backward()
        outs.append((cat.detach(), vcat.detach(), lap.detach(), x.grad, v.grad))
    for a, b in zip(*outs):
        assert rel_err(b.cpu(), a.cpu()) < 5e-6
