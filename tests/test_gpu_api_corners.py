"""GPU tests of the public-API corners outside the shipped models' path ("What's missing" #5, ADVICE):
DeltaConv with aggr in {sum, mean, min}, operators that only support `@`, edge_index that is not a fixed-k graph,
VectorNonLin with another activation, knn_graph(loop=False) on coincident points, BatchNorm statistics of a column whose
mean dwarfs its spread, optimizer state loaded before the first step.  Run with `-m gpu`."""
import copy

import pytest
import torch

from helpers import rel_err
from oracle import geometry as og
from oracle import network as on

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def dc():
    import deltaconv_b200
    return deltaconv_b200


def _layer_inputs(N=600, k=12, C=8):
    D = dc()
    torch.manual_seed(3)
    data = D.utils.synthetic_batch(2, N // 2, config=6)
    pos, batch = data.pos.to(DEV), data.batch.to(DEV)
    ei = D.geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
    xb, yb = D.geometry.build_tangent_basis(data.norm.to(DEV))
    grad, div = D.geometry.build_grad_div(pos, data.norm.to(DEV), xb, yb, ei, batch)
    x = torch.randn(N, C, device=DEV)
    v = torch.randn(2 * N, C, device=DEV)
    return D, ei, grad, div, x, v


def _reference_layer(conv, x, v, grad_d, div_d, ei, aggr):
    """deltaconv.py:44-70 in fp64 on the CPU with dense operators and torch's scatter_reduce."""
    ref = on.DeltaConv(conv.in_channels, conv.out_channels, depth=1, centralized=conv.centralized, vector=conv.v_mlp is not None).double()
    ref.load_state_dict({k_: (t_.double() if t_.is_floating_point() else t_) for k_, t_ in conv.state_dict().items()})
    ref.train()
    x, v = x.double().cpu(), v.double().cpu()
    row, col = ei.cpu()
    N = x.size(0)
    feat = ref.s_mlp_max(x[col] - x[row]) if conv.centralized else ref.s_mlp_max(x)[col]
    red = {'sum': 'sum', 'mean': 'mean', 'min': 'amin', 'max': 'amax'}[aggr]
    x_max = torch.zeros(N, feat.size(1), dtype=torch.float64).scatter_reduce(0, row[:, None].expand_as(feat), feat, red, include_self=False)
    div_v = div_d @ v
    curl_v = -(div_d @ og.rot90(v))
    xo = x_max + ref.s_mlp(torch.cat([x, div_v, curl_v, og.vec_norm(v)], 1))
    hodge = -(grad_d @ div_v + og.rot90(grad_d @ curl_v))
    vo = ref.v_mlp(og.with_rot90(torch.cat([v, hodge, grad_d @ xo], 1)))
    return xo, vo


@pytest.mark.parametrize('aggr', ['sum', 'mean', 'min'])
@pytest.mark.parametrize('centralized', [False, True])
def test_deltaconv_other_aggregations(aggr, centralized):
    D, ei, grad, div, x, v = _layer_inputs()
    conv = D.nn.DeltaConv(8, 16, depth=1, centralized=centralized, aggr=aggr).to(DEV).train()
    xo, vo = conv(x, v, grad, div, ei)
    xr, vr = _reference_layer(conv, x, v, grad.to_dense().double().cpu(), div.to_dense().double().cpu(), ei, aggr)
    assert rel_err(xo.detach().cpu(), xr.detach()) < 1e-4
    assert rel_err(vo.detach().cpu(), vr.detach()) < 1e-4
    (xo.sum() + (vo * vo).sum()).backward()
    assert all(torch.isfinite(p.grad).all() for p in conv.parameters() if p.grad is not None)


def test_deltaconv_accepts_plain_matmul_operators_and_ragged_graphs():
    """grad / div given as dense matrices (anything with `@`), and an edge_index with a varying number of neighbours per
    point (the max aggregation of deltaconv.py:54 is defined for any graph): same result as the fused path restricted to the
    same edges."""
    D, ei, grad, div, x, v = _layer_inputs()
    conv = D.nn.DeltaConv(8, 16, depth=1).to(DEV).train()
    fused = [t_.detach() for t_ in copy.deepcopy(conv)(x, v, grad, div, ei)]
    dense = [t_.detach() for t_ in copy.deepcopy(conv)(x, v, grad.to_dense(), div.to_dense(), ei)]
    assert rel_err(dense[0].cpu(), fused[0].cpu()) < 1e-5 and rel_err(dense[1].cpu(), fused[1].cpu()) < 1e-5
    keep = torch.ones(ei.size(1), dtype=torch.bool, device=DEV)
    keep[5::7] = False                                   # drop every 7th edge: no longer a fixed-k graph
    keep[::12] = True                                    # (keep every self edge)
    ragged = ei[:, keep]
    xo, vo = copy.deepcopy(conv)(x, v, grad, div, ragged)
    xr, vr = _reference_layer(conv, x, v, grad.to_dense().double().cpu(), div.to_dense().double().cpu(), ragged, 'max')
    assert rel_err(xo.detach().cpu(), xr.detach()) < 1e-4 and rel_err(vo.detach().cpu(), vr.detach()) < 1e-4
    with pytest.raises(D._lib.DeltaConvLibraryError):
        D.geometry.build_grad_div(x[:, :3].contiguous(), x[:, :3].contiguous(), x[:, :3].contiguous(), x[:, :3].contiguous(), ragged)
    with pytest.raises(D._lib.DeltaConvLibraryError):     # a source_to_target graph is not silently read as ELL
        D.ops.graph_of(ei.flip(0).contiguous(), x.size(0))


@pytest.mark.parametrize('with_bn', [True, False])
def test_vector_nonlin_other_activations(with_bn):
    D = dc()
    torch.manual_seed(0)
    N, C = 300, 12
    v0 = torch.randn(2 * N, C)
    mod = D.nn.VectorNonLin(C, nonlin=torch.nn.Softplus(), batchnorm=D.nn.BatchNorm1d(C) if with_bn else None).to(DEV).train()
    with torch.no_grad():
        mod.bias.copy_(0.1 * torch.randn(C))
    ref = on.VectorNonLin(C, nonlin=torch.nn.Softplus(), batchnorm=on.BatchNorm1d(C) if with_bn else None).double().train()
    ref.load_state_dict({k_: (t_.double() if t_.is_floating_point() else t_) for k_, t_ in mod.state_dict().items()})
    v = v0.to(DEV).requires_grad_(True)
    out = mod(v)
    vr = v0.double().requires_grad_(True)
    want = ref(vr)
    assert rel_err(out.detach().cpu(), want.detach()) < 1e-5
    (out * out).sum().backward()
    (want * want).sum().backward()
    assert rel_err(v.grad.cpu(), vr.grad) < 1e-4


def test_knn_graph_without_self_loops_on_coincident_points():
    """torch_cluster removes the self edge by INDEX: with duplicated points the query is not necessarily in slot 0."""
    D = dc()
    torch.manual_seed(1)
    base = torch.randn(40, 3)
    pos = torch.cat([base, base[:10], base[:5]])          # points 40..49 duplicate 0..9, 50..54 duplicate 0..4
    ei = D.geometry.knn_graph(pos.to(DEV), 6, loop=False, flow='target_to_source').cpu()
    row, col = ei
    assert (row != col).all()                              # no self edge anywhere
    idx = col.view(-1, 6)
    want = og.knn_ell(pos, 7)                              # (d2, index) order, self included
    for i in range(pos.size(0)):
        expect = [j for j in want[i].tolist() if j != i][:6]
        assert idx[i].tolist() == expect, i
    flipped = D.geometry.knn_graph(pos.to(DEV), 6, loop=False, flow='source_to_target')
    assert torch.equal(flipped.cpu(), ei.flip(0)) and getattr(flipped, '_dc_graph', None) is None


def test_batchnorm_statistics_with_a_large_column_mean():
    """mean ~ 1e3, std ~ 1e-2: E[x^2] - mean^2 in fp32 would have no correct digit; the pivoted sums keep them."""
    D = dc()
    torch.manual_seed(5)
    M, C = 20000, 8
    x0 = 1000.0 + 0.01 * torch.randn(M, C)
    x0[:, 1] = -250.0 + 3.0 * torch.randn(M)
    bn = D.nn.BatchNorm1d(C).to(DEV).train()
    out = bn(x0.to(DEV))
    xd = x0.double()
    want = (xd - xd.mean(0)) / torch.sqrt(xd.var(0, unbiased=False) + 1e-5)
    assert rel_err(out.detach().cpu(), want) < 2e-3        # the fp32 inputs themselves carry 6e-5 / 1e-2 of relative noise
    assert rel_err(bn.bn.running_var.cpu(), 0.9 + 0.1 * xd.var(0, unbiased=True)) < 1e-4
    v = torch.zeros(2 * M, C)
    v[0::2] = x0
    nl = D.nn.VectorNonLin(C, batchnorm=D.nn.BatchNorm1d(C)).to(DEV).train()
    got = nl(v.to(DEV))
    n = x0.abs().double()
    gain = torch.relu((n - n.mean(0)) / torch.sqrt(n.var(0, unbiased=False) + 1e-5)) / n.clamp(min=1e-8)
    assert rel_err(got.detach().cpu()[0::2], xd * gain) < 2e-3


@pytest.mark.parametrize('kind', ['sgd', 'adam'])
def test_flat_optimizer_state_loaded_before_the_first_step(kind):
    """build -> load_state_dict -> train (the usual resume order) continues exactly where the saved run would have."""
    D = dc()
    from deltaconv_b200.parallel import FlatGradients, FlatSGD, FlatAdam
    from deltaconv_b200.utils.losses import calc_loss
    data = D.utils.synthetic_batch(4, 128, config=2, num_classes=5).to(DEV)

    def make():
        torch.manual_seed(0)
        net = D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 5, conv_channels=[8, 16], num_neighbors=8), 2).to(DEV).train()
        for m in net.modules():
            if isinstance(m, torch.nn.Dropout):
                m.p = 0.0
        flat = FlatGradients(net.parameters())
        opt = FlatSGD(flat, lr=0.05, momentum=0.9, weight_decay=1e-4) if kind == 'sgd' else FlatAdam(flat, lr=1e-2)
        return net, flat, opt

    def step(net, flat, opt):
        flat.zero()
        loss = calc_loss(net(data), data.y)
        loss.backward()
        opt.step()
        return loss.item()

    net, flat, opt = make()
    for _ in range(3):
        step(net, flat, opt)
    saved_model, saved_opt = copy.deepcopy(net.state_dict()), opt.state_dict()
    want = [step(net, flat, opt) for _ in range(2)]
    net2, flat2, opt2 = make()
    net2.load_state_dict(saved_model)
    opt2.load_state_dict(saved_opt)                       # before any step: the flat buffers do not exist yet
    got = [step(net2, flat2, opt2) for _ in range(2)]
    assert got == pytest.approx(want, rel=1e-6)
    for p, q in zip(net.parameters(), net2.parameters()):
        assert torch.allclose(p, q, rtol=1e-6, atol=1e-7)


def test_morton_order_pretransform_is_a_permutation_inside_every_cloud():
    """transforms.MortonOrder: clouds keep their place, per-point fields move together, and the classifier's logits are those
    of the original batch up to fp32 summation order."""
    D = dc()
    data = D.utils.synthetic_batch(3, 0, config=8, sizes=[700, 300, 1024], num_classes=7).to(DEV)
    ref = D.utils.Data(**data.__dict__)
    out = D.transforms.MortonOrder()(D.utils.Data(**data.__dict__))
    perm = out.morton_perm
    assert torch.equal(torch.sort(perm).values, torch.arange(2024, device=DEV))
    assert torch.equal(data.batch[perm], data.batch)                       # a permutation inside every cloud
    assert torch.equal(out.pos, ref.pos[perm]) and torch.equal(out.norm, ref.norm[perm]) and torch.equal(out.y, ref.y)
    model = D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 7, conv_channels=[16, 32], num_neighbors=12), 4).to(DEV).eval()
    with torch.no_grad():
        a, b = model(ref), model(out)
    assert rel_err(b.cpu(), a.cpu()) < 1e-4
