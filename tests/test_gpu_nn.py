"""GPU parity: Linear / BatchNorm / VectorNonLin / aggregation kernels, DeltaConv layers and whole
networks (forward AND backward) against torch fp64, the oracle and the golden fixtures produced
by the unmodified reference.  Run with `-m gpu`."""
import copy

import pytest
import torch
import torch.nn.functional as F

from helpers import load, t, rel_err, report, LOGIT_REL, GRAD_REL
from oracle import geometry as og
from oracle import network as on

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def dc():
    import deltaconv_b200
    return deltaconv_b200


def no_dropout(m):
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
    return m


# ----------------------------------------------------------------------------------------------
# kernels against torch fp64
# ----------------------------------------------------------------------------------------------

@pytest.mark.parametrize('M,K,N,bias', [(1, 1, 1, False), (37, 3, 64, False), (1000, 12, 1, True), (4096, 256, 64, False),
                                        (333, 517, 129, True), (32, 2048, 512, False), (20000, 24, 16, False),
                                        (4, 64, 40, True), (64, 36, 20, True), (33, 100, 7, True), (17, 30, 12, True), (63, 512, 256, False),
                                        (4096, 12, 64, False), (2000, 3, 130, True), (5000, 16, 512, False), (32768, 7, 64, False)])
def test_linear_forward_backward(M, K, N, bias):
    x0, w0 = torch.randn(M, K), torch.randn(N, K) / K ** 0.5
    b0 = torch.randn(N) if bias else None
    g0 = torch.randn(M, N)
    x, w = x0.to(DEV).requires_grad_(True), w0.to(DEV).requires_grad_(True)
    b = b0.to(DEV).requires_grad_(True) if bias else None
    y = dc().ops.linear(x, w, b)
    (y * g0.to(DEV)).sum().backward()
    xr, wr = x0.double().requires_grad_(True), w0.double().requires_grad_(True)
    br = b0.double().requires_grad_(True) if bias else None
    yr = F.linear(xr, wr, br)
    (yr * g0.double()).sum().backward()
    assert rel_err(y.detach().cpu(), yr.detach()) < 1e-5
    assert rel_err(x.grad.cpu(), xr.grad) < 1e-5
    assert rel_err(w.grad.cpu(), wr.grad) < 1e-5
    if bias:
        assert rel_err(b.grad.cpu(), br.grad) < 2e-6


@pytest.mark.parametrize('mode,tol', [(3, 1e-5), (1, 3e-3), (0, 2e-6)])
@pytest.mark.parametrize('M,K,N', [(4096, 256, 64), (1000, 512, 1024), (32768, 140, 64), (8192, 64, 256), (777, 36, 40),
                                   (130, 1024, 16), (65536, 24, 128)])
def test_tensor_core_linear_modes(M, K, N, mode, tol):
    """dc_linear_* through the tcgen05 kernel: 3xTF32 (default, fp32-grade), plain TF32, and tensor cores off."""
    from deltaconv_b200 import _lib
    lib = _lib.load()
    x0, w0, g0 = torch.randn(M, K), torch.randn(N, K) / K ** 0.5, torch.randn(M, N)
    prev = lib.dc_get_gemm_mode()
    try:
        assert lib.dc_set_gemm_mode(mode) == 0
        x, w = x0.to(DEV).requires_grad_(True), w0.to(DEV).requires_grad_(True)
        y = dc().ops.linear(x, w)
        (y * g0.to(DEV)).sum().backward()
        torch.cuda.synchronize()
    finally:
        lib.dc_set_gemm_mode(prev)
    xr, wr = x0.double().requires_grad_(True), w0.double().requires_grad_(True)
    yr = F.linear(xr, wr)
    (yr * g0.double()).sum().backward()
    e = [report(f'tc_linear.mode{mode}.{M}x{K}x{N}.{n_}', rel_err(a.cpu(), b))
         for n_, a, b in (('y', y.detach(), yr.detach()), ('dx', x.grad, xr.grad), ('dw', w.grad, wr.grad))]
    assert max(e) < tol, e


@pytest.mark.parametrize('M,C,slope,train', [(1000, 64, 0.2, True), (257, 5, 1.0, True), (4096, 130, 0.2, False),
                                             (50000, 16, 0.0, True), (3, 2, 0.2, True)])
def test_batchnorm_activation(M, C, slope, train):
    nn = dc().nn
    mod = nn.BatchNorm1d(C).to(DEV)
    ref = torch.nn.BatchNorm1d(C).double()
    with torch.no_grad():
        for p_, q_ in ((mod.bn.weight, ref.weight), (mod.bn.bias, ref.bias), (mod.bn.running_mean, ref.running_mean),
                       (mod.bn.running_var, ref.running_var)):
            val = torch.randn(C) * 0.3 + (1.0 if p_ is mod.bn.weight or p_ is mod.bn.running_var else 0.0)
            if p_ is mod.bn.running_var:
                val = val.abs() + 0.1
            p_.copy_(val)
            q_.copy_(val.double())
    mod.train(train)
    ref.train(train)
    x0 = torch.randn(M, C) * 2 + 3          # mean >> 0: exercises the variance accumulation
    add0, g0 = torch.randn(M, C), torch.randn(M, C)
    x = x0.to(DEV).requires_grad_(True)
    add = add0.to(DEV).requires_grad_(True)
    out = mod(x, slope, add)
    (out * g0.to(DEV)).sum().backward()
    xr, addr = x0.double().requires_grad_(True), add0.double().requires_grad_(True)
    outr = F.leaky_relu(ref(xr), slope) + addr
    (outr * g0.double()).sum().backward()
    assert rel_err(out.detach().cpu(), outr.detach()) < 5e-6
    assert rel_err(x.grad.cpu(), xr.grad) < 1e-4 or (x.grad.cpu() - xr.grad).abs().max() < 1e-6   # tiny M: dx ~ 0 by cancellation
    assert rel_err(add.grad.cpu(), addr.grad) < 1e-6
    assert rel_err(mod.bn.weight.grad.cpu(), ref.weight.grad) < 1e-4
    assert rel_err(mod.bn.bias.grad.cpu(), ref.bias.grad) < 1e-4
    assert rel_err(mod.bn.running_mean.cpu(), ref.running_mean) < 1e-5
    assert rel_err(mod.bn.running_var.cpu(), ref.running_var) < 1e-5
    assert int(mod.bn.num_batches_tracked) == int(ref.num_batches_tracked)


@pytest.mark.parametrize('M,C,with_bn,train', [(777, 32, True, True), (100, 7, False, True), (2048, 64, True, False)])
def test_vector_nonlin(M, C, with_bn, train):
    nn = dc().nn
    mod = nn.VectorNonLin(C, batchnorm=nn.BatchNorm1d(C) if with_bn else None).to(DEV)
    ref = on.VectorNonLin(C, batchnorm=on.BatchNorm1d(C) if with_bn else None).double()
    with torch.no_grad():
        mod.bias.copy_(torch.randn(C) * 0.5 - 0.3)
        if with_bn:
            mod.batchnorm.bn.weight.copy_(torch.randn(C) * 0.3 + 1)
            mod.batchnorm.bn.bias.copy_(torch.randn(C) * 0.3)
    ref.load_state_dict({k_: v_.double() if v_.is_floating_point() else v_ for k_, v_ in mod.state_dict().items()})
    mod.train(train)
    ref.train(train)
    v0 = torch.randn(2 * M, C)
    v0[:4] = 0                               # zero vectors: the 1e-8 clamp path
    g0 = torch.randn(2 * M, C)
    v = v0.to(DEV).requires_grad_(True)
    out = mod(v)
    (out * g0.to(DEV)).sum().backward()
    vr = v0.double().requires_grad_(True)
    outr = ref(vr)
    (outr * g0.double()).sum().backward()
    assert rel_err(out.detach().cpu(), outr.detach()) < 5e-6
    assert rel_err(v.grad.cpu(), vr.grad) < 1e-4
    if with_bn:
        assert rel_err(mod.batchnorm.bn.weight.grad.cpu(), ref.batchnorm.bn.weight.grad) < 1e-4
        assert rel_err(mod.batchnorm.bn.bias.grad.cpu(), ref.batchnorm.bn.bias.grad) < 1e-4
        assert mod.bias.grad is None                       # unused when BN is on
    else:
        assert rel_err(mod.bias.grad.cpu(), ref.bias.grad) < 1e-4


def test_gather_max_pool_broadcast_edge_ops():
    d = dc().utils.synthetic_batch(3, 0, config=8, sizes=[90, 130, 64])
    idx = og.knn_ell(d.pos, 9, d.batch)
    g = dc().ops.knn(d.pos.to(DEV), 9, d.batch.to(DEV))
    N, C = d.pos.size(0), 70
    y0, g0 = torch.randn(N, C), torch.randn(N, C)
    y = y0.to(DEV).requires_grad_(True)
    out = dc().ops.gather_max(y, g)
    (out * g0.to(DEV)).sum().backward()
    yr = y0.double().requires_grad_(True)
    outr = yr[idx].max(dim=1).values
    (outr * g0.double()).sum().backward()
    assert torch.equal(out.detach().cpu().double(), outr.detach()) and rel_err(y.grad.cpu(), yr.grad) < 1e-6
    # edge differences + segment max (centralized first layer, deltaconv.py:50-52)
    x = y0[:, :5].contiguous().to(DEV).requires_grad_(True)
    e = dc().ops.edge_diff(x, g)
    m = dc().ops.segment_max(e * 1.0, g)
    (m * g0[:, :5].to(DEV)).sum().backward()
    xr = y0[:, :5].double().requires_grad_(True)
    er = (xr[idx] - xr[:, None]).reshape(N * 9, 5)
    mr = er.reshape(N, 9, 5).max(dim=1).values
    (mr * g0[:, :5].double()).sum().backward()
    assert torch.equal(e.detach().cpu().double(), er.detach().float().double()) or rel_err(e.detach().cpu(), er.detach()) < 1e-6
    assert rel_err(m.detach().cpu(), mr.detach()) < 1e-6 and rel_err(x.grad.cpu(), xr.grad) < 1e-5
    # pools
    batch = d.batch.to(DEV)
    for mode in ('max', 'mean', 'maxmean'):
        z = y0.to(DEV).requires_grad_(True)
        p = dc().ops.pool(z, batch, mode)
        gp = torch.randn(p.shape)
        (p * gp.to(DEV)).sum().backward()
        zr = y0.double().requires_grad_(True)
        mx, mn = on._pool(zr, d.batch, 3, 'max'), on._pool(zr, d.batch, 3, 'mean')
        pr = {'max': mx, 'mean': mn, 'maxmean': torch.cat([mx, mn], 1)}[mode]
        (pr * gp.double()).sum().backward()
        assert rel_err(p.detach().cpu(), pr.detach()) < 1e-6 and rel_err(z.grad.cpu(), zr.grad) < 1e-6
    assert torch.equal(dc().geometry.global_max_pool(y0.to(DEV), batch).cpu(), on._pool(y0, d.batch, 3, 'max'))
    # broadcast of per-cloud rows
    s0 = torch.randn(3, 11)
    s = s0.to(DEV).requires_grad_(True)
    bc = dc().ops.broadcast_rows(s, batch, N)
    gb = torch.randn(N, 11)
    (bc * gb.to(DEV)).sum().backward()
    assert torch.equal(bc.detach().cpu(), s0[d.batch])
    assert rel_err(s.grad.cpu(), torch.zeros(3, 11).index_add(0, d.batch, gb)) < 1e-5
    s4 = torch.randn(3, 24)                                 # the float4 kernel (C % 4 == 0)
    assert torch.equal(dc().ops.broadcast_rows(s4.to(DEV), batch, N).cpu(), s4[d.batch])


# ----------------------------------------------------------------------------------------------
# layers and networks against the reference's own outputs (golden) and the oracle
# ----------------------------------------------------------------------------------------------

def test_deltaconv_layers_vs_reference_golden():
    g, z = load('layers_deltaconv')
    D = dc()
    pos, normal, batch, k = t(g.pos).to(DEV), t(g.normal).to(DEV), t(g.batch).to(DEV), int(g.k)
    ei = D.geometry.knn_graph(pos, k, batch, loop=True, flow='target_to_source')
    xb, yb = D.geometry.build_tangent_basis(normal)
    grad, div = D.geometry.build_grad_div(pos, normal, xb, yb, ei, batch)
    conv0 = D.utils.fill_deterministic(D.nn.DeltaConv(3, 16, depth=1, centralized=True, vector=True), 3).to(DEV)
    conv1 = D.utils.fill_deterministic(D.nn.DeltaConv(16, 8, depth=2, centralized=False, vector=True), 4).to(DEV)
    assert repr(conv0) == 'DeltaConv(3, 16)'
    x_in = pos.clone().requires_grad_(True)
    x1, v1 = conv0(x_in, grad @ x_in, grad, div, ei)
    x2, v2 = conv1(x1, v1, grad, div, ei)
    ((x2 * t(g.wx).to(DEV)).sum() + (v2 * t(g.wv).to(DEV)).sum()).backward()
    for ours, key in ((x1, 'x1'), (v1, 'v1'), (x2, 'x2'), (v2, 'v2')):
        assert rel_err(ours.detach().cpu(), t(z[f'ref64_{key}'])) < LOGIT_REL, key
    assert rel_err(x_in.grad.cpu(), t(z['ref64_dx_in'])) < GRAD_REL
    for cname, conv in (('conv0', conv0), ('conv1', conv1)):
        for pname, p in conv.named_parameters():
            key = f'ref64_grad.{cname}.{pname}'
            if key in z.files:
                assert rel_err(p.grad.cpu(), t(z[key])) < GRAD_REL, key
            else:
                assert p.grad is None, key
        for bname, b in conv.named_buffers():
            if 'running' in bname:
                assert rel_err(b.cpu(), t(z[f'ref64_buf.{cname}.{bname}'])) < 1e-4, bname
    # vector=False returns the input v object untouched (deltaconv.py:64-70, test_deltaconv.py:37-40)
    conv2 = D.nn.DeltaConv(16, 8, vector=False).to(DEV)
    _, v_same = conv2(x1.detach(), v1.detach(), grad, div, ei)
    assert v_same.data_ptr() == v1.data_ptr()


def _models(kind, k):
    D = dc()
    if kind == 'model_cls':
        return D.models.DeltaNetClassification(3, 7, conv_channels=[16, 16, 32], num_neighbors=k)
    if kind == 'model_cls_nonormal':
        return D.models.DeltaNetClassification(3, 5, conv_channels=[8, 16], num_neighbors=k)
    return D.models.DeltaNetSegmentation(3, 6, conv_channels=[8, 16, 16], mlp_depth=2, embedding_size=64,
                                         categorical_vector=True, num_neighbors=k)


@pytest.mark.parametrize('kind', ['model_cls', 'model_cls_nonormal', 'model_seg'])
def test_models_vs_reference_golden(kind):
    g, z = load(kind)
    D = dc()
    model = no_dropout(D.utils.fill_deterministic(_models(kind, int(g.k)), 2)).to(DEV)
    data = D.utils.Data(pos=t(g.pos), batch=t(g.batch), y=t(g.y))
    data.norm = t(g.normal) if 'normal' in z.files else None
    if 'category' in z.files:
        data.category = t(g.category)
    data = data.to(DEV)
    model.eval()
    with torch.no_grad():
        assert report(f'{kind}.logits_eval', rel_err(model(data).cpu(), t(g.ref64_logits_eval))) < LOGIT_REL
    model.train()
    logits = model(data)
    loss = F.cross_entropy(logits, data.y)
    loss.backward()
    assert report(f'{kind}.logits_train', rel_err(logits.detach().cpu(), t(g.ref64_logits_train))) < LOGIT_REL
    assert abs(loss.item() - float(g.ref64_loss)) < LOGIT_REL * max(1.0, abs(float(g.ref64_loss)))
    names = dict(model.named_parameters())
    # gradients that vanish analytically (a BN bias ahead of another BN) are compared on the scale
    # of the network's typical gradient, not on their own (round-off sized) scale
    scale = max(float(z[key]) for key in z.files if key.startswith('ref64_gnorm.'))
    worst = 0.0
    for key in z.files:
        if key.startswith('ref64_gnorm.'):
            p = names[key[len('ref64_gnorm.'):]]
            ref = float(z[key])
            assert p.grad is not None, key
            err = abs(p.grad.norm().item() - ref) / max(ref, 1e-3 * scale)
            worst = max(worst, err)
            assert err <= GRAD_REL, key
    report(f'{kind}.grad_norms', worst)
    for name, p in names.items():
        if f'ref64_gnorm.{name}' not in z.files:
            assert p.grad is None, name            # VectorNonLin.bias under BN


def test_full_width_classifier_vs_oracle():
    """The shipped architecture (convs [64,64,128,256], 2,045,928 parameters) on a 4 x 256 batch:
    logits and loss against the fp64 oracle with identical weights, training-mode BN."""
    D = dc()
    data = D.utils.synthetic_batch(4, 256, config=2)
    model = no_dropout(D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 40), 1))
    oracle = no_dropout(on.DeltaNetClassification(3, 40)).double()
    oracle.load_state_dict({k_: (v_.double() if v_.is_floating_point() else v_) for k_, v_ in model.state_dict().items()})
    model = model.to(DEV)
    model.train()
    oracle.train()
    logits = model(data.to(DEV))
    ref = oracle(data.double())
    assert report('full_width.logits_train', rel_err(logits.detach().cpu(), ref.detach())) < LOGIT_REL
    on.smoothed_cross_entropy(ref, data.y).backward()
    on.smoothed_cross_entropy(logits, data.y.to(DEV)).backward()
    ref_grads = dict(oracle.named_parameters())
    for name, p in model.named_parameters():
        if ref_grads[name].grad is None:
            assert p.grad is None, name
            continue
        assert report(f'full_width.grad.{name}', rel_err(p.grad.cpu(), ref_grads[name].grad)) < 5 * GRAD_REL, name
    # state dicts stay interchangeable with the reference layout
    assert set(model.state_dict()) == set(oracle.state_dict())


# ----------------------------------------------------------------------------------------------
# the reference's own property tests, run against the CUDA modules
# ----------------------------------------------------------------------------------------------

def test_vector_mlp_equivariance():
    """MLP(T v) = T MLP(v) for per-point rotations/reflections T (reference: test/nn/test_mlp.py:42-78)."""
    D = dc()
    N, C = 500, 16
    mlp = D.nn.VectorMLP([C, 32, 8]).to(DEV).eval()
    v = torch.randn(2 * N, C, device=DEV)
    ang = torch.rand(N, device=DEV) * 2 * torch.pi
    refl = torch.where(torch.rand(N, device=DEV) < 0.5, -1.0, 1.0)
    T = torch.stack([torch.cos(ang), -torch.sin(ang) * refl, torch.sin(ang), torch.cos(ang) * refl], 1).view(N, 2, 2)

    def apply(Tm, field):
        return torch.einsum('nab,nbc->nac', Tm, field.view(N, 2, -1)).reshape(2 * N, -1)
    assert torch.allclose(mlp(apply(T, v)), apply(T, mlp(v)), atol=1e-5)


def test_scalar_vector_mlp_conventions():
    D = dc()
    m = D.nn.ScalarVectorMLP([4, 8]).to(DEV)
    x, v = torch.randn(50, 4, device=DEV), torch.randn(100, 4, device=DEV)
    xo, vo = m((x, v))
    assert xo.shape == (50, 8) and vo.shape == (100, 8)
    with pytest.raises(AssertionError):
        m(x)
    m2 = D.nn.ScalarVectorMLP([4, 8], vector_stream=False).to(DEV)
    assert m2(x).shape == (50, 8)
    assert D.nn.ScalarVectorIdentity()((x, v))[0] is x
    assert repr(D.nn.BatchNorm1d(8)) == 'BatchNorm1d(8)'
    assert repr(D.nn.VectorNonLin(8, batchnorm=D.nn.BatchNorm1d(8))) == 'VectorNonLin(batchnorm=BatchNorm1d(8))'


def test_coordinate_independence_of_weight_gradients():
    """Same weight gradients with randomly rotated tangent frames (reference: test/nn/test_deltaconv.py:42-74)."""
    D = dc()
    torch.manual_seed(1)
    N = 1000
    x = torch.rand(N, 3, device=DEV)
    ei = D.geometry.knn_graph(x, 20, loop=True, flow='target_to_source')
    normal, xb, yb = D.geometry.estimate_basis(x, ei)
    # lambda -> 0 as in the reference's test: the regulariser (lambda on all six monomials) is itself
    # not rotation invariant, so gauge independence only holds in that limit
    grad, div = D.geometry.build_grad_div(x, normal, xb, yb, ei, regularizer=1e-8)
    xb_rot = D.geometry.rotate_around(xb, normal, torch.rand(N, device=DEV) * 2 * torch.pi)
    yb_rot = torch.linalg.cross(normal, xb_rot, dim=-1)
    grad_r, div_r = D.geometry.build_grad_div(x, normal, xb_rot, yb_rot, ei, regularizer=1e-8)
    target = torch.rand(N, 1, device=DEV)
    conv = D.nn.DeltaConv(3, 1, depth=1, centralized=False).to(DEV)
    grads = []
    for gr, dv in ((grad, div), (grad_r, div_r)):
        conv.zero_grad()
        conv2 = copy.deepcopy(conv)          # identical running stats for both passes
        out, _ = conv2(x, gr @ x, gr, dv, ei)
        F.l1_loss(out, target).backward()
        grads.append(torch.cat([p.grad.flatten() for p in conv2.parameters() if p.grad is not None]))
    assert report('coordinate_independence.abs', (grads[0] - grads[1]).abs().max().item()) < 1e-4


def test_cpu_tensors_fail_loudly():
    D = dc()
    with pytest.raises(D._lib.DeltaConvLibraryError):
        D.geometry.build_tangent_basis(torch.randn(10, 3))
    with pytest.raises(D._lib.DeltaConvLibraryError):
        D.nn.Linear(4, 4)(torch.randn(3, 4))


def test_graphed_train_step_matches_eager():
    """CUDA-graph replay of the step = the eager step (same kernels, same order): identical weights."""
    D = dc()
    from deltaconv_b200.parallel import FlatGradients
    from deltaconv_b200.utils.losses import calc_loss
    data = D.utils.synthetic_batch(4, 256, config=2, num_classes=10).to(DEV)
    data._num_clouds, data._max_points = 4, 256
    nets = []
    for _ in range(2):
        torch.manual_seed(0)
        nets.append(no_dropout(D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 10, conv_channels=[16, 32]), 5)).to(DEV).train())
    eager, graphed_net = nets
    flat = FlatGradients(eager.parameters())
    opt = torch.optim.SGD(eager.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    opt2 = torch.optim.SGD(graphed_net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    step = D.utils.GraphedTrainStep(graphed_net, opt2, calc_loss, data, warmup=2)
    losses_e, losses_g = [], []
    for _ in range(2 + 3):                       # the graphed net already took 2 warm-up steps at construction
        flat.zero()
        loss = calc_loss(eager(data), data.y)
        loss.backward()
        opt.step()
        losses_e.append(loss.item())
    host = data.to('cpu').pin_memory()
    for it in range(3):
        if it == 1:                                  # side-stream prefetch + device-to-device commit instead of load()
            step.prefetch(host)
            step.commit()
        else:
            step.load(data)
        losses_g.append(step().item())
    assert losses_g == pytest.approx(losses_e[2:], rel=1e-5)
    for (n1, p1), (n2, p2) in zip(eager.named_parameters(), graphed_net.named_parameters()):
        assert torch.allclose(p1, p2, rtol=1e-5, atol=1e-6), n1
    for (n1, b1), (n2, b2) in zip(eager.named_buffers(), graphed_net.named_buffers()):
        if n1.endswith('num_batches_tracked'):
            assert int(b1) == int(b2) == 5, n1                # incremented by the statistics kernels, also under replay
        else:
            assert torch.allclose(b1, b2, rtol=1e-5, atol=1e-6), n1


def test_config3_segmentation_shapenet_shape_vs_oracle():
    """BASELINE configs[2] architecture (convs [64,128,256], depth-2 MLPs, embedding 1024, categorical vector, 50
    classes, k=30, 2048 points) on 2 clouds: logits and loss against the fp64 oracle, training-mode BN."""
    D = dc()
    data = D.utils.synthetic_batch(2, 2048, config=3, num_classes=50, per_point_labels=True, categories=16)
    kw = dict(conv_channels=[64, 128, 256], mlp_depth=2, embedding_size=1024, categorical_vector=True, num_neighbors=30)
    model = no_dropout(D.utils.fill_deterministic(D.models.DeltaNetSegmentation(3, 50, **kw), 7))
    oracle = no_dropout(on.DeltaNetSegmentation(3, 50, **kw)).double()
    oracle.load_state_dict({k_: (v_.double() if v_.is_floating_point() else v_) for k_, v_ in model.state_dict().items()})
    model = model.to(DEV).train()
    oracle.train()
    logits = model(data.to(DEV))
    with torch.no_grad():
        ref = oracle(data.double())
    assert tuple(logits.shape) == (4096, 50)
    assert report('config3_seg.logits_train', rel_err(logits.detach().cpu(), ref)) < LOGIT_REL
    F.cross_entropy(logits, data.y.to(DEV)).backward()
    assert all(torch.isfinite(p.grad).all() for p in model.parameters() if p.grad is not None)


def test_full_size_configs_run_and_are_deterministic():
    """configs[2] (16 x 2048, k=30, segmentation) and configs[4] (32 x 2048, k=20, no normals -> estimated frames,
    15 classes) at full size: a training step is finite and bitwise reproducible run to run."""
    D = dc()
    cases = [
        (lambda: D.models.DeltaNetSegmentation(3, 50, conv_channels=[64, 128, 256], mlp_depth=2, embedding_size=1024,
                                               categorical_vector=True, num_neighbors=30),
         D.utils.synthetic_batch(16, 2048, config=3, num_classes=50, per_point_labels=True, categories=16)),
        (lambda: D.models.DeltaNetClassification(3, 15, conv_channels=[64, 64, 64, 128], num_neighbors=20),
         D.utils.synthetic_batch(32, 2048, config=5, num_classes=15, normals=False)),
    ]
    for build, data in cases:
        data = data.to(DEV)
        if not hasattr(data, 'norm'):
            data.norm = None
        outs = []
        for _ in range(2):
            model = no_dropout(D.utils.fill_deterministic(build(), 3)).to(DEV).train()
            logits = model(data)
            loss = F.cross_entropy(logits, data.y)
            loss.backward()
            g = torch.cat([p.grad.flatten() for p in model.parameters() if p.grad is not None])
            assert torch.isfinite(logits).all() and torch.isfinite(g).all()
            outs.append((logits.detach().clone(), g.clone()))
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize('M,K,C,leaky', [(32768, 256, 64, False), (5000, 64, 128, False), (4096, 512, 1024, False),
                                         (777, 12, 40, True), (100, 36, 20, True)])
def test_fused_linear_batchnorm_block(M, K, C, leaky):
    """One-node MLP block (Linear -> BN(train) -> activation + addend) vs torch fp64.  The large shapes use the
    identity activation: among millions of pre-activations a few fall within rounding distance of zero, where the
    LeakyReLU derivative jumps, which makes an fp32-vs-fp64 gradient comparison ill-posed."""
    D = dc()
    act = torch.nn.LeakyReLU(0.2) if leaky else torch.nn.Identity()
    mlp = D.nn.MLP([K, C], nonlin=act).to(DEV).train()
    with torch.no_grad():
        mlp[0][1].bn.weight.copy_(torch.randn(C) * 0.3 + 1)
        mlp[0][1].bn.bias.copy_(torch.randn(C) * 0.3)
    ref_lin = torch.nn.Linear(K, C, bias=False).double()
    ref_bn = torch.nn.BatchNorm1d(C).double()
    ref_lin.weight.data.copy_(mlp[0][0].weight.detach().cpu().double())
    ref_bn.weight.data.copy_(mlp[0][1].bn.weight.detach().cpu().double())
    ref_bn.bias.data.copy_(mlp[0][1].bn.bias.detach().cpu().double())
    x0, add0, g0 = torch.randn(M, K) + 0.5, torch.randn(M, C), torch.randn(M, C)
    x, add = x0.to(DEV).requires_grad_(True), add0.to(DEV).requires_grad_(True)
    out = mlp(x, add)
    (out * g0.to(DEV)).sum().backward()
    xr, addr = x0.double().requires_grad_(True), add0.double().requires_grad_(True)
    outr = act(ref_bn(ref_lin(xr))) + addr
    (outr * g0.double()).sum().backward()
    assert report(f'fused_lin_bn.{M}x{K}x{C}.out', rel_err(out.detach().cpu(), outr.detach())) < 1e-5
    assert report(f'fused_lin_bn.{M}x{K}x{C}.dx', rel_err(x.grad.cpu(), xr.grad)) < 1e-4 and rel_err(add.grad.cpu(), addr.grad) < 1e-6
    assert rel_err(mlp[0][0].weight.grad.cpu(), ref_lin.weight.grad) < 1e-4
    assert rel_err(mlp[0][1].bn.weight.grad.cpu(), ref_bn.weight.grad) < 1e-4
    assert rel_err(mlp[0][1].bn.running_mean.cpu(), ref_bn.running_mean) < 1e-5
    assert rel_err(mlp[0][1].bn.running_var.cpu(), ref_bn.running_var) < 1e-5
    assert int(mlp[0][1].bn.num_batches_tracked) == int(ref_bn.num_batches_tracked) == 1      # counted by the kernel


@pytest.mark.parametrize('sizes,K,C,mode', [([100, 37, 263], 36, 64, 'maxmean'), ([1024] * 8, 128, 1024, 'maxmean'),
                                            ([500], 20, 32, 'max'), ([64, 64, 1, 300], 16, 40, 'mean'),
                                            ([90, 10], 24, 30, 'maxmean')])
def test_embedding_pool_fusion_matches_unfused(sizes, K, C, mode):
    """lin_embedding + BatchNorm + LeakyReLU evaluated inside the pooling pass (dc_bn_act_pool_*) = the unfused
    sequence MLP -> pool: same output, same gradients for the input and every parameter, same running statistics.
    C = 30 is not a multiple of 4: forward_pooled must fall back to the unfused kernels by itself."""
    import copy
    D = dc()
    from deltaconv_b200 import ops
    torch.manual_seed(3)
    N = sum(sizes)
    batch = torch.repeat_interleave(torch.arange(len(sizes)), torch.tensor(sizes)).to(DEV)
    fused = D.nn.MLP([K, C]).to(DEV).train()
    with torch.no_grad():
        fused[0][1].bn.weight.copy_(torch.randn(C) * 0.5)            # both signs: the max can sit at the column minimum
        fused[0][1].bn.bias.copy_(torch.randn(C) * 0.3)
    plain = copy.deepcopy(fused)
    x0 = torch.randn(N, K)
    width = C * (2 if mode == 'maxmean' else 1)
    g0 = torch.randn(len(sizes), width, device=DEV)
    xa, xb = x0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    out_f = fused.forward_pooled(xa, batch, mode, len(sizes))
    out_p = ops.pool(plain(xb), batch, mode, len(sizes))
    (out_f * g0).sum().backward()
    (out_p * g0).sum().backward()
    assert report(f'embed_pool.{mode}.{C}.out', rel_err(out_f.detach().cpu(), out_p.detach().cpu())) < 1e-6
    assert report(f'embed_pool.{mode}.{C}.dx', rel_err(xa.grad.cpu(), xb.grad.cpu())) < 1e-5
    for (n1, p1), (_, p2) in zip(fused.named_parameters(), plain.named_parameters()):
        assert rel_err(p1.grad.cpu(), p2.grad.cpu()) < 1e-5, n1
    for (n1, b1), (_, b2) in zip(fused.named_buffers(), plain.named_buffers()):
        assert torch.allclose(b1.float(), b2.float(), rtol=1e-6, atol=1e-7), n1
    # eval mode uses the running statistics: unfused path, same numbers as MLP -> pool
    fused.eval(), plain.eval()
    with torch.no_grad():
        assert torch.equal(fused.forward_pooled(xa, batch, mode, len(sizes)), ops.pool(plain(xb), batch, mode, len(sizes)))


@pytest.mark.parametrize('switch', ['dc_set_gemm_pair', 'dc_set_gemm_tma_store'])
def test_gemm_kernel_variants_agree(switch):
    """The single-CTA kernel (dc_set_gemm_pair(0)) and the direct-store epilogue (dc_set_gemm_tma_store(0), single-CTA only)
    compute the same products as the default CTA-pair kernel (tcgen05 cta_group::2, TMA-store epilogue): all three layouts, ragged M / N / K, bias, split-K partials."""
    from deltaconv_b200 import _lib
    lib = _lib.load()
    torch.manual_seed(5)
    ws = torch.zeros(64 << 20, dtype=torch.uint8, device=DEV)
    on = 0                                                 # the non-default setting of either switch
    try:
        for (M, N, K, a_mn, b_mn, with_bias) in [(1300, 72, 100, 0, 0, True), (2049, 200, 64, 0, 1, False), (640, 136, 4000, 1, 1, False),
                                                 (4096, 64, 256, 0, 0, True), (1153, 128, 96, 0, 0, False)]:
            a, b = torch.randn(M, K, device=DEV), torch.randn(N, K, device=DEV)
            bias = torch.randn(N, device=DEV) if with_bias else None
            A = a.t().contiguous() if a_mn else a
            B = b.t().contiguous() if b_mn else b
            outs = []
            for setting in (1 - on, on):
                getattr(lib, switch)(setting)
                C = torch.full((M, N), float('nan'), device=DEV)
                _lib.call('dc_gemm_tc_raw', A.data_ptr(), A.stride(0), a_mn, B.data_ptr(), B.stride(0), b_mn,
                          bias.data_ptr() if with_bias else None, M, N, K, C.data_ptr(), N, ws.data_ptr(), ws.numel())
                outs.append(C)
            ref = a.double() @ b.double().t() + (bias.double() if with_bias else 0)
            scale = ref.abs().max()
            assert ((outs[1].double() - ref).abs().max() / scale) < 2e-5, (M, N, K)
            assert ((outs[1] - outs[0]).abs().max() / scale) < 4e-6, (M, N, K)      # same terms, different accumulation grouping at most
    finally:
        lib.dc_set_gemm_pair(1)
        lib.dc_set_gemm_tma_store(1)


def test_flat_sgd_matches_torch_sgd_eager_and_graphed():
    """FlatSGD (one dc_sgd_step launch on flat buffers) follows torch.optim.SGD(momentum, weight_decay) step for step, eagerly
    and inside a captured training step; a learning-rate change reaches the captured kernel through the device scalar."""
    D = dc()
    from deltaconv_b200.parallel import FlatGradients, FlatSGD
    from deltaconv_b200.utils.losses import calc_loss
    data = D.utils.synthetic_batch(4, 256, config=2, num_classes=10).to(DEV)
    data._num_clouds, data._max_points = 4, 256
    nets = []
    for _ in range(3):
        torch.manual_seed(0)
        nets.append(no_dropout(D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 10, conv_channels=[16, 32]), 5)).to(DEV).train())
    ref_net, flat_net, graph_net = nets
    ref_flat = FlatGradients(ref_net.parameters())
    ref_opt = torch.optim.SGD(ref_net.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    fg = FlatGradients(flat_net.parameters())
    opt = FlatSGD(fg, lr=0.05, momentum=0.9, weight_decay=1e-4)
    gg = FlatGradients(graph_net.parameters())
    gopt = FlatSGD(gg, lr=0.05, momentum=0.9, weight_decay=1e-4)
    step = D.utils.GraphedTrainStep(graph_net, gopt, calc_loss, data, flat=gg, warmup=2)
    lrs = [0.05, 0.05, 0.05, 0.02, 0.02]
    for it, lr in enumerate(lrs):
        ref_opt.param_groups[0]['lr'] = lr
        ref_flat.zero()
        calc_loss(ref_net(data), data.y).backward()
        ref_opt.step()
        opt.set_lr(lr)
        opt.zero_grad()
        calc_loss(flat_net(data), data.y).backward()
        opt.step()
        if it >= 2:                                   # the graphed net already took 2 warm-up steps at lr 0.05
            gopt.set_lr(lr)
            step()
    for (n1, p1), (_, p2), (_, p3) in zip(ref_net.named_parameters(), flat_net.named_parameters(), graph_net.named_parameters()):
        scale = p1.abs().max().item() + 1e-12
        assert (p1 - p2).abs().max().item() < 2e-5 * scale, n1
        assert (p1 - p3).abs().max().item() < 2e-5 * scale, n1
    sd = flat_net.state_dict()                        # parameters are views of one flat buffer: state dicts keep their shapes
    assert all(sd[k].shape == v.shape for k, v in ref_net.state_dict().items())


def test_flat_adam_matches_torch_adam():
    D = dc()
    from deltaconv_b200.parallel import FlatGradients, FlatAdam
    from deltaconv_b200.utils.losses import calc_loss
    data = D.utils.synthetic_batch(4, 256, config=2, num_classes=10).to(DEV)
    data._num_clouds, data._max_points = 4, 256
    nets = []
    for _ in range(2):
        torch.manual_seed(0)
        nets.append(no_dropout(D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 10, conv_channels=[16, 32]), 5)).to(DEV).train())
    ref_net, flat_net = nets
    ref_flat = FlatGradients(ref_net.parameters())
    ref_opt = torch.optim.Adam(ref_net.parameters(), lr=2e-3, weight_decay=1e-4)
    fg = FlatGradients(flat_net.parameters())
    opt = FlatAdam(fg, lr=2e-3, weight_decay=1e-4)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.5)          # schedulers attach as to torch.optim.Adam
    ref_sched = torch.optim.lr_scheduler.StepLR(ref_opt, step_size=2, gamma=0.5)
    for _ in range(4):
        ref_flat.zero()
        calc_loss(ref_net(data), data.y).backward()
        ref_opt.step(); ref_sched.step()
        opt.zero_grad()
        calc_loss(flat_net(data), data.y).backward()
        opt.step(); sched.step()
    for (n1, p1), (_, p2) in zip(ref_net.named_parameters(), flat_net.named_parameters()):
        scale = p1.abs().max().item() + 1e-12
        assert (p1 - p2).abs().max().item() < 5e-4 * scale, n1        # Adam divides by sqrt(v): early steps amplify fp32 rounding of the gradients


def test_large_cloud_mode_morton_order_is_transparent():
    """DeltaNetBase.reorder_min_points: clouds processed in Morton order give the same per-point logits (handed back in the
    caller's order) and the same gradients as the plain order; the permutation never mixes clouds."""
    D = dc()
    from deltaconv_b200 import ops
    torch.manual_seed(2)
    data = D.utils.synthetic_batch(3, 700, config=3, num_classes=6).to(DEV)
    perm = ops.spatial_order(data.pos, data.batch, 3)
    assert torch.equal(torch.sort(perm).values, torch.arange(2100, device=DEV)) and torch.equal(data.batch[perm], data.batch)
    cell = ((data.pos[perm] - data.pos.min(0).values) * 4).long()          # consecutive points are spatially close
    assert (cell[1:] - cell[:-1]).abs().sum(1).float().mean() < (cell[torch.randperm(2100, device=DEV)] - cell).abs().sum(1).float().mean() * 0.5
    nets = []
    for _ in range(2):
        torch.manual_seed(0)
        nets.append(no_dropout(D.utils.fill_deterministic(D.models.DeltaNetSegmentation(3, 6, conv_channels=[16, 32], embedding_size=64), 4)).to(DEV).train())
    plain, reordered = nets
    reordered.deltanet_base.reorder_min_points = 1
    y = torch.randint(0, 6, (2100,), device=DEV)
    outs = []
    for net in nets:
        logits = net(data)
        torch.nn.functional.cross_entropy(logits, y).backward()
        outs.append(logits.detach())
    assert data._dc_perm is not None
    assert rel_err(outs[1].cpu(), outs[0].cpu()) < 1e-4
    scale = max(p.grad.abs().max().item() for p in plain.parameters() if p.grad is not None)     # some gradients are pure rounding noise
    for (n1, p1), (_, p2) in zip(plain.named_parameters(), reordered.named_parameters()):
        if p1.grad is not None:
            assert (p1.grad - p2.grad).abs().max().item() < 2e-3 * scale, n1


def test_transforms_run_on_a_device_batch():
    """The augmentations act on a whole device-resident batch (per-cloud parameters, no host round trip) and the result feeds
    the model as is."""
    D = dc()
    import deltaconv_b200.transforms as T
    from deltaconv_b200.compat import Compose
    torch.manual_seed(1)
    data = D.utils.synthetic_batch(4, 256, config=2, num_classes=10).to(DEV)
    data._num_clouds, data._max_points = 4, 256
    p0 = data.pos.clone()
    aug = Compose([T.NormalizeScale(), T.RandomScale((0.8, 1.25)), T.RandomRotate(20, axis=1), T.RandomTranslateGlobal(0.1), T.RandomNormals(0.02)])
    out = aug(data)
    assert out.pos.is_cuda and out.norm.is_cuda and not torch.equal(out.pos, p0)
    assert (out.norm.norm(dim=1) - 1).abs().max() < 1e-5
    net = no_dropout(D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 10, conv_channels=[16, 32]), 5)).to(DEV).train()
    assert torch.isfinite(net(out)).all()


@pytest.mark.parametrize('train', [True, False])
def test_fused_gathermax_over_batchnorm_activation_matches_the_two_pass_form(train):
    """ops.linear_bn_act_gathermax (the activation is never written: the gather-max applies BatchNorm + LeakyReLU to the
    per-channel extreme of the pre-activations) against Linear -> BatchNorm -> LeakyReLU -> gather-max as separate kernels:
    bitwise the same values, the same arg routing, the same gradients — including channels with a negative gamma, where the
    maximum of the activation is the activation of the MINIMUM."""
    D = dc()
    from deltaconv_b200 import ops
    torch.manual_seed(4)
    d = D.utils.synthetic_batch(3, 0, config=12, sizes=[300, 700, 40])
    graph = ops.knn(d.pos.to(DEV), 12, d.batch.to(DEV), num_clouds=3)
    N, K, C = d.pos.size(0), 24, 32
    lin = D.nn.Linear(K, C, bias=False).to(DEV)
    bn = torch.nn.BatchNorm1d(C).to(DEV)
    with torch.no_grad():
        bn.weight.copy_(torch.randn(C))                   # about half of the channels get a negative gamma
        bn.bias.copy_(0.3 * torch.randn(C))
        bn.running_mean.copy_(0.2 * torch.randn(C)); bn.running_var.copy_(0.5 + torch.rand(C))
    x0 = torch.randn(N, K, device=DEV)
    g0 = torch.randn(N, C, device=DEV)
    results = []
    for fused in (False, True):
        bn_i = copy.deepcopy(bn).train(train)
        lin_i = copy.deepcopy(lin)
        x = x0.clone().requires_grad_(True)
        if fused:
            out = ops.linear_bn_act_gathermax(x, lin_i.weight, bn_i, 0.2, graph)
        else:
            out = ops.gather_max(ops.bn_act(ops.linear(x, lin_i.weight), bn_i, 0.2), graph)
        (out * g0).sum().backward()
        results.append((out.detach(), x.grad, lin_i.weight.grad, bn_i.weight.grad, bn_i.bias.grad, bn_i.running_mean.clone()))
    assert (bn.weight < 0).any() and (bn.weight > 0).any()
    for a, b in zip(*results):
        assert torch.equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize('train', [True, False])
def test_fused_segmax_over_batchnorm_activation_matches_the_two_pass_form(train):
    """ops.linear_bn_act_segmax (last block of a deep centralized edge MLP folded into the max over each point's k edge rows;
    backward writes the gradient of the pre-activation directly, the BatchNorm sums run over the arg-max rows only) against
    Linear -> BatchNorm -> LeakyReLU -> segment-max as separate kernels.  Values and routing are bitwise the same; the gradients
    agree to rounding (the sums are formed over N rows instead of N*k rows with zeros: another summation tree)."""
    D = dc()
    from deltaconv_b200 import ops
    torch.manual_seed(5)
    k = 12
    d = D.utils.synthetic_batch(3, 0, config=12, sizes=[300, 700, 8])          # the 8-point cloud has padded (invalid) slots
    graph = ops.knn(d.pos.to(DEV), k, d.batch.to(DEV), num_clouds=3)
    N, K, C = d.pos.size(0), 24, 32
    lin = D.nn.Linear(K, C, bias=False).to(DEV)
    bn = torch.nn.BatchNorm1d(C).to(DEV)
    with torch.no_grad():
        bn.weight.copy_(torch.randn(C))
        bn.bias.copy_(0.3 * torch.randn(C))
        bn.running_mean.copy_(0.2 * torch.randn(C)); bn.running_var.copy_(0.5 + torch.rand(C))
    h0 = torch.randn(N * k, K, device=DEV)
    g0 = torch.randn(N, C, device=DEV)
    results = []
    for fused in (False, True):
        bn_i = copy.deepcopy(bn).train(train)
        lin_i = copy.deepcopy(lin)
        h = h0.clone().requires_grad_(True)
        if fused:
            out = ops.linear_bn_act_segmax(h, lin_i.weight, bn_i, 0.2, graph)
        else:
            out = ops.segment_max(ops.bn_act(ops.linear(h, lin_i.weight), bn_i, 0.2), graph)
        (out * g0).sum().backward()
        results.append((out.detach(), bn_i.running_mean.clone(), h.grad, lin_i.weight.grad, bn_i.weight.grad, bn_i.bias.grad))
    assert (bn.weight < 0).any() and (bn.weight > 0).any()
    assert torch.equal(results[0][0], results[1][0]) and torch.equal(results[0][1], results[1][1])
    for name, a, b in zip(('dh', 'dW', 'dgamma', 'dbeta'), results[0][2:], results[1][2:]):
        err = rel_err(b, a)
        report(f'fused_segmax_bn.train{int(train)}.{name}_vs_two_pass', err)
        assert err < 2e-5, (name, err)


@pytest.mark.gpu
@pytest.mark.parametrize('train', [True, False])
@pytest.mark.parametrize('depth', [2, 3])
def test_deep_centralized_edge_mlp_matches_the_plain_sequence_and_the_oracle(train, depth):
    """DeltaConv(centralized=True, depth >= 2) — the ShapeNet configuration's first layer.  The scheduled form (first block
    written from the coordinates with moment statistics, last block folded into the max) against the plain sequence
    edge_diff -> MLP -> segment_max on the same parameters: output, running statistics and every parameter gradient; and the
    output against the fp64 oracle's edge MLP."""
    D = dc()
    from deltaconv_b200 import ops
    torch.manual_seed(6)
    k = 10
    d = D.utils.synthetic_batch(3, 0, config=12, sizes=[500, 260, 7])
    pos = d.pos.to(DEV)
    graph = ops.knn(pos, k, d.batch.to(DEV), num_clouds=3)
    conv = D.nn.DeltaConv(3, 32, depth=depth, centralized=True).to(DEV)
    with torch.no_grad():
        for m in conv.s_mlp_max.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.weight.copy_(torch.randn_like(m.weight)); m.bias.copy_(0.3 * torch.randn_like(m.bias))
                m.running_mean.copy_(0.1 * torch.randn_like(m.running_mean)); m.running_var.copy_(0.5 + torch.rand_like(m.running_var))
    g0 = torch.randn(pos.size(0), 32, device=DEV)
    results = []
    for fused in (False, True):
        c = copy.deepcopy(conv).train(train)
        if fused:
            out = c._centralized_deep(pos, graph)
        else:
            out = ops.segment_max(c.s_mlp_max(ops.edge_diff(pos, graph)), graph)
        (out * g0).sum().backward()
        params = dict(c.s_mlp_max.named_parameters())
        stats = {n: b.clone() for n, b in c.s_mlp_max.named_buffers() if 'running' in n}
        results.append((out.detach(), {n: p.grad for n, p in params.items()}, stats))
    (out_a, grads_a, stats_a), (out_b, grads_b, stats_b) = results
    err = rel_err(out_b, out_a)
    report(f'deep_centralized.depth{depth}.train{int(train)}.out_vs_plain', err)
    assert err < 1e-5, err
    for n in stats_a:
        assert rel_err(stats_b[n], stats_a[n]) < 1e-5, n
    for n in grads_a:
        assert grads_b[n] is not None, n
        e = rel_err(grads_b[n], grads_a[n])
        report(f'deep_centralized.depth{depth}.train{int(train)}.d{n}', e)
        assert e < 2e-4, (n, e)            # arg-max ties / kinks between two fp32 evaluations of the first block
    # fp64 restatement of deltaconv.py:50-52 (torch, CPU)
    ref = copy.deepcopy(conv).train(train).cpu().double()
    idx = graph.idx.cpu().long()
    N = pos.size(0)
    valid = (idx >= 0) & (idx < N)
    nb = torch.where(valid, idx, torch.arange(N)[:, None].expand_as(idx))
    p64 = pos.cpu().double()
    t = (p64[nb] - p64[:, None, :]).reshape(N * k, 3)
    for block in ref.s_mlp_max:
        lin, bnw, act = block[0], block[1].bn, block[2]
        t = t @ lin.weight.T
        if train:
            mu, var = t.mean(0), t.var(0, unbiased=False)
        else:
            mu, var = bnw.running_mean, bnw.running_var
        t = torch.nn.functional.leaky_relu((t - mu) / torch.sqrt(var + bnw.eps) * bnw.weight + bnw.bias, act.negative_slope)
    t = t.reshape(N, k, -1).masked_fill(~valid[:, :, None], float('-inf')).amax(1)
    err = rel_err(out_b.cpu().double(), t)
    report(f'deep_centralized.depth{depth}.train{int(train)}.out_vs_fp64', err)
    assert err < 2e-5, err


@pytest.mark.gpu
@pytest.mark.parametrize('train', [True, False])
def test_segmentation_head_without_the_broadcast_matches_the_materialised_form(train):
    """DeltaNetSegmentation computes the head's first Linear over [x_max[batch] | categorical[batch] | conv features] as a per-point
    product plus a per-cloud product added per cloud (ops.linear_bcast_bn_act).  Against the materialised concatenation
    (decompose_head = False) on the same parameters: logits and every parameter gradient."""
    D = dc()
    torch.manual_seed(8)
    d = D.utils.synthetic_batch(3, 0, config=13, sizes=[400, 250, 350], per_point_labels=True, num_classes=50, categories=16)
    data = d.to(DEV)
    net = D.models.DeltaNetSegmentation(3, 50, conv_channels=[32, 64], mlp_depth=2, embedding_size=128, categorical_vector=True,
                                        num_neighbors=12).to(DEV)
    net = D.utils.fill_deterministic(net, 3).to(DEV)
    results = []
    for decompose in (False, True):
        m = copy.deepcopy(net).train(train)
        m.decompose_head = decompose
        for mod in m.modules():
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
        out = m(data)
        F.cross_entropy(out, data.y).backward()
        results.append((out.detach(), {n: p.grad for n, p in m.named_parameters()}))
    (out_a, ga), (out_b, gb) = results
    err = report(f'seg_head_decomposed.train{int(train)}.logits', rel_err(out_b, out_a))
    assert err < 2e-5, err
    # Gradients.  The two forms agree to ~2e-6 in the head's first pre-activation, which is enough to put a handful of its
    # ~640 K LeakyReLU inputs (three head blocks x 1000 points) on the other side of zero; ONE such flip changes the gradient of
    # one point by O(1) of itself, i.e. every upstream tensor by ~1/N = 1e-3 here (measured on B200: training mode no flip,
    # median 2.5e-6, worst 2.7e-5; evaluation mode one flip, median 1.6e-3, worst 9.9e-3).  A wrong block offset or a missing
    # term shows as O(0.1 - 1): the bounds sit between.  A tensor's error is relative to max(its own scale, 1e-3 of the
    # network's largest gradient) — in training mode some gradients are analytically zero (a shift of lin_global's output is
    # removed by the head's BatchNorm).  Parity of the gradients with the fp64 oracle is test_gpu_configs.py's job.
    scale = max(float(g.abs().max()) for g in ga.values() if g is not None)
    errs = {}
    for n in ga:
        if ga[n] is None:
            assert gb[n] is None, n
            continue
        errs[n] = float((gb[n] - ga[n]).abs().max()) / max(float(ga[n].abs().max()), 1e-3 * scale)
    ordered = sorted(errs.values())
    median, worst = ordered[len(ordered) // 2], ordered[-1]
    report(f'seg_head_decomposed.train{int(train)}.median_grad', median)
    report(f'seg_head_decomposed.train{int(train)}.worst_grad', worst)
    assert median < 1e-2, median
    assert worst < 1e-1, max(errs, key=errs.get)
