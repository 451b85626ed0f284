"""Pins oracle/ (the CPU restatement) against outputs of the unmodified reference stored in
tests/golden/*.npz by oracle/gen_golden.py.  CPU only."""
import numpy as np
import pytest
import torch

from helpers import load, t, row_scaled_err, rel_err, check_frames, OPERATOR_ROW_REL
from oracle import geometry as og
from oracle import network as on
from deltaconv_b200.utils import Data, fill_deterministic

GEOM = ['geom_normals_k10', 'geom_ragged_k20', 'geom_estimated_k16', 'geom_unnormalized_shape']


def _fit(g, dtype):
    pos = t(g.pos, dtype)
    batch = t(g.batch)
    k = int(g.k)
    idx = og.knn_ell(pos, k, batch)
    if int(g.given_normals):
        normal = t(g.input_normal, dtype)
        xb, yb = og.tangent_basis_from_normal(normal)
    else:                      # frames are gauge-dependent (SVD sign): take the fp64 reference's frames
        normal, xb, yb = (t(getattr(g, f'ref64_{n}'), dtype) for n in ('normal', 'xb', 'yb'))
    lam_shape = None if float(g.lam_shape) < 0 else float(g.lam_shape)
    G, D, parts = og.fit_operators(pos, normal, xb, yb, idx, batch, float(g.width), float(g.lam),
                                   bool(int(g.normalized)), lam_shape, return_parts=True)
    return pos, batch, idx, normal, xb, yb, G, D, parts


@pytest.mark.parametrize('name', GEOM)
def test_knn_matches_reference_graph(name):
    g, _ = load(name)
    for tag, dt in (('ref64', torch.float64), ('ref32', torch.float32)):
        idx = og.knn_ell(t(g.pos, dt), int(g.k), t(g.batch))
        ei = og.edge_index_from_ell(idx)
        assert torch.equal(ei, t(getattr(g, f'{tag}_edge_index')))
        assert torch.equal(og.ell_from_edge_index(ei, idx.size(0)), idx)
        assert (idx[:, 0] == torch.arange(idx.size(0))).all()       # self first


@pytest.mark.parametrize('name', GEOM)
def test_tangent_basis(name):
    g, _ = load(name)
    if not int(g.given_normals):
        pytest.skip('frames estimated')
    for tag, dt, tol in (('ref64', torch.float64, 1e-13), ('ref32', torch.float32, 1e-6)):
        xb, yb = og.tangent_basis_from_normal(t(g.input_normal, dt))
        assert (xb - t(getattr(g, f'{tag}_xb'))).abs().max() < tol
        assert (yb - t(getattr(g, f'{tag}_yb'))).abs().max() < tol


def test_estimate_basis_gauge_invariant():
    g, _ = load('geom_estimated_k16')
    pos = t(g.pos, torch.float64)
    idx10 = og.ell_from_edge_index(t(g.ref64_edge_index10), pos.size(0))
    assert torch.equal(idx10, og.knn_ell(pos, 10, t(g.batch)))
    normal, xb, yb = og.estimate_basis(pos, idx10, orientation=pos)
    check_frames(normal, xb, yb, 1e-9)
    n_ref, x_ref = t(g.ref64_normal), t(g.ref64_xb)
    assert ((normal * n_ref).sum(1) - 1).abs().max() < 1e-9           # oriented the same way
    assert ((xb * x_ref).sum(1).abs() - 1).abs().max() < 1e-7          # same axis up to sign


@pytest.mark.parametrize('name', GEOM)
def test_operator_entries(name):
    g, _ = load(name)
    # fp64 oracle vs fp64 reference: agreement far below the parity tolerance
    pos, batch, idx, *_rest, G, D, parts = _fit(g, torch.float64)
    N = pos.size(0)
    (gr, gc, gv), (dr, dc, dv) = og.ell_to_coo(idx, G, D)
    assert torch.equal(gr, t(g.ref64_grad_row)) and torch.equal(gc, t(g.ref64_grad_col))
    assert torch.equal(dr, t(g.ref64_div_row)) and torch.equal(dc, t(g.ref64_div_col))
    assert row_scaled_err(gv, t(g.ref64_grad_val), gr, 2 * N) < 1e-9
    assert row_scaled_err(dv, t(g.ref64_div_val), dr, N) < 1e-9
    E = N * int(g.k)
    assert (parts['coords'].reshape(E, 2) - t(g.ref64_coords)).abs().max() < 1e-12
    assert (parts['weights'].reshape(E) - t(g.ref64_weights)).abs().max() < 1e-12
    assert rel_err(parts['wls'].reshape(E, 6), t(g.ref64_wls)) < 1e-9
    assert (parts['mapping'].reshape(E, 2, 2) - t(g.ref64_mapping)).abs().max() < 1e-9
    # fp32 oracle vs fp64 reference: inside the stated parity tolerance
    pos, batch, idx, *_rest, G, D, _ = _fit(g, torch.float32)
    (gr, gc, gv), (dr, dc, dv) = og.ell_to_coo(idx, G, D)
    assert row_scaled_err(gv, t(g.ref64_grad_val), gr, 2 * N) < OPERATOR_ROW_REL
    assert row_scaled_err(dv, t(g.ref64_div_val), dr, N) < OPERATOR_ROW_REL


@pytest.mark.parametrize('name', GEOM)
def test_operator_application(name):
    g, _ = load(name)
    pos, batch, idx, *_rest, G, D, _ = _fit(g, torch.float64)
    x, v = t(g.x, torch.float64), t(g.v, torch.float64)
    assert rel_err(og.apply_grad(idx, G, x), t(g.ref64_grad_x)) < 1e-9
    assert rel_err(og.apply_div(idx, D, v), t(g.ref64_div_v)) < 1e-9
    assert rel_err(og.curl(idx, D, v), t(g.ref64_curl_v)) < 1e-9
    assert rel_err(og.laplacian(idx, G, D, x), t(g.ref64_lap_x)) < 1e-9
    assert rel_err(og.hodge_laplacian(idx, G, D, v), t(g.ref64_hodge_v)) < 1e-9
    assert rel_err(og.vec_norm(v), t(g.ref64_norm_v)) < 1e-12
    assert torch.equal(og.rot90(v), t(g.ref64_J_v))


def test_deltaconv_layers_forward_backward():
    g, z = load('layers_deltaconv')
    dt = torch.float64
    pos, normal, batch, k = t(g.pos, dt), t(g.normal, dt), t(g.batch), int(g.k)
    idx = og.knn_ell(pos, k, batch)
    xb, yb = og.tangent_basis_from_normal(normal)
    G, D = og.fit_operators(pos, normal, xb, yb, idx, batch)
    conv0 = fill_deterministic(on.DeltaConv(3, 16, depth=1, centralized=True, vector=True), 3).to(dt)
    conv1 = fill_deterministic(on.DeltaConv(16, 8, depth=2, centralized=False, vector=True), 4).to(dt)
    x_in = pos.clone().requires_grad_(True)
    x1, v1 = conv0(x_in, og.apply_grad(idx, G, x_in), idx, G, D)
    x2, v2 = conv1(x1, v1, idx, G, D)
    ((x2 * t(g.wx, dt)).sum() + (v2 * t(g.wv, dt)).sum()).backward()
    for ours, key in ((x1, 'x1'), (v1, 'v1'), (x2, 'x2'), (v2, 'v2'), (x_in.grad, 'dx_in')):
        assert rel_err(ours, t(z[f'ref64_{key}'])) < 1e-8, key
    for cname, conv in (('conv0', conv0), ('conv1', conv1)):
        for pname, p in conv.named_parameters():
            key = f'ref64_grad.{cname}.{pname}'
            if key in z.files:
                assert rel_err(p.grad, t(z[key])) < 1e-7, key
            else:
                assert p.grad is None or p.grad.abs().max() == 0, key
        for bname, b in conv.named_buffers():
            if 'running' in bname:
                assert rel_err(b, t(z[f'ref64_buf.{cname}.{bname}'])) < 1e-9, bname


def _model(kind, k):
    if kind == 'model_cls':
        return on.DeltaNetClassification(3, 7, conv_channels=[16, 16, 32], num_neighbors=k)
    if kind == 'model_cls_nonormal':
        return on.DeltaNetClassification(3, 5, conv_channels=[8, 16], num_neighbors=k)
    return on.DeltaNetSegmentation(3, 6, conv_channels=[8, 16, 16], mlp_depth=2, embedding_size=64,
                                   categorical_vector=True, num_neighbors=k)


@pytest.mark.parametrize('kind', ['model_cls', 'model_cls_nonormal', 'model_seg'])
def test_models(kind):
    g, z = load(kind)
    dt = torch.float64
    model = fill_deterministic(_model(kind, int(g.k)), 2).to(dt)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    data = Data(pos=t(g.pos, dt), batch=t(g.batch), y=t(g.y))
    data.norm = t(g.normal, dt) if 'normal' in z.files else None
    if 'category' in z.files:
        data.category = t(g.category, dt)
    model.eval()
    with torch.no_grad():
        assert rel_err(model(data), t(g.ref64_logits_eval)) < 1e-8
    model.train()
    logits = model(data)
    loss = torch.nn.functional.cross_entropy(logits, data.y)
    loss.backward()
    # estimated frames are gauge-dependent (SVD sign) but logits are gauge-invariant only up to
    # BN-in-training round-off, hence the same tight bound holds
    assert rel_err(logits, t(g.ref64_logits_train)) < 1e-7
    assert abs(loss.item() - float(g.ref64_loss)) < 1e-9
    names = dict(model.named_parameters())
    for key in z.files:
        if key.startswith('ref64_gnorm.'):
            p = names[key[len('ref64_gnorm.'):]]
            ref = float(z[key])
            assert abs(p.grad.norm().item() - ref) <= 1e-6 * max(ref, 1e-12) + 1e-12, key


def test_state_dict_keys_match_reference_layout():
    keys = set(on.DeltaNetClassification(3, 40).state_dict())
    for probe in ('deltanet_base.convs.0.s_mlp_max.0.0.weight', 'deltanet_base.convs.0.s_mlp.0.1.bn.running_mean',
                  'deltanet_base.convs.0.v_mlp.0.1.bias', 'deltanet_base.convs.0.v_mlp.0.1.batchnorm.bn.weight',
                  'classification_head.4.bias', 'lin_embedding.0.0.weight'):
        assert probe in keys
    assert len(keys) == 89
    assert sum(p.numel() for p in on.DeltaNetClassification(3, 40).parameters()) == 2045928
