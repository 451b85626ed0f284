"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) on seeded
inputs.  Runs only in the build container (the reference tree does not travel to the GPU box):

    python oracle/gen_golden.py            # writes tests/golden/*.npz

The reference's third-party imports are satisfied by oracle/ref_standins (pure torch, semantics
per SURVEY.md App. B).  Each fixture stores the inputs and the reference's outputs in fp64
("ref64": the value the parity tolerance is stated against) and in fp32 ("ref32": what the
reference itself produces in its native precision).  TEST INFRASTRUCTURE.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REFERENCE = os.environ.get('DELTACONV_REFERENCE', '/root/reference')
sys.path[:0] = [os.path.join(HERE, 'ref_standins'), REFERENCE, REPO]

import deltaconv as ref                                             # noqa: E402  (the reference)
from deltaconv.geometry import grad_div_mls as ref_g                # noqa: E402
from deltaconv.geometry import operators as ref_op                  # noqa: E402
from torch_geometric.nn import knn_graph                            # noqa: E402  (stand-in)

from deltaconv_b200.utils import Data, synthetic_batch, fill_deterministic   # noqa: E402

assert ref.__file__.startswith(REFERENCE), ref.__file__
OUT = os.path.join(REPO, 'tests', 'golden')


def npy(t):
    return t.detach().cpu().numpy()


def coo(sp):
    r, c, v = sp.coo()
    return npy(r), npy(c), npy(v)


def geometry_case(name, sizes, k, cfg, given_normals, lam=1e-3, width=1.0, normalized=True, lam_shape=None):
    data = synthetic_batch(len(sizes), 0, config=cfg, sizes=sizes)
    store = dict(pos=npy(data.pos), batch=npy(data.batch), k=k, lam=lam, width=width,
                 normalized=int(normalized), lam_shape=-1.0 if lam_shape is None else lam_shape,
                 given_normals=int(given_normals), input_normal=npy(data.norm))
    for tag, dt in (('ref64', torch.float64), ('ref32', torch.float32)):
        pos = data.pos.to(dt)
        edge_index = knn_graph(pos, k, data.batch, loop=True, flow='target_to_source')
        if given_normals:
            normal = data.norm.to(dt)
            xb, yb = ref_g.build_tangent_basis(normal)
        else:
            ei10 = knn_graph(pos, 10, data.batch, loop=True, flow='target_to_source')
            normal, xb, yb = ref_g.estimate_basis(pos, ei10, orientation=pos)
            store[f'{tag}_edge_index10'] = npy(ei10)
        grad, div = ref_g.build_grad_div(pos, normal, xb, yb, edge_index.clone(), data.batch,
                                         kernel_width=width, regularizer=lam, normalized=normalized,
                                         shape_regularizer=lam_shape)
        # the public helpers, called the way build_grad_div calls them (grad_div_mls.py:230-268)
        row, col = edge_index
        coords = ref_g.coords_projected(pos, normal, xb, yb, edge_index, k)
        dist = torch.linalg.norm(pos[col] - pos[row], dim=1)
        weights = ref_g.gaussian_weights(dist, k, data.batch, width)
        if lam_shape is None:
            wls = wls_surface = ref_g.weighted_least_squares(coords, weights, k, lam)
        else:                                                       # grad_div_mls.py:244,266-267
            wls, wls_surface = ref_g.weighted_least_squares(coords, weights, k, lam, lam_shape)
        mapping = ref_g.fit_vector_mapping(pos, normal, xb, yb, (row, col), wls_surface, coords)
        gen = torch.Generator().manual_seed(7)
        x = torch.randn(pos.size(0), 5, generator=gen).to(dt)
        v = torch.randn(2 * pos.size(0), 5, generator=gen).to(dt)
        gr, gc, gv = coo(grad)
        dr, dc, dv = coo(div)
        store.update({
            f'{tag}_edge_index': npy(edge_index), f'{tag}_normal': npy(normal),
            f'{tag}_xb': npy(xb), f'{tag}_yb': npy(yb),
            f'{tag}_grad_row': gr, f'{tag}_grad_col': gc, f'{tag}_grad_val': gv,
            f'{tag}_div_row': dr, f'{tag}_div_col': dc, f'{tag}_div_val': dv,
            f'{tag}_coords': npy(coords), f'{tag}_weights': npy(weights), f'{tag}_wls': npy(wls),
            f'{tag}_mapping': npy(mapping),
            f'{tag}_grad_x': npy(grad @ x), f'{tag}_div_v': npy(div @ v),
            f'{tag}_curl_v': npy(ref_op.curl(v, div)), f'{tag}_lap_x': npy(ref_op.laplacian(x, grad, div)),
            f'{tag}_hodge_v': npy(ref_op.hodge_laplacian(v, grad, div)),
            f'{tag}_norm_v': npy(ref_op.norm(v)), f'{tag}_J_v': npy(ref_op.J(v)),
        })
        if tag == 'ref64':
            store['x'], store['v'] = npy(x.float()), npy(v.float())
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **store)
    print('wrote', name)


def no_dropout(model):
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    return model


def sampled(t, limit=4096):
    flat = t.detach().reshape(-1)
    if flat.numel() <= limit:
        return npy(flat)
    step = flat.numel() // limit + 1
    return npy(flat[::step])


def layer_case(name):
    """One centralized DeltaConv (3 -> 16) feeding a plain one (16 -> 8), training-mode BN, with
    gradients w.r.t. every parameter and the inputs (reference: deltaconv.py:44-70)."""
    sizes, k = [80, 112], 12
    data = synthetic_batch(len(sizes), 0, config=11, sizes=sizes)
    store = dict(pos=npy(data.pos), batch=npy(data.batch), normal=npy(data.norm), k=k)
    for tag, dt in (('ref64', torch.float64), ('ref32', torch.float32)):
        pos, normal = data.pos.to(dt), data.norm.to(dt)
        edge_index = knn_graph(pos, k, data.batch, loop=True, flow='target_to_source')
        xb, yb = ref_g.build_tangent_basis(normal)
        grad, div = ref_g.build_grad_div(pos, normal, xb, yb, edge_index.clone(), data.batch)
        conv0 = fill_deterministic(ref.nn.DeltaConv(3, 16, depth=1, centralized=True, vector=True), 3).to(dt)
        conv1 = fill_deterministic(ref.nn.DeltaConv(16, 8, depth=2, centralized=False, vector=True), 4).to(dt)
        x_in = pos.clone().requires_grad_(True)
        v0 = grad @ x_in
        x1, v1 = conv0(x_in, v0, grad, div, edge_index)
        x2, v2 = conv1(x1, v1, grad, div, edge_index)
        gen = torch.Generator().manual_seed(5)
        wx = torch.randn(x2.shape, generator=gen).to(dt)
        wv = torch.randn(v2.shape, generator=gen).to(dt)
        loss = (x2 * wx).sum() + (v2 * wv).sum()
        loss.backward()
        store.update({f'{tag}_x1': npy(x1), f'{tag}_v1': npy(v1), f'{tag}_x2': npy(x2), f'{tag}_v2': npy(v2),
                      f'{tag}_dx_in': npy(x_in.grad)})
        for cname, conv in (('conv0', conv0), ('conv1', conv1)):
            for pname, p in conv.named_parameters():
                if p.grad is not None:
                    store[f'{tag}_grad.{cname}.{pname}'] = npy(p.grad)
            for bname, b in conv.named_buffers():
                if 'running' in bname:
                    store[f'{tag}_buf.{cname}.{bname}'] = npy(b)
        if tag == 'ref64':
            store['wx'], store['wv'] = npy(wx.float()), npy(wv.float())
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **store)
    print('wrote', name)


def model_case(name, kind):
    """Whole networks with small conv widths: logits in eval mode, logits + loss + parameter
    gradients in training mode (dropout probability set to 0 so the pass is deterministic)."""
    if kind == 'cls':
        sizes, k, cfg = [96, 96, 96], 10, 21
        data = synthetic_batch(len(sizes), 0, config=cfg, sizes=sizes, num_classes=7)
        build = lambda: ref.models.DeltaNetClassification(3, 7, conv_channels=[16, 16, 32], num_neighbors=k)   # noqa: E731
    elif kind == 'cls_nonormal':
        sizes, k, cfg = [120, 100, 90, 110], 12, 22      # >= 4 clouds: the head's BatchNorm runs over B rows
        data = synthetic_batch(len(sizes), 0, config=cfg, sizes=sizes, num_classes=5, normals=False)
        build = lambda: ref.models.DeltaNetClassification(3, 5, conv_channels=[8, 16], num_neighbors=k)       # noqa: E731
    else:
        sizes, k, cfg = [128, 128], 14, 23
        data = synthetic_batch(len(sizes), 0, config=cfg, sizes=sizes, num_classes=6, per_point_labels=True,
                               categories=16)
        build = lambda: ref.models.DeltaNetSegmentation(3, 6, conv_channels=[8, 16, 16], mlp_depth=2,          # noqa: E731
                                                        embedding_size=64, categorical_vector=True,
                                                        num_neighbors=k)
    store = dict(pos=npy(data.pos), batch=npy(data.batch), y=npy(data.y), k=k)
    if hasattr(data, 'norm'):
        store['normal'] = npy(data.norm)
    if hasattr(data, 'category'):
        store['category'] = npy(data.category)
    for tag, dt in (('ref64', torch.float64), ('ref32', torch.float32)):
        model = no_dropout(fill_deterministic(build(), 2)).to(dt)
        d = data.double() if dt == torch.float64 else data
        if not hasattr(d, 'norm'):
            d.norm = None
        model.eval()
        with torch.no_grad():
            store[f'{tag}_logits_eval'] = npy(model(d))
        model.train()
        logits = model(d)
        loss = torch.nn.functional.cross_entropy(logits, data.y)
        loss.backward()
        store[f'{tag}_logits_train'] = npy(logits)
        store[f'{tag}_loss'] = npy(loss)
        for pname, p in model.named_parameters():
            if p.grad is not None:
                store[f'{tag}_gnorm.{pname}'] = npy(p.grad.norm())
                store[f'{tag}_gsample.{pname}'] = sampled(p.grad)
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **store)
    print('wrote', name)


if __name__ == '__main__':
    torch.manual_seed(0)
    os.makedirs(OUT, exist_ok=True)
    geometry_case('geom_normals_k10', [128, 128], 10, 1, True)
    geometry_case('geom_ragged_k20', [50, 64, 80], 20, 2, True)
    geometry_case('geom_estimated_k16', [96, 160], 16, 3, False)
    geometry_case('geom_unnormalized_shape', [100], 12, 4, True, lam=1e-4, width=1.5, normalized=False,
                  lam_shape=1e-2)
    layer_case('layers_deltaconv')
    model_case('model_cls', 'cls')
    model_case('model_cls_nonormal', 'cls_nonormal')
    model_case('model_seg', 'seg')
