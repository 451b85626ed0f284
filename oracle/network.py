"""Oracle (CPU torch) for the learned half of the hot path: DeltaConv, the scalar/vector MLPs,
DeltaNetBase and the two heads, computing on the ELL operators of oracle.geometry.

TEST INFRASTRUCTURE — see oracle/__init__.py.  Module/attribute names reproduce the reference's
state-dict keys (`deltanet_base.convs.0.s_mlp_max.0.0.weight`,
`...s_mlp.0.1.bn.running_mean`, `...v_mlp.0.1.batchnorm.bn.weight`, `classification_head.4.bias`)
so that one set of weights loads into the reference, the oracle and the CUDA product.
"""
import torch
from torch import nn

from . import geometry as og


class BatchNorm1d(nn.Module):
    """nonlin.py:11-35 — the reference reshapes [N,C] -> [1,C,N]; the statistics are the same
    as nn.BatchNorm1d on [N,C].  State-dict prefix `.bn.`."""

    def __init__(self, channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.bn = nn.BatchNorm1d(channels, eps, momentum, affine, track_running_stats)

    def forward(self, x):
        return self.bn(x)


class VectorNonLin(nn.Module):
    """nonlin.py:38-86 — non-linearity on the per-point norm of a [2N,C] vector field."""

    def __init__(self, channels, nonlin=None, batchnorm=None):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(channels))
        self.nonlin = nn.ReLU() if nonlin is None else nonlin
        self.batchnorm = batchnorm

    def forward(self, v):
        rows, C = v.shape
        vv = v.reshape(-1, 2, C)
        n = vv.norm(dim=1)
        shifted = n + self.bias[None, :] if self.batchnorm is None else self.batchnorm(n)
        gain = self.nonlin(shifted) / n.clamp(1e-8)
        return (vv * gain[:, None, :]).reshape(rows, C)


def MLP(channels, bias=False, nonlin=None):
    """mlp.py:7-11."""
    act = nn.LeakyReLU(negative_slope=0.2) if nonlin is None else nonlin
    return nn.Sequential(*[
        nn.Sequential(nn.Linear(channels[i - 1], channels[i], bias=bias), BatchNorm1d(channels[i]), act)
        for i in range(1, len(channels))])


def VectorMLP(channels, batchnorm=True):
    """mlp.py:13-17."""
    return nn.Sequential(*[
        nn.Sequential(nn.Linear(channels[i - 1], channels[i], bias=False),
                      VectorNonLin(channels[i], batchnorm=BatchNorm1d(channels[i]) if batchnorm else None))
        for i in range(1, len(channels))])


class DeltaConv(nn.Module):
    """deltaconv.py:8-72 on ELL operators: forward(x, v, idx, G, D)."""

    def __init__(self, in_channels, out_channels, depth=1, centralized=False, vector=True):
        super().__init__()
        self.in_channels, self.out_channels, self.centralized = in_channels, out_channels, centralized
        self.s_mlp_max = MLP([in_channels] + [out_channels] * depth)
        self.s_mlp = MLP([in_channels * 4] + [out_channels] * depth)
        self.v_mlp = VectorMLP([in_channels * 4 + out_channels * 2] + [out_channels] * depth) if vector else None

    def forward(self, x, v, idx, G, D):
        N, k = idx.shape
        if self.centralized:                                        # deltaconv.py:50-52
            edge_feat = (x[idx] - x[:, None, :]).reshape(N * k, -1)
            x_max = self.s_mlp_max(edge_feat).reshape(N, k, -1).max(dim=1).values
        else:                                                       # deltaconv.py:54
            x_max = self.s_mlp_max(x)[idx].max(dim=1).values
        div_v = og.apply_div(idx, D, v)
        curl_v = og.curl(idx, D, v)
        x = x_max + self.s_mlp(torch.cat([x, div_v, curl_v, og.vec_norm(v)], dim=1))   # :57-59
        if self.v_mlp is not None:                                  # :64-68
            hodge = -(og.apply_grad(idx, G, div_v) + og.rot90(og.apply_grad(idx, G, curl_v)))
            v_cat = torch.cat([v, hodge, og.apply_grad(idx, G, x)], dim=1)
            v = self.v_mlp(og.with_rot90(v_cat))
        return x, v


def _fast_knn(pos, k, ptr):
    """cdist + topk kNN used ONLY when timing the CPU baseline (the timed
    stand-in must not be a strawman); tie order is whatever topk gives."""
    out = torch.empty(pos.size(0), k, dtype=torch.long)
    for s, e in zip(ptr[:-1], ptr[1:]):
        d = torch.cdist(pos[s:e], pos[s:e])
        out[s:e] = d.topk(k, dim=1, largest=False).indices + s
    return out


class DeltaNetBase(nn.Module):
    """deltanet_base.py:8-87."""

    def __init__(self, in_channels, conv_channels, mlp_depth, num_neighbors, grad_regularizer,
                 grad_kernel_width, centralize_first=True):
        super().__init__()
        self.k, self.lam, self.width = num_neighbors, grad_regularizer, grad_kernel_width
        chans = [in_channels] + list(conv_channels)
        self.convs = nn.ModuleList([
            DeltaConv(chans[i], chans[i + 1], depth=mlp_depth, centralized=(centralize_first and i == 0),
                      vector=(i != len(chans) - 2)) for i in range(len(chans) - 1)])
        self.fast_knn = False

    def operators(self, data):
        pos, batch = data.pos, data.batch
        with torch.no_grad():
            if self.fast_knn:
                ptr = og.cloud_ptr(batch, pos.size(0)).tolist()
                idx = _fast_knn(pos, self.k, ptr)
            else:
                idx = og.knn_ell(pos, self.k, batch)
            normal = getattr(data, 'norm', None)
            if normal is not None:
                xb, yb = og.tangent_basis_from_normal(normal)
            else:
                if self.fast_knn:
                    idx_n = _fast_knn(pos, 10, ptr)
                else:
                    idx_n = og.knn_ell(pos, 10, batch)
                normal, xb, yb = og.estimate_basis(pos, idx_n, orientation=pos)
            G, D = og.fit_operators(pos, normal, xb, yb, idx, batch, self.width, self.lam)
        return idx, G, D

    def forward(self, data):
        idx, G, D = self.operators(data)
        x = data.x if getattr(data, 'x', None) is not None else data.pos
        v = og.apply_grad(idx, G, x)
        outs = []
        for conv in self.convs:
            x, v = conv(x, v, idx, G, D)
            outs.append(x)
        return outs


def _pool(x, batch, num_clouds, mode):
    if mode == 'max':
        init = x.new_full((num_clouds, x.size(1)), float('-inf'))
        return init.scatter_reduce(0, batch[:, None].expand_as(x), x, reduce='amax', include_self=True)
    total = x.new_zeros(num_clouds, x.size(1)).index_add(0, batch, x)
    count = torch.bincount(batch, minlength=num_clouds).clamp(min=1).to(x.dtype)
    return total / count[:, None]


class DeltaNetClassification(nn.Module):
    """deltanet_classification.py:9-51."""

    def __init__(self, in_channels, num_classes, conv_channels=(64, 64, 128, 256), num_neighbors=20,
                 grad_regularizer=1e-3, grad_kernel_width=1):
        super().__init__()
        conv_channels = list(conv_channels)
        self.deltanet_base = DeltaNetBase(in_channels, conv_channels, 1, num_neighbors, grad_regularizer,
                                          grad_kernel_width)
        self.lin_embedding = MLP([sum(conv_channels), 1024])
        self.classification_head = nn.Sequential(
            MLP([2048, 512]), nn.Dropout(0.5), MLP([512, 256]), nn.Dropout(0.5), nn.Linear(256, num_classes))

    def forward(self, data):
        x = self.lin_embedding(torch.cat(self.deltanet_base(data), dim=1))
        B = int(data.batch.max()) + 1
        pooled = torch.cat([_pool(x, data.batch, B, 'max'), _pool(x, data.batch, B, 'mean')], dim=1)
        return self.classification_head(pooled)


class DeltaNetSegmentation(nn.Module):
    """deltanet_segmentation.py:9-69."""

    def __init__(self, in_channels, num_classes, conv_channels=(64, 128, 256), mlp_depth=2,
                 embedding_size=1024, categorical_vector=False, num_neighbors=20,
                 grad_regularizer=0.001, grad_kernel_width=1):
        super().__init__()
        conv_channels = list(conv_channels)
        self.categorical_vector = categorical_vector
        self.deltanet_base = DeltaNetBase(in_channels, conv_channels, mlp_depth, num_neighbors,
                                          grad_regularizer, grad_kernel_width)
        self.lin_global = MLP([sum(conv_channels), embedding_size])
        width = embedding_size + sum(conv_channels)
        if categorical_vector:
            self.lin_categorical = MLP([16, 64])
            width += 64
        self.segmentation_head = nn.Sequential(
            MLP([width, 256]), nn.Dropout(0.5), MLP([256, 256]), nn.Dropout(0.5),
            nn.Linear(256, 128), nn.LeakyReLU(negative_slope=0.2), nn.Linear(128, num_classes))

    def forward(self, data):
        conv_out = self.deltanet_base(data)
        x = self.lin_global(torch.cat(conv_out, dim=1))
        B = int(data.batch.max()) + 1
        x_glob = _pool(x, data.batch, B, 'max')[data.batch]
        if self.categorical_vector:
            x_glob = torch.cat([x_glob, self.lin_categorical(data.category)[data.batch]], dim=1)
        return self.segmentation_head(torch.cat([x_glob] + conv_out, dim=1))


def smoothed_cross_entropy(pred, target, eps=0.2):
    """experiments/utils.py:7-24 (label smoothing on, the scripts' default)."""
    n_class = pred.size(1)
    soft = torch.full_like(pred, eps / (n_class - 1)).scatter(1, target.view(-1, 1), 1 - eps)
    return -(soft * torch.log_softmax(pred, dim=1)).sum(dim=1).mean()
