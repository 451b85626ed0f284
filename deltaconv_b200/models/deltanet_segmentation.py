"""DeltaNet part-segmentation network (reference: deltaconv/models/deltanet_segmentation.py:9-69)."""
import torch
from torch.nn import Sequential as Seq, Dropout, LeakyReLU

from .. import ops
from ..nn import MLP, Linear
from ..nn.mlp import _slope_of
from .deltanet_base import DeltaNetBase


class DeltaNetSegmentation(torch.nn.Module):
    # False: the head's input [x_max[batch] | categorical[batch] | conv features] is materialised as in the reference (tests compare)
    decompose_head = True

    def __init__(self, in_channels, num_classes, conv_channels=[64, 128, 256], mlp_depth=2, embedding_size=1024,
                 categorical_vector=False, num_neighbors=20, grad_regularizer=0.001, grad_kernel_width=1):
        super().__init__()
        self.categorical_vector = categorical_vector
        self.deltanet_base = DeltaNetBase(in_channels, conv_channels, mlp_depth, num_neighbors, grad_regularizer,
                                          grad_kernel_width)
        self.lin_global = MLP([sum(conv_channels), embedding_size])
        width = embedding_size + sum(conv_channels)
        if categorical_vector:
            self.lin_categorical = MLP([16, 64])
            width += 64
        self.segmentation_head = Seq(
            MLP([width, 256]), Dropout(0.5), MLP([256, 256]), Dropout(0.5),
            Linear(256, 128), LeakyReLU(negative_slope=0.2), Linear(128, num_classes))

    def forward(self, data):
        batch, N = data.batch, data.pos.size(0)
        clouds = getattr(data, '_num_clouds', None)
        first = self.segmentation_head[0][0]                  # Linear -> BatchNorm1d -> LeakyReLU over [x_max | categorical | conv features]
        slope = _slope_of(first[2])
        conv_width = sum(c.out_channels for c in self.deltanet_base.convs)
        if self.decompose_head and len(self.segmentation_head[0]) == 1 and first[0].bias is None and slope is not None:
            # deltanet_segmentation.py:58-66 without the broadcast: x_max[batch] and the category features are per-cloud rows, so
            # their share of the head's first Linear is computed on the B cloud rows and added per cloud (ops.linear_bcast_bn_act);
            # the conv features sit side by side in one buffer and are read in place by lin_global and by the head
            buffer = ops.ColumnBuffer(N, conv_width, data.pos.device)
            conv_out = self.deltanet_base(data, feature_slot=(buffer, 0))
            feats = ops.join_columns(conv_out)
            per_cloud = self.lin_global.forward_pooled(feats, batch, 'max', clouds)
            if self.categorical_vector:
                per_cloud = torch.cat([per_cloud, self.lin_categorical(data.category)], dim=1)
            out = ops.linear_bcast_bn_act(feats, per_cloud, first[0].weight, first[1].bn, slope, batch, clouds)
            for module in list(self.segmentation_head)[1:]:
                out = module(out)
        else:
            # one buffer for the head's input [x_max | categorical | conv features] (deltanet_segmentation.py:66): every producer
            # writes its own block, the backbone's features are read in place for lin_global (:57)
            head_in = first[0].in_features
            buffer = ops.ColumnBuffer(N, head_in, data.pos.device)
            conv_out = self.deltanet_base(data, feature_slot=(buffer, head_in - conv_width))
            # global_max_pool(lin_global(x), batch)[batch]  (deltanet_segmentation.py:58-61); lin_global reads the features in place
            x_max = ops.broadcast_rows(self.lin_global.forward_pooled(ops.join_columns(conv_out), batch, 'max', clouds), batch, N,
                                       slot=(buffer, 0))
            parts = [x_max]
            if self.categorical_vector:
                parts.append(ops.broadcast_rows(self.lin_categorical(data.category), batch, N, slot=(buffer, x_max.size(1))))
            out = self.segmentation_head(ops.join_columns(parts + conv_out))
        perm = getattr(data, '_dc_perm', None)
        if perm is not None:                         # large-cloud mode: rows are in Morton order, hand them back in the caller's
            restored = torch.empty_like(out)
            restored[perm] = out
            out = restored
        return out
