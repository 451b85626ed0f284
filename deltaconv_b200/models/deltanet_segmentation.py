"""DeltaNet part-segmentation network (reference: deltaconv/models/deltanet_segmentation.py:9-69)."""
import torch
from torch.nn import Sequential as Seq, Dropout, LeakyReLU

from .. import ops
from ..nn import MLP, Linear
from .deltanet_base import DeltaNetBase


class DeltaNetSegmentation(torch.nn.Module):
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
        conv_out = self.deltanet_base(data)
        batch, N = data.batch, data.pos.size(0)
        clouds = getattr(data, '_num_clouds', None)
        # global_max_pool(lin_global(x), batch)[batch]  (deltanet_segmentation.py:58-61)
        x_max = ops.broadcast_rows(self.lin_global.forward_pooled(torch.cat(conv_out, dim=1), batch, 'max', clouds), batch, N)
        parts = [x_max]
        if self.categorical_vector:
            parts.append(ops.broadcast_rows(self.lin_categorical(data.category), batch, N))
        out = self.segmentation_head(torch.cat(parts + conv_out, dim=1))
        perm = getattr(data, '_dc_perm', None)
        if perm is not None:                         # large-cloud mode: rows are in Morton order, hand them back in the caller's
            restored = torch.empty_like(out)
            restored[perm] = out
            out = restored
        return out
