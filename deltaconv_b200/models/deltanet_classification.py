"""DeltaNet classifier (reference: deltaconv/models/deltanet_classification.py:9-51)."""
import torch
from torch.nn import Sequential as Seq, Dropout

from .. import ops
from ..nn import MLP, Linear
from .deltanet_base import DeltaNetBase


class DeltaNetClassification(torch.nn.Module):
    def __init__(self, in_channels, num_classes, conv_channels=[64, 64, 128, 256], num_neighbors=20,
                 grad_regularizer=1e-3, grad_kernel_width=1):
        super().__init__()
        self.deltanet_base = DeltaNetBase(in_channels, conv_channels, 1, num_neighbors, grad_regularizer,
                                          grad_kernel_width)
        self.lin_embedding = MLP([sum(conv_channels), 1024])
        self.classification_head = Seq(
            MLP([1024 * 2, 512]), Dropout(0.5), MLP([512, 256]), Dropout(0.5),
            Linear(256, num_classes))

    def forward(self, data):
        conv_out = self.deltanet_base(data)
        # lin_embedding, then [global max | global mean] in one segmented pass that also applies the layer's BatchNorm and
        # activation (deltanet_classification.py:45-49)
        x = self.lin_embedding.forward_pooled(ops.join_columns(conv_out), data.batch, 'maxmean', getattr(data, '_num_clouds', None))
        return self.classification_head(x)
