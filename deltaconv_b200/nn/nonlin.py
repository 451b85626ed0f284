"""BatchNorm1d wrapper and VectorNonLin with the reference's constructor signatures, reprs and
state-dict keys (deltaconv/nn/nonlin.py), computing through the CUDA library."""
import torch
from torch import Tensor

from .. import ops

__all__ = ['BatchNorm1d', 'VectorNonLin']


class BatchNorm1d(torch.nn.Module):
    """Batch normalisation over the rows of an [N,C] tensor (nonlin.py:11-35).  Parameters and
    running statistics live in `self.bn` (a torch.nn.BatchNorm1d used as a container, so the
    state-dict keys keep the `.bn.` prefix); the statistics and the normalisation are computed
    by `dc_bn_stats` / `dc_bn_act_fwd`."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.bn = torch.nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)
        self.reset_parameters()

    def reset_parameters(self):
        self.bn.reset_parameters()

    def forward(self, x: Tensor, slope: float = 1.0, addend: Tensor = None, slot=None) -> Tensor:
        return ops.bn_act(x, self.bn, slope, addend, slot)

    def __repr__(self):
        return f'{self.__class__.__name__}({self.bn.num_features})'


class VectorNonLin(torch.nn.Module):
    """Non-linearity on the per-point norm of vector features (nonlin.py:38-86):
    v <- v * nonlin(shift(|v|)) / max(|v|, 1e-8), shift = batchnorm or `+ bias`."""

    def __init__(self, in_channels, nonlin=torch.nn.ReLU(), batchnorm=None):
        super().__init__()
        self.bias = torch.nn.Parameter(torch.empty(in_channels))      # zero-filled by reset_parameters (nonlin.py:50,58)
        self.nonlin = nonlin
        self.batchnorm = batchnorm
        self.reset_parameters()

    def reset_parameters(self):
        self.bias.data.fill_(0)
        if self.batchnorm is not None:
            self.batchnorm.reset_parameters()

    def forward(self, x: Tensor) -> Tensor:
        if not isinstance(self.nonlin, torch.nn.ReLU):
            # any other activation (nonlin.py:48,78): the formula step by step — BatchNorm through the library, the rest as
            # torch device ops.  The fused kernel covers ReLU, the reference's own and only setting.
            rows, C = x.shape
            xv = x.view(-1, 2, C)
            n = torch.linalg.norm(xv, dim=1)
            shifted = self.batchnorm(n) if self.batchnorm is not None else n + self.bias[None, :]
            gain = self.nonlin(shifted) / n.clamp(min=1e-8)
            return (xv * gain[:, None, :]).reshape(rows, C)
        bn = self.batchnorm.bn if self.batchnorm is not None else None
        return ops.vec_nonlin(x, self.bias, bn)

    def __repr__(self):
        return f'{self.__class__.__name__}(batchnorm={self.batchnorm.__repr__()})'
