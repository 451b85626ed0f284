"""MLP builders with the reference's names, signatures and module trees (deltaconv/nn/mlp.py),
so state dicts are interchangeable; the Linear, BatchNorm and activation of each block run
through the CUDA library."""
import torch
from torch.nn import LeakyReLU

from .. import ops
from .nonlin import BatchNorm1d, VectorNonLin

__all__ = ['Linear', 'MLP', 'VectorMLP', 'ScalarVectorMLP', 'ScalarVectorIdentity']


class Linear(torch.nn.Linear):
    """torch.nn.Linear (same parameters, init and state-dict keys) whose forward/backward are
    the library's GEMM kernels."""

    def forward(self, x):
        if x.dim() != 2:
            return ops.linear(x.reshape(-1, x.size(-1)), self.weight, self.bias).reshape(*x.shape[:-1], self.out_features)
        return ops.linear(x, self.weight, self.bias)


def _slope_of(nonlin):
    if isinstance(nonlin, LeakyReLU):
        return float(nonlin.negative_slope)
    if isinstance(nonlin, torch.nn.Identity):
        return 1.0
    if isinstance(nonlin, torch.nn.ReLU):
        return 0.0
    return None


class _LinBNAct(torch.nn.Sequential):
    """Seq(Linear, BatchNorm1d, nonlin) — children 0,1,2 as in the reference (mlp.py:9) — with
    BatchNorm and the activation fused into one pass; `addend` is added to the result
    (DeltaConv's `x_max + s_mlp(...)`, deltaconv.py:59)."""

    def forward(self, x, addend=None, slot=None):
        lin, bn, act = self[0], self[1], self[2]
        slope = _slope_of(act)
        inner = bn.bn
        batch_stats = inner.training or not inner.track_running_stats or inner.running_mean is None
        if slope is not None and lin.bias is None and batch_stats and x.dim() == 2:
            # Linear + BatchNorm statistics + activation as one autograd node
            return ops.linear_bn_act(x, lin.weight, inner, slope, addend, slot)
        y = lin(x)
        if slope is None:                      # another activation: `slot` is not honoured, the caller concatenates instead
            out = act(bn(y))
            return out if addend is None else out + addend
        return bn(y, slope, addend, slot)


class _MLP(torch.nn.Sequential):
    def forward(self, x, addend=None, slot=None):
        """`addend` is added to the last block's output; `slot` = (ops.ColumnBuffer, column) lets the last block write its
        output straight into a block of a concatenation buffer."""
        last = len(self) - 1
        for i, block in enumerate(self):
            if isinstance(block, _LinBNAct):
                x = block(x, addend if i == last else None, slot if i == last else None)
            else:
                x = block(x)
        return x

    def forward_pooled(self, x, batch, mode, num_clouds=None):
        """global pool ('max' | 'mean' | 'maxmean') of this MLP's output over the clouds of `batch`; when the last block
        qualifies, its BatchNorm + activation are evaluated inside the pooling pass (ops.linear_bn_act_pool)."""
        blocks = list(self)
        for block in blocks[:-1]:
            x = block(x)
        tail = blocks[-1]
        if isinstance(tail, _LinBNAct) and x.dim() == 2:
            lin, inner, slope = tail[0], tail[1].bn, _slope_of(tail[2])
            if ops.pool_fusable(inner, lin.out_features, lin.bias, slope):
                return ops.linear_bn_act_pool(x, lin.weight, inner, slope, batch, mode, num_clouds)
            if not torch.is_grad_enabled() and ops.pool_fusable_inference(inner, lin.out_features, lin.bias, slope, x):
                return ops.linear_bn_act_pool_inference(x, lin.weight, inner, slope, batch, mode, num_clouds)
        return ops.pool(tail(x), batch, mode, num_clouds)


def MLP(channels, bias=False, nonlin=LeakyReLU(negative_slope=0.2)):
    """Stack of Linear -> BatchNorm1d -> nonlin blocks (mlp.py:7-11)."""
    return _MLP(*[
        _LinBNAct(Linear(channels[i - 1], channels[i], bias=bias), BatchNorm1d(channels[i]), nonlin)
        for i in range(1, len(channels))
    ])


def VectorMLP(channels, batchnorm=True):
    """Stack of Linear(no bias) -> VectorNonLin blocks on [2N,C] vector features (mlp.py:13-17)."""
    return torch.nn.Sequential(*[
        torch.nn.Sequential(Linear(channels[i - 1], channels[i], bias=False),
                            VectorNonLin(channels[i], batchnorm=BatchNorm1d(channels[i]) if batchnorm else None))
        for i in range(1, len(channels))
    ])


class ScalarVectorMLP(torch.nn.Module):
    """Scalar MLP and (optionally) vector MLP applied to an (x, v) tuple (mlp.py:19-39)."""

    def __init__(self, channels, nonlin=True, vector_stream=True):
        super().__init__()
        self.scalar_mlp = MLP(channels, nonlin=LeakyReLU(negative_slope=0.2) if nonlin else torch.nn.Identity())
        self.vector_mlp = VectorMLP(channels) if vector_stream else None

    def forward(self, x):
        assert self.vector_mlp is None or (self.vector_mlp is not None and type(x) is tuple)
        if type(x) is tuple:
            x, v = x
        x = self.scalar_mlp(x)
        if self.vector_mlp is not None:
            v = self.vector_mlp(v)
            x = (x, v)
        return x


class ScalarVectorIdentity(torch.nn.Module):
    def __init__(self, *args, **kwargs):
        super().__init__()

    def forward(self, input):
        return input
