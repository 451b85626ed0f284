"""Operator algebra on interleaved [2N,C] vector fields (reference: deltaconv/geometry/operators.py).
`norm`, `J` and `I_J` are tiny element-wise torch expressions kept for API parity — inside
DeltaConv they are never materialised (folded into the SpMM kernels and the weight layout).
`curl`, `laplacian`, `hodge_laplacian` run on the CUDA operators when given EllOperators."""
import torch

from .. import ops

__all__ = ['norm', 'J', 'I_J', 'curl', 'laplacian', 'hodge_laplacian']


def norm(v):
    """Per-point norm of a vector field: [2N,C] -> [N,C] (operators.py:4-7)."""
    C = v.size(1)
    return torch.linalg.norm(v.view(-1, 2, C), dim=1)


def J(v):
    """90-degree counter-clockwise rotation (operators.py:9-17)."""
    rows, C = v.shape
    vv = v.view(-1, 2, C)
    return torch.stack([-vv[:, 1], vv[:, 0]], dim=1).view(rows, C)


def I_J(v):
    """[v | J v] (operators.py:19-21)."""
    return torch.cat([v, J(v)], dim=1)


def curl(v, div):
    """curl = -div J v (operators.py:23-27)."""
    if isinstance(div, ops.EllOperator):
        return ops.curl_of(v, div)
    return -(div @ J(v))


def laplacian(x, grad, div):
    """-div grad x (operators.py:29-33)."""
    return -(div @ (grad @ x))


def hodge_laplacian(v, grad, div):
    """-(grad div + J grad curl) v (operators.py:35-46)."""
    if isinstance(grad, ops.EllOperator) and isinstance(div, ops.EllOperator):
        _, dc = ops.scalar_cat(v.new_zeros(v.size(0) // 2, v.size(1)), v, div)
        C = v.size(1)
        cat = ops.vector_cat(v, dc, v.new_zeros(v.size(0) // 2, 1), grad)
        return cat[:, C:2 * C]
    return -((grad @ (div @ v)) + J(grad @ curl(v, div)))
