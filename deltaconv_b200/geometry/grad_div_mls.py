"""Operator construction entry points with the reference's signatures
(deltaconv/geometry/grad_div_mls.py), backed by the CUDA library.

`estimate_basis`, `build_tangent_basis` and `build_grad_div` are the public entry points the
models use.  The reference's intermediate helpers (`coords_projected`, `gaussian_weights`,
`fit_vector_mapping`) are exposed from the same fused kernel through its debug outputs.
"""
import torch

from .. import ops
# the reference defines its four intermediate helpers in this module (grad_div_mls.py:72-194) and its tests reach them
# through `from deltaconv.geometry.grad_div_mls import *`
from .helpers import coords_projected, gaussian_weights, weighted_least_squares, fit_vector_mapping   # noqa: F401

DEFER_NORMALISATION_FROM = 1 << 22          # edges (N * k) from which build_grad_div defers the per-cloud normalisation

__all__ = ['estimate_basis', 'build_tangent_basis', 'build_grad_div', 'operator_parts',
           'coords_projected', 'gaussian_weights', 'weighted_least_squares', 'fit_vector_mapping']


def estimate_basis(pos, edge_index, k=None, orientation=None):
    """(normal, x_basis, y_basis) from the local 3 x k neighbourhood (grad_div_mls.py:10-47).
    `k` is accepted for signature parity; the graph itself carries it."""
    graph = ops.graph_of(edge_index, pos.size(0))
    if k is not None and int(k) != graph.k:
        raise ValueError(f"k={int(k)} does not match the graph's k={graph.k}")
    return ops.estimate_frames(pos, graph, orientation)


def build_tangent_basis(normal):
    """(x_basis, y_basis) orthonormal to `normal` (grad_div_mls.py:50-69)."""
    return ops.tangent_basis(normal)


def _prepare(pos, edge_index, batch):
    graph = ops.graph_of(edge_index, pos.size(0))
    if graph.cloud_ptr is not None and (batch is None or getattr(batch, '_dc_cloud', (None,))[0] is graph.cloud_ptr):
        return graph, graph.cloud_ptr, graph.B
    cp, B, maxp = ops.cloud_layout(batch, pos.size(0), pos.device)
    if graph.cloud_ptr is None:          # a foreign edge_index: remember the cloud layout for the SpMM kernels
        graph.cloud_ptr, graph.B, graph.max_cloud = cp, B, maxp
    return graph, cp, B


def build_grad_div(pos, normal, x_basis, y_basis, edge_index, batch=None, kernel_width=1, regularizer=0.001,
                   normalized=True, shape_regularizer=None):
    """(grad, div): the 2N x N gradient and N x 2N divergence of the weighted-least-squares fit
    (grad_div_mls.py:197-277) as `EllOperator`s sharing one fixed-k ELL storage."""
    graph, cp, B = _prepare(pos, edge_index, batch)
    with torch.no_grad():
        # Per-cloud normalisation: deferred to the operator kernels for large batches (it saves a read-modify-write pass over
        # 16 k bytes per point of entries: 0.11 of 0.56 ms at 2^20 points); for the model batches the entries are still in L2
        # when the scale pass runs (5 us) and the un-scaled operator kernels keep their leaner register allocation.
        deferred = graph.N * graph.k >= DEFER_NORMALISATION_FROM
        G, D, scale = ops.build_operators(pos, normal, x_basis, y_basis, graph, cp, B, kernel_width, regularizer,
                                          normalized, shape_regularizer, deferred=deferred)
    return ops.EllOperator('grad', graph, G, D, scale), ops.EllOperator('div', graph, G, D, scale)


def operator_parts(pos, normal, x_basis, y_basis, edge_index, batch=None, kernel_width=1, regularizer=0.001,
                   normalized=True, shape_regularizer=None):
    """dict(coords [E,2], weights [E], mapping [E,2,2]): what the reference's `coords_projected`
    (:72-97), `gaussian_weights` (:100-116) and `fit_vector_mapping` (:155-194) return for the same
    inputs, read out of the fused build kernel."""
    graph, cp, B = _prepare(pos, edge_index, batch)
    with torch.no_grad():
        _, _, (coords, weights, mapping) = ops.build_operators(pos, normal, x_basis, y_basis, graph, cp, B, kernel_width,
                                                               regularizer, normalized, shape_regularizer, debug=True)
    E = graph.N * graph.k
    return dict(coords=coords.view(E, 2), weights=weights.view(E), mapping=mapping.view(E, 2, 2))
