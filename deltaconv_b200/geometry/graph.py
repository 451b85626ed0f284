"""Graph construction and pooling entry points with the PyTorch-Geometric signatures the
reference imports (`torch_geometric.nn.knn_graph`, `global_max_pool`, `global_mean_pool`:
deltanet_base.py:2, deltanet_classification.py:3, deltanet_segmentation.py:3), backed by the
CUDA library."""
import torch

from .. import ops


def knn_graph(x, k, batch=None, loop=False, flow='source_to_target', num_clouds=None, max_points=None, **_unused):
    """[2, N*k] int64 edge_index.  With flow='target_to_source' row 0 holds the query (centre) and
    row 1 its neighbours in ascending (d2, index) order, self first when loop=True — the layout
    DeltaConv assumes (grad_div_mls.py:24-25).  The ELL form of the same graph rides along as
    `edge_index._dc_graph`, so downstream calls never re-derive it."""
    if flow not in ('source_to_target', 'target_to_source'):
        raise ValueError(f"unknown flow '{flow}'")
    if loop:
        graph = ops.knn(x, k, batch, num_clouds, max_points=max_points)
        ei = graph.edge_index()
    else:
        # (k+1)-graph minus the self edge.  torch_cluster removes self BY INDEX: with coincident points slot 0 is the lowest
        # index among the d2 = 0 ties, not necessarily the query, so the slot to drop is the one holding the row's own index
        # (the last slot if self is absent, which only happens when more than k coincident points precede it)
        wide = ops.knn(x, k + 1, batch, num_clouds, want_edge_index=False, max_points=max_points)
        own = wide.idx == torch.arange(wide.N, device=wide.idx.device, dtype=wide.idx.dtype)[:, None]
        drop = torch.where(own.any(1), own.to(torch.int8).argmax(1), torch.full_like(own[:, 0], k, dtype=torch.long))
        slots = torch.arange(k, device=wide.idx.device)[None, :]
        idx = wide.idx.gather(1, slots + (slots >= drop[:, None]).long()).contiguous()
        graph = ops.EllGraph(idx, wide.cloud_ptr, wide.B, wide.max_cloud)
        ei = graph.edge_index()
    if flow == 'source_to_target':
        # row 0 = neighbours, row 1 = centres.  No ELL graph rides on this tensor: the DeltaConv entry points expect the
        # target_to_source layout and ops.graph_of() refuses (or DeltaConv goes generic on) anything else.
        return ei.flip(0).contiguous()
    return ei


def global_max_pool(x, batch, size=None):
    return ops.pool(x, batch, 'max', size)


def global_mean_pool(x, batch, size=None):
    return ops.pool(x, batch, 'mean', size)
