"""Graph construction and pooling entry points with the PyTorch-Geometric signatures the
reference imports (`torch_geometric.nn.knn_graph`, `global_max_pool`, `global_mean_pool`:
deltanet_base.py:2, deltanet_classification.py:3, deltanet_segmentation.py:3), backed by the
CUDA library."""
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
    else:                                   # drop the self slot of a (k+1)-graph
        wide = ops.knn(x, k + 1, batch, num_clouds, want_edge_index=False, max_points=max_points)
        idx = wide.idx[:, 1:].contiguous()
        graph = ops.EllGraph(idx, wide.cloud_ptr, wide.B, wide.max_cloud)
        ei = graph.edge_index()
    if flow == 'source_to_target':
        flipped = ei.flip(0).contiguous()
        return flipped
    return ei


def global_max_pool(x, batch, size=None):
    return ops.pool(x, batch, 'max', size)


def global_mean_pool(x, batch, size=None):
    return ops.pool(x, batch, 'mean', size)
