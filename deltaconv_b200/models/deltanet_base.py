"""DeltaNet backbone (reference: deltaconv/models/deltanet_base.py:8-87)."""
import torch

from .. import ops
from ..geometry.graph import knn_graph
from ..geometry.grad_div_mls import build_grad_div, build_tangent_basis, estimate_basis
from ..nn import DeltaConv


class DeltaNetBase(torch.nn.Module):
    def __init__(self, in_channels, conv_channels, mlp_depth, num_neighbors, grad_regularizer, grad_kernel_width,
                 centralize_first=True):
        super().__init__()
        self.k = num_neighbors
        self.grad_regularizer = grad_regularizer
        self.grad_kernel_width = grad_kernel_width
        conv_channels = [in_channels] + list(conv_channels)
        self.convs = torch.nn.ModuleList()
        for i in range(len(conv_channels) - 1):
            last_layer = i == (len(conv_channels) - 2)
            self.convs.append(DeltaConv(conv_channels[i], conv_channels[i + 1], depth=mlp_depth,
                                        centralized=(centralize_first and i == 0), vector=not last_layer))

    # Large-cloud mode (no counterpart in the reference): when set, batches whose largest cloud has at least this many
    # points are processed in Morton order inside every cloud (ops.spatial_order) — the gathers of the operators then stay in
    # cache (2x on a 2^20-point cloud).  The returned features are in that order; `data._dc_perm` holds the permutation and
    # the segmentation head undoes it.  Off by default: exact distance ties in the kNN resolve by index, so a reordering can
    # change which of two equidistant neighbours is kept.
    reorder_min_points = None

    def forward(self, data, feature_slot=None):
        """Returns the list of per-layer scalar features (deltanet_base.py:87).  They are written side by side into one
        [N, sum(channels)] buffer — `feature_slot` = (ops.ColumnBuffer, first column) when the head provides a wider one — so
        that the heads' `torch.cat(conv_out, dim=1)` is `ops.join_columns(conv_out)`, without a copy."""
        pos, batch = data.pos, data.batch
        # optional hints (number of clouds, size of the largest one) that avoid a device sync
        clouds, maxp = getattr(data, '_num_clouds', None), getattr(data, '_max_points', None)
        perm = None
        if self.reorder_min_points is not None:
            if maxp is None:
                _, clouds, maxp = ops.cloud_layout(batch, pos.size(0), pos.device, clouds)
            if maxp >= self.reorder_min_points:
                perm = ops.spatial_order(pos, batch, clouds)
                pos = pos[perm]
        if perm is not None or getattr(data, '_dc_perm', None) is not None:
            data._dc_perm = perm                     # the batch object is only touched in large-cloud mode

        # operator construction (deltanet_base.py:52-69): kNN graph -> frames -> grad/div
        edge_index = knn_graph(pos, self.k, batch, loop=True, flow='target_to_source', num_clouds=clouds, max_points=maxp)
        if hasattr(data, 'norm') and data.norm is not None:
            normal = data.norm if perm is None else data.norm[perm]
            x_basis, y_basis = build_tangent_basis(normal)
        else:
            # the reference builds a second graph with k = 10 (deltanet_base.py:63).  Neighbours are ordered by (d2, index), so
            # the 10 nearest are exactly the first 10 slots of the k-nearest when k >= 10: no second search
            graph = ops.graph_of(edge_index, pos.size(0))
            if graph.k >= 10:
                edge_index_normal = ops.EllGraph(graph.idx[:, :10].contiguous(), graph.cloud_ptr, graph.B, graph.max_cloud)
            else:
                edge_index_normal = knn_graph(pos, 10, batch, loop=True, flow='target_to_source', num_clouds=clouds, max_points=maxp)
            normal, x_basis, y_basis = estimate_basis(pos, edge_index_normal, orientation=pos)
        grad, div = build_grad_div(pos, normal, x_basis, y_basis, edge_index, batch,
                                   kernel_width=self.grad_kernel_width, regularizer=self.grad_regularizer)

        # convolutions (deltanet_base.py:76-87)
        if hasattr(data, 'x') and data.x is not None:
            x = data.x if perm is None else data.x[perm]
        else:
            x = pos
        v = grad @ x
        if feature_slot is None:
            feature_slot = (ops.ColumnBuffer(pos.size(0), sum(c.out_channels for c in self.convs), pos.device), 0)
        buffer, col = feature_slot
        out = []
        for conv in self.convs:
            x, v = conv(x, v, grad, div, edge_index, x_slot=(buffer, col))
            col += conv.out_channels
            out.append(x)
        return out
