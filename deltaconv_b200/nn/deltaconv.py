"""DeltaConv layer with the reference's constructor, forward signature, repr and parameter tree
(deltaconv/nn/deltaconv.py:8-72), scheduled onto the fused CUDA operators:

  scalar stream   [x | div v | curl v | |v|]  one SpMM pass over v        (dc_divcurlnorm_fwd)
                  x' = max_nbr(s_mlp_max(.)) + s_mlp(cat)                  (dc_gathermax_fwd, GEMM, fused BN+act+add)
  vector stream   [v | hodge(v) | grad x']     from the same div/curl       (dc_hodge_fwd, dc_grad_fwd)
                  v' = v_mlp(I_J(cat))         I_J folded into the weights  (dc_expand_vec_weight, GEMM, dc_vecnonlin)
"""
import torch

from .. import ops
from .mlp import MLP, VectorMLP, _slope_of

__all__ = ['DeltaConv']


class DeltaConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, depth=1, centralized=False, vector=True, aggr='max'):
        super().__init__()
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.centralized = centralized
        self.aggr = aggr
        self.s_mlp_max = MLP([in_channels] + [out_channels] * depth)
        self.s_mlp = MLP([in_channels * 4] + [out_channels] * depth)
        self.v_mlp = VectorMLP([in_channels * 4 + out_channels * 2] + [out_channels] * depth) if vector else None

    def forward(self, x, v, grad, div, edge_index, x_slot=None):
        """The reference's signature (deltaconv.py:44).  `x_slot` = (ops.ColumnBuffer, column), an extension used by
        DeltaNetBase: the layer's scalar output is written into that block of the network's feature concatenation."""
        ell = isinstance(grad, ops.EllOperator) and isinstance(div, ops.EllOperator)
        graph = None
        if ell and self.aggr == 'max':
            graph = ops.graph_of(edge_index, x.size(0), strict=False) if edge_index is not None else grad.graph
        if graph is None:
            # anything the fused schedule does not cover — another aggregation (deltaconv.py:35,52,54 pass `aggr` to
            # torch_scatter), an edge_index that is not a fixed-k graph, operators that only support `@` — takes the
            # reference's own formulation step by step (correct, not fast; no shipped model comes here)
            return self._forward_generic(x, v, grad, div, edge_index if edge_index is not None else grad.graph.edge_index())

        # scalar stream (deltaconv.py:50-59)
        if self.centralized:
            block = self.s_mlp_max[0]
            slope = _slope_of(block[2])
            if (len(self.s_mlp_max) == 1 and slope is not None and block[0].bias is None
                    and ops.centralized_max_supported(x.size(1), graph.k)):
                # edge MLP + BatchNorm over the E edge rows + max, without any per-edge tensor
                x_max = ops.centralized_max(x, block[0].weight, block[1].bn, graph, slope)
            else:
                x_max = self._centralized_deep(x, graph)
        else:
            blocks = list(self.s_mlp_max)
            tail = blocks[-1]
            if ops.gathermax_fusable(tail[0], _slope_of(tail[2]), x):
                h = x
                for block in blocks[:-1]:
                    h = block(h)
                # Linear + BatchNorm statistics + (BatchNorm, activation, max over the neighbours) as one node
                x_max = ops.linear_bn_act_gathermax(h, tail[0].weight, tail[1].bn, _slope_of(tail[2]), graph)
            else:
                x_max = ops.gather_max(self.s_mlp_max(x), graph)
        x_cat, div_curl = ops.scalar_cat(x, v, div)
        x = self.s_mlp(x_cat, x_max, x_slot)

        # vector stream (deltaconv.py:64-68)
        if self.v_mlp is not None:
            v_cat = ops.vector_cat(v, div_curl, x, grad, pad=True)
            first = self.v_mlp[0]
            v = first[1](ops.vector_linear_with_rotation(v_cat, first[0].weight))
            for block in list(self.v_mlp)[1:]:
                v = block(v)
        return x, v

    def _centralized_deep(self, x, graph):
        """scatter_max(s_mlp_max(x_j - x_i)) for an edge MLP of depth >= 2 (deltaconv.py:50-52; train_shapenet.py builds depth 2).
        The [N*k, C] tensors between the blocks have to exist, the ones around them do not: the first block's activation is
        written straight from the coordinates with statistics from the offsets' moments (ops.central_edge_act), the last block's
        BatchNorm + activation are folded into the max (ops.linear_bn_act_segmax).  Anything those nodes do not cover (a
        Linear bias, another activation, wide inputs, a gradient w.r.t. x) takes the plain sequence."""
        blocks = list(self.s_mlp_max)
        head, tail = blocks[0], blocks[-1]
        head_slope, tail_slope = _slope_of(head[2]), _slope_of(tail[2])
        fuse_head = (len(blocks) >= 2 and head_slope is not None and head[0].bias is None and not x.requires_grad
                     and ops.centralized_max_supported(x.size(1), graph.k))
        fuse_tail = len(blocks) >= 2 and ops.segmax_fusable(tail[0], tail_slope)
        if fuse_head:
            h = ops.central_edge_act(x, head[0].weight, head[1].bn, graph, head_slope)
        else:
            h = head(ops.edge_diff(x, graph))
        for block in blocks[1:-1] if fuse_tail else blocks[1:]:
            h = block(h)
        if fuse_tail:
            return ops.linear_bn_act_segmax(h, tail[0].weight, tail[1].bn, tail_slope, graph)
        return ops.segment_max(h, graph)

    def _forward_generic(self, x, v, grad, div, edge_index):
        """deltaconv.py:44-70 literally: gather, MLP, scatter-reduce; operators through `@`.  The Linear / BatchNorm /
        VectorNonLin modules still run in the CUDA library; gathers, scatter-reduce and concatenations are torch device ops."""
        from ..geometry.operators import I_J, curl, hodge_laplacian, norm
        if self.aggr not in ('max', 'min', 'sum', 'add', 'mean'):
            raise ValueError(f"DeltaConv: unknown aggr '{self.aggr}'")
        N = x.size(0)
        row, col = edge_index[0], edge_index[1]
        keep = (col >= 0) & (col < N)                       # the padding sentinel of clouds with fewer than k points
        row, col = row[keep], col[keep]
        if self.centralized:
            feat = self.s_mlp_max(x[col] - x[row])
        else:
            feat = self.s_mlp_max(x)[col]
        x_max = _scatter_rows(feat, row, N, self.aggr)
        x = x_max + self.s_mlp(torch.cat([x, div @ v, curl(v, div), norm(v)], dim=1))
        if self.v_mlp is not None:
            v = self.v_mlp(I_J(torch.cat([v, hodge_laplacian(v, grad, div), grad @ x], dim=1)))
        return x, v

    def __repr__(self):
        return f'{self.__class__.__name__}({self.in_channels}, {self.out_channels})'


def _scatter_rows(src, index, num_rows, aggr):
    """torch_scatter.scatter(src, index, dim=0, reduce=aggr) with its conventions: empty groups give 0, mean divides by the
    group size."""
    out = src.new_zeros(num_rows, src.size(1))
    expand = index[:, None].expand_as(src)
    if aggr in ('sum', 'add'):
        return out.index_add(0, index, src)
    if aggr == 'mean':
        count = torch.bincount(index, minlength=num_rows).clamp(min=1).to(src.dtype)
        return out.index_add(0, index, src) / count[:, None]
    return out.scatter_reduce(0, expand, src, reduce='amax' if aggr == 'max' else 'amin', include_self=False)
