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

    def forward(self, x, v, grad, div, edge_index):
        if not (isinstance(grad, ops.EllOperator) and isinstance(div, ops.EllOperator)):
            raise TypeError("DeltaConv expects the grad/div operators returned by deltaconv_b200's build_grad_div")
        if self.aggr != 'max':
            raise NotImplementedError("DeltaConv: the CUDA path implements aggr='max' (every shipped model's setting)")
        graph = ops.graph_of(edge_index, x.size(0)) if edge_index is not None else grad.graph

        # scalar stream (deltaconv.py:50-59)
        if self.centralized:
            block = self.s_mlp_max[0]
            slope = _slope_of(block[2])
            if (len(self.s_mlp_max) == 1 and slope is not None and block[0].bias is None
                    and ops.centralized_max_supported(x.size(1), graph.k)):
                # edge MLP + BatchNorm over the E edge rows + max, without any per-edge tensor
                x_max = ops.centralized_max(x, block[0].weight, block[1].bn, graph, slope)
            else:
                x_max = ops.segment_max(self.s_mlp_max(ops.edge_diff(x, graph)), graph)
        else:
            x_max = ops.gather_max(self.s_mlp_max(x), graph)
        x_cat, div_curl = ops.scalar_cat(x, v, div)
        x = self.s_mlp(x_cat, x_max)

        # vector stream (deltaconv.py:64-68)
        if self.v_mlp is not None:
            v_cat = ops.vector_cat(v, div_curl, x, grad)
            first = self.v_mlp[0]
            v = first[1](ops.vector_linear_with_rotation(v_cat, first[0].weight))
            for block in list(self.v_mlp)[1:]:
                v = block(v)
        return x, v

    def __repr__(self):
        return f'{self.__class__.__name__}({self.in_channels}, {self.out_channels})'
