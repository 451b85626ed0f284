"""One operator build on 2^20 points (1024 clouds x 1024 points, k from argv) — the target of ncu captures of the build kernels."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops

k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
B = (1 << 20) // P
data = D.utils.synthetic_batch(B, P, config=4).to('cuda')
graph = ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False)
xb, yb = ops.tangent_basis(data.norm)
for _ in range(3):
    G, Dv = ops.build_operators(data.pos, data.norm, xb, yb, graph, graph.cloud_ptr, B)
torch.cuda.synchronize()
print('ok', float(G.abs().max()), float(Dv.abs().max()))
