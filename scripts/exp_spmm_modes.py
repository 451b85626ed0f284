"""SpMM scheduling modes (dc_set_spmm_staged 0 / 1 / 2) on the training batch shape: 32 clouds x 1024 points, k = 20.
Per-op device time, L2 flushed; results must be bitwise identical between modes."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops, _lib

lib = _lib.load()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
B, P, k = (int(a) for a in (sys.argv[1:4] if len(sys.argv) > 3 else (32, 1024, 20)))
data = D.utils.synthetic_batch(B, P, config=2).to('cuda')
graph = ops.knn(data.pos, k, data.batch, num_clouds=B, max_points=P, want_edge_index=False)
xb, yb = ops.tangent_basis(data.norm)
G, Dv = ops.build_operators(data.pos, data.norm, xb, yb, graph, graph.cloud_ptr, B)
grad, div = ops.EllOperator('grad', graph, G, Dv), ops.EllOperator('div', graph, G, Dv)
graph.transpose()
N = B * P


def timed(fn):
    ts = []
    for _ in range(7):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); out = fn(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[3], out


for C in (64, 128, 256):
    x = torch.randn(N, C, device='cuda')
    v = torch.randn(2 * N, C, device='cuda')
    gy = torch.randn(2 * N, C, device='cuda')
    ref = {}
    for mode in (0, 2, 1):
        lib.dc_set_spmm_staged(mode)
        with torch.no_grad():
            t_g, o_g = timed(lambda: grad @ x)
            t_d, o_d = timed(lambda: ops.scalar_cat(x, v, div)[0])
        xr = x.clone().requires_grad_(True)
        yv = grad @ xr
        t_gt, o_gt = timed(lambda: torch.autograd.grad(yv, xr, gy, retain_graph=True)[0])
        outs = (o_g, o_d, o_gt)
        if mode == 0:
            ref = outs
        same = all(torch.equal(a, b) for a, b in zip(outs, ref))
        print(f'C={C} mode={mode}: grad {t_g * 1e3:.1f} us  divcurlnorm+copy {t_d * 1e3:.1f} us  grad^T {t_gt * 1e3:.1f} us  bitwise={same}', flush=True)
lib.dc_set_spmm_staged(0)
