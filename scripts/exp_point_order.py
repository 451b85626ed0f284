import sys; sys.path.insert(0, '/root/repo')
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops
dev='cuda'
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def time_ms(fn, reps=5):
    out=[]
    for _ in range(reps):
        flush.zero_()
        s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); out.append(s.elapsed_time(e))
    return sorted(out)[len(out)//2]
P=1<<20; k=20
data = D.utils.synthetic_batch(1, P, config=4).to(dev)
for mode in ('as generated', 'cell order G=64', 'cell order G=256', 'morton 10 bits'):
    pos, nrm = data.pos, data.norm
    if mode != 'as generated':
        lo = pos.min(0).values; ext = (pos.max(0).values - lo).max()
        if mode.startswith('cell'):
            G = int(mode.split('=')[1])
            c = ((pos - lo) / ext * (G - 1e-3)).long().clamp(0, G - 1)
            key = (c[:, 2] * G + c[:, 1]) * G + c[:, 0]
        else:
            c = ((pos - lo) / ext * 1023.999).long().clamp(0, 1023)
            def spread(v):
                v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F; v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249; return v
            key = spread(c[:, 0]) | (spread(c[:, 1]) << 1) | (spread(c[:, 2]) << 2)
        perm = torch.argsort(key, stable=True)
        pos, nrm = pos[perm].contiguous(), nrm[perm].contiguous()
    graph = ops.knn(pos, k, data.batch, num_clouds=1, want_edge_index=False, max_points=P)
    xb, yb = ops.tangent_basis(nrm)
    G_, Dv = ops.build_operators(pos, nrm, xb, yb, graph, graph.cloud_ptr, 1)
    grad, div = ops.EllOperator('grad', graph, G_, Dv), ops.EllOperator('div', graph, G_, Dv)
    graph.transpose()
    res=[]
    for C in (16, 64, 128):
        x = torch.randn(P, C, device=dev); v = torch.randn(2*P, C, device=dev)
        with torch.no_grad():
            tg = time_ms(lambda: grad @ x); td = time_ms(lambda: ops.scalar_cat(x, v, div))
        res.append(f'C{C}: grad {P*(12*k+12*C)/tg/1e6:.0f} GB/s divcurl {P*(12*k+28*C)/td/1e6:.0f} GB/s')
    tb = time_ms(lambda: ops.build_operators(pos, nrm, xb, yb, graph, graph.cloud_ptr, 1))
    print(f'{mode}: build {tb:.2f} ms | ' + ' | '.join(res), flush=True)
