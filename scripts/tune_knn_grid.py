import os, sys
sys.path.insert(0, '/root/repo')
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops, _lib
lib = _lib.load()
def t_ms(fn, n=3):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
ops.KNN_METHOD = 'grid'
for lg in (12, 14, 20):
    P = 1 << lg; B = (1 << 20) // P
    data = D.utils.synthetic_batch(B, P, config=4).to('cuda')
    for k in (10, 20, 40):
        out = []
        for dens in (0.5, 1, 2, 4, 8, 16):
            lib.dc_set_knn_grid_density(dens)
            out.append(f'{dens}: {t_ms(lambda: ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False, max_points=P)):.2f}')
        print(f'P={P} k={k}  ' + '  '.join(out), flush=True)
