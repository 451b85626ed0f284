"""Brute-force against grid kNN (both exact) over cloud sizes: where is the crossover?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops
dev = 'cuda'
def t_ms(fn, n=3):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for lg in (10, 11, 12, 13, 14, 17, 20):
    P = 1 << lg; B = max(1, (1 << 20) // P)
    data = D.utils.synthetic_batch(B, P, config=4).to(dev)
    res = {}
    for method in ('brute', 'grid'):
        if method == 'brute' and lg > 17: continue
        ops.KNN_METHOD = method
        res[method] = t_ms(lambda: ops.knn(data.pos, 20, data.batch, num_clouds=B, want_edge_index=False, max_points=P), 1 if lg >= 17 else 3)
    ops.KNN_METHOD = 'auto'
    print(f'P={P} B={B} k=20: ' + '  '.join(f'{m} {v:.3f} ms' for m, v in res.items()), flush=True)
