"""Grid kNN over its density knob (points per bounding-box cell) for the model batch shapes and for large clouds; device time
by CUDA events over back-to-back calls after a warm-up (the host side is faster than these kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops, _lib
lib = _lib.load()
def t_us(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3
warm = D.utils.synthetic_batch(32, 1024, config=5).to('cuda')
for _ in range(20): ops.knn(warm.pos, 20, warm.batch, num_clouds=32, max_points=1024)
for B, P in ((32, 1024), (32, 2048), (16, 2048), (32, 4096), (64, 16384), (1, 1 << 20)):
    data = D.utils.synthetic_batch(B, P, config=5).to('cuda')
    for k in (10, 20, 30, 40):
        f = lambda: ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False, max_points=P)
        out = []
        if P <= 4096:
            ops.KNN_METHOD = 'brute'
            out.append(f'brute {t_us(f):.0f}')
        ops.KNN_METHOD = 'grid'
        for dens in (1.5, 2, 3, 4, 5, 6, 8):
            lib.dc_set_knn_grid_density(dens)
            out.append(f'@{dens} {t_us(f, 10 if P < 100000 else 3):.0f}')
        lib.dc_set_knn_grid_density(0)
        ops.KNN_METHOD = 'auto'
        print(f'B={B} P={P} k={k} (us):  ' + '  '.join(out), flush=True)
