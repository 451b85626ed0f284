"""kNN (brute force, 1024 clouds x 1024 points) for k = 30 / 40 / 48 / 64 over merge thresholds of the two-slot list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops, _lib
lib = _lib.load()
data = D.utils.synthetic_batch(1024, 1024, config=4).to('cuda')
def t_ms(k):
    f = lambda: ops.knn(data.pos, k, data.batch, num_clouds=1024, want_edge_index=False, max_points=1024)
    f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); f(); f(); f(); e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / 3
ref = {k: ops.knn(data.pos, k, data.batch, num_clouds=1024, want_edge_index=False, max_points=1024).idx.clone() for k in (40, 64)}
print('k=30 (one slot):', ' '.join(f'min{m}={(lib.dc_set_knn_merge_min(m, 0), t_ms(30))[1]:.3f}' for m in (8, 12, 16)))
lib.dc_set_knn_merge_min(12, 0)
for k in (40, 48, 64):
    out = []
    for m in (2, 3, 4, 6, 8, 12, 33):
        lib.dc_set_knn_merge_min(0, m)
        out.append(f'min{m}={t_ms(k):.3f}')
        if k in ref:
            assert torch.equal(ops.knn(data.pos, k, data.batch, num_clouds=1024, want_edge_index=False, max_points=1024).idx, ref[k])
    print(f'k={k}:', ' '.join(out), flush=True)
