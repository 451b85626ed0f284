"""Operator build on 2^20 points: time per call for several fit/normalise chunk sizes (dc_set_build_chunk) and cloud sizes."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deltaconv_b200 as D
from deltaconv_b200 import ops, _lib

lib = _lib.load()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for P, k in ((1024, 20), (1024, 10), (1024, 30), (1024, 40), (16384, 20), (1 << 20, 20)):
    B = (1 << 20) // P
    data = D.utils.synthetic_batch(B, P, config=4).to('cuda')
    graph = ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False)
    xb, yb = ops.tangent_basis(data.norm)
    for chunk in (0, 65536, 131072, 262144, 1 << 20):
        lib.dc_set_build_chunk(chunk)
        ts = []
        for _ in range(5):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            ops.build_operators(data.pos, data.norm, xb, yb, graph, graph.cloud_ptr, B)
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        t = sorted(ts)[2]
        print(f'P={P} k={k} chunk={chunk}: {t:.4f} ms  {(1 << 20) * (48 + 20 * k) / t / 1e6:.0f} GB/s', flush=True)
    lib.dc_set_build_chunk(0)
    del data, graph
