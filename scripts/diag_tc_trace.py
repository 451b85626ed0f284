"""Timeline of the pair kernel's pipeline in CTA 0 (clock64 stamps per k-block): where does a k-block wait?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deltaconv_b200 import _lib
lib = _lib.load()
dev = 'cuda'
ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
M, N, K = 32768, 1024, 512
a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev); C = torch.empty(M, N, device=dev)
f = lambda: _lib.call('dc_gemm_tc_raw', a.data_ptr(), K, 0, b.data_ptr(), K, 0, None, M, N, K, C.data_ptr(), N, ws.data_ptr(), ws.numel())
lib.dc_set_gemm_debug(int(os.environ.get('DC_REP', '0')), 0)
lib.dc_set_gemm_pair(1)
for _ in range(3): f()
trace = torch.zeros(2, 8, 256, dtype=torch.int64, device=dev)
lib.dc_set_gemm_trace(trace.data_ptr())
f(); torch.cuda.synchronize()
lib.dc_set_gemm_trace(None)
lib.dc_set_gemm_debug(0, 0)
tt = trace.cpu()
sync0, sync1 = int(tt[0, 7, 255]), int(tt[1, 7, 255])
tt[:, 7, 255] = 0
print('clock at the cluster barrier: CTA0', sync0, 'CTA1', sync1)
peer = tt[1].clone()
peer[peer > 0] += sync0 - sync1                      # peer stamps on CTA 0's clock (up to the skew of leaving the barrier)
t = tt[0]
t0 = int(t[t > 0].min())
names = ['MMA:top', 'MMA:waited', 'MMA:fenced', 'MMA:issued', 'cv:a_full', 'cv:ta_empty', 'cv:b_full', 'cv:arrive']
print('k-block ' + ' '.join(f'{n:>11s}' for n in names) + '   (clock cycles since the first stamp; converter columns: warp 2/6 of the owning group)')
for it in range(20, 30):
    print(f'{it:7d} ' + ' '.join(f'{int(t[e, it]) - t0:11d}' if t[e, it] > 0 else f'{"-":>11s}' for e in range(8)))
d = (t[3, 33:129] - t[3, 32:128]).float()
print('MMA issue interval per k-block: mean', d.mean().item(), 'cycles')
print('peer CTA (shifted to CTA 0 clock):')
print('k-block ' + ' '.join(f'{n:>11s}' for n in names))
for it in range(20, 30):
    print(f'{it:7d} ' + ' '.join(f'{int(peer[e, it]) - t0:11d}' if peer[e, it] > 0 else f'{"-":>11s}' for e in range(8)))
