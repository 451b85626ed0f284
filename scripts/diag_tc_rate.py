"""Tensor rate of the two 3xTF32 kernels: time of the 32768 x 1024 x 512 product with 0 / 2 / 4 extra copies of every k-block's
12 MMAs (dc_set_gemm_debug); the slope is the cost of one tcgen05.mma in the running pipeline."""
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
def t_ms(n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for pair in (0, 1):
    lib.dc_set_gemm_pair(pair)
    res = []
    for rep in (0, 2, 4):
        lib.dc_set_gemm_debug(rep, int(os.environ.get('DC_WIDE', '0')))
        res.append(t_ms())
    lib.dc_set_gemm_debug(0, 0)
    # k-blocks executed per SM: tiles per SM (ceil) x 16
    per_sm_kb = -(-(256 * 8) // 148) * 16
    slope_us = (res[2] - res[0]) * 1000 / (per_sm_kb * 48)
    print(f'pair={pair}: {res[0]*1000:.1f} / {res[1]*1000:.1f} / {res[2]*1000:.1f} us for 0/2/4 extra copies -> {slope_us*1000:.1f} ns = {slope_us*1.965e3:.0f} cycles per extra MMA; '
          f'base per k-block {res[0]*1000/per_sm_kb*1.965e3:.0f} cycles (12 MMAs)')
lib.dc_set_gemm_pair(0)
