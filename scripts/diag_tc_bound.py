"""What bounds the tcgen05 GEMM: plain TF32 (1 MMA per k-step, no converter warps, operands from shared memory) against the
default 3xTF32 (3 MMAs, converters, A from tensor memory) on the layer shapes of configs[1]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deltaconv_b200 import _lib
lib = _lib.load()
dev = 'cuda'
ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]

for (M, N, K, a_mn, b_mn) in [(32768, 1024, 512, 0, 0), (32768, 512, 1024, 0, 1), (1024, 512, 32768, 1, 1), (32768, 256, 512, 0, 0),
                              (32768, 128, 384, 0, 0), (32768, 128, 256, 0, 0), (32768, 64, 256, 0, 0), (32768, 256, 128, 0, 1)]:
    a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev)
    A = a.t().contiguous() if a_mn else a
    B = b.t().contiguous() if b_mn else b
    C = torch.empty(M, N, device=dev)
    out = []
    for mode in (1, 3):
        lib.dc_set_gemm_mode(mode)
        f = lambda: _lib.call('dc_gemm_tc_raw', A.data_ptr(), A.stride(0), a_mn, B.data_ptr(), B.stride(0), b_mn, None, M, N, K, C.data_ptr(), N, ws.data_ptr(), ws.numel())
        t = timeit(f)
        out.append(f'mode{mode} {t * 1000:7.1f} us {2 * M * N * K / t / 1e9:6.0f} TF/s')
    lib.dc_set_gemm_mode(3)
    hbm = 4 * (M * K + N * K + M * N) / 6.5e12 * 1e6
    print(f'M{M} N{N} K{K} a_mn{a_mn} b_mn{b_mn}: ' + ' | '.join(out) + f' | HBM floor {hbm:5.1f} us', flush=True)
