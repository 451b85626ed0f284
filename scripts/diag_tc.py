"""GPU experiment matrix for the tcgen05 GEMM: layouts x tile widths x precision modes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deltaconv_b200 import _lib
lib = _lib.load()
torch.manual_seed(0)
dev = 'cuda'
ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
def run(M, N, K, a_mn, b_mn, mode):
    a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev)
    A = a.t().contiguous() if a_mn else a          # stored [K,M] when a_mn
    B = b.t().contiguous() if b_mn else b
    C = torch.full((M, N), float('nan'), device=dev)
    lib.dc_set_gemm_mode(mode)
    try:
        _lib.call('dc_gemm_tc_raw', A.data_ptr(), A.stride(0), a_mn, B.data_ptr(), B.stride(0), b_mn, None, M, N, K, C.data_ptr(), N, ws.data_ptr(), ws.numel())
        torch.cuda.synchronize()
    except Exception as e:
        return 'EXC ' + str(e)[:60]
    finally:
        lib.dc_set_gemm_mode(3)
    ref = a.double() @ b.double().t()
    err = ((C.double() - ref).abs().max() / ref.abs().max()).item()
    nan = int(C.isnan().sum()); zero = int((C == 0).sum())
    return f'err {err:.2e} nan {nan} zeros {zero}'
for (M, N, K) in [(256, 64, 64), (256, 128, 64), (256, 256, 64), (256, 64, 256), (1024, 256, 512)]:
    for (a_mn, b_mn) in [(0, 0), (0, 1), (1, 1)]:
        for mode in (1, 3):
            print(f'M{M} N{N} K{K} a_mn{a_mn} b_mn{b_mn} mode{mode}:', run(M, N, K, a_mn, b_mn, mode), flush=True)

# ---- timing of representative layer shapes, tensor-memory A operand (default) vs shared-memory A operand
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
print('--- timing (ms), fwd-type product, mode 3')
for (M, N, K) in [(32768, 1024, 512), (32768, 256, 512), (32768, 64, 256), (32768, 128, 384), (65536, 128, 768)]:
    a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev); C = torch.empty(M, N, device=dev)
    ref = a.double() @ b.double().t()
    out = []
    for no_ts in (0, 2):
        lib.dc_set_gemm_no_ts(0)
        lib.dc_set_gemm_keep_raw_hi(1 if no_ts == 2 else 0)
        f = lambda: _lib.call('dc_gemm_tc_raw', a.data_ptr(), K, 0, b.data_ptr(), K, 0, None, M, N, K, C.data_ptr(), N, ws.data_ptr(), ws.numel())
        t = timeit(f)
        err = ((C.double() - ref).abs().max() / ref.abs().max()).item()
        out.append(f'{"TS+rawhi" if no_ts else "TS"} {t:.4f} ms {2*M*N*K/t/1e9:.0f} TF/s err {err:.1e}')
    lib.dc_set_gemm_no_ts(0); lib.dc_set_gemm_keep_raw_hi(0)
    print(f'M{M} N{N} K{K}:', ' | '.join(out), flush=True)
