"""Numerics (ragged edges, bias, split-K partials) and timing of the CTA-pair kernel against the single-CTA kernel; derived from diag_tc_epilogue: TMA store through shared memory
(default) against row-per-lane direct stores (dc_set_gemm_tma_store(0))."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deltaconv_b200 import _lib
lib = _lib.load()
dev = 'cuda'
ws = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
torch.manual_seed(0)

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

shapes = [(300, 72, 100, 0, 0, True), (257, 200, 64, 0, 1, False), (200, 136, 1000, 1, 1, False), (1200, 64, 256, 0, 0, True), (131, 20, 40, 0, 0, True),
          (32768, 1024, 512, 0, 0, False), (32768, 512, 1024, 0, 1, False), (1024, 512, 32768, 1, 1, False), (32768, 256, 512, 0, 0, True),
          (256, 512, 32768, 1, 1, False), (32768, 128, 384, 0, 0, False), (32768, 64, 256, 0, 0, False), (32768, 256, 128, 0, 1, False), (64, 64, 32768, 1, 1, False)]
for (M, N, K, a_mn, b_mn, with_bias) in shapes:
    a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev)
    bias = torch.randn(N, device=dev) if with_bias else None
    A = a.t().contiguous() if a_mn else a
    B = b.t().contiguous() if b_mn else b
    ref = a.double() @ b.double().t() + (bias.double() if with_bias else 0)
    out = []
    for tma in (0, 1):  # here: CTA pairs off / on
        lib.dc_set_gemm_pair(tma)
        ld = N + (4 if M < 2000 else 0)                       # a padded output pitch on the small cases
        Cbuf = torch.full((M, ld), float('nan'), device=dev)
        f = lambda: _lib.call('dc_gemm_tc_raw', A.data_ptr(), A.stride(0), a_mn, B.data_ptr(), B.stride(0), b_mn, bias.data_ptr() if with_bias else None,
                              M, N, K, Cbuf.data_ptr(), ld, ws.data_ptr(), ws.numel())
        f(); torch.cuda.synchronize()
        C = Cbuf[:, :N]
        err = ((C.double() - ref).abs().max() / ref.abs().max()).item()
        pad_ok = bool(Cbuf[:, N:].isnan().all())              # nothing written outside the N columns
        t = timeit(f) if M * N * K > 1e8 else 0.0
        out.append(f'pair{tma} err {err:.1e} pad_untouched {pad_ok} {t * 1000:6.1f} us {2 * M * N * K / max(t, 1e-9) / 1e9:5.0f} TF/s')
    lib.dc_set_gemm_pair(1)
    print(f'M{M} N{N} K{K} a_mn{a_mn} b_mn{b_mn}: ' + ' | '.join(out), flush=True)
