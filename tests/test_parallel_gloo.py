"""CPU, world_size 2, gloo: cloud sharding and the flat-gradient all-reduce give the same
parameter update as one process on the whole batch."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deltaconv_b200.parallel import FlatGradients, shard_range


def test_shard_range_partitions_everything():
    for n in (0, 1, 7, 32, 33):
        for w in (1, 2, 3, 8):
            ranges = [shard_range(n, r, w) for r in range(w)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            sizes = [e - b for b, e in ranges]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _model():
    torch.manual_seed(3)
    m = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 2))
    m.extra = torch.nn.Parameter(torch.zeros(4))      # never used: grad slot must stay zero on every rank
    return m


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    torch.manual_seed(0)
    x, y = torch.randn(8, 6), torch.randn(8, 2)
    model = _model()
    flat = FlatGradients(model.parameters())
    b, e = shard_range(8, rank, world)
    flat.zero()
    loss = ((model(x[b:e]) - y[b:e]) ** 2).sum() / 8 * world      # local mean scaled so the rank-average is the global mean
    loss.backward()
    assert not flat.packed
    reduced = flat.all_reduce()
    assert flat.packed and flat.flat.data_ptr() == reduced.data_ptr()   # what FlatSGD.step() reads instead of packing again
    assert model.extra.grad is None                                 # never used: skipped on every rank
    assert model[0].weight.grad.data_ptr() == reduced.data_ptr()    # grads are views of the reduced flat buffer
    if rank == 0:
        torch.save(reduced.clone(), out)
    address = reduced.data_ptr()
    flat.zero()
    assert not flat.packed                                          # a new step starts unpacked ...
    (((model(x[b:e]) - y[b:e]) ** 2).sum() / 8 * world).backward()
    assert flat.all_reduce().data_ptr() == address and flat.packed  # ... and packs into the same (stable) buffer
    mean = flat.flat.clone()
    # deferred averaging (what the captured step uses): the buffer keeps the SUM and the factor 1 / world is handed to the
    # optimizer kernel through `pending_scale`; zero() resets it
    flat.zero()
    assert flat.pending_scale == 1.0
    (((model(x[b:e]) - y[b:e]) ** 2).sum() / 8 * world).backward()
    summed = flat.all_reduce(defer_average=True)
    assert flat.pending_scale == 1.0 / world and torch.allclose(summed * flat.pending_scale, mean, atol=1e-7)
    flat.zero()
    assert flat.pending_scale == 1.0
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_flat_gradient_allreduce_matches_single_process(tmp_path):
    out = str(tmp_path / 'flat.pt')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    x, y = torch.randn(8, 6), torch.randn(8, 2)
    model = _model()
    flat = FlatGradients(model.parameters())
    (((model(x) - y) ** 2).sum() / 8).backward()
    assert flat.all_reduce() is None                                # single process: nothing to pack or reduce
    assert not flat.packed                                          # so an optimizer on the flat buffer packs itself
    assert flat.pack().numel() == sum(p.numel() for p in model.parameters() if p.grad is not None) and flat.packed
    want = torch.cat([p.grad.reshape(-1) for p in model.parameters() if p.grad is not None])
    assert torch.allclose(got, want, atol=1e-6)
    flat.zero()
    assert all(p.grad is None for p in model.parameters())
