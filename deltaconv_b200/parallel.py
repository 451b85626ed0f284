"""Multi-GPU plumbing for the hot path: clouds are independent, so work shards by cloud and the
only exchange is one all-reduce of the flat gradient per training step.  The
reference itself is single-process (experiments/train_modelnet.py:191); this module is what a
launcher puts beside its unchanged training loop.  Backend-agnostic (NCCL on GPUs, gloo in the
CPU tests)."""
import torch
import torch.distributed as dist


def shard_range(num_items, rank, world_size):
    """Contiguous [begin, end) slice of `num_items` clouds owned by `rank` (sizes differ by <= 1)."""
    base, extra = divmod(num_items, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


class FlatGradients:
    """All parameter gradients live in ONE contiguous fp32 buffer: `p.grad` is a view into it, so
    autograd accumulates in place, `zero()` is one memset and `all_reduce()` is one collective
    with no packing.  Parameters that never receive a gradient (VectorNonLin.bias under
    BatchNorm) keep a zero slot, identically on every rank."""

    def __init__(self, parameters):
        self.params = [p for p in parameters if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        total = sum(p.numel() for p in self.params)
        first = self.params[0]
        self.flat = torch.zeros(total, dtype=first.dtype, device=first.device)
        offset = 0
        for p in self.params:
            n = p.numel()
            p.grad = self.flat[offset:offset + n].view_as(p)
            offset += n

    def zero(self):
        self.flat.zero_()

    def all_reduce(self, group=None, average=True):
        """Sum (or mean) of the flat gradient over all ranks; no-op for a single process."""
        if not (dist.is_available() and dist.is_initialized()):
            return self.flat
        world = dist.get_world_size(group)
        if world == 1:
            return self.flat
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            self.flat.div_(world)
        return self.flat

    def nbytes(self):
        return self.flat.numel() * self.flat.element_size()
