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
    """One contiguous fp32 buffer holding every parameter gradient, for ONE collective per step.

    Gradients are produced by autograd as separate tensors (assigning is free; pre-allocated `.grad` views would
    cost one accumulate kernel per parameter per step).  `all_reduce()` packs them with a single concatenation,
    reduces the flat buffer, and re-points every `p.grad` at its slice of the result, so the optimizer reads the
    averaged gradients without a copy back.  Parameters that never receive a gradient (VectorNonLin.bias under
    BatchNorm) are skipped, identically on every rank.  With a single process nothing is packed."""

    def __init__(self, parameters):
        self.params = [p for p in parameters if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        self.flat = None

    def zero(self):
        for p in self.params:
            p.grad = None

    def _world(self, group):
        if not (dist.is_available() and dist.is_initialized()):
            return 1
        return dist.get_world_size(group)

    def pack(self, sources=None):
        """Concatenate the gradients into `self.flat` (re-used when the size is unchanged, so its address is stable)
        and make every `p.grad` a view of it.  `sources` overrides where the gradients are read from (a list
        parallel to the parameters that currently have a gradient) — used when they live in CUDA-graph memory."""
        live = [p for p in self.params if p.grad is not None]
        pieces = [g.reshape(-1) for g in (sources if sources is not None else [p.grad for p in live])]
        total = sum(t.numel() for t in pieces)
        if self.flat is not None and self.flat.numel() == total and self.flat.device == pieces[0].device:
            torch.cat(pieces, out=self.flat)
        else:
            self.flat = torch.cat(pieces)
        offset = 0
        for p in live:
            n = p.numel()
            p.grad = self.flat[offset:offset + n].view_as(p)
            offset += n
        return self.flat

    def all_reduce(self, group=None, average=True, sources=None):
        """Sum (or mean) of all gradients over the ranks with one collective; no-op for a single process."""
        world = self._world(group)
        if world == 1:
            return None
        flat = self.pack(sources)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat.div_(world)
        return flat

    def nbytes(self):
        return sum(p.numel() * p.element_size() for p in self.params)
