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
        self.packed = False          # True between pack() and the next zero(): `flat` holds this step's gradients
        self.pending_scale = 1.0     # factor the optimizer still has to apply to `flat` (1 / world after a SUM all-reduce)

    def zero(self):
        for p in self.params:
            p.grad = None
        self.packed = False
        self.pending_scale = 1.0

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
        self.packed = True
        return self.flat

    def all_reduce(self, group=None, average=True, sources=None, defer_average=False):
        """Sum (or mean) of all gradients over the ranks with one collective; no-op for a single process.
        defer_average=True leaves the SUM in the buffer and records 1 / world in `pending_scale`: FlatSGD / FlatAdam fold it
        into their update kernel (any other reader of `p.grad` sees the sum)."""
        world = self._world(group)
        if world == 1:
            return None
        flat = self.pack(sources)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            if defer_average:
                self.pending_scale = 1.0 / world
            else:
                flat.div_(world)
        return flat

    def nbytes(self):
        return sum(p.numel() * p.element_size() for p in self.params)


class FlatSGD(torch.optim.Optimizer):
    """torch.optim.SGD(lr, momentum, weight_decay) — the optimizer of experiments/train_*.py — as ONE kernel per step
    (`dc_sgd_step`) over flat parameter / gradient / momentum buffers instead of a few foreach launches over ~130
    tensors.  A `torch.optim.Optimizer` (learning-rate schedulers attach to it as to SGD).  The parameters that receive
    gradients become views of one flat buffer at the first step (state dicts are unaffected); gradients come from
    `FlatGradients.pack()` (already packed when they were all-reduced).  The learning rate lives in a device scalar, so
    a CUDA-graph-captured step follows schedulers: `sync_lr()` (called by GraphedTrainStep before every replay) copies
    `param_groups[0]['lr']` to the device when it changed."""

    def __init__(self, flat, lr, momentum=0.0, weight_decay=0.0):
        if not isinstance(flat, FlatGradients):
            raise TypeError("FlatSGD works on a FlatGradients object")
        super().__init__(flat.params, dict(lr=float(lr), momentum=float(momentum), weight_decay=float(weight_decay)))
        if len(self.param_groups) != 1:
            raise ValueError("FlatSGD supports one parameter group")
        self.flat = flat
        self._live = None
        self._p = self._buf = self._lr = None
        self._lr_value = None

    def _materialise(self):
        live = [p for p in self.flat.params if p.grad is not None]
        if not live or not live[0].is_cuda:
            from ._lib import DeltaConvLibraryError
            raise DeltaConvLibraryError("FlatSGD needs CUDA parameters with gradients (there is no CPU path)")
        with torch.no_grad():
            self._p = torch.cat([p.data.reshape(-1) for p in live])
            offset = 0
            for p in live:
                n = p.numel()
                p.data = self._p[offset:offset + n].view_as(p)
                offset += n
        self._buf = torch.zeros_like(self._p)
        self._lr = torch.empty(1, dtype=torch.float32, device=self._p.device)
        self._live = live
        self._apply_pending()

    def _apply_pending(self):
        """State loaded before the flat buffers existed (the usual resume order: build, load_state_dict, train)."""
        state, self._pending_state = getattr(self, '_pending_state', None), None
        if state is not None and state.get('momentum_buffer') is not None:
            if state['momentum_buffer'].numel() != self._buf.numel():
                raise ValueError(f"FlatSGD.load_state_dict: momentum buffer has {state['momentum_buffer'].numel()} elements, "
                                 f"the parameters with gradients have {self._buf.numel()}")
            self._buf.copy_(state['momentum_buffer'])

    def sync_lr(self):
        lr = float(self.param_groups[0]['lr'])
        if self._lr is not None and self._lr_value != lr:
            self._lr.fill_(lr)
            self._lr_value = lr

    def set_lr(self, lr):
        self.param_groups[0]['lr'] = float(lr)
        self.sync_lr()

    def zero_grad(self, set_to_none=True):
        self.flat.zero()

    @torch.no_grad()
    def step(self, closure=None):
        from ._lib import call, DeltaConvLibraryError
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g = self.flat.flat if self.flat.packed else self.flat.pack()
        if self._live is None:
            self._materialise()
        if g.numel() != self._p.numel():
            raise DeltaConvLibraryError("FlatSGD: the set of parameters with gradients changed between steps")
        self.sync_lr()
        group = self.param_groups[0]
        call('dc_sgd_step', self._p.data_ptr(), g.data_ptr(), self._buf.data_ptr(), self._p.numel(), self._lr.data_ptr(),
             float(group['momentum']), float(group['weight_decay']), float(self.flat.pending_scale))
        self.flat.pending_scale = 1.0
        return loss

    def state_dict(self):
        return dict(param_groups=[{k: v for k, v in self.param_groups[0].items() if k != 'params'}],
                    momentum_buffer=None if self._buf is None else self._buf.clone())

    def load_state_dict(self, state):
        self.param_groups[0].update(state['param_groups'][0])
        if self._buf is None:
            self._pending_state = state          # applied when the first step() creates the flat buffers
        else:
            self._pending_state = state
            self._apply_pending()


class FlatAdam(FlatSGD):
    """torch.optim.Adam(lr, betas, eps, weight_decay) (experiments/train_shapeseg.py:82) on the flat buffers: one
    `dc_adam_step` call per step; step counter and learning rate are device scalars, so captured steps stay correct."""

    def __init__(self, flat, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if not isinstance(flat, FlatGradients):
            raise TypeError("FlatAdam works on a FlatGradients object")
        torch.optim.Optimizer.__init__(self, flat.params, dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), eps=float(eps),
                                                               weight_decay=float(weight_decay)))
        self.flat = flat
        self._live = None
        self._p = self._buf = self._lr = None
        self._lr_value = None
        self._v = self._step = None

    @torch.no_grad()
    def step(self, closure=None):
        from ._lib import call, DeltaConvLibraryError
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g = self.flat.flat if self.flat.packed else self.flat.pack()
        if self._live is None:
            self._materialise()          # creates the moment buffers and applies a state loaded before the first step
        if g.numel() != self._p.numel():
            raise DeltaConvLibraryError("FlatAdam: the set of parameters with gradients changed between steps")
        self.sync_lr()
        group = self.param_groups[0]
        call('dc_adam_step', self._p.data_ptr(), g.data_ptr(), self._buf.data_ptr(), self._v.data_ptr(), self._p.numel(),
             self._lr.data_ptr(), self._step.data_ptr(), group['betas'][0], group['betas'][1], group['eps'], group['weight_decay'],
             float(self.flat.pending_scale))
        self.flat.pending_scale = 1.0
        return loss

    def state_dict(self):
        return dict(param_groups=[{k: v for k, v in self.param_groups[0].items() if k != 'params'}],
                    exp_avg=None if self._buf is None else self._buf.clone(), exp_avg_sq=None if self._v is None else self._v.clone(),
                    step=None if self._step is None else int(self._step))

    def _apply_pending(self):
        if self._v is None:
            self._v = torch.zeros_like(self._p)
            self._step = torch.zeros(1, dtype=torch.int32, device=self._p.device)
        state, self._pending_state = getattr(self, '_pending_state', None), None
        if state is not None and state.get('exp_avg') is not None:
            if state['exp_avg'].numel() != self._buf.numel():
                raise ValueError(f"FlatAdam.load_state_dict: state has {state['exp_avg'].numel()} elements, "
                                 f"the parameters with gradients have {self._buf.numel()}")
            self._buf.copy_(state['exp_avg'])
            self._v.copy_(state['exp_avg_sq'])
            self._step.fill_(int(state['step']))
