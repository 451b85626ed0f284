"""Whole training step captured in CUDA graphs.

At 1K-point clouds one step is a few hundred short kernels: launched one by one from Python the
host takes longer than the GPU.  Shapes are static once (clouds, points, k) are fixed, so the step
is captured once and replayed; new batches are copied into the static input tensors.

The reference's loop (experiments/train_modelnet.py:98-108) maps onto it as
    step.load(batch)        # data = data.to(device)
    loss = step()           # zero_grad / forward / loss / backward / (all-reduce) / optimizer.step
"""
import torch
import torch.distributed as dist

from .. import ops
from ..parallel import FlatGradients


class _StaticInputs:
    """Static input tensors of a captured graph and the two ways of refilling them: `load()` (a plain copy) and
    `prefetch()` / `commit()` (host -> staging buffers on a side stream while the previous replay runs, then device to
    device on the replay's stream)."""

    static = None
    _stage = None

    def _keep_workspaces(self):
        # The captured kernels have the addresses of the library's scratch buffers baked in.  `ops.workspace()` replaces
        # a buffer when a later eager call asks for a larger one; holding the ones that were live during capture keeps
        # their memory from going back to the allocator for as long as this object (and its graphs) exists.
        self._workspaces = ops.live_workspaces()

    def load(self, batch):
        """Copy a (pinned-host or device) batch with the captured shapes into the static inputs."""
        for name, dst in self.static.__dict__.items():
            if torch.is_tensor(dst):
                dst.copy_(getattr(batch, name), non_blocking=True)

    def prefetch(self, batch):
        """Start copying the NEXT batch (pinned host memory) into device staging buffers on a side stream, so that the
        transfer overlaps the step that is running; `commit()` then moves it into the static inputs device-to-device."""
        if self._stage is None:
            self._stage = {name: torch.empty_like(t) for name, t in self.static.__dict__.items() if torch.is_tensor(t)}
            self._copy_stream = torch.cuda.Stream()
            self._staged = torch.cuda.Event()
            self._consumed = torch.cuda.Event()
            self._consumed.record()
        self._copy_stream.wait_event(self._consumed)                 # the previous commit() has read the staging buffers
        with torch.cuda.stream(self._copy_stream):
            for name, dst in self._stage.items():
                dst.copy_(getattr(batch, name), non_blocking=True)
            self._staged.record()

    def commit(self):
        """Make the prefetched batch the input of the next replay (waits for its transfer on the device, not the host)."""
        cur = torch.cuda.current_stream()
        cur.wait_event(self._staged)
        for name, src in self._stage.items():
            getattr(self.static, name).copy_(src, non_blocking=True)
        self._consumed.record(cur)


class GraphedInference(_StaticInputs):
    """Evaluation forward (`model.eval()`, no autograd) captured in one CUDA graph: `load(batch)` (or `prefetch` /
    `commit`), then `logits = runner()`.  The returned tensor is the graph's static output — copy it before the next
    replay.  Same contract on `example` as GraphedTrainStep (same shapes and cloud layout on every batch)."""

    def __init__(self, model, example, warmup=2):
        self.model, self.static = model, example
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):
                model(self.static)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = model(self.static)
        self._keep_workspaces()

    def __call__(self):
        self.graph.replay()
        return self.out


class GraphedTrainStep(_StaticInputs):
    """forward + loss + backward (+ gradient all-reduce) + optimizer step as CUDA graph replays.

    `example` is a device-resident `Data` batch whose tensors become the static inputs; every later
    batch must have the same shapes and the same cloud layout (same points per cloud), because the
    per-cloud offsets are baked into the captured kernels' arguments."""

    def __init__(self, model, optimizer, loss_fn, example, flat=None, warmup=3):
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        self.static = example
        self.flat = flat if flat is not None else FlatGradients(model.parameters())
        self.distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._forward_backward()
                self.flat.all_reduce()
                self.optimizer.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph_fb = torch.cuda.CUDAGraph()
        self.graph_opt = None
        self._graph_grads = None
        self._stage = None
        from ..parallel import FlatSGD
        self._fold = self.distributed and isinstance(self.optimizer, FlatSGD)      # FlatSGD / FlatAdam apply 1 / world themselves
        if self.distributed:
            # The collective stays between two graphs: fwd+bwd graph -> pack + NCCL all-reduce (SUM) -> optimizer graph, which
            # also applies 1 / world.  (Capturing the all-reduce into one graph with the rest hung on 2 GPUs in this
            # environment — torch 2.11 / NCCL 2.28, capture_error_mode thread_local — and is not used.)
            # The gradients a replay produces live in graph memory: keep the tensors so every step can pack them.
            with torch.cuda.graph(self.graph_fb):
                self.loss = self._forward_backward()
            self._graph_grads = [p.grad for p in self.flat.params if p.grad is not None]
            self.flat.all_reduce(sources=self._graph_grads, defer_average=self._fold)   # p.grad now are views of the (static) flat buffer
            self.graph_opt = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_opt, pool=self.graph_fb.pool()):
                self.optimizer.step()
        else:
            with torch.cuda.graph(self.graph_fb):
                self.loss = self._forward_backward()
                self.optimizer.step()
        self._keep_workspaces()

    def _forward_backward(self):
        self.flat.zero()
        loss = self.loss_fn(self.model(self.static), self.static.y)
        loss.backward()
        return loss.detach()

    def __call__(self):
        sync_lr = getattr(self.optimizer, 'sync_lr', None)
        if sync_lr is not None:
            sync_lr()                    # a scheduler may have changed the learning rate: the captured kernel reads it from the device
        self.graph_fb.replay()
        if self.graph_opt is not None:
            self.flat.all_reduce(sources=self._graph_grads, defer_average=self._fold)
            self.graph_opt.replay()
        return self.loss
