"""Minimal stand-in for `torch_geometric.data.Batch` — the duck type the models read
(reference: deltanet_base.py:43-44,59,76; deltanet_segmentation.py:64)."""
import torch


class Data:
    """Attribute bag with `.to(device)`; the models touch `pos`, `batch`, and optionally
    `norm`, `x`, `category`, `y`."""

    def __init__(self, **fields):
        for name, value in fields.items():
            setattr(self, name, value)

    def _tensors(self):
        return {k: v for k, v in self.__dict__.items() if torch.is_tensor(v)}

    def to(self, device, non_blocking=False):
        out = Data(**self.__dict__)
        for name, t in self._tensors().items():
            setattr(out, name, t.to(device, non_blocking=non_blocking))
        return out

    def pin_memory(self):
        out = Data(**self.__dict__)
        for name, t in self._tensors().items():
            setattr(out, name, t.pin_memory())
        return out

    def double(self):
        out = Data(**self.__dict__)
        for name, t in self._tensors().items():
            if t.is_floating_point():
                setattr(out, name, t.double())
        return out

    @property
    def num_graphs(self):
        return int(self.batch.max()) + 1 if self.batch.numel() else 0

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self._tensors().values())
