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


class Batch(Data):
    """Several clouds concatenated along the point axis with a `batch` vector, as `torch_geometric.data.Batch`
    lays them out (what the reference's DataLoader hands to the models, experiments/train_*.py).  `_num_clouds` / `_max_points` are the size hints that let the CUDA path skip a device
    read-back of the cloud layout."""

    @staticmethod
    def from_data_list(clouds):
        if not clouds:
            raise ValueError("Batch.from_data_list: empty list")
        names = [k for k, v in clouds[0].__dict__.items() if torch.is_tensor(v)]
        sizes = [int(c.pos.size(0)) for c in clouds]
        # like PyG's collate: every tensor field is concatenated along dim 0 (per-point fields become [N, ...], per-cloud
        # fields such as a label of shape [1] become [B, ...]); 0-dim tensors are stacked
        fields = {name: torch.cat([getattr(c, name).reshape(1) if getattr(c, name).dim() == 0 else getattr(c, name) for c in clouds], dim=0)
                  for name in names}
        fields['batch'] = torch.repeat_interleave(torch.arange(len(clouds)), torch.tensor(sizes))
        out = Batch(**fields)
        out._num_clouds, out._max_points = len(clouds), max(sizes)
        return out

    def to(self, device, non_blocking=False):
        out = Batch(**{k: (v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})
        return out

    def pin_memory(self):
        return Batch(**{k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})


class DataLoader(torch.utils.data.DataLoader):
    """torch DataLoader over a dataset of `Data` clouds that collates with `Batch.from_data_list` — the role of
    `torch_geometric.loader.DataLoader` in the reference's training scripts."""

    def __init__(self, dataset, batch_size=1, shuffle=False, **kwargs):
        kwargs.pop('collate_fn', None)
        super().__init__(dataset, batch_size=batch_size, shuffle=shuffle, collate_fn=Batch.from_data_list, **kwargs)
