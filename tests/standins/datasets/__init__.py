"""Stand-in for the reference's `experiments/datasets` package (test infrastructure): `ModelNet` with the constructor the
training script calls (experiments/train_modelnet.py:43-44, datasets/modelnet.py:53-54) over synthetic meshes, because the
real class downloads ModelNet and reads .off files through torch_geometric.io.  Like an InMemoryDataset it applies
`pre_transform` once per shape at construction and `transform` on every access."""
import copy

import torch

from deltaconv_b200.utils import synthetic_mesh

SHAPES_PER_SPLIT = {True: 16, False: 8}


class ModelNet(torch.utils.data.Dataset):
    def __init__(self, root, n_per_class=None, name='10', train=True, transform=None, pre_transform=None, pre_filter=None):
        assert name in ['10', '40']
        self.root, self.name, self.transform = root, name, transform
        gen = torch.Generator().manual_seed(10 + int(train))
        classes = int(name)
        self.items = []
        for i in range(SHAPES_PER_SPLIT[bool(train)]):
            data = synthetic_mesh(i % classes, gen)
            if pre_filter is not None and not pre_filter(data):
                continue
            self.items.append(pre_transform(data) if pre_transform is not None else data)

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        data = copy.copy(self.items[i])
        return self.transform(data) if self.transform is not None else data
