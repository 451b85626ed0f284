"""The training script of the reference (experiments/train_modelnet.py), line for line where it matters, on this package and
a synthetic mesh dataset — the datasets of the reference need files and readers that are not in this image.

    python examples/train_synthetic.py [--shapes 64] [--num_points 1024] [--epochs 2]

What is the same as in the reference: the pre-transforms (NormalizeScale, SamplePoints with a sampling margin of 8,
GeodesicFPS), the training transforms, `DataLoader`, `DeltaNetClassification`, SGD + cosine schedule, label-smoothed loss.
What differs: the imports resolve through `deltaconv_b200.compat.install()` (no torch_geometric / torch_cluster /
torch_sparse / torch_scatter), and GeodesicFPS samples every shape of the dataset in ONE launch instead of shape by shape.
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deltaconv_b200 import compat           # noqa: E402

compat.install()
import deltaconv.transforms as T            # noqa: E402  (= deltaconv_b200.transforms)
from deltaconv.models import DeltaNetClassification          # noqa: E402
from torch_geometric.data import Data, Batch                  # noqa: E402  (stand-ins unless the real package is installed)
from torch_geometric.loader import DataLoader                 # noqa: E402
from torch_geometric.transforms import Compose, SamplePoints  # noqa: E402


def ellipsoid_mesh(label, gen):
    from deltaconv_b200.utils import synthetic_mesh
    return synthetic_mesh(label, gen)


def calc_loss(pred, gold, eps=0.2):
    """label-smoothed cross entropy (experiments/utils.py:7-24)"""
    n_class = pred.size(1)
    one_hot = torch.zeros_like(pred).scatter(1, gold.view(-1, 1), 1)
    one_hot = one_hot * (1 - eps) + (1 - one_hot) * eps / (n_class - 1)
    return -(one_hot * torch.nn.functional.log_softmax(pred, dim=1)).sum(dim=1).mean()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--shapes', type=int, default=64)
    ap.add_argument('--num_classes', type=int, default=4)
    ap.add_argument('--num_points', type=int, default=1024)
    ap.add_argument('--sampling_margin', type=int, default=8)
    ap.add_argument('--batch_size', type=int, default=16)
    ap.add_argument('--epochs', type=int, default=2)
    ap.add_argument('--k', type=int, default=20)
    ap.add_argument('--lr', type=float, default=0.001)
    args = ap.parse_args()
    device = torch.device('cuda')
    gen = torch.Generator().manual_seed(0)

    # pre-transform, as train_modelnet.py:30-34: the two mesh steps shape by shape on the host ...
    per_shape = Compose((T.NormalizeScale(), SamplePoints(args.num_points * args.sampling_margin, include_normals=True)))
    shapes = [per_shape(ellipsoid_mesh(i % args.num_classes, gen)) for i in range(args.shapes)]
    # ... and the geodesic sampling for the whole dataset at once on the GPU
    t0 = time.time()
    sampled = T.GeodesicFPS(args.num_points)(Batch.from_data_list(shapes).to(device))
    torch.cuda.synchronize()
    print(f'GeodesicFPS: {args.shapes} shapes x {args.num_points * args.sampling_margin} -> {args.num_points} points in '
          f'{1000 * (time.time() - t0):.1f} ms (first call, includes loading the library)')
    P = args.num_points
    dataset = [Data(pos=sampled.pos[i * P:(i + 1) * P].cpu(), norm=sampled.norm[i * P:(i + 1) * P].cpu(), y=shapes[i].y)
               for i in range(args.shapes)]

    transform = Compose((T.RandomScale((4 / 5, 5 / 4)), T.RandomTranslateGlobal(0.1)))        # train_modelnet.py:37-40
    loader = DataLoader(dataset, batch_size=args.batch_size, shuffle=True, drop_last=True)
    model = DeltaNetClassification(in_channels=3, num_classes=args.num_classes, num_neighbors=args.k).to(device)
    optimizer = torch.optim.SGD(model.parameters(), lr=100 * args.lr, momentum=0.9, weight_decay=1e-4)
    scheduler = torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, args.epochs, eta_min=args.lr)
    for epoch in range(1, args.epochs + 1):
        model.train()
        total, correct, seen = 0.0, 0, 0
        for data in loader:
            data = transform(data.to(device))             # the augmentations act on the whole device batch, per cloud
            optimizer.zero_grad()
            out = model(data)
            loss = calc_loss(out, data.y)
            loss.backward()
            optimizer.step()
            total += loss.item() * data.num_graphs
            correct += int((out.argmax(1) == data.y).sum())
            seen += data.num_graphs
        scheduler.step()
        print(f'epoch {epoch}: loss {total / seen:.4f}  accuracy {correct / seen:.3f}')


if __name__ == '__main__':
    main()
