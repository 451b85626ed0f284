"""CPU: import-level stand-ins for the reference package name and the PyG modules its scripts import."""
import subprocess
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_compat_install_resolves_reference_imports():
    code = '''
import sys
sys.path.insert(0, %r)
from deltaconv_b200 import compat
provided = compat.install()
import deltaconv
from deltaconv.models import DeltaNetClassification, DeltaNetSegmentation
from deltaconv.nn import DeltaConv, MLP
from deltaconv.geometry import build_grad_div, estimate_basis, geodesic_fps
import deltaconv.transforms as T
# the nine transforms of the reference (deltaconv/transforms/__init__.py)
for name in ('NormalizeScale', 'NormalizeArea', 'NormalizeAxes', 'RandomScale', 'RandomTranslateGlobal', 'RandomRotate',
             'RandomNormals', 'SamplePoints', 'GeodesicFPS'):
    assert hasattr(T, name), name
assert repr(T.GeodesicFPS(1024)) == 'GeodesicFPS()'
assert repr(T.RandomScale((0.8, 1.2))) == 'RandomScale((0.8, 1.2))'
assert deltaconv.__name__ == 'deltaconv_b200'
if provided:
    from torch_geometric.data import Data, Batch
    from torch_geometric.loader import DataLoader
    from torch_geometric.transforms import Compose, SamplePoints
    from torch_geometric.nn import knn_graph, global_max_pool, global_mean_pool
    import torch
    class Shift:
        def __call__(self, d):
            d.pos = d.pos + 1
            return d
    d = Compose([Shift(), Shift()])(Data(pos=torch.zeros(4, 3)))
    assert float(d.pos.sum()) == 24.0
    b = next(iter(DataLoader([Data(pos=torch.zeros(n, 3), y=torch.tensor([n])) for n in (3, 5)], batch_size=2)))
    assert b.batch.tolist() == [0, 0, 0, 1, 1, 1, 1, 1]
net = DeltaNetClassification(3, 40)
assert sum(p.numel() for p in net.parameters()) == 2045928
print("ok", provided)
''' % ROOT
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip().startswith('ok')
