"""Import-level compatibility for code written against the reference and PyTorch Geometric.

`install()` makes these imports resolve on a box without PyG, to this package's CUDA implementations:

    import deltaconv                      ->  deltaconv_b200 (geometry, nn, models, transforms too)
    from torch_geometric.data import Data, Batch
    from torch_geometric.loader import DataLoader
    from torch_geometric.transforms import Compose, BaseTransform, SamplePoints
    from torch_geometric.nn import knn_graph, global_max_pool, global_mean_pool

A real `torch_geometric`, if importable, is left alone (its `Batch` objects work with the models as they are).
Datasets are not provided; deltaconv_b200/transforms.py has all nine transforms of the reference (`GeodesicFPS` on the GPU)."""
import importlib
import sys
import types


class BaseTransform:
    def __call__(self, data):
        return data

    def __repr__(self):
        return f'{self.__class__.__name__}()'


class Compose(BaseTransform):
    """Apply a list of transforms in order (torch_geometric.transforms.Compose)."""

    def __init__(self, transforms):
        self.transforms = list(transforms)

    def __call__(self, data):
        for t in self.transforms:
            data = [t(d) for d in data] if isinstance(data, (list, tuple)) else t(data)
        return data

    def __repr__(self):
        return 'Compose([\n' + ',\n'.join(f'  {t}' for t in self.transforms) + '\n])'


def _module(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


def install(alias_reference=True, provide_pyg=None):
    """Register the stand-in modules.  provide_pyg=None: only when the real torch_geometric is absent."""
    import deltaconv_b200 as pkg
    from .utils import Data, Batch, DataLoader
    from . import geometry, transforms
    if alias_reference:
        sys.modules.setdefault('deltaconv', pkg)
        for sub in ('geometry', 'nn', 'models', 'utils', 'transforms'):
            sys.modules.setdefault(f'deltaconv.{sub}', importlib.import_module(f'deltaconv_b200.{sub}'))
    if provide_pyg is None:
        try:
            importlib.import_module('torch_geometric')
            provide_pyg = False
        except Exception:
            provide_pyg = True
    if provide_pyg:
        root = _module('torch_geometric', __version__='0.0-deltaconv_b200-standin')
        root.data = _module('torch_geometric.data', Data=Data, Batch=Batch)
        root.loader = _module('torch_geometric.loader', DataLoader=DataLoader)
        root.transforms = _module('torch_geometric.transforms', Compose=Compose, BaseTransform=BaseTransform,
                                  SamplePoints=transforms.SamplePoints)
        root.nn = _module('torch_geometric.nn', knn_graph=geometry.knn_graph, global_max_pool=geometry.global_max_pool,
                          global_mean_pool=geometry.global_mean_pool)
    return provide_pyg
