from torch_cluster import knn_graph  # noqa: F401
from torch_scatter import scatter as _scatter
from . import inits  # noqa: F401


def global_max_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return _scatter(x, batch, dim=0, dim_size=size, reduce='max')


def global_mean_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return _scatter(x, batch, dim=0, dim_size=size, reduce='mean')
