"""Timing stand-in for torch_scatter: the same results as ../ref_standins/torch_scatter through index_add /
scatter_reduce, without the arg-max bookkeeping (the reference never reads the arg).  Bench infrastructure only."""
import torch


def _size(index, dim_size):
    if dim_size is not None:
        return int(dim_size)
    return int(index.max()) + 1 if index.numel() else 0


def _bcast(index, src, dim):
    if dim < 0:
        dim += src.dim()
    if index.dim() == 1 and src.dim() > 1:
        shape = [1] * src.dim()
        shape[dim] = -1
        index = index.view(shape).expand_as(src)
    return index, dim


def scatter_add(src, index, dim=-1, out=None, dim_size=None):
    if index.dim() == 1:
        d = dim if dim >= 0 else dim + src.dim()
        shape = list(src.shape)
        shape[d] = _size(index, dim_size)
        return src.new_zeros(shape).index_add(d, index, src)
    index, dim = _bcast(index, src, dim)
    shape = list(src.shape)
    shape[dim] = _size(index, dim_size)
    return src.new_zeros(shape).scatter_add(dim, index, src)


scatter_sum = scatter_add


def scatter_mean(src, index, dim=-1, out=None, dim_size=None):
    total = scatter_add(src, index, dim, dim_size=dim_size)
    count = scatter_add(torch.ones_like(src), index, dim, dim_size=dim_size).clamp(min=1)
    return total / count


def _extreme(src, index, dim, dim_size, mode):
    index, dim = _bcast(index, src, dim)
    shape = list(src.shape)
    shape[dim] = _size(index, dim_size)
    return src.new_zeros(shape).scatter_reduce(dim, index, src, reduce=mode, include_self=False)


def scatter_max(src, index, dim=-1, out=None, dim_size=None):
    return _extreme(src, index, dim, dim_size, 'amax'), None


def scatter_min(src, index, dim=-1, out=None, dim_size=None):
    return _extreme(src, index, dim, dim_size, 'amin'), None


def scatter(src, index, dim=-1, out=None, dim_size=None, reduce='sum'):
    if reduce in ('sum', 'add'):
        return scatter_add(src, index, dim, dim_size=dim_size)
    if reduce == 'mean':
        return scatter_mean(src, index, dim, dim_size=dim_size)
    if reduce == 'max':
        return scatter_max(src, index, dim, dim_size=dim_size)[0]
    if reduce == 'min':
        return scatter_min(src, index, dim, dim_size=dim_size)[0]
    raise ValueError(reduce)
